"""volkit-b200: B200-native (sm_100a) backend for the StructuredVolume core of Kniggi/volkit.

The product is the C-ABI shared library volkit_b200/lib/libvktb200.so (include/vktb/backend.h) plus the
C++ host mirror of the vkt:: API (volkit_b200/lib/libvkt_b200.so, include/vkt/*.hpp).  This Python
package only binds them for tests, the benchmark and multi-GPU orchestration.
"""
from . import capi  # noqa: F401
from .device import DeviceVolume, launch_count, sync  # noqa: F401
