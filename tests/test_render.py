"""Headless ray-marcher (last row): the reference's RayMarchingKernel
(src/vkt/Render_kernel.hpp:95-162) as CUDA arithmetic writing RGBA into a device buffer -- no window, no GL.

Chain of evidence:  oracle/_ref/libvkt_ref_render.so = the reference's OWN kernel + visionaray's own camera, compiled
here (oracle/ref_render_shim.cpp)  ==  oracle/vkt_render_oracle.c (plain-C restatement, float for float)  ~  the
CUDA kernel through vktb_render_raymarch (within 1 LSB per RGBA8 channel: powf / tanf differ between libm and CUDA)."""
import numpy as np
import pytest

from oracle import pyoracle as po


def lut_ramp(n=64):
    """a colourful transfer function with a soft opacity ramp"""
    x = np.linspace(0.0, 1.0, n, dtype=np.float32)
    return np.stack([x, 1.0 - x, 0.5 + 0.5 * np.sin(6.0 * x), 0.05 + 0.25 * x * x], axis=1).astype(np.float32)


def blob(dims, fmt):
    """a smooth blob plus counter-based noise, so rays see structure and hard voxel edges"""
    x, y, z = dims
    zz, yy, xx = np.meshgrid(np.arange(z), np.arange(y), np.arange(x), indexing="ij")
    r = np.sqrt(((xx - x / 2) / x) ** 2 + ((yy - y / 2) / y) ** 2 + ((zz - z / 2) / z) ** 2)
    v = np.clip(1.0 - 2.2 * r, 0.0, 1.0)
    top = {po.UINT8: 255, po.UINT16: 65535, po.UINT32: 4294967295, po.INT16: 32767}[fmt]
    noise = po.synth(dims, po.UINT8, 3).astype(np.float64) / 255.0
    return (v * 0.85 * top + noise * 0.1 * top).astype(po.DTYPES[fmt])


CAMERAS = [  # eye, center, up, fovy
    ((40.0, 30.0, 130.0), (24.0, 20.0, 16.0), (0.0, 1.0, 0.0), 45.0),
    ((-60.0, 70.0, -50.0), (24.0, 20.0, 16.0), (0.1, 1.0, 0.0), 30.0),
    ((24.0, 20.0, 16.0), (60.0, 20.0, 16.0), (0.0, 0.0, 1.0), 70.0),     # eye INSIDE the volume: tnear < 0
]


@pytest.mark.needs_ref
@pytest.mark.parametrize("fmt", [po.UINT8, po.UINT16, po.UINT32, po.INT16])
def test_restatement_equals_the_reference_kernel(fmt):
    if not po.have_reference_render():
        pytest.skip("oracle/_ref/libvkt_ref_render.so not built")
    dims = (48, 40, 32)
    vol = po.Vol(blob(dims, fmt), fmt)
    ref, mine = po.render_reference(), po.render_oracle()
    for eye, center, up, fovy in CAMERAS:
        for lut in (None, lut_ramp()):
            for dt, srgb, dist in ((1.0, True, (1.0, 1.0, 1.0)), (0.37, False, (1.0, 1.5, 0.8))):
                a32, a8 = ref.raymarch(vol, lut, eye, center, up, fovy, 40, 28, dt, srgb, dist)
                b32, b8 = mine.raymarch(vol, lut, eye, center, up, fovy, 40, 28, dt, srgb, dist)
                assert a32.tobytes() == b32.tobytes(), (fmt, eye, lut is None, dt)
                assert a8.tobytes() == b8.tobytes()
                assert a32[..., 3].max() > 0.05        # the rays do hit the volume


@pytest.mark.needs_ref
def test_default_camera_matches_visionarays_view_all():
    if not po.have_reference_render():
        pytest.skip("oracle/_ref/libvkt_ref_render.so not built")
    for dims, dist in (((64, 64, 64), (1.0, 1.0, 1.0)), ((100, 37, 250), (1.0, 2.0, 0.5))):
        assert po.default_camera(dims, 512, 512, dist, reference=True) == po.default_camera(dims, 512, 512, dist)
