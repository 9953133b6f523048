"""ctypes binding of the C-ABI (include/vktb/backend.h) -- the boundary the reference's shims bind.

There is no Python or CPU implementation behind this module: if libvktb200.so is missing or no B200
is usable, calls raise.  Build the library with `python -c "import __graft_entry__ as g; g.build()"`
or `make -C volkit_b200/csrc`.
"""
from __future__ import annotations

import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libvktb200.so")

# vkt::DataFormat values (reference include/cpp/vkt/common.hpp:53-66)
INT8, INT16, INT32, UINT8, UINT16, UINT32, FLOAT32 = 1, 2, 3, 4, 5, 6, 7
H2H, H2D, D2H, D2D = 0, 1, 2, 3
SUM, DIFF, PROD, QUOT, ABSDIFF, SAFESUM, SAFEDIFF, SAFEPROD, SAFEQUOT, SAFEABSDIFF = range(10)
NEAREST, LINEAR, TRILINEAR = 0, 1, 2
BYTES_PER_VOXEL = {INT8: 1, UINT8: 1, INT16: 2, UINT16: 2, INT32: 4, UINT32: 4, FLOAT32: 4}


class VktbVol(ctypes.Structure):
    _fields_ = [("data", ctypes.c_void_p), ("dims", ctypes.c_int32 * 3), ("format", ctypes.c_int32),
                ("lo", ctypes.c_float), ("hi", ctypes.c_float)]


I3 = ctypes.c_int32 * 3

class VktbAggregates(ctypes.Structure):
    """vktb_aggregates = vkt::Aggregates (reference include/cpp/vkt/Aggregates.hpp:14-25)"""
    _fields_ = [("min", ctypes.c_float), ("max", ctypes.c_float), ("mean", ctypes.c_float), ("stddev", ctypes.c_float),
                ("var", ctypes.c_float), ("sum", ctypes.c_float), ("prod", ctypes.c_float),
                ("argmin", ctypes.c_int32 * 3), ("argmax", ctypes.c_int32 * 3)]


class VktbAggPartial(ctypes.Structure):
    """vktb_agg_partial: combinable moments of one z-slab"""
    _fields_ = [("sum_d", ctypes.c_double), ("sum_d2", ctypes.c_double), ("prod_mant", ctypes.c_double),
                ("prod_exp", ctypes.c_int64), ("count", ctypes.c_uint64), ("min", ctypes.c_float),
                ("max", ctypes.c_float), ("shift", ctypes.c_float), ("has_min", ctypes.c_int32),
                ("has_max", ctypes.c_int32), ("argmin", ctypes.c_int32 * 3), ("argmax", ctypes.c_int32 * 3),
                ("pad", ctypes.c_int32)]


class VktbCamera(ctypes.Structure):
    _fields_ = [("eye", ctypes.c_float * 3), ("center", ctypes.c_float * 3), ("up", ctypes.c_float * 3),
                ("fovy_degrees", ctypes.c_float)]


READ_FN = ctypes.CFUNCTYPE(ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t)
WRITE_FN = ctypes.CFUNCTYPE(ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t)

# name -> (restype, argtypes); every symbol include/vktb/backend.h declares
_P = ctypes.POINTER
_VP = _P(VktbVol)
SIGNATURES = {
    "vktb_abi_version": (ctypes.c_int, []),
    "vktb_last_error": (ctypes.c_char_p, []),
    "vktb_device_count": (ctypes.c_int, [_P(ctypes.c_int)]),
    "vktb_set_device": (ctypes.c_int, [ctypes.c_int]),
    "vktb_get_device": (ctypes.c_int, [_P(ctypes.c_int)]),
    "vktb_set_devices": (ctypes.c_int, [_P(ctypes.c_int), ctypes.c_int]),
    "vktb_get_devices": (ctypes.c_int, [_P(ctypes.c_int), ctypes.c_int, _P(ctypes.c_int)]),
    "vktb_is_striped": (ctypes.c_int, [ctypes.c_void_p, _P(ctypes.c_int)]),
    "vktb_launch_count": (ctypes.c_uint64, []),
    "vktb_get_stream": (ctypes.c_void_p, []),
    "vktb_set_stream": (ctypes.c_int, [ctypes.c_void_p]),
    "vktb_sync": (ctypes.c_int, []),
    "vktb_alloc": (ctypes.c_int, [_P(ctypes.c_void_p), ctypes.c_size_t]),
    "vktb_free": (ctypes.c_int, [ctypes.c_void_p]),
    "vktb_memcpy": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]),
    "vktb_memset_range": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t]),
    "vktb_host_alloc": (ctypes.c_int, [_P(ctypes.c_void_p), ctypes.c_size_t]),
    "vktb_host_free": (ctypes.c_int, [ctypes.c_void_p]),
    "vktb_host_owns": (ctypes.c_int, [ctypes.c_void_p]),
    "vktb_mem_stats": (ctypes.c_int, [_P(ctypes.c_size_t)] * 4),
    "vktb_trim_pools": (ctypes.c_int, []),
    "vktb_fill_range": (ctypes.c_int, [_VP, I3, I3, ctypes.c_float]),
    "vktb_copy_range": (ctypes.c_int, [_VP, _VP, I3, I3, I3]),
    "vktb_copy_range_batch": (ctypes.c_int, [ctypes.c_int, _VP, _VP, ctypes.POINTER(ctypes.c_int32),
                                             ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]),
    "vktb_copy_range_batch_noalloc": (ctypes.c_int, [ctypes.c_int, _VP, _VP, ctypes.POINTER(ctypes.c_int32),
                                                     ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]),
    "vktb_copy_range_batch_noalloc_max": (ctypes.c_int, []),
    "vktb_scan_range": (ctypes.c_int, [_VP, _VP, I3, I3, I3]),
    "vktb_arith_range": (ctypes.c_int, [ctypes.c_int, _VP, _VP, _VP, I3, I3, I3]),
    "vktb_transform_range1": (ctypes.c_int, [_VP, I3, I3, ctypes.c_int]),
    "vktb_transform_range2": (ctypes.c_int, [_VP, _VP, I3, I3, ctypes.c_int]),
    "vktb_resample": (ctypes.c_int, [_VP, _VP, ctypes.c_int]),
    "vktb_resample_window": (ctypes.c_int, [_VP, _VP, ctypes.c_int, ctypes.c_int32, I3, ctypes.c_int32, I3]),
    "vktb_resample_window_src_range": (ctypes.c_int, [ctypes.c_int, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                                      ctypes.c_int32, _P(ctypes.c_int32), _P(ctypes.c_int32)]),
    "vktb_resample_window_src_range3": (ctypes.c_int, [ctypes.c_int, ctypes.c_int32, ctypes.c_int32, I3, I3,
                                                       _P(ctypes.c_int32), _P(ctypes.c_int32)]),
    "vktb_flip_range": (ctypes.c_int, [_VP, _VP, I3, I3, I3, ctypes.c_int]),
    "vktb_stream_in": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_size_t, READ_FN, ctypes.c_void_p, _P(ctypes.c_size_t)]),
    "vktb_stream_out": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_size_t, WRITE_FN, ctypes.c_void_p, _P(ctypes.c_size_t)]),
    "vktb_affine_range": (ctypes.c_int, [_VP, _VP, I3, I3, _P(ctypes.c_float), _P(ctypes.c_float)]),
    "vktb_apply_filter": (ctypes.c_int, [_VP, _VP, _P(ctypes.c_float), I3, ctypes.c_int]),
    "vktb_aggregates_range": (ctypes.c_int, [_VP, I3, I3, _P(VktbAggregates)]),
    "vktb_aggregates_partial": (ctypes.c_int, [_VP, I3, I3, _P(VktbAggPartial)]),
    "vktb_aggregates_finish": (ctypes.c_int, [_P(VktbAggPartial), ctypes.c_int, ctypes.c_uint64, _P(VktbAggregates)]),
    "vktb_render_raymarch": (ctypes.c_int, [_VP, _P(ctypes.c_float), _P(VktbCamera), ctypes.c_void_p, ctypes.c_int,
                                            ctypes.c_float, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                            ctypes.c_void_p]),
    "vktb_render_default_camera": (ctypes.c_int, [I3, _P(ctypes.c_float), _P(VktbCamera)]),
    "vktb_histogram_range": (ctypes.c_int, [_VP, I3, I3, ctypes.c_void_p, ctypes.c_size_t, _P(ctypes.c_uint64)]),
    "vktb_hosted_arith_range": (ctypes.c_int, [ctypes.c_int, _VP, _VP, _VP, I3, I3, I3]),
    "vktb_hosted_transform_range1": (ctypes.c_int, [_VP, I3, I3, ctypes.c_int]),
    "vktb_hosted_resample": (ctypes.c_int, [_VP, _VP, ctypes.c_int]),
    "vktb_unary_op_by_name": (ctypes.c_int, [ctypes.c_char_p, _P(ctypes.c_int)]),
    "vktb_binary_op_by_name": (ctypes.c_int, [ctypes.c_char_p, _P(ctypes.c_int)]),
    "vktb_register_unary_op_image": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_char_p, _P(ctypes.c_int)]),
    "vktb_register_binary_op_image": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_char_p, _P(ctypes.c_int)]),
    "vktb_register_unary_op_source": (ctypes.c_int, [ctypes.c_char_p, ctypes.c_char_p, _P(ctypes.c_int)]),
    "vktb_register_binary_op_source": (ctypes.c_int, [ctypes.c_char_p, ctypes.c_char_p, _P(ctypes.c_int)]),
    "vktb_nvrtc_available": (ctypes.c_int, []),
    "vktb_bind_unary_op": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "vktb_bind_binary_op": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "vktb_lookup_unary_op": (ctypes.c_int, [ctypes.c_void_p, _P(ctypes.c_int)]),
    "vktb_lookup_binary_op": (ctypes.c_int, [ctypes.c_void_p, _P(ctypes.c_int)]),
    "vktb_debug_div_proven": (ctypes.c_int, [ctypes.c_float, ctypes.c_float, _P(ctypes.c_int)]),
    "vktb_debug_map_unmap": (ctypes.c_int, [ctypes.c_int, ctypes.c_float, ctypes.c_float, ctypes.c_void_p,
                                            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t]),
}


class VktbError(RuntimeError):
    pass


_lib = None


def load() -> ctypes.CDLL:
    """Loads libvktb200.so; raises (never falls back) when it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise VktbError(f"{LIB_PATH} is missing: build it first (make -C volkit_b200/csrc); "
                            "volkit_b200 has no CPU fallback")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)   # AttributeError if the library does not export the symbol
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().vktb_last_error()
        raise VktbError(f"{what or 'vktb call'} failed ({rc}): {msg.decode() if msg else '?'}")


def vol(ptr: int, dims, fmt: int, lo: float = 0.0, hi: float = 1.0) -> VktbVol:
    v = VktbVol()
    v.data = ptr
    v.dims[:] = dims
    v.format = fmt
    v.lo = lo
    v.hi = hi
    return v
