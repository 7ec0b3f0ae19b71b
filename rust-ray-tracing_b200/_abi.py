"""ctypes binding of include/rtb200.h (librtb200.so).

The library is the product: if it is missing or does not load, importing this module raises — there is no
Python or CPU fallback for any compute entry point.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librtb200.so")

from ._abi_types import (OK, ERR_INVALID, ERR_CUDA, ERR_STATE, LAMBERTIAN, METAL, DIELECTRIC, Material, Sphere, Scene, Camera, Params, Stats,
                         STATIC_SCRATCH_BYTES)  # noqa: F401


# every symbol include/rtb200.h declares: name -> (restype, argtypes)
_P = C.c_void_p
_F = C.POINTER(C.c_float)
SYMBOLS = {
    "rtb200_create": (C.c_int, [C.c_int32, C.POINTER(_P)]),
    "rtb200_destroy": (None, [_P]),
    "rtb200_last_error": (C.c_char_p, []),
    "rtb200_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(_P)]),
    "rtb200_host_free": (None, [_P]),
    "rtb200_device_info": (C.c_int, [_P, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "rtb200_camera_init": (C.c_int, [C.POINTER(Camera), C.c_int32, C.c_int32, C.c_double, C.c_double,
                                     C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_double]),
    "rtb200_upload_scene": (C.c_int, [_P, C.POINTER(Scene)]),
    "rtb200_render": (C.c_int, [_P, C.POINTER(Camera), C.POINTER(Params), _P, _P, C.POINTER(Stats)]),
    "rtb200_intersect_batch": (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _P, C.c_int64, _P, _P]),
    "rtb200_intersect_batch_f64": (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _P, C.c_int64, _P, _P]),
    "rtb200_trace_device": (C.c_int, [_P, C.POINTER(Camera), C.POINTER(Params), _P, _P]),
    "rtb200_trace_device_static": (C.c_int, [_P, C.POINTER(Camera), C.POINTER(Params), _P, _P, C.POINTER(C.c_uint64)]),
    "rtb200_resolve_device": (C.c_int, [_P, _P, C.c_int64, C.c_int32, _P, _P]),
    "rtb200_intersect_batch_device": (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _P, C.c_int64, _P, _P, _P]),
    "rtb200_device_alloc": (C.c_int, [_P, C.c_size_t, C.POINTER(_P)]),
    "rtb200_device_free": (C.c_int, [_P, _P]),
    "rtb200_ipc_export": (C.c_int, [_P, _P, _P]),
    "rtb200_ipc_open": (C.c_int, [_P, _P, C.POINTER(_P)]),
    "rtb200_ipc_close": (C.c_int, [_P, _P]),
    "rtb200_resolve_peers_device": (C.c_int, [_P, C.POINTER(_P), C.c_int32, C.c_int64, C.c_int64, C.c_int32, _P, _P]),
    "rtb200_resolve_grid_device": (C.c_int, [_P, C.POINTER(_P), C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_int64, C.c_int32, _P, _P]),
    "rtb200_resolve_grid_sum_device": (C.c_int, [_P, C.POINTER(_P), C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_int64, C.c_int32, _P, _P, _P]),
    "rtb200_peer_barrier_device": (C.c_int, [_P, C.POINTER(_P), C.c_int32, C.c_int32, C.c_uint32, C.c_double, _P]),
    "rtb200_multi_create": (C.c_int, [C.POINTER(C.c_int32), C.c_int32, C.POINTER(_P)]),
    "rtb200_multi_destroy": (None, [_P]),
    "rtb200_multi_info": (C.c_int, [_P, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "rtb200_multi_upload_scene": (C.c_int, [_P, C.POINTER(Scene)]),
    "rtb200_multi_set_hit_mode": (C.c_int, [_P, C.c_int32]),
    "rtb200_multi_set_progress": (C.c_int, [_P, C.c_void_p, C.c_void_p, C.c_int32]),
    "rtb200_multi_render": (C.c_int, [_P, C.POINTER(Camera), C.POINTER(Params), C.c_int32, _P, _P, C.POINTER(Stats)]),
    "rtb200_read_counters": (C.c_int, [_P, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "rtb200_launch_count": (C.c_int64, [_P]),
    "rtb200_trace_kernel_stats": (C.c_int, [_P, C.c_int32, C.POINTER(C.c_double), C.POINTER(C.c_int32), C.POINTER(C.c_char_p), C.POINTER(C.c_int32)]),
    "rtb200_set_hit_mode": (C.c_int32, [_P, C.c_int32]),
    "rtb200_set_progress": (C.c_int32, [_P, C.c_void_p, C.c_void_p, C.c_int32]),
    "rtb200_fp32_peak": (C.c_int, [_P, C.c_int32, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "rtb200_filter_check": (C.c_int, [_P, C.c_uint64, C.c_int32, C.c_uint32, C.POINTER(C.c_double), C.POINTER(C.c_uint64), C.POINTER(C.c_double)]),
}


def load(path: str = LIB_PATH) -> C.CDLL:
    if not os.path.exists(path):
        raise ImportError(f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(rtb200 has no CPU fallback)")
    lib = C.CDLL(path)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)   # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    return lib


lib = load()


class Rtb200Error(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"rtb200 error {code}: {message}")
        self.code = code
        self.message = message


def check(rc: int) -> None:
    if rc != OK:
        raise Rtb200Error(rc, (lib.rtb200_last_error() or b"").decode("utf-8", "replace"))
