"""ctypes wrapper over oracle/liboracle.so — the CPU restatement used as the parity checker.
Test infrastructure only: nothing under rust-ray-tracing_b200/ imports this."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "liboracle.so")


def build(force: bool = False) -> None:
    src = [os.path.join(ORACLE_DIR, f) for f in ("oracle_capi.cpp", "rt_oracle.hpp", "cpu_ref.cpp")]
    stale = not os.path.exists(LIB) or any(os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(LIB) for s in src)
    if force or stale:
        subprocess.run(["make", "-C", ORACLE_DIR], check=True, capture_output=True)


build()
_lib = C.CDLL(LIB)
_D = C.POINTER(C.c_double)
_I = C.POINTER(C.c_int32)
_U = C.POINTER(C.c_uint32)


def _mode(f32) -> int:
    """The oracle's scalar mode.  False: f64, reference-faithful.  True: the GPU comparand = MIXED (FP32 path state, the
    closest-hit sweep in f64 on the f64 scene, the root rounded once to FP32 — what the CUDA kernels do).  "pure": FP32
    throughout (what FP32 hit arithmetic would cost; no product path runs it)."""
    if f32 == "pure":
        return 2
    return 1 if f32 else 0


def _d(a):
    return a.ctypes.data_as(_D)


def _i(a):
    return a.ctypes.data_as(_I)


def philox(ctr, key):
    c = (C.c_uint32 * 4)(*ctr)
    k = (C.c_uint32 * 2)(*key)
    o = (C.c_uint32 * 4)()
    _lib.oracle_philox4x32_10(c, k, o)
    return [int(x) for x in o]


def cossin_2pi(v: float, f32: bool):
    cs, sn = C.c_double(), C.c_double()
    _lib.oracle_cossin_2pi(C.c_double(v), _mode(f32), C.byref(cs), C.byref(sn))
    return cs.value, sn.value


def unit_vector_from(u: float, v: float, f32: bool):
    out = (C.c_double * 3)()
    _lib.oracle_unit_vector_from(C.c_double(u), C.c_double(v), _mode(f32), out)
    return np.array(out[:])


def camera_new(width, height, focal, vfov, center, look_at, up, defocus):
    out = (C.c_double * 18)()
    d3 = C.c_double * 3
    _lib.oracle_camera_new(int(width), int(height), C.c_double(focal), C.c_double(vfov), d3(*center), d3(*look_at), d3(*up),
                           C.c_double(defocus), out)
    return np.array(out[:])


def camera_get_ray(width, height, cam18, col, row, rnd4, f32: bool):
    out = (C.c_double * 6)()
    cam = np.ascontiguousarray(cam18, dtype=np.float64)
    r = (C.c_uint32 * 4)(*rnd4)
    _lib.oracle_camera_get_ray(int(width), int(height), _d(cam), int(col), int(row), r, _mode(f32), out)
    return np.array(out[:])


def set_both_roots(on: bool) -> None:
    """Hit mode of later intersect()/render() calls: False = the reference's packed path (near root only)."""
    _lib.oracle_set_both_roots(1 if on else 0)


def intersect(spheres, rays, enabled=None, f32: bool = True):
    """spheres [n,4] (cx,cy,cz,r) f64; rays [m,6] f64 -> (idx[m] i32, t[m] f64)."""
    sp = np.ascontiguousarray(spheres, dtype=np.float64).reshape(-1, 4)
    ry = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 6)
    m = ry.shape[0]
    idx = np.empty(m, dtype=np.int32)
    t = np.empty(m, dtype=np.float64)
    en = None
    if enabled is not None:
        en = np.ascontiguousarray(enabled, dtype=np.int32)
    _lib.oracle_intersect(_d(sp), int(sp.shape[0]), _d(ry), _i(en) if en is not None else None, C.c_int64(m), _mode(f32), _i(idx), _d(t))
    return idx, t


def hit_and_scatter(sphere4, kind, hollow, albedo, param, ray6, rnd4, f32: bool):
    out = (C.c_double * 18)()
    d4 = (C.c_double * 4)(*sphere4)
    d3 = (C.c_double * 3)(*albedo)
    d6 = (C.c_double * 6)(*ray6)
    r = (C.c_uint32 * 4)(*rnd4)
    _lib.oracle_hit_and_scatter(d4, int(kind), int(hollow), d3, C.c_double(param), d6, r, _mode(f32), out)
    o = np.array(out[:])
    return {"hit": bool(o[0]), "t": o[1], "p": o[2:5], "n": o[5:8], "front": bool(o[8]), "attenuation": o[9:12],
            "origin": o[12:15], "direction": o[15:18]}


def sky(d3, f32: bool):
    out = (C.c_double * 3)()
    _lib.oracle_sky((C.c_double * 3)(*d3), _mode(f32), out)
    return np.array(out[:])


_lib.oracle_quantise.restype = C.c_int32
_lib.oracle_reflectance.restype = C.c_double


def quantise(c: float, f32: bool) -> int:
    return int(_lib.oracle_quantise(C.c_double(c), _mode(f32)))


def reflectance(cosine: float, ior: float, f32: bool) -> float:
    return float(_lib.oracle_reflectance(C.c_double(cosine), C.c_double(ior), _mode(f32)))


def render(scene_arrays, width, height, cam18, spp, depth, seed=1, sample_range=None, row_range=None, threads=None,
           f32: bool = True, want_rgb8: bool = True):
    """Per-path (v1/v3) render.  Returns (sum[H,W,3] f64 over the sample range, rgb8[H,W,3] u8, segments)."""
    spheres, mat_index, kind, hollow, albedo, param = scene_arrays
    s0, s1 = sample_range if sample_range is not None else (0, spp)
    r0, r1 = row_range if row_range is not None else (0, height)
    if threads is None:
        threads = os.cpu_count() or 1
    sp = np.ascontiguousarray(spheres, dtype=np.float64)
    mi = np.ascontiguousarray(mat_index, dtype=np.int32)
    kd = np.ascontiguousarray(kind, dtype=np.int32)
    hl = np.ascontiguousarray(hollow, dtype=np.int32)
    al = np.ascontiguousarray(albedo, dtype=np.float64)
    pr = np.ascontiguousarray(param, dtype=np.float64)
    cam = np.ascontiguousarray(cam18, dtype=np.float64)
    out_sum = np.zeros((height, width, 3), dtype=np.float64)
    out_rgb = np.zeros((height, width, 3), dtype=np.uint8)
    seg = C.c_uint64(0)
    _lib.oracle_render(_d(sp), _i(mi), int(sp.shape[0]), _i(kd), _i(hl), _d(al), _d(pr), int(kd.shape[0]), int(width), int(height),
                       _d(cam), int(s0), int(s1), int(spp), int(depth), C.c_uint32(seed & 0xFFFFFFFF), C.c_uint32((seed >> 32) & 0xFFFFFFFF),
                       int(r0), int(r1), int(threads), _mode(f32), _d(out_sum),
                       out_rgb.ctypes.data_as(C.POINTER(C.c_uint8)) if want_rgb8 else None, C.byref(seg))
    return out_sum, out_rgb, int(seg.value)


# ---- CPU baseline port of the reference's live renderer (oracle/cpu_ref.cpp) ---------------------------------
CPUREF = os.path.join(ORACLE_DIR, "libcpuref.so")
_cpuref = C.CDLL(CPUREF)
_cpuref.cpuref_render.restype = C.c_double


def cpuref_render(scene_arrays, width, height, cam18, spp, depth, threads=0, block_size=128, seed=1, row_range=None, want_rgb8=True):
    """TileRenderer + render_vectorized2 + trace_vectorized2 (4 x f64 AVX2 lanes, own PRNG).
    Returns (rgb8[H,W,3], seconds, lane_tests, segments)."""
    spheres, mat_index, kind, hollow, albedo, param = scene_arrays
    r0, r1 = row_range if row_range is not None else (0, height)
    sp = np.ascontiguousarray(spheres, dtype=np.float64)
    mi = np.ascontiguousarray(mat_index, dtype=np.int32)
    kd = np.ascontiguousarray(kind, dtype=np.int32)
    hl = np.ascontiguousarray(hollow, dtype=np.int32)
    al = np.ascontiguousarray(albedo, dtype=np.float64)
    pr = np.ascontiguousarray(param, dtype=np.float64)
    cam = np.ascontiguousarray(cam18, dtype=np.float64)
    rgb = np.zeros((height, width, 3), dtype=np.uint8)
    tests, seg = C.c_uint64(0), C.c_uint64(0)
    sec = _cpuref.cpuref_render(_d(sp), _i(mi), int(sp.shape[0]), _i(kd), _i(hl), _d(al), _d(pr), int(kd.shape[0]), int(width), int(height),
                                _d(cam), int(spp), int(depth), int(threads), int(block_size), C.c_uint64(seed), int(r0), int(r1),
                                rgb.ctypes.data_as(C.POINTER(C.c_uint8)) if want_rgb8 else None, C.byref(tests), C.byref(seg))
    return rgb, float(sec), int(tests.value), int(seg.value)
