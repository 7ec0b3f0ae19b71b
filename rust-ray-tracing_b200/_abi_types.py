"""Plain-ctypes mirrors of the structs and constants of include/rtb200.h.  Pure host code: importing this module does NOT
load librtb200.so (the scene generators and the TOML loader need only these; everything that computes goes through _abi)."""
from __future__ import annotations

import ctypes as C

OK, ERR_INVALID, ERR_CUDA, ERR_STATE = 0, -1, -2, -3
LAMBERTIAN, METAL, DIELECTRIC = 0, 1, 2
STATIC_SCRATCH_BYTES = 33554432   # RTB200_STATIC_SCRATCH_BYTES: capacity of the static scratch of rtb200_trace_device_static


class Material(C.Structure):
    _fields_ = [("kind", C.c_int32), ("hollow", C.c_int32), ("albedo", C.c_double * 3), ("param", C.c_double)]


class Sphere(C.Structure):
    _fields_ = [("center", C.c_double * 3), ("radius", C.c_double), ("material", C.c_int32), ("_pad", C.c_int32)]


class Scene(C.Structure):
    _fields_ = [("spheres", C.POINTER(Sphere)), ("n_spheres", C.c_int32), ("materials", C.POINTER(Material)),
                ("n_materials", C.c_int32)]


class Camera(C.Structure):
    _fields_ = [("image_width", C.c_int32), ("image_height", C.c_int32), ("center", C.c_double * 3),
                ("viewport_upper_left_corner", C.c_double * 3), ("viewport_u", C.c_double * 3),
                ("viewport_v", C.c_double * 3), ("defocus_u", C.c_double * 3), ("defocus_v", C.c_double * 3)]


class Params(C.Structure):
    _fields_ = [("max_bounces", C.c_int32), ("samples_per_pixel", C.c_int32), ("seed", C.c_uint64),
                ("sample_begin", C.c_int32), ("sample_end", C.c_int32), ("row_begin", C.c_int32),
                ("row_end", C.c_int32), ("flags", C.c_int32), ("row_stride", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("seconds", C.c_double), ("pixels_rendered", C.c_uint64), ("pixels_per_second", C.c_double),
                ("paths", C.c_uint64), ("segments", C.c_uint64), ("sphere_tests", C.c_uint64),
                ("trace_ms", C.c_double), ("resolve_ms", C.c_double), ("h2d_ms", C.c_double), ("d2h_ms", C.c_double)]
