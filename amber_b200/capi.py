"""ctypes binding of the C ABI in ``include/amber_cuda.h``.

This is plumbing: it declares the POD structs and argument types of
``libamber_cuda.so`` so that the Python host mirror (``amber_b200.render``),
the tests and ``bench.py`` call the product exactly as a foreign host would.
There is no fallback: if the shared library is missing, loading raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libamber_cuda.so")

ABI_VERSION = 2
SAMPLES_IN_FILM = 0xFFFFFFFFFFFFFFFF
CAST_STATIC_MAX = 65536

PRIM_TRIANGLE, PRIM_SPHERE = 0, 1
MAT_LAMBERTIAN, MAT_PHONG, MAT_SPECULAR, MAT_REFRACTION, MAT_DIFFUSE_LIGHT, MAT_EYE = range(6)
ESTIMATOR_REFERENCE, ESTIMATOR_NEE, ESTIMATOR_LIGHT_TRACING, ESTIMATOR_BDPT = 0, 1, 2, 3
LENS_THIN, LENS_PINHOLE = 0, 1

# every symbol include/amber_cuda.h declares (checked by tests/test_abi.py)
SYMBOLS = (
    "amber_cuda_abi_version",
    "amber_cuda_last_error",
    "amber_cuda_create",
    "amber_cuda_destroy",
    "amber_cuda_upload_scene",
    "amber_cuda_set_sensor",
    "amber_cuda_film_clear",
    "amber_cuda_render",
    "amber_cuda_render_async",
    "amber_cuda_wait",
    "amber_cuda_pending",
    "amber_cuda_wait_pending",
    "amber_cuda_film_reduce",
    "amber_cuda_resolve",
    "amber_cuda_resolve_ldr",
    "amber_cuda_film_device",
    "amber_cuda_film_set_samples",
    "amber_cuda_cast",
    "amber_cuda_cast_device",
    "amber_cuda_cast_static",
    "amber_cuda_sample_light",
    "amber_cuda_eye_rays",
    "amber_cuda_sample_importance",
    "amber_cuda_light_rays",
    "amber_cuda_lens_response",
    "amber_cuda_normalize",
    "amber_cuda_get_stats",
    "amber_cuda_set_profiling",
)


class Primitive(C.Structure):
    _fields_ = [("kind", C.c_int32), ("material", C.c_int32), ("data", C.c_float * 9)]


class Material(C.Structure):
    _fields_ = [("kind", C.c_int32), ("rgb", C.c_float * 3), ("param", C.c_float)]


class Lens(C.Structure):
    _fields_ = [
        ("transform", C.c_float * 16),
        ("focal_length", C.c_float),
        ("focus_distance", C.c_float),
        ("radius", C.c_float),
        ("n_blades", C.c_int32),
        ("kind", C.c_int32),
        ("sensor_distance", C.c_float),
    ]


class SceneDesc(C.Structure):
    _fields_ = [
        ("n_primitives", C.c_uint32),
        ("primitives", C.POINTER(Primitive)),
        ("n_materials", C.c_uint32),
        ("materials", C.POINTER(Material)),
        ("lens", Lens),
    ]


class RenderParams(C.Structure):
    _fields_ = [
        ("sample_begin", C.c_uint64),
        ("sample_count", C.c_uint32),
        ("max_depth", C.c_uint32),
        ("seed", C.c_uint64),
        ("estimator", C.c_int32),
        ("pool_paths", C.c_uint32),
        ("pixel_begin", C.c_uint32),
        ("pixel_count", C.c_uint32),
        ("sample_stride", C.c_uint32),
        ("reserved", C.c_uint32),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("paths", C.c_uint64),
        ("closest_rays", C.c_uint64),
        ("shadow_rays", C.c_uint64),
        ("truncated_paths", C.c_uint64),
        ("phong_capped", C.c_uint64),
        ("kernel_launches", C.c_uint64),
        ("last_render_ms", C.c_double),
        ("extend_ms", C.c_double),
        ("shade_ms", C.c_double),
        ("connect_ms", C.c_double),
        ("raygen_ms", C.c_double),
        ("extend_launches", C.c_uint64),
        ("accel", C.c_uint32),
        ("bvh_nodes", C.c_uint32),
        ("light_splats", C.c_uint64),
        ("bvh_node_visits", C.c_uint64),
        ("bvh_triangle_tests", C.c_uint64),
        ("bvh_node_visits_connect", C.c_uint64),
        ("bvh_triangle_tests_connect", C.c_uint64),
        ("bvh_node_bytes", C.c_uint32),
        ("bvh_triangle_bytes", C.c_uint32),
        ("chained_renders", C.c_uint64),
        ("scene_build_ms", C.c_double),
        ("scene_upload_ms", C.c_double),
    ]


class AmberCudaError(RuntimeError):
    """A C-ABI call returned non-zero (the message is amber_cuda_last_error)."""


_lib = None


def load(path: str | None = None) -> C.CDLL:
    """Load libamber_cuda.so and declare prototypes.  Raises if it is not built."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or LIB_PATH
    if not os.path.exists(path):
        raise AmberCudaError(
            f"{path} is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)"
        )
    lib = C.CDLL(path)
    vp, u64, u32, i32, f32p = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int32, C.POINTER(C.c_float)
    lib.amber_cuda_abi_version.restype = C.c_int
    lib.amber_cuda_abi_version.argtypes = []
    lib.amber_cuda_last_error.restype = C.c_char_p
    lib.amber_cuda_last_error.argtypes = [vp]
    protos = {
        "amber_cuda_create": [C.POINTER(vp), C.c_int],
        "amber_cuda_destroy": [vp],
        "amber_cuda_upload_scene": [vp, C.POINTER(SceneDesc)],
        "amber_cuda_set_sensor": [vp, u32, u32, C.c_float, C.c_float],
        "amber_cuda_film_clear": [vp],
        "amber_cuda_render": [vp, C.POINTER(RenderParams)],
        "amber_cuda_render_async": [vp, C.POINTER(RenderParams)],
        "amber_cuda_wait": [vp],
        "amber_cuda_pending": [vp],
        "amber_cuda_wait_pending": [vp, C.c_int],
        "amber_cuda_film_reduce": [C.POINTER(vp), C.c_int],
        "amber_cuda_resolve": [vp, vp, C.POINTER(u64)],
        "amber_cuda_resolve_ldr": [vp, vp, C.c_int, C.POINTER(u64)],
        "amber_cuda_film_device": [vp, C.POINTER(vp), C.POINTER(u64), C.POINTER(u64)],
        "amber_cuda_film_set_samples": [vp, u64],
        "amber_cuda_cast": [vp, vp, vp, u64, vp, vp],
        "amber_cuda_cast_device": [vp, vp, vp, u64, vp, f32p],
        "amber_cuda_cast_static": [vp, vp, vp, u64, vp, vp],
        "amber_cuda_sample_light": [vp, i32, u64, vp, vp, vp, vp, vp, vp],
        "amber_cuda_eye_rays": [vp, u64, vp, vp, vp, vp, vp, vp],
        "amber_cuda_sample_importance": [vp, i32, u64, vp, vp, vp, vp, vp, vp],
        "amber_cuda_light_rays": [vp, u64, vp, vp, vp, vp],
        "amber_cuda_lens_response": [vp, u64, vp, vp, vp, vp],
        "amber_cuda_normalize": [vp, u64, vp, vp, vp],
        "amber_cuda_get_stats": [vp, C.POINTER(Stats)],
        "amber_cuda_set_profiling": [vp, C.c_int],
    }
    for name, argtypes in protos.items():
        fn = getattr(lib, name)
        fn.restype = C.c_int
        fn.argtypes = argtypes
    if lib.amber_cuda_abi_version() != ABI_VERSION:
        raise AmberCudaError("libamber_cuda.so ABI version mismatch")
    if path == LIB_PATH:
        _lib = lib
    return lib


def check(lib: C.CDLL, ctx, code: int) -> None:
    if code != 0:
        msg = lib.amber_cuda_last_error(ctx)
        raise AmberCudaError(f"amber_cuda error {code}: {msg.decode() if msg else '?'}")
