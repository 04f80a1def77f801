"""TEST INFRASTRUCTURE: ctypes wrappers for the two checkers under oracle/.

  * ``Reference`` -- oracle/_ref/libamber_ref.so, the UNMODIFIED reference compiled
    from /root/reference (oracle/Makefile, ``make ref``) behind oracle/ref_harness.cc.
  * ``Oracle``    -- oracle/libamber_oracle.so, the plain-C restatement
    (oracle/amber_oracle.c).

Only tests/, ``__graft_entry__.smoke()`` and bench.py's cpu_baseline / --impl
reference legs may import this module.  The product (amber_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(_HERE, "_ref", "libamber_ref.so")
ORACLE_LIB = os.path.join(_HERE, "libamber_oracle.so")
REFERENCE_ROOT = "/root/reference"


def build(ref: bool = True) -> None:
    """Compile the C restatement and, when /root/reference is present, the reference."""
    subprocess.check_call(["make", "-C", _HERE, "oracle"], stdout=subprocess.DEVNULL)
    if ref and os.path.isdir(REFERENCE_ROOT):
        subprocess.check_call(["make", "-C", _HERE, "ref", "-j8"], stdout=subprocess.DEVNULL)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Reference:
    """The unmodified reference behind a flat scene description."""

    def __init__(self, flat_scene=None, builtin_cornel_box=False):
        if not os.path.exists(REF_LIB):
            raise FileNotFoundError(REF_LIB)
        self.lib = C.CDLL(REF_LIB)
        self.lib.ref_scene_create.restype = C.c_void_p
        self.lib.ref_scene_create_cornel_box.restype = C.c_void_p
        self.flat = flat_scene
        if builtin_cornel_box:
            self.handle = C.c_void_p(self.lib.ref_scene_create_cornel_box())
        else:
            self.handle = C.c_void_p(self.lib.ref_scene_create(C.byref(flat_scene.desc)))

    def close(self):
        if self.handle:
            self.lib.ref_scene_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def cast(self, origins, directions, accel=0, threads=8):
        o, d = _f32(origins), _f32(directions)
        n = o.shape[0]
        prim = np.empty(n, dtype=np.int32)
        t = np.empty(n, dtype=np.float32)
        rc = self.lib.ref_cast(self.handle, C.c_int(accel), C.c_uint64(n), _ptr(o), _ptr(d),
                               _ptr(prim), _ptr(t), C.c_int(threads))
        assert rc == 0
        return prim, t

    def sample_light(self, material, normals, dirs_out, xi8):
        n_, wo, xi = _f32(normals), _f32(dirs_out), _f32(xi8)
        n = n_.shape[0]
        wi = np.empty((n, 3), dtype=np.float32)
        w = np.empty((n, 3), dtype=np.float32)
        draws = np.empty(n, dtype=np.int32)
        rc = self.lib.ref_sample_light(C.byref(material), C.c_uint64(n), _ptr(n_), _ptr(wo), _ptr(xi),
                                       _ptr(wi), _ptr(w), _ptr(draws))
        assert rc == 0
        return wi, w, draws

    def bsdf(self, material, normals, dirs_out, dirs_in):
        n_, wo, wi = _f32(normals), _f32(dirs_out), _f32(dirs_in)
        n = n_.shape[0]
        f = np.empty((n, 3), dtype=np.float32)
        le = np.empty((n, 3), dtype=np.float32)
        rc = self.lib.ref_bsdf(C.byref(material), C.c_uint64(n), _ptr(n_), _ptr(wo), _ptr(wi), _ptr(f), _ptr(le))
        assert rc == 0
        return f, le

    def triangle_sample(self, v9, xi2):
        v, xi = _f32(v9), _f32(xi2)
        n = xi.shape[0]
        p = np.empty((n, 3), dtype=np.float32)
        nn = np.empty((n, 3), dtype=np.float32)
        area = C.c_float()
        rc = self.lib.ref_triangle_sample(_ptr(v), C.c_uint64(n), _ptr(xi), _ptr(p), _ptr(nn), C.byref(area))
        assert rc == 0
        return p, nn, area.value

    def eye_rays(self, w, h, film_w, film_h, px, py, xi5):
        px_ = np.ascontiguousarray(px, dtype=np.uint32)
        py_ = np.ascontiguousarray(py, dtype=np.uint32)
        xi = _f32(xi5)
        n = px_.shape[0]
        o = np.empty((n, 3), dtype=np.float32)
        d = np.empty((n, 3), dtype=np.float32)
        wt = np.empty(n, dtype=np.float32)
        rc = self.lib.ref_eye_rays(self.handle, C.c_uint32(w), C.c_uint32(h), C.c_float(film_w),
                                   C.c_float(film_h), C.c_uint64(n), _ptr(px_), _ptr(py_), _ptr(xi),
                                   _ptr(o), _ptr(d), _ptr(wt))
        assert rc == 0
        return o, d, wt

    def render(self, w, h, film_w, film_h, spp, threads, algo=0, want_image=True):
        """The reference's own PT (algo 0) / BDPT (algo 1).  Returns (rgb, casts, seconds)."""
        rgb = np.empty((h, w, 3), dtype=np.float32) if want_image else None
        casts = C.c_uint64()
        seconds = C.c_double()
        rc = self.lib.ref_render(self.handle, C.c_int(algo), C.c_uint32(w), C.c_uint32(h),
                                 C.c_float(film_w), C.c_float(film_h), C.c_uint32(spp),
                                 C.c_uint32(threads), _ptr(rgb) if want_image else None,
                                 C.byref(casts), C.byref(seconds))
        assert rc == 0
        return rgb, casts.value, seconds.value

    def render_stream(self, w, h, film_w, film_h, sample_begin, sample_count, seed, max_depth=0,
                      threads=8):
        rgb = np.empty((h, w, 3), dtype=np.float32)
        stats = (C.c_uint64 * 2)()
        rc = self.lib.ref_render_stream(self.handle, C.c_uint32(w), C.c_uint32(h), C.c_float(film_w),
                                        C.c_float(film_h), C.c_uint64(sample_begin),
                                        C.c_uint32(sample_count), C.c_uint64(seed),
                                        C.c_uint32(max_depth), C.c_uint32(threads), _ptr(rgb), stats)
        assert rc == 0
        return rgb, {"rays": stats[0], "truncated": stats[1]}

    # ---- light tracing (row N2) ----
    def sample_importance(self, material, normals, dirs_out, xi8):
        n_, wo, xi = _f32(normals), _f32(dirs_out), _f32(xi8)
        n = n_.shape[0]
        wi = np.empty((n, 3), dtype=np.float32)
        w = np.empty((n, 3), dtype=np.float32)
        draws = np.empty(n, dtype=np.int32)
        rc = self.lib.ref_sample_importance(C.byref(material), C.c_uint64(n), _ptr(n_), _ptr(wo), _ptr(xi),
                                            _ptr(wi), _ptr(w), _ptr(draws))
        assert rc == 0
        return wi, w, draws

    def lens_response(self, w, h, film_w, film_h, positions, dirs_out):
        p, d = _f32(positions), _f32(dirs_out)
        n = p.shape[0]
        pixel = np.empty(n, dtype=np.int32)
        value = np.empty(n, dtype=np.float32)
        rc = self.lib.ref_lens_response(self.handle, C.c_uint32(w), C.c_uint32(h), C.c_float(film_w),
                                        C.c_float(film_h), C.c_uint64(n), _ptr(p), _ptr(d), _ptr(pixel), _ptr(value))
        assert rc == 0
        return pixel, value

    def light_rays(self, xi5):
        xi = _f32(xi5)
        n = xi.shape[0]
        o = np.empty((n, 3), dtype=np.float32)
        d = np.empty((n, 3), dtype=np.float32)
        wt = np.empty((n, 3), dtype=np.float32)
        rc = self.lib.ref_light_rays(self.handle, C.c_uint64(n), _ptr(xi), _ptr(o), _ptr(d), _ptr(wt))
        assert rc == 0
        return o, d, wt

    def render_lt_stream(self, w, h, film_w, film_h, sample_begin, sample_count, seed, max_depth=0,
                         threads=8):
        """algorithm_lt.cc's loop over the reference Scene API on the device's Philox stream."""
        rgb = np.empty((h, w, 3), dtype=np.float32)
        stats = (C.c_uint64 * 3)()
        rc = self.lib.ref_render_lt_stream(self.handle, C.c_uint32(w), C.c_uint32(h), C.c_float(film_w),
                                           C.c_float(film_h), C.c_uint64(sample_begin),
                                           C.c_uint32(sample_count), C.c_uint64(seed),
                                           C.c_uint32(max_depth), C.c_uint32(threads), _ptr(rgb), stats)
        assert rc == 0
        return rgb, {"rays": stats[0], "truncated": stats[1], "splats": stats[2]}

    def render_bdpt_stream(self, w, h, film_w, film_h, sample_begin, sample_count, seed, max_vertices=0,
                           threads=8):
        """The reference's own GenerateLightSubpath / GenerateEyeSubpath / Combine driven per pixel
        as algorithm_bdpt.cc:116-148 does, each subpath on its sequential Philox stream."""
        rgb = np.empty((h, w, 3), dtype=np.float32)
        stats = (C.c_uint64 * 3)()
        rc = self.lib.ref_render_bdpt_stream(self.handle, C.c_uint32(w), C.c_uint32(h), C.c_float(film_w),
                                             C.c_float(film_h), C.c_uint64(sample_begin),
                                             C.c_uint32(sample_count), C.c_uint64(seed),
                                             C.c_uint32(max_vertices), C.c_uint32(threads), _ptr(rgb), stats)
        assert rc == 0
        return rgb, {"light_events": stats[0], "eye_events": stats[1], "splats": stats[2]}


def reference_tonemap(rgb):
    """Gamma()(Filmic()(image)) of the unmodified reference; rgb (h, w, 3) float32 -> uint8."""
    lib = C.CDLL(REF_LIB)
    rgb = _f32(rgb)
    h, w = rgb.shape[:2]
    out = np.empty((h, w, 3), dtype=np.uint8)
    assert lib.ref_tonemap(C.c_uint32(w), C.c_uint32(h), _ptr(rgb), _ptr(out)) == 0
    return out


def oracle_tonemap(rgb):
    """The C restatement of the same post-process (oracle_tonemap)."""
    if not os.path.exists(ORACLE_LIB):
        build(ref=False)
    lib = C.CDLL(ORACLE_LIB)
    rgb = _f32(rgb)
    out = np.empty(rgb.shape, dtype=np.uint8)
    lib.oracle_tonemap(C.c_uint64(rgb.size), _ptr(rgb), _ptr(out))
    return out


def reference_available() -> bool:
    return os.path.exists(REF_LIB)


class Oracle:
    """The plain-C restatement (oracle/amber_oracle.c) behind a flat scene description."""

    def __init__(self, flat_scene):
        if not os.path.exists(ORACLE_LIB):
            build(ref=False)
        self.lib = C.CDLL(ORACLE_LIB)
        self.lib.oracle_scene_create.restype = C.c_void_p
        self.flat = flat_scene
        self.handle = C.c_void_p(self.lib.oracle_scene_create(C.byref(flat_scene.desc)))

    def close(self):
        if self.handle:
            self.lib.oracle_scene_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def cast(self, origins, directions, threads=8):
        o, d = _f32(origins), _f32(directions)
        n = o.shape[0]
        prim = np.empty(n, dtype=np.int32)
        t = np.empty(n, dtype=np.float32)
        self.lib.oracle_cast(self.handle, C.c_uint64(n), _ptr(o), _ptr(d), _ptr(prim), _ptr(t), C.c_int(threads))
        return prim, t

    def sample_light(self, material, normals, dirs_out, xi8):
        n_, wo, xi = _f32(normals), _f32(dirs_out), _f32(xi8)
        n = n_.shape[0]
        wi = np.empty((n, 3), dtype=np.float32)
        w = np.empty((n, 3), dtype=np.float32)
        draws = np.empty(n, dtype=np.int32)
        self.lib.oracle_sample_light(C.byref(material), C.c_uint64(n), _ptr(n_), _ptr(wo), _ptr(xi),
                                     _ptr(wi), _ptr(w), _ptr(draws))
        return wi, w, draws

    def bsdf(self, material, normals, dirs_out, dirs_in):
        n_, wo, wi = _f32(normals), _f32(dirs_out), _f32(dirs_in)
        n = n_.shape[0]
        f = np.empty((n, 3), dtype=np.float32)
        le = np.empty((n, 3), dtype=np.float32)
        self.lib.oracle_bsdf(C.byref(material), C.c_uint64(n), _ptr(n_), _ptr(wo), _ptr(wi), _ptr(f), _ptr(le))
        return f, le

    def eye_rays(self, w, h, film_w, film_h, px, py, xi5):
        px_ = np.ascontiguousarray(px, dtype=np.uint32)
        py_ = np.ascontiguousarray(py, dtype=np.uint32)
        xi = _f32(xi5)
        n = px_.shape[0]
        o = np.empty((n, 3), dtype=np.float32)
        d = np.empty((n, 3), dtype=np.float32)
        wt = np.empty(n, dtype=np.float32)
        self.lib.oracle_eye_rays(self.handle, C.c_uint32(w), C.c_uint32(h), C.c_float(film_w),
                                 C.c_float(film_h), C.c_uint64(n), _ptr(px_), _ptr(py_), _ptr(xi),
                                 _ptr(o), _ptr(d), _ptr(wt))
        return o, d, wt

    def render(self, w, h, film_w, film_h, sample_begin, sample_count, seed, max_depth=0,
               estimator=0, threads=8):
        rgb = np.empty((h, w, 3), dtype=np.float32)
        stats = (C.c_uint64 * 4)()
        self.lib.oracle_render(self.handle, C.c_uint32(w), C.c_uint32(h), C.c_float(film_w),
                               C.c_float(film_h), C.c_uint64(sample_begin), C.c_uint32(sample_count),
                               C.c_uint64(seed), C.c_uint32(max_depth), C.c_int(estimator),
                               C.c_int(threads), _ptr(rgb), stats)
        return rgb, {"closest_rays": stats[0], "shadow_rays": stats[1], "truncated": stats[2],
                     "paths": stats[3]}

    # ---- light tracing (row N2) ----
    def sample_importance(self, material, normals, dirs_out, xi8):
        n_, wo, xi = _f32(normals), _f32(dirs_out), _f32(xi8)
        n = n_.shape[0]
        wi = np.empty((n, 3), dtype=np.float32)
        w = np.empty((n, 3), dtype=np.float32)
        draws = np.empty(n, dtype=np.int32)
        self.lib.oracle_sample_importance(C.byref(material), C.c_uint64(n), _ptr(n_), _ptr(wo), _ptr(xi),
                                          _ptr(wi), _ptr(w), _ptr(draws))
        return wi, w, draws

    def lens_response(self, w, h, film_w, film_h, positions, dirs_out):
        p, d = _f32(positions), _f32(dirs_out)
        n = p.shape[0]
        pixel = np.empty(n, dtype=np.int32)
        value = np.empty(n, dtype=np.float32)
        self.lib.oracle_lens_response(self.handle, C.c_uint32(w), C.c_uint32(h), C.c_float(film_w),
                                      C.c_float(film_h), C.c_uint64(n), _ptr(p), _ptr(d), _ptr(pixel), _ptr(value))
        return pixel, value

    def light_rays(self, xi5):
        xi = _f32(xi5)
        n = xi.shape[0]
        o = np.empty((n, 3), dtype=np.float32)
        d = np.empty((n, 3), dtype=np.float32)
        wt = np.empty((n, 3), dtype=np.float32)
        self.lib.oracle_light_rays(self.handle, C.c_uint64(n), _ptr(xi), _ptr(o), _ptr(d), _ptr(wt))
        return o, d, wt

    def render_lt(self, w, h, film_w, film_h, sample_begin, sample_count, seed, max_depth=0, threads=8):
        rgb = np.empty((h, w, 3), dtype=np.float32)
        stats = (C.c_uint64 * 4)()
        self.lib.oracle_render_lt(self.handle, C.c_uint32(w), C.c_uint32(h), C.c_float(film_w),
                                  C.c_float(film_h), C.c_uint64(sample_begin), C.c_uint32(sample_count),
                                  C.c_uint64(seed), C.c_uint32(max_depth), C.c_int(threads), _ptr(rgb), stats)
        return rgb, {"closest_rays": stats[0], "splats": stats[1], "truncated": stats[2], "paths": stats[3]}

    def render_bdpt(self, w, h, film_w, film_h, sample_begin, sample_count, seed, max_vertices=0, threads=8):
        rgb = np.empty((h, w, 3), dtype=np.float32)
        stats = (C.c_uint64 * 3)()
        self.lib.oracle_render_bdpt(self.handle, C.c_uint32(w), C.c_uint32(h), C.c_float(film_w),
                                    C.c_float(film_h), C.c_uint64(sample_begin), C.c_uint32(sample_count),
                                    C.c_uint64(seed), C.c_uint32(max_vertices), C.c_int(threads), _ptr(rgb), stats)
        return rgb, {"light_events": stats[0], "eye_events": stats[1], "splats": stats[2]}

    def ray_set(self, w, h, film_w, film_h, seed, n):
        """2n rays: n camera rays then n secondary rays (see oracle_ray_set)."""
        o = np.empty((2 * n, 3), dtype=np.float32)
        d = np.empty((2 * n, 3), dtype=np.float32)
        self.lib.oracle_ray_set(self.handle, C.c_uint32(w), C.c_uint32(h), C.c_float(film_w), C.c_float(film_h),
                                C.c_uint64(seed), C.c_uint64(n), _ptr(o), _ptr(d))
        return o, d

    def philox(self, ctr, key):
        c = (C.c_uint32 * 4)(*ctr)
        k = (C.c_uint32 * 2)(*key)
        out = (C.c_uint32 * 4)()
        self.lib.oracle_philox4x32_10(c, k, out)
        return list(out)

    def stream(self, seed, pixel, sample, n):
        out = np.empty(n, dtype=np.float32)
        self.lib.oracle_stream(C.c_uint64(seed), C.c_uint32(pixel), C.c_uint64(sample), C.c_uint32(n), _ptr(out))
        return out

    def trace_path(self, w, h, film_w, film_h, px, py, sample, seed, max_depth=0, estimator=0, max_vertices=64):
        """Debug: per-vertex (prim, t, x, y, z, draws, throughput.r, p_rr) of one path."""
        out = np.zeros((max_vertices, 8), dtype=np.float32)
        n = self.lib.oracle_trace_path(self.handle, C.c_uint32(w), C.c_uint32(h), C.c_float(film_w),
                                       C.c_float(film_h), C.c_uint32(px), C.c_uint32(py), C.c_uint64(sample),
                                       C.c_uint64(seed), C.c_uint32(max_depth or (1 << 16)), C.c_int(estimator),
                                       _ptr(out), C.c_int(max_vertices))
        return out[:n]
