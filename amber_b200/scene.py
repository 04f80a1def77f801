"""Flat scene description for the C ABI, with the reference's factory verbs.

``FlatSceneBuilder`` mirrors the way the reference assembles a scene
(``src/amber/etude/cornel_box.cc:38-204``): ``make_*`` material factories,
``make_triangle`` / ``make_sphere`` primitive factories, ``make_thin_lens`` and
``add_object(primitive, material)``; primitive ids are creation order.  The result
is the POD ``amber_cuda_scene_desc`` of ``include/amber_cuda.h`` -- the same bytes
are handed to the CUDA library and to the reference harness (oracle/ref_harness.cc),
which rebuilds the scene with the reference's own factories.

``cornel_box()`` is the reference's built-in scene as data.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from . import capi

_libm = C.CDLL("libm.so.6")
for _name in ("cosf", "sinf"):
    getattr(_libm, _name).restype = C.c_float
    getattr(_libm, _name).argtypes = [C.c_float]

f32 = np.float32
# static_cast<float>(kPI), include/amber/constants.h:25
PI_F = f32(3.141592653589793238462643383279503)


class FlatScene:
    """Owns the ctypes arrays behind an ``amber_cuda_scene_desc``."""

    def __init__(self, primitives, materials, lens):
        self.primitives = (capi.Primitive * len(primitives))(*primitives)
        self.materials = (capi.Material * len(materials))(*materials)
        self.lens = lens
        self.desc = capi.SceneDesc(
            len(primitives),
            C.cast(self.primitives, C.POINTER(capi.Primitive)),
            len(materials),
            C.cast(self.materials, C.POINTER(capi.Material)),
            lens,
        )

    @property
    def n_primitives(self) -> int:
        return len(self.primitives)

    def material_kind(self, prim: int) -> int:
        return self.materials[self.primitives[prim].material].kind


class FlatSceneBuilder:
    def __init__(self):
        self._materials: list[capi.Material] = []
        self._primitives: list[capi.Primitive] = []
        self._lens: capi.Lens | None = None
        self._eye_material: int | None = None

    # -- materials (include/amber/scene/material_*.h factories) -----------------
    def _material(self, kind, rgb, param=0.0) -> int:
        m = capi.Material(kind, (C.c_float * 3)(*[float(c) for c in rgb]), float(param))
        self._materials.append(m)
        return len(self._materials) - 1

    def make_lambertian(self, kd) -> int:
        return self._material(capi.MAT_LAMBERTIAN, kd)

    def make_phong(self, ks, exponent) -> int:
        return self._material(capi.MAT_PHONG, ks, exponent)

    def make_specular(self, ks) -> int:
        return self._material(capi.MAT_SPECULAR, ks)

    def make_refraction(self, ior) -> int:
        # BasicMaterialForwarder multiplies by rho = Radiant(1):
        # include/amber/scene/material_refraction.h:94-101
        return self._material(capi.MAT_REFRACTION, (1.0, 1.0, 1.0), ior)

    def make_diffuse_light(self, radiance) -> int:
        return self._material(capi.MAT_DIFFUSE_LIGHT, radiance)

    def make_eye(self) -> int:
        if self._eye_material is None:
            self._eye_material = self._material(capi.MAT_EYE, (1.0, 1.0, 1.0))
        return self._eye_material

    # -- primitives (include/amber/scene/primitive_{triangle,sphere}.h) ---------
    def add_triangle(self, v0, v1, v2, material: int) -> int:
        data = (C.c_float * 9)(*[float(c) for c in (*v0, *v1, *v2)])
        self._primitives.append(capi.Primitive(capi.PRIM_TRIANGLE, material, data))
        return len(self._primitives) - 1

    def add_sphere(self, center, radius, material: int) -> int:
        data = (C.c_float * 9)(*[float(c) for c in (*center, radius)], 0, 0, 0, 0, 0)
        self._primitives.append(capi.Primitive(capi.PRIM_SPHERE, material, data))
        return len(self._primitives) - 1

    def add_triangles(self, vertices9: np.ndarray, material: int) -> None:
        """Bulk add: vertices9 is (n, 9) float32."""
        for row in np.asarray(vertices9, dtype=np.float32):
            self._primitives.append(
                capi.Primitive(capi.PRIM_TRIANGLE, material, (C.c_float * 9)(*row.tolist()))
            )

    # -- lens (include/amber/scene/lens_thin.h:30-38) ---------------------------
    def make_thin_lens(self, transform, focal_length, focus_distance, radius, n_blades,
                       add_aperture_objects=True) -> None:
        """``transform`` is a row-major 4x4.  With ``add_aperture_objects`` the blade
        triangles are emplaced as Eye-material objects first, as
        src/amber/etude/cornel_box.cc:62-64 does with lens->ApertureObjects()."""
        t = [float(f32(v)) for v in np.asarray(transform, dtype=np.float64).reshape(16)]
        self._lens = capi.Lens((C.c_float * 16)(*t), float(focal_length), float(focus_distance),
                               float(radius), int(n_blades))
        if add_aperture_objects:
            eye = self.make_eye()
            for tri in thin_lens_blades(t, radius, n_blades):
                self.add_triangle(tri[0:3], tri[3:6], tri[6:9], eye)

    def make_pinhole_lens(self, transform, sensor_distance) -> None:
        """MakePinholeLens (include/amber/scene/lens_pinhole.h:32-37).  Its aperture object is a
        degenerate triangle at the lens origin (src/amber/scene/lens_pinhole.cc:38) that no ray
        can hit, so nothing is added to the primitives."""
        t = [float(f32(v)) for v in np.asarray(transform, dtype=np.float64).reshape(16)]
        self._lens = capi.Lens((C.c_float * 16)(*t), 0.0, 0.0, 0.0, 0, capi.LENS_PINHOLE, float(sensor_distance))

    def build(self) -> FlatScene:
        if self._lens is None:
            raise ValueError("scene has no lens")
        return FlatScene(self._primitives, self._materials, self._lens)


def thin_lens_blades(transform16, radius, n_blades) -> np.ndarray:
    """Aperture blade triangles in float32, op for op as BasicThin's constructor
    (src/amber/scene/lens_thin.cc:45-55): (origin + G*(r*(cos a_i, sin a_i, 0)),
    origin + G*(r*(cos a_{i+1}, sin a_{i+1}, 0)), origin)."""
    t = np.asarray(transform16, dtype=np.float32).reshape(4, 4)
    origin = t[:3, 3].copy()
    g = t[:3, :3]
    r = f32(radius)
    nb = f32(n_blades)

    def rim(i):
        angle = f32(2) * PI_F / nb * f32(i)
        v = np.array([r * f32(_libm.cosf(angle)), r * f32(_libm.sinf(angle)), r * f32(0)], dtype=np.float32)
        out = np.empty(3, dtype=np.float32)
        for row in range(3):
            out[row] = f32(f32(g[row, 0] * v[0]) + f32(g[row, 1] * v[1])) + f32(g[row, 2] * v[2])
        return origin + out

    tris = np.empty((int(n_blades), 9), dtype=np.float32)
    for i in range(int(n_blades)):
        tris[i, 0:3] = rim(i)
        tris[i, 3:6] = rim(i + 1)
        tris[i, 6:9] = origin
    return tris


def cornel_box(focal_length=0.050, aperture_radius=0.050, aperture_n_blades=6) -> FlatScene:
    """The reference's built-in scene, src/amber/etude/cornel_box.cc:38-204, called as
    CornelBox(0.050, 0.050, 6) by src/amber/cli/application.cc:69-73.  Primitive ids:
    0-5 aperture, 6-7 light, 8-9 left, 10-11 right, 12-13 back (Phong), 14-15 floor,
    16-17 ceiling, 18-21 water (Refraction 1.333), 22-24 spheres."""
    b = FlatSceneBuilder()
    b.make_thin_lens(
        [1, 0, 0, 0,
         0, 1, 0, 0,
         0, 0, 1, 4,
         0, 0, 0, 1],
        focal_length, 4.0, aperture_radius, aperture_n_blades)

    def quad(material, a, bb, c, d, e, f):
        b.add_triangle(a, bb, c, material)
        b.add_triangle(d, e, f, material)

    light = b.make_diffuse_light((1e11, 1e11, 1e11))
    quad(light, (0.01, 0.99, 0.01), (-0.01, 0.99, 0.01), (-0.01, 0.99, -0.01),
         (-0.01, 0.99, -0.01), (0.01, 0.99, -0.01), (0.01, 0.99, 0.01))
    left = b.make_lambertian((.5, 0, 0))
    quad(left, (-1, 1, 1), (-1, -1, 1), (-1, -1, -1), (-1, -1, -1), (-1, 1, -1), (-1, 1, 1))
    right = b.make_lambertian((0, .5, 0))
    quad(right, (1, 1, 1), (1, 1, -1), (1, -1, -1), (1, -1, -1), (1, -1, 1), (1, 1, 1))
    back = b.make_phong((.95, .95, .95), 256)
    quad(back, (1, 1, -1), (-1, 1, -1), (-1, -1, -1), (-1, -1, -1), (1, -1, -1), (1, 1, -1))
    floor = b.make_lambertian((.5, .5, .5))
    quad(floor, (1, -1, 1), (1, -1, -1), (-1, -1, -1), (-1, -1, -1), (-1, -1, 1), (1, -1, 1))
    ceiling = b.make_lambertian((.5, .5, .5))
    quad(ceiling, (1, 1, 1), (-1, 1, 1), (-1, 1, -1), (-1, 1, -1), (1, 1, -1), (1, 1, 1))
    water = b.make_refraction(1.333)
    quad(water, (1, -0.5, 1), (1, -0.5, -1), (-1, -0.5, -1), (-1, -0.5, -1), (-1, -0.5, 1), (1, -0.5, 1))
    quad(water, (1, -0.5, 1), (-1, -0.5, 1), (-1, -1, 1), (-1, -1, 1), (1, -1, 1), (1, -0.5, 1))
    b.add_sphere((0.4, -0.6, -0.5), 0.4, b.make_lambertian((.5, .5, .5)))
    b.add_sphere((-0.4, -0.7, 0.1), 0.3, b.make_specular((.95, .95, .95)))
    b.add_sphere((0.1, -0.8, 0.6), 0.2, b.make_refraction(1.125))
    return b.build()


def _hash32(x: np.ndarray, seed: int) -> np.ndarray:
    """32-bit integer hash (lowbias32-style) used to displace the config-4 mesh."""
    x = (x.astype(np.uint64) ^ np.uint64(seed)) & np.uint64(0xFFFFFFFF)
    x ^= x >> np.uint64(16)
    x = (x * np.uint64(0x7FEB352D)) & np.uint64(0xFFFFFFFF)
    x ^= x >> np.uint64(15)
    x = (x * np.uint64(0x846CA68B)) & np.uint64(0xFFFFFFFF)
    x ^= x >> np.uint64(16)
    return x.astype(np.uint32)


def icosphere(subdivisions: int, center=(0.0, -0.3, 0.0), radius=0.5, displacement=0.02,
              seed=0x5EED) -> np.ndarray:
    """(20*4^s, 9) float32 triangles: icosahedron subdivided ``subdivisions`` times,
    each vertex displaced radially by displacement*(hash32(vertex_id, seed)/2^32 - 0.5)
    (s=8 -> 1 310 720 triangles)."""
    t = (1.0 + math.sqrt(5.0)) / 2.0
    verts = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t],
                      [0, -1, -t], [0, 1, -t], [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]],
                     dtype=np.float64)
    verts /= np.linalg.norm(verts, axis=1, keepdims=True)
    faces = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9],
                      [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2],
                      [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10],
                      [8, 6, 7], [9, 8, 1]], dtype=np.int64)
    for _ in range(subdivisions):
        nv = len(verts)
        edges = np.concatenate([faces[:, [0, 1]], faces[:, [1, 2]], faces[:, [2, 0]]])
        edges_sorted = np.sort(edges, axis=1)
        keys = edges_sorted[:, 0] * nv + edges_sorted[:, 1]
        uniq, inverse = np.unique(keys, return_inverse=True)
        mid = verts[uniq // nv] + verts[uniq % nv]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        verts = np.concatenate([verts, mid])
        nf = len(faces)
        ab = nv + inverse[:nf]
        bc = nv + inverse[nf:2 * nf]
        ca = nv + inverse[2 * nf:]
        a, b_, c = faces[:, 0], faces[:, 1], faces[:, 2]
        faces = np.concatenate([
            np.stack([a, ab, ca], 1), np.stack([b_, bc, ab], 1),
            np.stack([c, ca, bc], 1), np.stack([ab, bc, ca], 1)])
    ids = np.arange(len(verts), dtype=np.uint32)
    jitter = _hash32(ids, seed).astype(np.float64) / 4294967296.0 - 0.5
    r = radius + displacement * jitter
    pts = (np.asarray(center, dtype=np.float64)[None, :] + verts * r[:, None]).astype(np.float32)
    return pts[faces].reshape(-1, 9)


def cornel_mesh(subdivisions=8) -> FlatScene:
    """Config 4: Cornell walls + light + lens; spheres and water
    replaced by a displaced icosphere, Lambertian(.5)."""
    b = FlatSceneBuilder()
    b.make_thin_lens([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 4, 0, 0, 0, 1], 0.050, 4.0, 0.050, 6)

    def quad(material, *v):
        b.add_triangle(v[0], v[1], v[2], material)
        b.add_triangle(v[3], v[4], v[5], material)

    quad(b.make_diffuse_light((1e11, 1e11, 1e11)),
         (0.01, 0.99, 0.01), (-0.01, 0.99, 0.01), (-0.01, 0.99, -0.01),
         (-0.01, 0.99, -0.01), (0.01, 0.99, -0.01), (0.01, 0.99, 0.01))
    quad(b.make_lambertian((.5, 0, 0)), (-1, 1, 1), (-1, -1, 1), (-1, -1, -1), (-1, -1, -1), (-1, 1, -1), (-1, 1, 1))
    quad(b.make_lambertian((0, .5, 0)), (1, 1, 1), (1, 1, -1), (1, -1, -1), (1, -1, -1), (1, -1, 1), (1, 1, 1))
    quad(b.make_phong((.95, .95, .95), 256), (1, 1, -1), (-1, 1, -1), (-1, -1, -1), (-1, -1, -1), (1, -1, -1), (1, 1, -1))
    quad(b.make_lambertian((.5, .5, .5)), (1, -1, 1), (1, -1, -1), (-1, -1, -1), (-1, -1, -1), (-1, -1, 1), (1, -1, 1))
    quad(b.make_lambertian((.5, .5, .5)), (1, 1, 1), (-1, 1, 1), (-1, 1, -1), (-1, 1, -1), (1, 1, -1), (1, 1, 1))
    mesh_material = b.make_lambertian((.5, .5, .5))
    tris = icosphere(subdivisions)
    n0 = len(b._primitives)
    prims = (capi.Primitive * (n0 + len(tris)))()
    for i, p in enumerate(b._primitives):
        prims[i] = p
    # bulk fill through a numpy view (a Python loop over 1.3 M ctypes structs is slow)
    view = np.frombuffer(prims, dtype=np.dtype([("kind", "<i4"), ("material", "<i4"), ("data", "<f4", (9,))]))
    view["kind"][n0:] = capi.PRIM_TRIANGLE
    view["material"][n0:] = mesh_material
    view["data"][n0:] = tris
    scene = FlatScene.__new__(FlatScene)
    scene.primitives = prims
    scene.materials = (capi.Material * len(b._materials))(*b._materials)
    scene.lens = b._lens
    scene.desc = capi.SceneDesc(len(prims), C.cast(prims, C.POINTER(capi.Primitive)),
                                len(scene.materials), C.cast(scene.materials, C.POINTER(capi.Material)),
                                b._lens)
    return scene
