"""Scene ingestion for triangle meshes (row N1): the counterpart of the reference's
``ImportScene`` (``src/amber/cli/import.cc:49-167``).

The reference reads scene files through assimp (absent here): triangulated, pre-transformed
meshes; one material per mesh chosen by the heuristics of ``import.cc:67-109``; the first
camera as a ``Matrix4`` (``:133-149``); a thin lens 0.050 / 4 / 0.010 / 6 (``:150-156``) whose
aperture blades are appended after the scene objects (``:157-159``).  This module does the same
for Wavefront OBJ + MTL text, feeding ``FlatSceneBuilder``:

  material mapping, in the reference's order of precedence
    emissive colour ``Ke`` not black            -> DiffuseLight(Ke)
    reflective colour * reflectivity not black  -> Specular     (``Kr`` / ``reflectivity``: not in
                                                   standard MTL; accepted as extensions)
    ``illum`` is a Phong model (assimp maps MTL ``illum 2`` to aiShadingMode_Phong) with ``Ks``
      and ``Ns``                                -> Phong(Ks, Ns)
    diffuse colour ``Kd``                       -> Lambertian(Kd)
    otherwise                                   -> Lambertian(.5)
    (like the reference -- "TODO refraction", ``import.cc:79`` -- dielectrics are not imported)

  faces: fan-triangulated (assimp's aiProcess_Triangulate does the same for convex polygons);
  a mesh without ``usemtl`` gets the default material.

OBJ files carry no camera, so the camera arrives as arguments (position, look-at direction, up),
turned into the lens transform exactly as ``import.cc:137-148``.
"""
from __future__ import annotations

import os

import numpy as np

from .scene import FlatScene, FlatSceneBuilder

f32 = np.float32


def camera_transform(position, look_at, up) -> np.ndarray:
    """``import.cc:137-148``: zaxis = -lookAt, xaxis = lookAt x up, yaxis = zaxis x xaxis; the
    columns of the rotation are (xaxis, yaxis, zaxis), the translation is the position.  float32
    arithmetic, no normalisation (assimp hands over what the file holds)."""
    look = np.asarray(look_at, dtype=f32)
    upv = np.asarray(up, dtype=f32)
    pos = np.asarray(position, dtype=f32)
    zaxis = -look

    def cross(a, b):
        return np.array([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]], dtype=f32)

    xaxis = cross(look, upv)
    yaxis = cross(zaxis, xaxis)
    m = np.zeros((4, 4), dtype=f32)
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = xaxis, yaxis, zaxis, pos
    m[3, 3] = 1
    return m


def parse_mtl(path: str) -> dict[str, dict]:
    """MTL text -> {name: {key: value}} for the keys the mapping looks at."""
    materials: dict[str, dict] = {}
    current = None
    with open(path) as f:
        for line in f:
            tok = line.split("#", 1)[0].split()
            if not tok:
                continue
            key = tok[0]
            if key == "newmtl":
                current = materials.setdefault(" ".join(tok[1:]), {})
            elif current is None:
                continue
            elif key in ("Kd", "Ks", "Ke", "Kr"):
                vals = [float(v) for v in tok[1:4]]
                current[key] = tuple(vals if len(vals) == 3 else vals[:1] * 3)
            elif key in ("Ns", "reflectivity"):
                current[key] = float(tok[1])
            elif key == "illum":
                current[key] = int(float(tok[1]))
    return materials


def map_material(builder: FlatSceneBuilder, m: dict | None) -> int:
    """The heuristics of ``import.cc:67-109``, in their order."""
    def black(c):
        return c is None or all(v == 0 for v in c)

    if m is not None:
        if not black(m.get("Ke")):
            return builder.make_diffuse_light(m["Ke"])
        if "Kr" in m and "reflectivity" in m:
            c = tuple(f32(v) * f32(m["reflectivity"]) for v in m["Kr"])
            if not black(c):
                return builder.make_specular(c)
        if m.get("illum") == 2 and "Ks" in m and "Ns" in m:
            return builder.make_phong(m["Ks"], m["Ns"])
        if "Kd" in m:
            return builder.make_lambertian(m["Kd"])
    return builder.make_lambertian((.5, .5, .5))


def load_obj(path: str, camera_position, camera_look_at, camera_up=(0.0, 1.0, 0.0)) -> FlatScene:
    """OBJ (+ the MTL files it names) -> FlatScene with the importer's thin lens."""
    vertices: list[tuple[float, float, float]] = []
    groups: list[tuple[str | None, list[tuple[int, int, int]]]] = []
    mtl: dict[str, dict] = {}
    faces: list[tuple[int, int, int]] = []
    current: str | None = None
    groups.append((current, faces))
    with open(path) as f:
        for line in f:
            tok = line.split("#", 1)[0].split()
            if not tok:
                continue
            if tok[0] == "v":
                vertices.append((float(tok[1]), float(tok[2]), float(tok[3])))
            elif tok[0] == "f":
                idx = []
                for ref in tok[1:]:
                    i = int(ref.split("/")[0])
                    idx.append(i - 1 if i > 0 else len(vertices) + i)
                for k in range(1, len(idx) - 1):  # fan
                    faces.append((idx[0], idx[k], idx[k + 1]))
            elif tok[0] == "usemtl":
                current = " ".join(tok[1:])
                faces = []
                groups.append((current, faces))
            elif tok[0] == "mtllib":
                for name in tok[1:]:
                    mtl.update(parse_mtl(os.path.join(os.path.dirname(path), name)))
    if not vertices:
        raise ValueError(f"{path}: no vertices")
    v = np.asarray(vertices, dtype=f32)
    builder = FlatSceneBuilder()
    material_ids: dict[str | None, int] = {}
    for name, tris in groups:
        if not tris:
            continue
        if name not in material_ids:
            if name is not None and name not in mtl:
                raise ValueError(f"{path}: material {name!r} is not defined")
            material_ids[name] = map_material(builder, mtl.get(name) if name is not None else None)
        rows = v[np.asarray(tris, dtype=np.int64)].reshape(-1, 9)
        builder.add_triangles(rows, material_ids[name])
    # thin lens 0.050 / 4 / 0.010 / 6, aperture objects after the scene's (import.cc:150-159)
    builder.make_thin_lens(camera_transform(camera_position, camera_look_at, camera_up), 0.050, 4.0, 0.010, 6)
    return builder.build()
