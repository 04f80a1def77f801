"""Row N1: OBJ/MTL ingestion mirroring the reference's ImportScene (src/amber/cli/import.cc:49-167)."""
import numpy as np
import pytest

from conftest import FILM, image_agreement

OBJ = """# a small room: floor, back wall, emitter, shiny block, default-material triangle
mtllib room.mtl
v -1 -1 1
v 1 -1 1
v 1 -1 -1
v -1 -1 -1
v -1 1 -1
v 1 1 -1
v -0.3 0.95 0.3
v 0.3 0.95 0.3
v 0.3 0.95 -0.3
v -0.3 0.95 -0.3
v -0.5 -1 -0.2
v 0.1 -1 -0.2
v 0.1 -0.2 -0.2
v -0.5 -0.2 -0.2
f 11 12 13
usemtl floor
f 1 2 3 4
usemtl wall
f 4/1 3/2 6/3 5/4
usemtl lamp
f 7 10 9 8
usemtl block
f -4 -3 -2 -1
usemtl mirror
f 1 4 5
"""
MTL = """newmtl floor
Kd 0.6 0.6 0.6
illum 1
newmtl wall
Kd 0.2 0.5 0.2
Ks 0.5 0.5 0.5
Ns 32
illum 1
newmtl lamp
Kd 1 1 1
Ke 9 8 7
newmtl block
Kd 0.1 0.1 0.1
Ks 0.8 0.8 0.7
Ns 40
illum 2
newmtl mirror
Kd 0.3 0.3 0.3
Kr 1 1 1
reflectivity 0.9
"""
CAMERA = dict(camera_position=(0.3, 0.1, 2.5), camera_look_at=(-0.1, -0.05, -1.0), camera_up=(0.0, 1.0, 0.0))


@pytest.fixture(scope="module")
def room(tmp_path_factory):
    from amber_b200.importer import load_obj
    d = tmp_path_factory.mktemp("room")
    (d / "room.obj").write_text(OBJ)
    (d / "room.mtl").write_text(MTL)
    return load_obj(str(d / "room.obj"), **CAMERA)


def test_material_mapping_and_triangulation(room):
    from amber_b200 import capi
    kinds = [room.material_kind(i) for i in range(room.n_primitives)]
    # default (1 tri), floor quad (2), wall quad (2), lamp quad (2), block quad (2), mirror (1), 6 aperture blades last
    assert kinds == ([capi.MAT_LAMBERTIAN] * 5 + [capi.MAT_DIFFUSE_LIGHT] * 2 + [capi.MAT_PHONG] * 2
                     + [capi.MAT_SPECULAR] + [capi.MAT_EYE] * 6)
    default = room.materials[room.primitives[0].material]
    assert tuple(default.rgb) == (0.5, 0.5, 0.5)
    wall = room.materials[room.primitives[3].material]       # Ks + Ns but illum 1: Lambertian(Kd)
    assert wall.kind == capi.MAT_LAMBERTIAN and np.allclose(tuple(wall.rgb), (0.2, 0.5, 0.2))
    block = room.materials[room.primitives[7].material]
    assert block.kind == capi.MAT_PHONG and block.param == 40 and np.allclose(tuple(block.rgb), (0.8, 0.8, 0.7))
    mirror = room.materials[room.primitives[9].material]
    assert np.allclose(tuple(mirror.rgb), (0.9, 0.9, 0.9))
    # fan triangulation of `f 1 2 3 4`, relative indices of the block face
    assert np.allclose(room.primitives[1].data[:], [-1, -1, 1, 1, -1, 1, 1, -1, -1])
    assert np.allclose(room.primitives[2].data[:], [-1, -1, 1, 1, -1, -1, -1, -1, -1])
    assert np.allclose(room.primitives[7].data[:3], [-0.5, -1, -0.2])
    # the importer's lens: 0.050 / 4 / 0.010 / 6
    lens = room.desc.lens
    assert (lens.focal_length, lens.focus_distance, lens.n_blades) == (np.float32(0.05), 4.0, 6)
    assert lens.radius == np.float32(0.010)


def test_camera_transform_follows_the_importer():
    from amber_b200.importer import camera_transform
    m = camera_transform((1, 2, 3), (0, 0, -1), (0, 1, 0))
    # lookAt = -z, up = y: xaxis = lookAt x up = +x, zaxis = +z, yaxis = z x x = +y
    assert np.array_equal(m, np.array([[1, 0, 0, 1], [0, 1, 0, 2], [0, 0, 1, 3], [0, 0, 0, 1]], dtype=np.float32))
    m = camera_transform(**{k.replace("camera_", ""): v for k, v in CAMERA.items()})
    x, y, z = m[:3, 0], m[:3, 1], m[:3, 2]
    assert abs(np.dot(x, y)) < 1e-6 and abs(np.dot(x, z)) < 1e-6 and np.allclose(z, [0.1, 0.05, 1.0])


def test_imported_scene_oracle_is_pinned_to_the_reference(room, reference, oracle_lib):
    """A tilted, non-orthonormal camera matrix through BasicThin: camera rays, closest hits and a
    rendered frame of the C restatement against the compiled reference, bit for bit."""
    ref = oracle_lib.Reference(room)
    oracle = oracle_lib.Oracle(room)
    rng = np.random.default_rng(4)
    n = 4096
    px, py = rng.integers(0, 96, n), rng.integers(0, 64, n)
    xi = rng.random((n, 5), dtype=np.float32)
    film = (0.036, 0.024)
    a = ref.eye_rays(96, 64, *film, px, py, xi)
    b = oracle.eye_rays(96, 64, *film, px, py, xi)
    for u, v in zip(a, b):
        assert np.array_equal(u, v)
    o, d = oracle.ray_set(96, 64, *film, 12, 20000)
    pa, ta = ref.cast(o, d)
    pb, tb = oracle.cast(o, d)
    assert np.array_equal(pa, pb) and np.array_equal(ta, tb, equal_nan=True)
    ia, _ = ref.render_stream(48, 32, *film, 0, 16, 3, 8)
    ib, _ = oracle.render(48, 32, *film, 0, 16, 3, 8, estimator=0)
    assert np.array_equal(ia, ib)


@pytest.mark.gpu
def test_imported_scene_on_the_gpu(room, device, oracle_lib):
    from amber_b200 import Sensor
    oracle = oracle_lib.Oracle(room)
    film = (0.036, 0.024)
    device.upload_scene(room)
    device.set_sensor(Sensor(96, 64, *film))
    rng = np.random.default_rng(5)
    n = 4096
    px, py = rng.integers(0, 96, n), rng.integers(0, 64, n)
    xi = rng.random((n, 5), dtype=np.float32)
    o, d, w = device.eye_rays(px, py, xi)
    oo, od, ow = oracle.eye_rays(96, 64, *film, px, py, xi)
    assert np.array_equal(o, oo) and np.array_equal(d, od) and np.allclose(w, ow, rtol=2e-5)
    ro, rd = oracle.ray_set(96, 64, *film, 12, 50000)
    want_prim, want_t = oracle.cast(ro, rd, threads=16)
    prim, t = device.cast(ro, rd)
    assert (prim == want_prim).mean() >= 0.9999
    both = (prim == want_prim) & (want_prim >= 0)
    assert np.all(np.abs(t[both] - want_t[both]) <= 1e-4 * want_t[both] + 1e-7)
    for estimator in (0, 1):
        device.film_clear()
        device.render(0, 32, seed=3, max_depth=16, estimator=estimator)
        image, _ = device.resolve()
        want, _ = oracle.render(96, 64, *film, 0, 32, 3, 16, estimator=estimator, threads=16)
        assert image_agreement(image, want) >= 0.99
