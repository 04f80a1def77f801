"""Host-side logic that needs no GPU: scene builder, Context/Sensor semantics, sample
partition, BVH builder, multi-rank film reduction over gloo."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import FILM, ROOT


def test_cornel_box_is_the_reference_scene(cornell):
    from amber_b200 import capi
    assert cornell.n_primitives == 25
    kinds = [cornell.material_kind(i) for i in range(25)]
    assert kinds[:6] == [capi.MAT_EYE] * 6
    assert kinds[6:8] == [capi.MAT_DIFFUSE_LIGHT] * 2
    assert kinds[12:14] == [capi.MAT_PHONG] * 2
    assert kinds[18:22] == [capi.MAT_REFRACTION] * 4
    assert [cornell.primitives[i].kind for i in (22, 23, 24)] == [capi.PRIM_SPHERE] * 3
    assert kinds[23] == capi.MAT_SPECULAR and kinds[24] == capi.MAT_REFRACTION
    # aperture blades: a hexagon of radius 0.05 in the plane z = 4
    for i in range(6):
        v = np.array(cornell.primitives[i].data[:]).reshape(3, 3)
        assert np.allclose(v[:, 2], 4.0)
        assert np.allclose(np.hypot(v[:2, 0], v[:2, 1]), 0.05, atol=1e-7)
        assert np.allclose(v[2], (0, 0, 4))


def test_blades_match_the_reference_constructor(cornell, oracle):
    # the oracle builds its blades inside oracle_scene_create; camera-ray origins must lie
    # on the builder's blade triangles (same float arithmetic as BasicThin's constructor)
    rng = np.random.default_rng(0)
    n = 600
    xi = rng.random((n, 5), dtype=np.float32)
    xi[:, 1:3] = 0.0  # u = v = 0 -> the blade's first vertex
    o, _, _ = oracle.eye_rays(64, 64, *FILM, np.zeros(n, np.uint32), np.zeros(n, np.uint32), xi)
    blade = np.minimum(5, np.floor(xi[:, 0] * np.float32(6)).astype(int))
    first_vertices = np.array([cornell.primitives[i].data[:3] for i in range(6)], dtype=np.float32)
    assert np.array_equal(o, first_vertices[blade])


def test_context_matches_cli_context():
    from amber_b200 import Context
    ctx = Context(4, 3)
    assert ctx.thread_count() == 4
    assert [ctx.iterate() for _ in range(5)] == [True, True, True, False, False]
    assert ctx.iteration_count() == 3
    unlimited = Context(1, 0)  # spp 0 = until expired (src/amber/cli/context.cc:48-60)
    assert all(unlimited.iterate() for _ in range(100))
    unlimited.expire()
    assert not unlimited.iterate() and unlimited.iteration_count() == 100


def test_sensor_defaults():
    from amber_b200 import Sensor
    s = Sensor(1024, 512)
    assert s.film_width == pytest.approx(0.036) and s.film_height == pytest.approx(0.018)
    assert s.size() == 1024 * 512


def test_partition_covers_every_sample_once():
    from amber_b200.dist import partition_samples
    for total in (0, 1, 7, 1024, 16384):
        for world in (1, 2, 3, 8):
            spans = [partition_samples(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == total
            for (b0, c0), (b1, _) in zip(spans, spans[1:]):
                assert b0 + c0 == b1
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1


def test_icosphere_mesh():
    from amber_b200.scene import icosphere
    tris = icosphere(3)
    assert tris.shape == (20 * 4 ** 3, 9) and tris.dtype == np.float32
    r = np.linalg.norm(tris.reshape(-1, 3) - np.array([0, -0.3, 0], np.float32), axis=1)
    assert r.min() >= 0.5 - 0.0101 and r.max() <= 0.5 + 0.0101
    assert np.ptp(r) > 0.01  # displaced


def test_render_rejects_a_foreign_scene():
    from amber_b200 import Context, CudaPathTracing, Sensor
    with pytest.raises(TypeError):
        CudaPathTracing().render(object(), Sensor(8, 8), Context(1, 1))


@pytest.mark.parametrize("width", [4, 8])
def test_bvh_builder_invariants(tmp_path, width):
    """Both node formats (bvh.cuh): 4-wide 64-byte nodes and 8-wide 96-byte nodes."""
    exe = tmp_path / "bvh_check"
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-pthread", f"-DAMBER_BVH_WIDTH={width}",
                           os.path.join(ROOT, "tests", "cpp", "bvh_check.cc"), "-o", str(exe)])
    out = subprocess.run([str(exe), "200000"], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    assert out.stdout.count("ok") == 3


_WORKER = r"""
import os, sys
sys.path.insert(0, {root!r})
import numpy as np, torch, torch.distributed as dist
from amber_b200 import cornel_box
from amber_b200.dist import partition_samples, all_reduce_film
from oracle import pyoracle
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:{port}", rank=rank, world_size=world)
oracle = pyoracle.Oracle(cornel_box())
W, S, SEED = 24, 10, 9
begin, count = partition_samples(S, rank, world)
image, _ = oracle.render(W, W, 0.036, 0.036, begin, count, SEED, 8, 1, threads=2)
film = torch.from_numpy(image.astype(np.float64) * count)          # un-normalised partial sum
total = all_reduce_film(film, count)
assert total == S
if rank == 0:
    whole, _ = oracle.render(W, W, 0.036, 0.036, 0, S, SEED, 8, 1, threads=2)
    np.testing.assert_allclose(film.numpy() / total, whole, rtol=1e-5, atol=1e-9)
# the same partition for the film-splatting algorithms (rows N2, N3): a rank's light paths land
# on any pixel, the reduce still only sums films
for render in (lambda b, c: oracle.render_lt(64, 64, 0.036, 0.036, b, c, SEED, threads=2)[0],
               lambda b, c: oracle.render_bdpt(W, W, 0.036, 0.036, b, c, SEED, 8, threads=2)[0]):
    S2 = 24
    begin, count = partition_samples(S2, rank, world)
    film = torch.from_numpy(render(begin, count).astype(np.float64) * count)
    assert all_reduce_film(film, count) == S2
    if rank == 0:
        np.testing.assert_allclose(film.numpy() / S2, render(0, S2), rtol=1e-5, atol=1e-9)
if rank == 0:
    print("PARTITION_OK")
dist.destroy_process_group()
"""


def test_sample_partition_and_film_reduce_over_gloo(tmp_path, oracle_lib):
    """World size 2 on CPU: each rank renders its slice of the samples (with the oracle,
    standing in for the device), one sum all-reduce of the film, and the result equals a
    single-rank render of all samples -- the image does not depend on the rank count."""
    script = tmp_path / "worker.py"
    port = 29500 + os.getpid() % 2000
    script.write_text(_WORKER.format(root=ROOT, port=port))
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=300)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    assert "PARTITION_OK" in outs[0]


def test_image_files_round_trip(tmp_path):
    """PNG / PFM writers of amber_b200.image_io (the reference's exporters need OpenCV)."""
    from amber_b200 import image_io
    rng = np.random.default_rng(2)
    ldr = rng.integers(0, 256, size=(37, 53, 3), dtype=np.uint8)
    hdr = rng.random((37, 53, 3), dtype=np.float32) * 1e3
    image_io.write_png(str(tmp_path / "a.png"), ldr)
    image_io.write_pfm(str(tmp_path / "a.pfm"), hdr)
    assert np.array_equal(image_io.read_png(str(tmp_path / "a.png")), ldr)
    assert np.array_equal(image_io.read_pfm(str(tmp_path / "a.pfm")), hdr)
