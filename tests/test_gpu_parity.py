"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle on the same
seeded inputs, against the golden fixtures of the unmodified reference, and -- at the
benchmark's full size -- through size-independent properties.

Tolerances (floating point; the north star asks for hit ids >= 99.99 % and hit distance
within 1e-4 relative): the device uses FMA contraction, MUFU reciprocal / rsqrt and its
own libm, so results differ from the host's by a few ulp; where a path takes a different
branch because of that (a rejection-loop retry, a Russian-roulette threshold) a pixel can
differ, so image checks require >= 99 % of pixels to agree to 2e-3 relative.
"""
import numpy as np
import pytest

from conftest import FILM, diffuse_box, golden, image_agreement, light_tracing_box, pinhole_box, unit_vectors

pytestmark = pytest.mark.gpu

ID_AGREEMENT = 0.9999
T_RTOL = 1e-4


def check_cast(prim, t, want_prim, want_t):
    agree = prim == want_prim
    assert agree.mean() >= ID_AGREEMENT, f"hit ids agree on {agree.mean():.6f}"
    both = agree & (want_prim >= 0)
    rel = np.abs(t[both] - want_t[both]) / want_t[both]
    # 1e-4 relative, plus an absolute floor of 0.1 kEPS = 1e-7: hits at t ~ 1e-5 are a
    # surface re-hitting itself, where the reference's own t is cancellation noise
    # (c = |c-o|^2 - r^2 in primitive_sphere.cc:79) and no two float evaluations agree
    # to 1e-4.  The pure 1e-4 bound must still hold for >= 99.99 % of the hits.
    assert np.all(rel <= T_RTOL + 1e-7 / want_t[both]), f"max rel t error {rel.max():.3e}"
    assert (rel <= T_RTOL).mean() >= 0.9999, f"pure relative bound holds on {(rel <= T_RTOL).mean():.6f}"
    assert np.isnan(t[agree & (want_prim < 0)]).all()
    return float(agree.mean()), float(rel.max()) if rel.size else 0.0


@pytest.fixture
def cornell_device(device, cornell):
    """The shared context with the Cornell box bound (other tests upload other scenes)."""
    from amber_b200 import Sensor
    if device.scene is not cornell:
        device.upload_scene(cornell)
    device.set_sensor(Sensor(64, 64))
    return device


def test_cast_matches_reference_golden(cornell_device):
    g = golden("cornell_cast.npz")
    prim, t = cornell_device.cast(g["origins"], g["directions"])
    check_cast(prim, t, g["prim"], g["t"])


def test_cast_matches_oracle_on_the_fixed_ray_set(cornell_device, oracle):
    """The 10 M-ray set of SURVEY.md section 8d: 5 M camera rays + 5 M secondary rays, seed 20261104."""
    o, d = oracle.ray_set(1024, 1024, *FILM, 20261104, 5_000_000)
    want_prim, want_t = oracle.cast(o, d, threads=32)
    prim, t = cornell_device.cast(o, d)
    agreement, max_rel = check_cast(prim, t, want_prim, want_t)
    print(f"10M-ray set: id agreement {agreement:.7f}, max rel t error {max_rel:.3e}")


def test_cast_edge_cases(cornell_device, oracle):
    rng = np.random.default_rng(11)
    # empty input
    prim, t = cornell_device.cast(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32))
    assert prim.shape == (0,) and t.shape == (0,)
    # rays that start on surfaces and leave along them / axis-parallel rays / rays from outside
    o = rng.uniform(-0.99, 0.99, size=(20000, 3)).astype(np.float32)
    d = unit_vectors(rng, 20000)
    o[:5000, 1] = -1.0                      # on the floor
    d[:2500, 1] = np.abs(d[:2500, 1])       # heading up
    d[5000:7500] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, 2500)] * rng.choice([-1, 1], (2500, 1)).astype(np.float32)
    o[7500:10000] *= 5.0                    # outside the box
    d[:10000] /= np.linalg.norm(d[:10000], axis=1, keepdims=True)
    want_prim, want_t = oracle.cast(o, d)
    prim, t = cornell_device.cast(o, d)
    check_cast(prim, t, want_prim, want_t)
    # a single ray (ragged block) straight at the back wall
    prim, t = cornell_device.cast(np.array([[0, 0, 0.5]], np.float32), np.array([[0, 0, -1]], np.float32))
    assert prim[0] in (12, 13) and abs(t[0] - 1.5) < 1e-5


def test_camera_rays(cornell_device, oracle):
    from amber_b200 import Sensor
    g = golden("eye_rays.npz")
    for tag, (w, h) in {"64x64": (64, 64), "1024x512": (1024, 512)}.items():
        cornell_device.set_sensor(Sensor(w, h))
        o, d, wt = cornell_device.eye_rays(g[f"{tag}.px"], g[f"{tag}.py"], g[f"{tag}.xi"])
        # origin and direction are computed with the host's operations: bit-identical
        assert np.array_equal(o, g[f"{tag}.origin"])
        assert np.array_equal(d, g[f"{tag}.direction"])
        assert np.allclose(wt, g[f"{tag}.weight"], rtol=2e-5)
    cornell_device.set_sensor(Sensor(64, 64))


def test_exact_normalisation_is_ieee(device):
    """vec.cuh normalize_exact (one shared reciprocal, zero numerators by select) must be bit for bit
    Normalize of include/amber/prelude/vector3.h:309-316 in IEEE binary32: sqrt((x*x + y*y) + z*z), then
    three divisions -- numpy float32 arithmetic is exactly that.  Covers the magnitudes the short
    sequence takes, the ones it hands to div.rn (tiny, huge, denormal), signed zeros, inf and NaN."""
    rng = np.random.default_rng(7)
    n = 1 << 22
    parts = [rng.standard_normal((n, 3))]                                       # directions
    v = rng.standard_normal((n, 3)); v[rng.random((n, 3)) < 0.4] = 0.0; parts.append(v)   # axis-aligned: zeros
    v = rng.standard_normal((n, 3)); v[rng.random((n, 3)) < 0.3] *= -0.0; parts.append(v)  # signed zeros
    parts.append(rng.standard_normal((n, 3)) * np.exp2(rng.integers(-30, 30, (n, 1))))    # any scale, one vector
    parts.append(rng.standard_normal((n, 3)) * np.exp2(rng.integers(-60, 20, (n, 3))))    # mixed scales in one vector
    parts.append(rng.standard_normal((n // 4, 3)) * np.exp2(rng.integers(-149, 127, (n // 4, 3))))  # ends of the range
    parts.append(np.array([[0, 0, 0], [np.inf, 1, 0], [np.nan, 1, 2], [1e-45, 0, 0], [3e38, 3e38, 0], [1, -0.0, 0.0],
                           [1e-30, 1, 0], [1e-39, 1e-39, 1e-39], [-0.0, -0.0, -1e-20]]))
    for part in parts:
        with np.errstate(all="ignore"):
            v = np.ascontiguousarray(part, dtype=np.float32)
            length = np.sqrt((v[:, 0] * v[:, 0] + v[:, 1] * v[:, 1]) + v[:, 2] * v[:, 2], dtype=np.float32)
            want = (v / length[:, None]).astype(np.float32)
        got = device.normalize(v)
        same = (got.view(np.uint32) == want.view(np.uint32)) | (np.isnan(got) & np.isnan(want))
        assert same.all(), f"{(~same).sum()} of {same.size} components differ, e.g. {v[~same.all(axis=1)][:3]}"
        # the same sequence as a plain division (barycentrics over det): either sign of the denominator
        with np.errstate(all="ignore"):
            den = (rng.standard_normal(v.shape[0]) * np.exp2(rng.integers(-45, 45, v.shape[0]))).astype(np.float32)
            den[:: 97] = np.float32(0.0)
            want = (v / den[:, None]).astype(np.float32)
        got = device.normalize(v, den)
        same = (got.view(np.uint32) == want.view(np.uint32)) | (np.isnan(got) & np.isnan(want))
        assert same.all(), f"division: {(~same).sum()} of {same.size} quotients differ"


def test_materials(device, oracle_lib):
    """SampleLight for every material kind against the reference golden vectors."""
    from amber_b200 import FlatSceneBuilder, Sensor, capi
    g = golden("materials.npz")
    names = sorted({k.split(".")[0] for k in g.files})
    b = FlatSceneBuilder()
    b.make_thin_lens(np.eye(4), 0.05, 4.0, 0.05, 6, add_aperture_objects=False)
    index = {}
    for name in names:
        kind, r, gg, bb, param = g[f"{name}.desc"]
        index[name] = b._material(int(kind), (r, gg, bb), param)
    b.add_triangle((0, 0, 0), (1, 0, 0), (0, 1, 0), index[names[0]])
    device.upload_scene(b.build())
    device.set_sensor(Sensor(8, 8))
    for name in names:
        dir_in, weight, draws = device.sample_light(index[name], g[f"{name}.normal"], g[f"{name}.wo"], g[f"{name}.xi"])
        same = draws == g[f"{name}.draws"]
        assert same.mean() >= 0.995, (name, same.mean())
        atol = 1e-4 if name.startswith("phong") else 5e-6
        assert np.allclose(dir_in[same], g[f"{name}.dir_in"][same], atol=atol), name
        assert np.allclose(weight[same], g[f"{name}.weight"][same], rtol=2e-4, atol=1e-7), name


@pytest.mark.parametrize("estimator", [0, 1])
def test_image_matches_oracle_on_the_same_stream(cornell_device, oracle, estimator):
    from amber_b200 import Sensor
    w, spp, seed, depth = 64, 32, 5, 32
    cornell_device.set_sensor(Sensor(w, w))
    cornell_device.render(0, spp, seed=seed, max_depth=depth, estimator=estimator)
    image, done = cornell_device.resolve()
    want, stats = oracle.render(w, w, *FILM, 0, spp, seed, depth, estimator=estimator)
    assert done == spp
    frac = image_agreement(image, want)
    print(f"estimator {estimator}: {frac:.4f} of pixels agree with the oracle at {spp} spp")
    assert frac >= 0.99, f"only {frac:.4f} of pixels agree"
    got = cornell_device.stats()
    assert got["paths"] == w * w * spp
    assert abs(got["closest_rays"] - stats["closest_rays"]) <= 2e-3 * stats["closest_rays"]
    if estimator == 1:
        assert abs(got["shadow_rays"] - stats["shadow_rays"]) <= 2e-3 * stats["shadow_rays"]
    else:
        assert got["shadow_rays"] == 0


def test_image_matches_reference_golden(cornell_device):
    """Whole images rendered by the unmodified reference on the shared Philox stream."""
    from amber_b200 import Sensor
    g = golden("cornell_stream.npz")
    for tag in sorted({k.rsplit(".", 1)[0] for k in g.files}):
        size, seed, depth = tag.split(".")
        w, _, spp = (int(x) for x in size.split("x"))
        cornell_device.set_sensor(Sensor(w, w))
        cornell_device.render(0, spp, seed=int(seed[4:]), max_depth=int(depth[5:]), estimator=0)
        image, _ = cornell_device.resolve()
        assert image_agreement(image, g[f"{tag}.image"]) >= 0.99, tag
        rays = int(g[f"{tag}.rays"][0])
        assert abs(cornell_device.stats()["closest_rays"] - rays) <= 2e-3 * rays


def test_specular_chain_stress(cornell_device, oracle):
    """Config 3: mirror + glass + water, depth cap 32 with Russian roulette; the cap is
    applied identically by the oracle and the truncated fraction is reported."""
    from amber_b200 import Sensor
    cornell_device.set_sensor(Sensor(48, 48))
    for depth in (2, 32):
        cornell_device.film_clear()
        cornell_device.render(0, 16, seed=8, max_depth=depth, estimator=1)
        image, _ = cornell_device.resolve()
        want, stats = oracle.render(48, 48, *FILM, 0, 16, 8, depth, estimator=1)
        assert image_agreement(image, want) >= 0.99
        got = cornell_device.stats()
        assert abs(got["truncated_paths"] - stats["truncated"]) <= max(8, 0.02 * stats["truncated"])
        assert got["phong_capped"] == 0


def test_accumulation_is_independent_of_batching_and_pool_size(cornell_device):
    """Sample partition invariance: [0,8) in one call == [0,3)+[3,8) in two == a tiny pool."""
    from amber_b200 import Sensor
    cornell_device.set_sensor(Sensor(96, 64))
    cornell_device.render(0, 8, seed=3, max_depth=16, estimator=1)
    whole, _ = cornell_device.resolve()
    cornell_device.film_clear()
    cornell_device.render(0, 3, seed=3, max_depth=16, estimator=1)
    cornell_device.render(3, 5, seed=3, max_depth=16, estimator=1)
    split, done = cornell_device.resolve()
    assert done == 8
    cornell_device.film_clear()
    cornell_device.render(0, 8, seed=3, max_depth=16, estimator=1, pool_paths=1024)
    small_pool, _ = cornell_device.resolve()
    for other in (split, small_pool):
        assert np.allclose(other, whole, rtol=1e-4, atol=1e-7 * whole.mean())


@pytest.mark.parametrize("size", [(1, 1), (37, 19), (5, 131)])
def test_ragged_frames_and_pools(cornell_device, size):
    """Frame sizes and pools that are no multiple of anything (a 1 x 1 frame, pools below one shade tile
    and off its boundary), every estimator: exactly W * H * spp paths, a finite image, and -- for the
    wavefront estimators -- the image of the default pool.  A render of zero samples is a no-op."""
    from amber_b200 import Sensor, capi
    w, h = size
    cornell_device.set_sensor(Sensor(w, h))
    spp = 6
    for estimator, depth in ((capi.ESTIMATOR_REFERENCE, 12), (capi.ESTIMATOR_NEE, 12),
                             (capi.ESTIMATOR_LIGHT_TRACING, 12), (capi.ESTIMATOR_BDPT, 0)):
        images = []
        for pool in (0, 300, 777):
            cornell_device.film_clear()
            before = cornell_device.stats()["paths"]
            cornell_device.render(0, 0, seed=5, max_depth=depth, estimator=estimator, pool_paths=pool)
            cornell_device.render(2, spp, seed=5, max_depth=depth, estimator=estimator, pool_paths=pool)
            image, done = cornell_device.resolve()
            assert done == spp
            assert cornell_device.stats()["paths"] - before == w * h * spp
            assert image.shape == (h, w, 3) and np.isfinite(image).all() and (image >= 0).all()
            images.append(image)
        if estimator != capi.ESTIMATOR_LIGHT_TRACING:  # its splats land in float atomics in any order
            for other in images[1:]:
                assert np.allclose(other, images[0], rtol=1e-4, atol=1e-6 * max(images[0].mean(), 1e-6))


def test_algorithm_interface_mirrors_the_reference(cornell, oracle):
    """Render(scene, sensor, context): one Iterate() = one sample per pixel; expiry ends early."""
    from amber_b200 import Context, CudaPathTracing, Sensor
    algo = CudaPathTracing(estimator="reference", max_depth=8, seed=2, batch_spp=4)
    context = Context(1, 10)
    image = algo.render(cornell, Sensor(32, 32), context)
    assert image.shape == (32, 32, 3) and image.dtype == np.float32
    assert context.iteration_count() == 10 and algo.last_stats["paths"] == 32 * 32 * 10
    want, _ = oracle.render(32, 32, *FILM, 0, 10, 2, 8, estimator=0)
    assert image_agreement(image, want) >= 0.99
    expired = Context(1, 0)
    expired.expire()
    assert np.all(algo.render(cornell, Sensor(8, 8), expired) == 0)


def test_estimators_agree_in_expectation_on_the_device(device):
    """NEE vs the reference's estimator, low-variance scene, millions of paths: means agree."""
    from amber_b200 import Sensor
    device.upload_scene(diffuse_box())
    device.set_sensor(Sensor(32, 32))
    means = {}
    for estimator, spp in ((0, 4096), (1, 1024)):
        runs = []
        for seed in range(4):
            device.film_clear()
            device.render(0, spp, seed=50 + seed, estimator=estimator)
            image, _ = device.resolve()
            runs.append(image.mean(axis=(0, 1)))
        means[estimator] = np.array(runs)
    diff = means[0].mean(0) - means[1].mean(0)
    sigma = np.sqrt(means[0].var(0, ddof=1) / 4 + means[1].var(0, ddof=1) / 4)
    assert np.all(np.abs(diff) < 4 * sigma + 2e-3 * means[0].mean(0)), (diff, sigma)


def test_bvh_path_matches_oracle(device, oracle_lib):
    """Scenes above the constant-memory limit go through the 4-wide BVH: closest hits and
    images against the brute-force oracle on a displaced icosphere in the Cornell room."""
    from amber_b200 import Sensor, cornel_mesh
    scene = cornel_mesh(subdivisions=4)  # 5120 + 18 triangles
    oracle = oracle_lib.Oracle(scene)
    device.upload_scene(scene)
    device.set_sensor(Sensor(32, 32))
    stats = device.stats()
    assert stats["accel"] == 1 and stats["bvh_nodes"] > 100
    o, d = oracle.ray_set(256, 256, *FILM, 77, 60000)
    want_prim, want_t = oracle.cast(o, d, threads=32)
    prim, t = device.cast(o, d)
    check_cast(prim, t, want_prim, want_t)
    for estimator in (0, 1):
        device.film_clear()
        device.render(0, 8, seed=6, max_depth=16, estimator=estimator)
        image, _ = device.resolve()
        want, _ = oracle.render(32, 32, *FILM, 0, 8, 6, 16, estimator=estimator, threads=32)
        assert image_agreement(image, want) >= 0.985


def test_bvh_far_from_the_coordinate_origin(device, oracle_lib):
    """The quantised child boxes are padded for the rounding of the device's slab arithmetic, which scales with
    the scene's largest coordinate: a mesh pushed 1000 units away (and one corner of it left near the origin)
    must not lose hits -- a culled box shows up as a miss where the brute-force oracle hits."""
    from amber_b200 import FlatSceneBuilder
    from amber_b200.scene import icosphere
    rng = np.random.default_rng(5)
    offset = np.array([1000.0, -700.0, 850.0])
    b = FlatSceneBuilder()
    b.make_thin_lens([1, 0, 0, 1000, 0, 1, 0, -700, 0, 0, 1, 854, 0, 0, 0, 1], 0.05, 4.0, 0.05, 6, add_aperture_objects=False)
    grey = b.make_lambertian((.5, .5, .5))
    b.add_triangles(icosphere(4, center=tuple(offset), radius=2.0, displacement=0.1), grey)       # 5120 triangles far out
    b.add_triangles(icosphere(2, center=(0.0, 0.0, 0.0), radius=1.0, displacement=0.05), grey)   # 320 near the origin
    b.add_triangle((0, 0, 0), (1000, -700, 850), (1000, -700, 852), b.make_diffuse_light((1, 1, 1)))  # spans both
    scene = b.build()
    oracle = oracle_lib.Oracle(scene)
    device.upload_scene(scene)
    assert device.stats()["accel"] == 1
    n = 40000
    # rays from around either cluster towards it, and rays from one cluster to the other
    centres = np.where(rng.random((n, 1)) < 0.5, offset[None, :], np.zeros((1, 3)))
    o = (centres + unit_vectors(rng, n) * rng.uniform(3.0, 6.0, (n, 1))).astype(np.float32)
    target = centres + rng.normal(0, 0.8, (n, 3))
    target[: n // 8] = (offset[None, :] - centres[: n // 8]) + rng.normal(0, 0.8, (n // 8, 3))   # across the scene
    d = (target - o).astype(np.float64)
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    want_prim, want_t = oracle.cast(o, d, threads=32)
    prim, t = device.cast(o, d)
    lost = (want_prim >= 0) & (prim < 0)
    print(f"far scene: oracle hits {(want_prim >= 0).mean():.3f}, ids agree {(prim == want_prim).mean():.5f}, lost hits {lost.sum()}")
    assert (want_prim >= 0).mean() > 0.3
    # at 1000 units one float ulp is 6e-5: grazing hits may go to a neighbouring triangle or across a silhouette on
    # either side; a conservative traversal loses (almost) none of the oracle's hits
    assert lost.mean() <= 2e-4 and (prim == want_prim).mean() >= 0.995


def test_bvh_built_on_the_device(device, oracle_lib, monkeypatch):
    """AMBER_CUDA_BVH_BUILDER=lbvh: the tree comes from the device builder (bvh_build_gpu.cuh: Morton order, binary
    radix tree, collapsed to the same 4-wide quantised nodes).  Another tree over the same triangles: closest hits
    against the brute-force oracle, hit for hit against the host-built tree, and images against the oracle."""
    from amber_b200 import Sensor, cornel_mesh
    scene = cornel_mesh(subdivisions=5)  # 20 480 + 18 triangles
    oracle = oracle_lib.Oracle(scene)
    o, d = oracle.ray_set(256, 256, *FILM, 78, 60000)
    want_prim, want_t = oracle.cast(o, d, threads=32)
    results = {}
    for builder in ("host", "lbvh"):
        monkeypatch.setenv("AMBER_CUDA_BVH_BUILDER", builder)
        device.upload_scene(scene)
        device.set_sensor(Sensor(32, 32))
        stats = device.stats()
        assert stats["accel"] == 1 and stats["bvh_nodes"] > 1000
        if stats["bvh_node_bytes"] != 64:
            pytest.skip("library built with the 8-wide node option: the device builder writes 4-wide nodes only")
        results[builder] = (stats["bvh_nodes"],) + tuple(device.cast(o, d))
        check_cast(results[builder][1], results[builder][2], want_prim, want_t)
    monkeypatch.delenv("AMBER_CUDA_BVH_BUILDER")
    assert results["host"][0] != results["lbvh"][0], "the two builders gave the same node count: was the device builder used?"
    # the primitive ids reported are the caller's (creation order), whatever order a builder put the triangles in
    same = results["host"][1] == results["lbvh"][1]
    assert same.mean() >= ID_AGREEMENT
    assert np.allclose(results["host"][2][same], results["lbvh"][2][same], rtol=T_RTOL, atol=2e-6, equal_nan=True)
    for estimator in (0, 1):  # (the scene of the last upload is the device-built one)
        device.film_clear()
        device.render(0, 8, seed=6, max_depth=16, estimator=estimator)
        image, _ = device.resolve()
        want, _ = oracle.render(32, 32, *FILM, 0, 8, 6, 16, estimator=estimator, threads=32)
        assert image_agreement(image, want) >= 0.985


def test_bvh_full_size_mesh_against_brute_force(device, oracle_lib):
    """Config 4's geometry at full size (1 310 720-triangle displaced icosphere + room): the
    persistent BVH traversal against the oracle's brute-force scan on a ray set small enough
    for the CPU (4000 rays x 1.3 M triangles), plus whole-render invariants."""
    from amber_b200 import Sensor, cornel_mesh
    scene = cornel_mesh(subdivisions=8)
    oracle = oracle_lib.Oracle(scene)
    device.upload_scene(scene)
    device.set_sensor(Sensor(256, 256))
    stats = device.stats()
    assert stats["accel"] == 1 and stats["bvh_nodes"] > 100000
    # 2000 camera rays + 2000 secondary rays leaving the surfaces those hit (the recipe of the
    # fixed ray set, with the brute-force casts spread over the host threads)
    rng = np.random.default_rng(99)
    n = 2000
    co, cd, _ = oracle.eye_rays(256, 256, *FILM, rng.integers(0, 256, n), rng.integers(0, 256, n),
                                rng.random((n, 5), dtype=np.float32))
    cprim, ct = oracle.cast(co, cd, threads=32)
    hit = cprim >= 0
    so = (co + ct[:, None] * cd)[hit].astype(np.float32)
    sd = unit_vectors(rng, int(hit.sum()))
    o, d = np.concatenate([co, so]), np.concatenate([cd, sd])
    want_prim, want_t = oracle.cast(o, d, threads=32)
    prim, t = device.cast(o, d)
    # 4000 rays cannot resolve 99.99 %: at most one disagreement, and t within tolerance
    agree = prim == want_prim
    assert (~agree).sum() <= 1, f"{(~agree).sum()} hit ids differ"
    both = agree & (want_prim >= 0)
    # 1e-4 relative + an absolute floor of 2e-6 (a few ulp of the unit-scale coordinates): the
    # secondary rays start a rounding error away from a 3 mm triangle of a bumpy surface and
    # re-hit its neighbours at t ~ 1e-6..1e-3, where both sides' t is cancellation noise
    err = np.abs(t[both] - want_t[both])
    assert np.all(err <= T_RTOL * want_t[both] + 2e-6), f"max abs t error {err.max():.3e}"
    assert (err <= T_RTOL * want_t[both]).mean() >= 0.99
    assert (want_prim[:n] >= 18).mean() > 0.05  # the mesh itself is being hit, not only the room
    # a render through the wavefront: path count exact, batching invariance
    device.render(0, 4, seed=2, max_depth=16, estimator=1)
    a, _ = device.resolve()
    assert device.stats()["paths"] == 4 * 256 * 256
    device.film_clear()
    device.render(0, 1, seed=2, max_depth=16, estimator=1)
    device.render(1, 3, seed=2, max_depth=16, estimator=1, pool_paths=4096)
    b, _ = device.resolve()
    assert np.isfinite(a).all() and np.allclose(a, b, rtol=1e-4, atol=1e-7 * a.mean())


def test_bvh_full_size_mesh_against_the_reference_bvh(device):
    """Config 4's geometry at full size against golden hits produced by the reference's OWN BVH
    (include/amber/raytracer/acceleration_bvh.h, built over the 1.3 M triangles by
    scripts/make_golden.py --only mesh): 20 000 camera rays + their secondary rays."""
    from amber_b200 import cornel_mesh
    g = golden("mesh8_cast.npz")
    if device.scene is None or device.scene.n_primitives != 1310738:
        device.upload_scene(cornel_mesh(subdivisions=8))
    prim, t = device.cast(g["origins"], g["directions"])
    agree = prim == g["prim"]
    assert agree.mean() >= ID_AGREEMENT, f"hit ids agree on {agree.mean():.6f}"
    both = agree & (g["prim"] >= 0)
    err = np.abs(t[both] - g["t"][both])
    assert np.all(err <= T_RTOL * g["t"][both] + 2e-6), f"max abs t error {err.max():.3e}"
    assert (err <= T_RTOL * g["t"][both]).mean() >= 0.99
    assert np.isnan(t[agree & (g["prim"] < 0)]).all()
    print(f"mesh8 golden: id agreement {agree.mean():.6f} on {len(prim)} rays")


def test_device_built_bvh_at_full_size_against_the_reference_bvh(device, monkeypatch):
    """The same golden hits (the reference's own BVH over the 1.3 M triangles) through the tree built on the device."""
    from amber_b200 import cornel_mesh
    g = golden("mesh8_cast.npz")
    monkeypatch.setenv("AMBER_CUDA_BVH_BUILDER", "lbvh")
    device.upload_scene(cornel_mesh(subdivisions=8))
    monkeypatch.delenv("AMBER_CUDA_BVH_BUILDER")
    stats = device.stats()
    if stats["bvh_node_bytes"] != 64:
        pytest.skip("library built with the 8-wide node option: the device builder writes 4-wide nodes only")
    assert stats["bvh_nodes"] != 281134, "the host builder's node count: was the device builder used?"
    prim, t = device.cast(g["origins"], g["directions"])
    agree = prim == g["prim"]
    assert agree.mean() >= ID_AGREEMENT, f"hit ids agree on {agree.mean():.6f}"
    both = agree & (g["prim"] >= 0)
    err = np.abs(t[both] - g["t"][both])
    assert np.all(err <= T_RTOL * g["t"][both] + 2e-6), f"max abs t error {err.max():.3e}"
    assert np.isnan(t[agree & (g["prim"] < 0)]).all()
    print(f"mesh8 golden, device-built tree ({stats['bvh_nodes']} nodes, scene prepared in {stats['scene_build_ms']:.0f} ms): "
          f"id agreement {agree.mean():.6f}")
    device.scene = None  # the next test that wants this mesh uploads it again with the default builder


def test_full_size_properties(cornell_device):
    """Config 2 geometry (Cornell 1024x1024) at a few spp: properties that do not need an
    oracle run of that size."""
    from amber_b200 import Sensor
    cornell_device.set_sensor(Sensor(1024, 1024))
    cornell_device.render(0, 4, seed=1, max_depth=32, estimator=1)
    a, _ = cornell_device.resolve()
    stats = cornell_device.stats()
    assert stats["paths"] == 4 * 1024 * 1024
    # the reference measures 2.094 closest-hit rays per path on this scene
    assert 1.9 < stats["closest_rays"] / stats["paths"] < 2.3
    assert np.isfinite(a).all() and (a >= 0).all()
    # idempotence: the same samples again give the same film (up to float summation order)
    cornell_device.film_clear()
    cornell_device.render(0, 4, seed=1, max_depth=32, estimator=1)
    b, _ = cornell_device.resolve()
    assert np.allclose(a, b, rtol=1e-4, atol=1e-7 * a.mean())
    # linearity of the film: 2 + 2 samples == 4 samples
    cornell_device.film_clear()
    cornell_device.render(0, 2, seed=1, max_depth=32, estimator=1)
    cornell_device.render(2, 2, seed=1, max_depth=32, estimator=1)
    c, _ = cornell_device.resolve()
    assert np.allclose(a, c, rtol=1e-4, atol=1e-7 * a.mean())
    # pixel values scale with 1/(W*H) (lens_thin.cc:139-148): 4x dimmer than 512x512
    cornell_device.set_sensor(Sensor(512, 512))
    cornell_device.render(0, 16, seed=1, max_depth=32, estimator=1)
    small, _ = cornell_device.resolve()
    def robust_mean(image):  # firefly-free: mean of the luminance clipped at its 90th percentile
        lum = image.sum(-1)
        return float(np.minimum(lum, np.percentile(lum, 90)).mean())
    ratio = robust_mean(small) / robust_mean(a)
    assert 3.6 < ratio < 4.4, ratio
    cornell_device.set_sensor(Sensor(64, 64))


def test_tonemapped_resolve_matches_the_reference_postprocess(cornell_device, oracle_lib, tmp_path):
    """Row N4: resolve fused with Filmic + Gamma against the oracle's restatement (pinned bit for
    bit to the reference) applied to the float image of the same film; and the exported files."""
    from amber_b200 import Sensor, image_io
    cornell_device.set_sensor(Sensor(96, 64))
    cornell_device.render(0, 64, seed=4, max_depth=32, estimator=1)
    hdr, _ = cornell_device.resolve()
    ldr, done = cornell_device.resolve_ldr()
    want = oracle_lib.oracle_tonemap(hdr)
    assert done == 64 and ldr.shape == (64, 96, 3)
    diff = np.abs(ldr.astype(np.int16) - want.astype(np.int16))
    # device powf vs glibc powf: at most one 8-bit level, on values that sit on a level boundary
    assert diff.max() <= 1 and (diff > 0).mean() < 2e-3, (diff.max(), (diff > 0).mean())
    assert np.unique(want).size > 100  # the frame spans the curve, not just black and white
    flipped, _ = cornell_device.resolve_ldr(flip_x=True)
    assert np.array_equal(flipped, ldr[:, ::-1])
    png, pfm = image_io.export(cornell_device, str(tmp_path / "render"))
    assert np.array_equal(image_io.read_png(png), flipped)
    assert np.array_equal(image_io.read_pfm(pfm), hdr[:, ::-1])
    cornell_device.set_sensor(Sensor(64, 64))


def test_two_contexts_share_a_device(cornell, oracle):
    """Constant memory is per device: contexts must re-bind their scene."""
    from amber_b200 import Sensor
    from amber_b200.render import Device
    with Device(0) as a, Device(0) as b:
        a.upload_scene(cornell)
        a.set_sensor(Sensor(32, 32))
        b.upload_scene(diffuse_box())
        b.set_sensor(Sensor(16, 16))
        b.render(0, 2, seed=1)
        a.render(0, 4, seed=9, max_depth=8, estimator=0)
        image, _ = a.resolve()
        want, _ = oracle.render(32, 32, *FILM, 0, 4, 9, 8, estimator=0)
        assert image_agreement(image, want) >= 0.99


def test_errors_are_reported_not_swallowed(device):
    from amber_b200 import FlatSceneBuilder, capi
    from amber_b200.render import Device
    with Device(0) as fresh:
        with pytest.raises(capi.AmberCudaError, match="upload_scene"):
            fresh.render(0, 1)
    b = FlatSceneBuilder()
    b.make_thin_lens(np.eye(4), 0.05, 4.0, 0.05, 6)
    b.add_triangle((0, 0, 0), (1, 0, 0), (0, 1, 0), 7)  # material out of range
    with pytest.raises(capi.AmberCudaError, match="material"):
        device.upload_scene(b.build())


# ---- light tracing (row N2): src/amber/rendering/algorithm_lt.cc on the same kernels ------
def test_light_tracing_unit_functions(cornell_device):
    """GenerateLightRay, SampleImportance and the lens response against the compiled reference's
    golden vectors."""
    from amber_b200 import FlatSceneBuilder, Sensor
    g = golden("light_tracing.npz")
    o, d, w = cornell_device.light_rays(g["light_rays.xi"])
    assert np.array_equal(o, g["light_rays.origin"])  # host operations one by one: bit-identical
    assert np.allclose(d, g["light_rays.direction"], atol=5e-6)
    assert np.array_equal(w, g["light_rays.weight"])
    for (w_, h_) in ((64, 64), (1024, 512)):
        cornell_device.set_sensor(Sensor(w_, h_))
        pixel, value = cornell_device.lens_response(g["response.position"], g["response.dir_out"])
        want_pixel, want_value = g[f"response.{w_}x{h_}.pixel"], g[f"response.{w_}x{h_}.value"]
        same = pixel == want_pixel
        assert same.mean() >= 0.9995, same.mean()
        assert np.allclose(value[same], want_value[same], rtol=1e-5)
    cornell_device.set_sensor(Sensor(64, 64))
    names = sorted({k.split(".")[1] for k in g.files if k.startswith("importance.")})
    b = FlatSceneBuilder()
    b.make_thin_lens(np.eye(4), 0.05, 4.0, 0.05, 6, add_aperture_objects=False)
    index = {}
    for name in names:
        kind, r, gg, bb, param = g[f"importance.{name}.desc"]
        index[name] = b._material(int(kind), (r, gg, bb), param)
    b.add_triangle((0, 0, 0), (1, 0, 0), (0, 1, 0), index[names[0]])
    cornell_device.upload_scene(b.build())
    for name in names:
        key = f"importance.{name}"
        dir_in, weight, draws = cornell_device.sample_importance(index[name], g[f"{key}.normal"], g[f"{key}.wo"], g[f"{key}.xi"])
        same = draws == g[f"{key}.draws"]
        assert same.mean() >= 0.995, (name, same.mean())
        atol = 1e-4 if name.startswith("phong") else 5e-6
        assert np.allclose(dir_in[same], g[f"{key}.dir_in"][same], atol=atol), name
        assert np.allclose(weight[same], g[f"{key}.weight"][same], rtol=2e-4, atol=1e-7), name


def light_tracing_agreement(image, want):
    """Light-tracing images are sums of a few large splats per pixel: compare the pixels either
    image lights up."""
    lit = (want.sum(-1) > 0) | (image.sum(-1) > 0)
    ok = np.abs(image - want) <= 2e-3 * np.abs(want) + 1e-6 * want[lit].mean()
    return float(ok.all(-1)[lit].mean()), int(lit.sum())


def test_light_tracing_matches_reference_golden_and_oracle(device, oracle_lib):
    from amber_b200 import Sensor, capi
    g = golden("light_tracing.npz")
    box = light_tracing_box()
    device.upload_scene(box)
    device.set_sensor(Sensor(96, 64, 0.036, 0.024))
    for depth in (0, 3):  # images of the unmodified reference on the shared stream
        device.film_clear()
        device.render(0, 64, seed=24, max_depth=depth, estimator=capi.ESTIMATOR_LIGHT_TRACING)
        image, done = device.resolve()
        rays, truncated, splats = (int(v) for v in g[f"box.96x64.spp64.seed24.depth{depth}.stats"])
        frac, lit = light_tracing_agreement(image, g[f"box.96x64.spp64.seed24.depth{depth}.image"])
        got = device.stats()
        print(f"depth {depth}: {frac:.4f} of {lit} lit pixels agree; splats {got['light_splats']} vs {splats}")
        assert done == 64 and frac >= 0.98
        assert got["paths"] == 96 * 64 * 64 and got["shadow_rays"] == 0
        assert abs(got["light_splats"] - splats) <= max(3, 0.01 * splats)
        assert abs(got["closest_rays"] - rays) <= 2e-3 * rays
        assert abs(got["truncated_paths"] - truncated) <= max(8, 0.02 * truncated)
    # a larger run against the oracle: 4.2 M light paths
    oracle = oracle_lib.Oracle(box)
    device.set_sensor(Sensor(256, 256))
    device.render(5, 64, seed=77, estimator=capi.ESTIMATOR_LIGHT_TRACING)
    image, _ = device.resolve()
    want, stats = oracle.render_lt(256, 256, *FILM, 5, 64, 77, threads=32)
    frac, lit = light_tracing_agreement(image, want)
    got = device.stats()
    print(f"256x256x64: {frac:.4f} of {lit} lit pixels agree; splats {got['light_splats']} vs {stats['splats']}")
    assert frac >= 0.98 and lit > 2000
    assert abs(got["light_splats"] - stats["splats"]) <= 0.01 * stats["splats"]


def test_light_tracing_on_the_cornell_box(cornell_device, oracle):
    """The built-in scene: the aperture is 3 to 5 units from everything, so splats are rare
    (3 in 1e5 paths) -- every one of them must land where the oracle puts it."""
    from amber_b200 import Sensor, capi
    cornell_device.set_sensor(Sensor(128, 128))
    cornell_device.render(0, 256, seed=9, estimator=capi.ESTIMATOR_LIGHT_TRACING)
    image, _ = cornell_device.resolve()
    want, stats = oracle.render_lt(128, 128, *FILM, 0, 256, 9, threads=32)
    frac, lit = light_tracing_agreement(image, want)
    got = cornell_device.stats()
    print(f"cornell: {frac:.4f} of {lit} lit pixels agree; splats {got['light_splats']} vs {stats['splats']}")
    assert lit > 50 and frac >= 0.97
    assert abs(got["light_splats"] - stats["splats"]) <= max(2, 0.02 * stats["splats"])
    # sample partition invariance, as for the path tracer
    cornell_device.film_clear()
    cornell_device.render(0, 100, seed=9, estimator=capi.ESTIMATOR_LIGHT_TRACING)
    cornell_device.render(100, 156, seed=9, estimator=capi.ESTIMATOR_LIGHT_TRACING, pool_paths=4096)
    split, done = cornell_device.resolve()
    assert done == 256 and np.allclose(split, image, rtol=1e-4, atol=1e-7 * image.mean())


def test_light_tracing_agrees_with_path_tracing_in_expectation(device):
    """Both algorithms measure the same image: mean radiance per channel over the film, wide
    aperture, tens of millions of paths each."""
    from amber_b200 import Sensor, capi
    device.upload_scene(light_tracing_box())
    device.set_sensor(Sensor(32, 32))
    means = {}
    for estimator, spp in ((capi.ESTIMATOR_LIGHT_TRACING, 32768), (capi.ESTIMATOR_NEE, 2048)):
        runs = []
        for seed in range(4):
            device.film_clear()
            device.render(0, spp, seed=90 + seed, estimator=estimator)
            image, _ = device.resolve()
            runs.append(image.mean(axis=(0, 1)))
        means[estimator] = np.array(runs)
    lt, pt = means[capi.ESTIMATOR_LIGHT_TRACING], means[capi.ESTIMATOR_NEE]
    diff = lt.mean(0) - pt.mean(0)
    sigma = np.sqrt(lt.var(0, ddof=1) / 4 + pt.var(0, ddof=1) / 4)
    print("light tracing", lt.mean(0), "path tracing", pt.mean(0), "sigma", sigma)
    assert np.all(np.abs(diff) < 4 * sigma + 5e-3 * pt.mean(0)), (diff, sigma)


def test_light_tracing_refuses_an_empty_light_set(device):
    from amber_b200 import FlatSceneBuilder, Sensor, capi
    from amber_b200.capi import AmberCudaError
    b = FlatSceneBuilder()
    b.make_thin_lens(np.eye(4), 0.05, 4.0, 0.05, 6, add_aperture_objects=True)
    b.add_triangle((0, 0, -2), (1, 0, -2), (0, 1, -2), b.make_lambertian((.5, .5, .5)))
    device.upload_scene(b.build())
    device.set_sensor(Sensor(8, 8))
    for estimator in (capi.ESTIMATOR_LIGHT_TRACING, capi.ESTIMATOR_BDPT):
        with pytest.raises(AmberCudaError, match="emitter"):  # LightSet::GenerateRay on an empty set is undefined
            device.render(0, 1, estimator=estimator)
    device.render(0, 1, estimator=capi.ESTIMATOR_REFERENCE)  # the path tracer just sees black


def test_emitting_spheres(device, oracle_lib):
    """Sphere emitters in the light set (Sphere::SampleSurfacePoint, primitive_sphere.cc:115-122; SphereSA as
    the reference wrote it): light rays, light tracing and BDPT against the compiled reference's golden
    vectors; the path tracers (spheres are not light-sampled by NEE) against the oracle and each other."""
    from amber_b200 import Sensor, capi
    from conftest import sphere_light_box
    g = golden("sphere_lights.npz")
    scene = sphere_light_box()
    device.upload_scene(scene)
    device.set_sensor(Sensor(96, 64, 0.036, 0.024))
    o, d, w = device.light_rays(g["light_rays.xi"])
    # points on the triangle emitter use the host's operations one by one (bit-identical); points on a
    # sphere go through sincosf, which may differ from glibc's by an ulp
    on_triangle = np.abs(g["light_rays.origin"][:, 1] - np.float32(0.98)) < 1e-6
    assert on_triangle.any() and not on_triangle.all()
    assert np.array_equal(o[on_triangle], g["light_rays.origin"][on_triangle])
    assert np.allclose(o, g["light_rays.origin"], atol=2e-7) and np.array_equal(w, g["light_rays.weight"])
    assert np.allclose(d, g["light_rays.direction"], atol=5e-6)
    device.render(0, 64, seed=25, estimator=capi.ESTIMATOR_LIGHT_TRACING)
    image, _ = device.resolve()
    frac, lit = light_tracing_agreement(image, g["lt.96x64.begin0.spp64.seed25.image"])
    splats = int(g["lt.96x64.begin0.spp64.seed25.stats"][2])
    got = device.stats()
    print(f"sphere lights, light tracing: {frac:.4f} of {lit} lit pixels agree; splats {got['light_splats']} vs {splats}")
    assert frac >= 0.98 and abs(got["light_splats"] - splats) <= max(3, 0.01 * splats)
    device.set_sensor(Sensor(48, 32, 0.036, 0.024))
    device.render(0, 16, seed=33, max_depth=10, estimator=capi.ESTIMATOR_BDPT)
    image, _ = device.resolve()
    frac = image_agreement(image, g["bdpt.48x32.begin0.spp16.seed33.cap10.image"], rtol=5e-3)
    print(f"sphere lights, bdpt: {frac:.4f} of pixels agree")
    assert frac >= 0.98
    device.film_clear()
    device.render(1, 32, seed=7, estimator=capi.ESTIMATOR_REFERENCE)
    image, _ = device.resolve()
    assert image_agreement(image, g["pt.48x32.begin1.spp32.seed7.image"]) >= 0.99
    oracle = oracle_lib.Oracle(scene)
    device.film_clear()
    device.render(0, 16, seed=9, max_depth=12, estimator=capi.ESTIMATOR_NEE)
    image, _ = device.resolve()
    want, _ = oracle.render(48, 32, 0.036, 0.024, 0, 16, 9, 12, estimator=1)
    assert image_agreement(image, want) >= 0.99
    # NEE vs the reference estimator in expectation, with sphere and triangle emitters together
    device.set_sensor(Sensor(32, 32))
    means = {}
    for estimator, spp in ((0, 8192), (1, 2048)):
        runs = []
        for seed in range(4):
            device.film_clear()
            device.render(0, spp, seed=80 + seed, estimator=estimator)
            runs.append(device.resolve()[0].mean(axis=(0, 1)))
        means[estimator] = np.array(runs)
    diff = means[0].mean(0) - means[1].mean(0)
    sigma = np.sqrt(means[0].var(0, ddof=1) / 4 + means[1].var(0, ddof=1) / 4)
    assert np.all(np.abs(diff) < 4 * sigma + 2e-3 * means[0].mean(0)), (diff, sigma)


# ---- pinhole lens (row N1): src/amber/scene/lens_pinhole.cc --------------------------------
def test_pinhole_lens(device, oracle_lib):
    from amber_b200 import Sensor
    g = golden("pinhole.npz")
    scene = pinhole_box()
    device.upload_scene(scene)
    for (w, h) in ((64, 48), (1024, 512)):
        tag = f"{w}x{h}"
        device.set_sensor(Sensor(w, h))
        o, d, wt = device.eye_rays(g[f"{tag}.px"], g[f"{tag}.py"], g[f"{tag}.xi"])
        assert np.array_equal(o, g[f"{tag}.origin"]) and np.array_equal(d, g[f"{tag}.direction"])
        assert np.allclose(wt, g[f"{tag}.weight"], rtol=2e-5)
    device.set_sensor(Sensor(64, 48))
    pixel, value = device.lens_response(g["response.position"], g["response.dir_out"])
    same = pixel == g["response.64x48.pixel"]
    assert same.mean() >= 0.9995 and np.array_equal(value[same], g["response.64x48.value"][same])
    # the reference's image on the shared stream, then the NEE estimator against the oracle
    device.set_sensor(Sensor(48, 32, 0.036, 0.024))
    device.render(2, 32, seed=5, estimator=0)
    image, _ = device.resolve()
    assert image_agreement(image, g["48x32.begin2.spp32.seed5.image"]) >= 0.99
    rays = int(g["48x32.begin2.spp32.seed5.rays"][0])
    assert abs(device.stats()["closest_rays"] - rays) <= 2e-3 * rays
    oracle = oracle_lib.Oracle(scene)
    device.film_clear()
    device.render(0, 16, seed=6, max_depth=12, estimator=1)
    image, _ = device.resolve()
    want, _ = oracle.render(48, 32, 0.036, 0.024, 0, 16, 6, 12, estimator=1)
    assert image_agreement(image, want) >= 0.99


# ---- bidirectional path tracing (row N3) ------------------------------------------------------
def test_bidirectional_path_tracing_matches_reference_golden(device):
    """bdpt_kernel against images of the reference's own subpath generators + Combine on the shared
    sequential streams: Cornell box (two subpath caps), the wide-aperture box with every material
    kind, and the pinhole box."""
    from amber_b200 import Sensor, capi
    from test_oracle_golden import bdpt_cases
    for tag, make, w, h, fw, fh, begin, spp, seed, cap, want, stats in bdpt_cases():
        device.upload_scene(make())
        device.set_sensor(Sensor(w, h, fw, fh))
        device.render(begin, spp, seed=seed, max_depth=cap, estimator=capi.ESTIMATOR_BDPT)
        image, done = device.resolve()
        frac = image_agreement(image, want, rtol=5e-3)
        got = device.stats()
        print(f"{tag}: {frac:.4f} of pixels agree; mean {image.mean():.6g} vs {want.mean():.6g}; splats {got['light_splats']} vs {stats[2]}")
        assert done == spp and got["paths"] == w * h * spp
        assert frac >= 0.98, tag
        assert abs(got["light_splats"] - stats[2]) <= max(2, 0.05 * stats[2])


def test_bidirectional_path_tracing_matches_oracle_and_path_tracing(device, oracle_lib):
    from amber_b200 import Sensor, capi
    from amber_b200.capi import AmberCudaError
    scene = diffuse_box()
    oracle = oracle_lib.Oracle(scene)
    device.upload_scene(scene)
    device.set_sensor(Sensor(64, 64))
    device.render(0, 32, seed=12, estimator=capi.ESTIMATOR_BDPT)
    image, _ = device.resolve()
    want, stats = oracle.render_bdpt(64, 64, *FILM, 0, 32, 12, 16, threads=32)
    frac = image_agreement(image, want, rtol=5e-3)
    print(f"diffuse box 64x64x32: {frac:.4f} of pixels agree with the oracle")
    assert frac >= 0.98
    # same measurement as the path tracer: mean radiance per channel, many samples
    device.set_sensor(Sensor(32, 32))
    means = {}
    for estimator, spp in ((capi.ESTIMATOR_BDPT, 512), (capi.ESTIMATOR_NEE, 1024)):
        runs = []
        for seed in range(4):
            device.film_clear()
            device.render(0, spp, seed=70 + seed, estimator=estimator)
            runs.append(device.resolve()[0].mean(axis=(0, 1)))
        means[estimator] = np.array(runs)
    bd, pt = means[capi.ESTIMATOR_BDPT], means[capi.ESTIMATOR_NEE]
    diff = bd.mean(0) - pt.mean(0)
    sigma = np.sqrt(bd.var(0, ddof=1) / 4 + pt.var(0, ddof=1) / 4)
    print("bdpt", bd.mean(0), "pt", pt.mean(0), "sigma", sigma)
    assert np.all(np.abs(diff) < 4 * sigma + 2e-2 * pt.mean(0)), (diff, sigma)
    with pytest.raises(AmberCudaError):  # more events per subpath than the kernel stores
        device.render(0, 1, max_depth=17, estimator=capi.ESTIMATOR_BDPT)


def test_bdpt_wavefront_equals_the_one_kernel_version(device, cornell, monkeypatch):
    """Two implementations of the same estimator (bdpt.cuh): the wavefront (pool of pixel samples; grow, connect and
    evaluate kernels; the default) and round 1's one-kernel state machine.  Same random streams, same arithmetic: the
    images differ only by the order of the film's float additions; on the constant-memory list and on the BVH."""
    from amber_b200 import Sensor, capi, cornel_mesh
    for scene, size, spp in ((cornell, 96, 24), (cornel_mesh(subdivisions=3), 48, 8)):
        device.upload_scene(scene)
        device.set_sensor(Sensor(size, size))
        images, stats = {}, {}
        for mode in ("megakernel", "wavefront"):
            monkeypatch.setenv("AMBER_CUDA_BDPT", mode)
            device.film_clear()
            before = device.stats()
            device.render(0, spp, seed=21, estimator=capi.ESTIMATOR_BDPT)
            images[mode], done = device.resolve()
            after = device.stats()
            stats[mode] = tuple(after[k] - before[k] for k in ("paths", "closest_rays", "light_splats"))
            assert done == spp
        monkeypatch.delenv("AMBER_CUDA_BDPT")
        assert stats["wavefront"] == stats["megakernel"], stats           # pixel samples, casts, light-image splats
        assert stats["wavefront"][0] == spp * size * size
        a, b = images["megakernel"], images["wavefront"]
        assert np.isfinite(b).all()
        frac = image_agreement(b, a, rtol=1e-4)
        print(f"bdpt wavefront vs one kernel, {size}x{size}x{spp}: {frac:.5f} of pixels agree to 1e-4")
        assert frac >= 0.999
        assert abs(float(b.mean()) - float(a.mean())) <= 1e-5 * float(a.mean())
    # a pool smaller than the job: items take several pairs in turn
    monkeypatch.setenv("AMBER_CUDA_BDPT_POOL", "2048")
    monkeypatch.setenv("AMBER_CUDA_BDPT", "wavefront")
    device.film_clear()
    device.render(0, spp, seed=21, estimator=capi.ESTIMATOR_BDPT)
    small_pool, _ = device.resolve()
    assert image_agreement(small_pool, images["wavefront"], rtol=1e-4) >= 0.999


def test_light_and_bidirectional_tracing_through_the_bvh(device, oracle_lib):
    """Rows N2 / N3 on a scene above the constant-memory limit: the light tracer's wavefront and the
    BDPT kernel's per-thread walk of the 4-wide BVH against the brute-force oracle."""
    from amber_b200 import FlatSceneBuilder, Sensor, capi
    from amber_b200.scene import icosphere
    b = FlatSceneBuilder()
    b.make_thin_lens([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0.9, 0, 0, 0, 1], 0.05, 1.5, 0.25, 5, add_aperture_objects=True)

    def quad(m, p0, p1, p2, p3):
        b.add_triangle(p0, p1, p2, m)
        b.add_triangle(p2, p3, p0, m)

    grey = b.make_lambertian((.6, .6, .6))
    quad(grey, (-1, -1, 1), (1, -1, 1), (1, -1, -1), (-1, -1, -1))
    quad(grey, (-1, 1, 1), (-1, 1, -1), (1, 1, -1), (1, 1, 1))
    quad(b.make_lambertian((.6, .2, .2)), (-1, -1, 1), (-1, -1, -1), (-1, 1, -1), (-1, 1, 1))
    quad(grey, (1, -1, 1), (1, 1, 1), (1, 1, -1), (1, -1, -1))
    quad(b.make_phong((.7, .7, .7), 16), (-1, -1, -1), (1, -1, -1), (1, 1, -1), (-1, 1, -1))
    quad(grey, (-1, -1, 1), (-1, 1, 1), (1, 1, 1), (1, -1, 1))
    quad(b.make_diffuse_light((4.0, 4.0, 3.0)), (-0.5, 0.98, 0.5), (-0.5, 0.98, -0.5), (0.5, 0.98, -0.5), (0.5, 0.98, 0.5))
    b.add_triangles(icosphere(3, center=(0.1, -0.4, -0.2), radius=0.45), b.make_lambertian((.3, .5, .7)))  # 1280 triangles
    b.add_sphere((-0.5, -0.7, 0.3), 0.25, b.make_refraction(1.5))
    scene = b.build()
    oracle = oracle_lib.Oracle(scene)
    device.upload_scene(scene)
    assert device.stats()["accel"] == 1
    device.set_sensor(Sensor(64, 64))
    device.render(0, 64, seed=41, estimator=capi.ESTIMATOR_LIGHT_TRACING)
    image, _ = device.resolve()
    want, stats = oracle.render_lt(64, 64, *FILM, 0, 64, 41, threads=32)
    frac, lit = light_tracing_agreement(image, want)
    got = device.stats()
    print(f"lt through the BVH: {frac:.4f} of {lit} lit pixels agree; splats {got['light_splats']} vs {stats['splats']}")
    assert frac >= 0.98 and lit > 100 and abs(got["light_splats"] - stats["splats"]) <= max(3, 0.02 * stats["splats"])
    device.set_sensor(Sensor(48, 48))
    device.render(0, 8, seed=42, max_depth=10, estimator=capi.ESTIMATOR_BDPT)
    image, _ = device.resolve()
    want, _ = oracle.render_bdpt(48, 48, *FILM, 0, 8, 42, 10, threads=32)
    frac = image_agreement(image, want, rtol=5e-3)
    print(f"bdpt through the BVH: {frac:.4f} of pixels agree")
    assert frac >= 0.98
