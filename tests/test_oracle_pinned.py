"""The CPU oracle against the compiled UNMODIFIED reference (oracle/_ref/libamber_ref.so),
on fresh random inputs.  Skipped where the reference was not built; the committed golden
fixtures (test_oracle_golden.py) cover that case."""
import numpy as np

from conftest import FILM, light_tracing_box, pinhole_box, unit_vectors


def test_flat_cornell_is_the_builtin_cornel_box(reference, oracle_lib):
    builtin = oracle_lib.Reference(builtin_cornel_box=True)
    a, sa = reference.render_stream(32, 32, *FILM, 0, 32, 5)
    b, sb = builtin.render_stream(32, 32, *FILM, 0, 32, 5)
    assert np.array_equal(a, b) and sa == sb


def test_cast_bit_exact(reference, oracle):
    o, d = oracle.ray_set(512, 512, *FILM, 99, 50000)
    for accel in (0, 1):  # reference BVH and reference List
        prim, t = reference.cast(o, d, accel=accel)
        oprim, ot = oracle.cast(o, d)
        assert np.array_equal(prim, oprim)
        assert np.array_equal(t, ot, equal_nan=True)


def test_random_rays_inside_the_box(reference, oracle):
    rng = np.random.default_rng(3)
    o = rng.uniform(-0.99, 0.99, size=(50000, 3)).astype(np.float32)
    d = unit_vectors(rng, 50000)
    prim, t = reference.cast(o, d)
    oprim, ot = oracle.cast(o, d)
    assert np.array_equal(prim, oprim) and np.array_equal(t, ot, equal_nan=True)


def test_camera_rays_bit_exact(reference, oracle):
    rng = np.random.default_rng(4)
    for (w, h) in ((37, 91), (1024, 1024)):
        n = 4000
        px, py = rng.integers(0, w, n), rng.integers(0, h, n)
        xi = rng.random((n, 5), dtype=np.float32)
        film = (0.036, 0.036 * h / w)
        a = reference.eye_rays(w, h, *film, px, py, xi)
        b = oracle.eye_rays(w, h, *film, px, py, xi)
        for x, y in zip(a, b):
            assert np.array_equal(x, y)


def test_materials_bit_exact(reference, oracle):
    from amber_b200 import capi
    rng = np.random.default_rng(5)
    n = 5000
    for kind, rgb, param in [(0, (.5, .4, .3), 0), (1, (.95, .9, .8), 256), (1, (.9, .9, .9), 2), (2, (.9, .9, .9), 0),
                             (3, (1, 1, 1), 1.333), (3, (1, 1, 1), 2.4), (4, (5, 5, 5), 0), (5, (1, 1, 1), 0)]:
        m = capi.Material(kind, (capi.C.c_float * 3)(*rgb), param)
        nrm, wo, wi = unit_vectors(rng, n), unit_vectors(rng, n), unit_vectors(rng, n)
        xi = rng.random((n, 8), dtype=np.float32)
        a = reference.sample_light(m, nrm, wo, xi)
        b = oracle.sample_light(m, nrm, wo, xi)
        for x, y in zip(a, b):
            assert np.array_equal(x, y), (kind, param)
        fa, la = reference.bsdf(m, nrm, wo, wi)
        fb, lb = oracle.bsdf(m, nrm, wo, wi)
        assert np.array_equal(la, lb)
        if kind in (0, 1, 4):
            assert np.array_equal(fa, fb)


def test_images_bit_exact_including_depth_cap(reference, oracle):
    for (w, spp, seed, depth) in ((40, 48, 21, 0), (24, 64, 22, 3)):
        a, sa = reference.render_stream(w, w, *FILM, 8, spp, seed, depth)
        b, sb = oracle.render(w, w, *FILM, 8, spp, seed, depth, estimator=0)
        assert np.array_equal(a, b)
        assert sa["rays"] == sb["closest_rays"] and sa["truncated"] == sb["truncated"]
        if depth:
            assert sb["truncated"] > 0


def test_tonemap_bit_exact(reference, oracle_lib):
    """Gamma()(Filmic()(image)) of the compiled reference against the C restatement."""
    rng = np.random.default_rng(8)
    hdr = (10.0 ** rng.uniform(-10, 3, size=(64, 96, 3))).astype(np.float32)
    hdr[0, :8] = 0.0
    assert np.array_equal(oracle_lib.reference_tonemap(hdr), oracle_lib.oracle_tonemap(hdr))


# ---- light tracing (row N2) -------------------------------------------------------------
def test_light_rays_bit_exact(reference, oracle):
    rng = np.random.default_rng(31)
    xi = rng.random((20000, 5), dtype=np.float32)
    for x, y in zip(reference.light_rays(xi), oracle.light_rays(xi)):
        assert np.array_equal(x, y)


def test_sample_importance_bit_exact(reference, oracle):
    from amber_b200 import capi
    rng = np.random.default_rng(32)
    n = 5000
    for kind, rgb, param in [(0, (.5, .4, .3), 0), (1, (.95, .9, .8), 64), (2, (.9, .9, .9), 0),
                             (3, (1, 1, 1), 1.333), (3, (1, 1, 1), 1.125), (3, (1, 1, 1), 2.4),
                             (4, (5, 5, 5), 0), (5, (1, 1, 1), 0)]:
        m = capi.Material(kind, (capi.C.c_float * 3)(*rgb), param)
        nrm, wo = unit_vectors(rng, n), unit_vectors(rng, n)
        xi = rng.random((n, 8), dtype=np.float32)
        for x, y in zip(reference.sample_importance(m, nrm, wo, xi), oracle.sample_importance(m, nrm, wo, xi)):
            assert np.array_equal(x, y), (kind, param)
        if kind == 3:  # the one material whose importance sampling is not its radiance sampling
            wl = oracle.sample_light(m, nrm, wo, xi)[1]
            wi = oracle.sample_importance(m, nrm, wo, xi)[1]
            assert not np.array_equal(wl, wi)


def lens_response_inputs(rng, n, lens_z, radius):
    """Points on (and around) the aperture disc at z = lens_z, directions mostly towards -z."""
    r = radius * np.sqrt(rng.random(n)) * 1.05
    phi = rng.uniform(0, 2 * np.pi, n)
    pos = np.stack([r * np.cos(phi), r * np.sin(phi), np.full(n, lens_z)], axis=1).astype(np.float32)
    d = np.stack([rng.normal(0, 0.3, n), rng.normal(0, 0.3, n), -np.abs(rng.normal(1, 0.3, n))], axis=1)
    d[: n // 20, 2] *= -1  # a few arrive from behind the lens
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    return pos, d


def test_lens_response_bit_exact(reference, oracle):
    rng = np.random.default_rng(33)
    pos, d = lens_response_inputs(rng, 20000, 4.0, 0.05)
    for (w, h) in ((64, 64), (1024, 512), (37, 91)):
        film = (0.036, 0.036 * h / w)
        pa, va = reference.lens_response(w, h, *film, pos, d)
        pb, vb = oracle.lens_response(w, h, *film, pos, d)
        assert np.array_equal(pa, pb) and np.array_equal(va, vb)
        assert (pa >= 0).mean() > 0.05


def test_light_tracing_images_bit_exact(reference, oracle, oracle_lib):
    a, sa = reference.render_lt_stream(64, 64, *FILM, 3, 32, 23)
    b, sb = oracle.render_lt(64, 64, *FILM, 3, 32, 23)
    assert np.array_equal(a, b) and sa["splats"] == sb["splats"]
    # a scene where the lens is hit often, with every material kind, and a depth cap
    box = light_tracing_box()
    ref_box, orc_box = oracle_lib.Reference(box), oracle_lib.Oracle(box)
    for depth in (0, 3):
        a, sa = ref_box.render_lt_stream(96, 64, 0.036, 0.024, 0, 64, 24, depth)
        b, sb = orc_box.render_lt(96, 64, 0.036, 0.024, 0, 64, 24, depth)
        assert sa["splats"] == sb["splats"] and sa["splats"] > 100
        assert sa["rays"] == sb["closest_rays"] and sa["truncated"] == sb["truncated"]
        assert np.array_equal(a, b)


# ---- pinhole lens (row N1): src/amber/scene/lens_pinhole.cc ---------------------------------
def test_pinhole_lens_bit_exact(oracle_lib):
    if not oracle_lib.reference_available():
        import pytest
        pytest.skip("oracle/_ref/libamber_ref.so not built (needs /root/reference)")
    flat = pinhole_box()
    ref, orc = oracle_lib.Reference(flat), oracle_lib.Oracle(flat)
    rng = np.random.default_rng(41)
    for (w, h) in ((64, 48), (1024, 512)):
        n = 5000
        px, py = rng.integers(0, w, n), rng.integers(0, h, n)
        xi = rng.random((n, 5), dtype=np.float32)
        film = (0.036, 0.036 * h / w)
        for x, y in zip(ref.eye_rays(w, h, *film, px, py, xi), orc.eye_rays(w, h, *film, px, py, xi)):
            assert np.array_equal(x, y)
    for estimator_depth in (0, 4):
        a, sa = ref.render_stream(48, 32, 0.036, 0.024, 2, 32, 5, estimator_depth)
        b, sb = orc.render(48, 32, 0.036, 0.024, 2, 32, 5, estimator_depth, estimator=0)
        assert np.array_equal(a, b) and sa["rays"] == sb["closest_rays"] and a.mean() > 0
    d = unit_vectors(rng, 5000)
    pos = np.tile(np.array([[0.2, 0.1, 0.9]], np.float32), (5000, 1))
    pa, va = ref.lens_response(64, 48, 0.036, 0.027, pos, d)
    pb, vb = orc.lens_response(64, 48, 0.036, 0.027, pos, d)
    assert np.array_equal(pa, pb) and np.array_equal(va, vb) and (pa >= 0).any()


# ---- bidirectional path tracing (row N3) ----------------------------------------------------
def test_bidirectional_path_tracing_bit_exact(reference, oracle, oracle_lib):
    """The C restatement of bidirectional_path_sampling.h against the reference's own
    GenerateLightSubpath / GenerateEyeSubpath / Combine (called unmodified by the harness): images,
    subpath lengths and light-image splats, with and without a tight cap on the subpath length."""
    for cap in (12, 3):
        a, sa = reference.render_bdpt_stream(40, 40, *FILM, 2, 16, 31, cap)
        b, sb = oracle.render_bdpt(40, 40, *FILM, 2, 16, 31, cap)
        assert sa == sb and np.array_equal(a, b) and a.mean() > 0
    for flat, film in ((light_tracing_box(), (0.036, 0.024)), (pinhole_box(), (0.036, 0.024))):
        ref_box, orc_box = oracle_lib.Reference(flat), oracle_lib.Oracle(flat)
        a, sa = ref_box.render_bdpt_stream(48, 32, *film, 0, 16, 32, 10)
        b, sb = orc_box.render_bdpt(48, 32, *film, 0, 16, 32, 10)
        assert sa == sb and np.array_equal(a, b) and a.mean() > 0


# ---- emitting spheres: Sphere::SampleSurfacePoint in LightSet::GenerateRay ------------------
def test_sphere_lights_bit_exact(oracle_lib):
    """A scene whose light set holds a triangle pair and two emitting spheres: light rays, the path tracer,
    the light tracer and BDPT of the C restatement against the compiled reference, bit for bit."""
    if not oracle_lib.reference_available():
        import pytest
        pytest.skip("oracle/_ref/libamber_ref.so not built (needs /root/reference)")
    from conftest import sphere_light_box
    flat = sphere_light_box()
    ref, orc = oracle_lib.Reference(flat), oracle_lib.Oracle(flat)
    xi = np.random.default_rng(51).random((20000, 5), dtype=np.float32)
    got, want = orc.light_rays(xi), ref.light_rays(xi)
    for x, y in zip(got, want):
        assert np.array_equal(x, y)
    on_sphere = np.abs(np.linalg.norm(want[0] - np.array([0.45, 0.1, -0.3], np.float32), axis=1) - 0.2) < 1e-5
    assert 0.2 < on_sphere.mean() < 0.9 and (want[0][on_sphere][:, 2] >= -0.3 - 1e-6).all()  # SphereSA: z >= 0 only
    a, sa = ref.render_stream(48, 32, 0.036, 0.024, 1, 16, 7, 0)
    b, sb = orc.render(48, 32, 0.036, 0.024, 1, 16, 7, 0, estimator=0)
    assert np.array_equal(a, b) and sa["rays"] == sb["closest_rays"]
    a, sa = ref.render_lt_stream(96, 64, 0.036, 0.024, 0, 64, 25, 0)
    b, sb = orc.render_lt(96, 64, 0.036, 0.024, 0, 64, 25, 0)
    assert sa["splats"] == sb["splats"] and sa["splats"] > 100 and np.array_equal(a, b)
    a, sa = ref.render_bdpt_stream(48, 32, 0.036, 0.024, 0, 16, 33, 10)
    b, sb = orc.render_bdpt(48, 32, 0.036, 0.024, 0, 16, 33, 10)
    assert sa == sb and np.array_equal(a, b) and a.mean() > 0
