"""The CPU oracle (oracle/amber_oracle.c) against the golden fixtures generated from the
UNMODIFIED reference by scripts/make_golden.py.  Bit-exact: the oracle restates the
reference float for float."""
import json
import os

import numpy as np
import pytest

from conftest import FILM, GOLDEN, golden


def test_philox_known_answers(oracle):
    # Random123 kat_vectors, philox4x32 10 rounds
    assert oracle.philox([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert oracle.philox([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert oracle.philox([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]
    u = oracle.stream(5, 17, 3, 4096)
    assert u.min() >= 0.0 and u.max() < 1.0 and abs(u.mean() - 0.5) < 0.02


def test_closest_hits_match_reference(oracle):
    g = golden("cornell_cast.npz")
    o, d = oracle.ray_set(1024, 1024, *FILM, int(g["seed"]), int(g["n"]))
    assert np.array_equal(o, g["origins"]) and np.array_equal(d, g["directions"])
    prim, t = oracle.cast(g["origins"], g["directions"])
    assert np.array_equal(prim, g["prim"])
    assert np.array_equal(t, g["t"], equal_nan=True)
    # the set exercises every primitive class
    hit = g["prim"][g["prim"] >= 0]
    assert set(np.unique(hit)) >= set(range(8, 25))


def test_materials_match_reference(oracle):
    from amber_b200 import capi
    g = golden("materials.npz")
    names = sorted({k.split(".")[0] for k in g.files})
    assert len(names) == 8
    for name in names:
        kind, r, gg, b, param = g[f"{name}.desc"]
        m = capi.Material(int(kind), (capi.C.c_float * 3)(r, gg, b), param)
        dir_in, weight, draws = oracle.sample_light(m, g[f"{name}.normal"], g[f"{name}.wo"], g[f"{name}.xi"])
        assert np.array_equal(dir_in, g[f"{name}.dir_in"]), name
        assert np.array_equal(weight, g[f"{name}.weight"]), name
        assert np.array_equal(draws, g[f"{name}.draws"]), name
        bsdf, radiance = oracle.bsdf(m, g[f"{name}.normal"], g[f"{name}.wo"], g[f"{name}.wi"])
        assert np.array_equal(radiance, g[f"{name}.radiance"]), name
        if int(kind) in (capi.MAT_LAMBERTIAN, capi.MAT_PHONG, capi.MAT_DIFFUSE_LIGHT):
            assert np.array_equal(bsdf, g[f"{name}.bsdf"]), name


def test_camera_rays_match_reference(oracle):
    g = golden("eye_rays.npz")
    for tag, (w, h) in {"64x64": (64, 64), "1024x512": (1024, 512)}.items():
        o, d, wt = oracle.eye_rays(w, h, 0.036, 0.036 * h / w, g[f"{tag}.px"], g[f"{tag}.py"], g[f"{tag}.xi"])
        assert np.array_equal(o, g[f"{tag}.origin"])
        assert np.array_equal(d, g[f"{tag}.direction"])
        assert np.array_equal(wt, g[f"{tag}.weight"])
    # pixel values scale with 1/(W*H): lens_thin.cc:139-148
    assert g["1024x512.weight"].mean() < g["64x64.weight"].mean() / 50


def test_rendered_images_match_reference(oracle):
    g = golden("cornell_stream.npz")
    tags = sorted({k.rsplit(".", 1)[0] for k in g.files})
    assert len(tags) == 2
    for tag in tags:
        size, seed, depth = tag.split(".")
        w, _, spp = (int(x) for x in size.split("x"))
        image, stats = oracle.render(w, w, *FILM, 0, spp, int(seed[4:]), int(depth[5:]), estimator=0)
        assert np.array_equal(image, g[f"{tag}.image"]), tag
        assert stats["closest_rays"] == int(g[f"{tag}.rays"][0])
        assert stats["truncated"] == int(g[f"{tag}.rays"][1])


def test_estimators_agree_in_expectation(oracle_lib):
    """NEE is not in the reference's PT; it must have the same expectation.  Low-variance
    scene (large light, diffuse + Phong): the two oracle estimators agree within 4 sigma."""
    from conftest import diffuse_box
    box = oracle_lib.Oracle(diffuse_box())
    w, spp = 16, 512
    ref_means, nee_means = [], []
    for seed in range(6):
        a, _ = box.render(w, w, *FILM, 0, spp, 100 + seed, 0, estimator=0)
        b, _ = box.render(w, w, *FILM, 0, spp // 8, 200 + seed, 0, estimator=1)
        ref_means.append(a.mean(axis=(0, 1)))
        nee_means.append(b.mean(axis=(0, 1)))
    ref_means, nee_means = np.array(ref_means), np.array(nee_means)
    diff = ref_means.mean(0) - nee_means.mean(0)
    sigma = np.sqrt(ref_means.var(0, ddof=1) / 6 + nee_means.var(0, ddof=1) / 6)
    assert np.all(np.abs(diff) < 4 * sigma + 1e-3 * ref_means.mean(0)), (diff, sigma)
    assert np.all(ref_means.mean(0) > 0)


def test_oracle_estimators_match_the_reference_renderer_on_the_cornell_box(oracle):
    """The Cornell box (mirror, glass, water: emission after non-connectable vertices) -- mean luminance of
    both oracle estimators against the reference's OWN renderer (MakeRGBPathTracing()->Render, independent
    runs recorded by scripts/make_golden.py) at two frame sizes (pixel values scale with 1 / (W * H)).  The
    2 cm light is found by BSDF sampling only, so a mean over 3e7 paths is still heavy-tailed and its sample
    standard error is an underestimate: the CPU check uses a 99 % interval (2.58 standard errors); the GPU
    repeats it with 4e9 paths per estimator and the 95 % interval (tests/test_image_parity.py)."""
    stats = json.load(open(os.path.join(GOLDEN, "cornell_luminance.json")))
    assert stats["runs"] == 16 and len(stats["pt_means"]) == 16
    conv = golden("cornell_converged.npz")
    for size, spp in ((64, 1024), (128, 384)):
        runs = conv[f"{size}x{size}.run_means"]
        cpu_mean, cpu_sem = runs.mean(), runs.std(ddof=1) / np.sqrt(len(runs))
        for estimator in (0, 1):
            means = np.array([oracle.render(size, size, *FILM, 0, spp, 500 + 10 * estimator + seed, 0, estimator=estimator)[0].mean()
                              for seed in range(8)])
            sem = means.std(ddof=1) / np.sqrt(len(means))
            assert abs(means.mean() - cpu_mean) <= 2.58 * np.hypot(cpu_sem, sem), (size, estimator, means.mean(), sem, cpu_mean, cpu_sem)
    # the two sizes of the golden agree with each other once scaled by the pixel count
    a, b = conv["64x64.run_means"], 4.0 * conv["128x128.run_means"]
    assert abs(a.mean() - b.mean()) <= 1.96 * np.hypot(a.std(ddof=1) / np.sqrt(len(a)), b.std(ddof=1) / np.sqrt(len(b)))
    assert abs(a.mean() - stats["pt_mean"]) <= 1.96 * np.hypot(a.std(ddof=1) / np.sqrt(len(a)), stats["pt_std_of_run"] / 4.0)


def test_image_metrics():
    from amber_b200.metrics import box_downsample, luminance, rrmse, water_line_rows
    rng = np.random.default_rng(3)
    image = rng.random((32, 48, 3))
    assert box_downsample(luminance(image), 4).shape == (8, 12)
    assert rrmse(image, image) == 0.0
    assert rrmse(1.1 * image, image) == pytest.approx(0.1 * np.sqrt(np.mean(box_downsample(luminance(image)) ** 2)) /
                                                      box_downsample(luminance(image)).mean())
    above, below = water_line_rows(1024)
    assert (above.start, above.stop) == (0, 158) and (below.start, below.stop) == (195, 256)
    above, below = water_line_rows(1024, 320, 32)
    assert (above.start, above.stop) == (0, 8) and below.stop - below.start <= 0 or below.start >= 8
    above, below = water_line_rows(1024, 800, 32)
    assert above.stop == 0 and (below.start, below.stop) == (0, 8)


def test_tonemap_matches_reference(oracle_lib):
    """Row N4: Filmic + Gamma + 8-bit truncation, bit for bit on the reference's golden ramp."""
    g = golden("tonemap.npz")
    ldr = oracle_lib.oracle_tonemap(g["hdr"])
    assert np.array_equal(ldr, g["ldr"])
    assert g["ldr"].min() == 0 and g["ldr"].max() == 255 and np.unique(g["ldr"]).size == 256


def test_light_tracing_matches_reference(oracle, oracle_lib):
    """Row N2: light rays, SampleImportance, lens response and light-tracing images of the
    compiled reference (tests/golden/light_tracing.npz), bit for bit."""
    from amber_b200 import capi
    from conftest import light_tracing_box
    g = golden("light_tracing.npz")
    o, d, w = oracle.light_rays(g["light_rays.xi"])
    assert np.array_equal(o, g["light_rays.origin"]) and np.array_equal(d, g["light_rays.direction"])
    assert np.array_equal(w, g["light_rays.weight"])
    # Irradiance / PDFArea of the Cornell light: pi * 1e11 * (total area 4e-4)
    assert np.allclose(w, np.pi * 1e11 * 4e-4, rtol=1e-5)
    names = sorted({k.split(".")[1] for k in g.files if k.startswith("importance.")})
    assert len(names) == 8
    for name in names:
        kind, r, gg, b, param = g[f"importance.{name}.desc"]
        m = capi.Material(int(kind), (capi.C.c_float * 3)(r, gg, b), param)
        dir_in, weight, draws = oracle.sample_importance(m, g[f"importance.{name}.normal"], g[f"importance.{name}.wo"],
                                                         g[f"importance.{name}.xi"])
        assert np.array_equal(dir_in, g[f"importance.{name}.dir_in"]), name
        assert np.array_equal(weight, g[f"importance.{name}.weight"]), name
        assert np.array_equal(draws, g[f"importance.{name}.draws"]), name
    for (w_, h_) in ((64, 64), (1024, 512)):
        pixel, value = oracle.lens_response(w_, h_, 0.036, 0.036 * h_ / w_, g["response.position"], g["response.dir_out"])
        assert np.array_equal(pixel, g[f"response.{w_}x{h_}.pixel"])
        assert np.array_equal(value, g[f"response.{w_}x{h_}.value"])
        assert 0.05 < (pixel >= 0).mean() < 0.95
    image, stats = oracle.render_lt(64, 64, *FILM, 3, 64, 23)
    assert np.array_equal(image, g["cornell.64x64.begin3.spp64.seed23.image"])
    assert stats["splats"] == int(g["cornell.64x64.begin3.spp64.seed23.stats"][2]) > 0
    box = oracle_lib.Oracle(light_tracing_box())
    for depth in (0, 3):
        image, stats = box.render_lt(96, 64, 0.036, 0.024, 0, 64, 24, depth)
        want = g[f"box.96x64.spp64.seed24.depth{depth}.stats"]
        assert np.array_equal(image, g[f"box.96x64.spp64.seed24.depth{depth}.image"]), depth
        assert (stats["closest_rays"], stats["truncated"], stats["splats"]) == tuple(int(v) for v in want)
    assert int(g["box.96x64.spp64.seed24.depth3.stats"][1]) > 0


def test_pinhole_lens_matches_reference(oracle_lib):
    """Row N1's second lens, MakePinholeLens (tests/golden/pinhole.npz)."""
    from conftest import pinhole_box
    g = golden("pinhole.npz")
    orc = oracle_lib.Oracle(pinhole_box())
    for (w, h) in ((64, 48), (1024, 512)):
        tag = f"{w}x{h}"
        o, d, wt = orc.eye_rays(w, h, 0.036, 0.036 * h / w, g[f"{tag}.px"], g[f"{tag}.py"], g[f"{tag}.xi"])
        assert np.array_equal(o, g[f"{tag}.origin"]) and np.array_equal(d, g[f"{tag}.direction"])
        assert np.array_equal(wt, g[f"{tag}.weight"])
        assert np.all(o == np.array([0.2, 0.1, 0.9], np.float32))  # every ray leaves the lens origin
    image, stats = orc.render(48, 32, 0.036, 0.024, 2, 32, 5, 0, estimator=0)
    assert np.array_equal(image, g["48x32.begin2.spp32.seed5.image"])
    assert stats["closest_rays"] == int(g["48x32.begin2.spp32.seed5.rays"][0])
    pixel, value = orc.lens_response(64, 48, 0.036, 0.027, g["response.position"], g["response.dir_out"])
    assert np.array_equal(pixel, g["response.64x48.pixel"]) and np.array_equal(value, g["response.64x48.value"])


def bdpt_cases():
    from amber_b200 import cornel_box
    from conftest import light_tracing_box, pinhole_box
    g = golden("bdpt.npz")
    scenes = {"cornell": cornel_box, "ltbox": light_tracing_box, "pinhole": pinhole_box}
    for tag in sorted({k.rsplit(".", 1)[0] for k in g.files}):
        name, size, begin, spp, seed, cap = tag.split(".")
        w, h = (int(v) for v in size.split("x"))
        yield (tag, scenes[name], w, h, 0.036, 0.036 * h / w, int(begin[5:]), int(spp[3:]), int(seed[4:]), int(cap[3:]),
               g[f"{tag}.image"], tuple(int(v) for v in g[f"{tag}.stats"]))


def test_bidirectional_path_tracing_matches_reference(oracle_lib):
    """Row N3: images of the reference's own subpath generators and Combine (tests/golden/bdpt.npz)."""
    n = 0
    for tag, make, w, h, fw, fh, begin, spp, seed, cap, want, stats in bdpt_cases():
        image, got = oracle_lib.Oracle(make()).render_bdpt(w, h, fw, fh, begin, spp, seed, cap)
        assert np.array_equal(image, want), tag
        assert (got["light_events"], got["eye_events"], got["splats"]) == stats, tag
        n += 1
    assert n == 4


def test_sphere_lights_match_reference(oracle_lib):
    """Emitting spheres: light rays (Sphere::SampleSurfacePoint -> SphereSA as written) and the images of
    the three algorithms on tests/golden/sphere_lights.npz, bit for bit."""
    from conftest import sphere_light_box
    g = golden("sphere_lights.npz")
    orc = oracle_lib.Oracle(sphere_light_box())
    o, d, w = orc.light_rays(g["light_rays.xi"])
    assert np.array_equal(o, g["light_rays.origin"]) and np.array_equal(d, g["light_rays.direction"])
    assert np.array_equal(w, g["light_rays.weight"])
    image, stats = orc.render(48, 32, 0.036, 0.024, 1, 32, 7, 0, estimator=0)
    assert np.array_equal(image, g["pt.48x32.begin1.spp32.seed7.image"])
    assert stats["closest_rays"] == int(g["pt.48x32.begin1.spp32.seed7.rays"][0])
    image, stats = orc.render_lt(96, 64, 0.036, 0.024, 0, 64, 25, 0)
    assert np.array_equal(image, g["lt.96x64.begin0.spp64.seed25.image"])
    assert stats["splats"] == int(g["lt.96x64.begin0.spp64.seed25.stats"][2]) > 100
    image, stats = orc.render_bdpt(48, 32, 0.036, 0.024, 0, 16, 33, 10)
    assert np.array_equal(image, g["bdpt.48x32.begin0.spp16.seed33.cap10.image"])
    assert [stats["light_events"], stats["eye_events"], stats["splats"]] == [int(v) for v in g["bdpt.48x32.begin0.spp16.seed33.cap10.stats"]]
    # NEE (not in the reference: spheres are not light-sampled, their emission counts on every hit) has the
    # reference estimator's expectation
    means = {}
    for estimator, spp in ((0, 1024), (1, 256)):
        means[estimator] = np.array([orc.render(24, 16, 0.036, 0.024, 0, spp, 60 + 7 * estimator + seed, 0, estimator=estimator)[0].mean(axis=(0, 1))
                                     for seed in range(6)])
    diff = means[0].mean(0) - means[1].mean(0)
    sigma = np.sqrt(means[0].var(0, ddof=1) / 6 + means[1].var(0, ddof=1) / 6)
    assert np.all(np.abs(diff) < 4 * sigma + 1e-3 * means[0].mean(0)), (diff, sigma)
