#!/usr/bin/env python
"""Generate tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libamber_ref.so).

Run in the build container (where /root/reference exists):
    make -C oracle ref && python scripts/make_golden.py
The fixtures are what keeps the CPU oracle pinned on machines without /root/reference
(the GPU box): tests/test_oracle_golden.py replays these inputs through
oracle/amber_oracle.c and requires bit-identical outputs.  The reference has no tests or
golden vectors of its own -- every expected value here is produced by
executing its code.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from amber_b200 import capi, scene  # noqa: E402
from amber_b200.metrics import box_downsample, luminance  # noqa: E402
from oracle import pyoracle  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
FILM = (0.036, 0.036)

# (name, kind, rgb, param)
MATERIALS = [
    ("lambertian", capi.MAT_LAMBERTIAN, (.5, .4, .3), 0.0),
    ("phong256", capi.MAT_PHONG, (.95, .95, .95), 256.0),
    ("phong8", capi.MAT_PHONG, (.9, .8, .7), 8.0),
    ("specular", capi.MAT_SPECULAR, (.95, .95, .95), 0.0),
    ("water", capi.MAT_REFRACTION, (1, 1, 1), 1.333),
    ("glass", capi.MAT_REFRACTION, (1, 1, 1), 1.125),
    ("light", capi.MAT_DIFFUSE_LIGHT, (7.0, 5.0, 3.0), 0.0),
    ("eye", capi.MAT_EYE, (1, 1, 1), 0.0),
]


def unit(rng, n):
    v = rng.normal(size=(n, 3)).astype(np.float32)
    return (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(np.float32)


def tonemap_fixture():
    """6. post-process (row N4): Gamma()(Filmic()(image)) of the reference on a value ramp that
    covers black, the toe, the shoulder, level boundaries and overflow."""
    rng = np.random.default_rng(20261104)
    values = np.concatenate([np.zeros(6), 10.0 ** rng.uniform(-9, 2, 6000), np.linspace(0, 0.25, 2988),
                             [1e30, 3.0e38, 1e-30, 1e-38, 0.5, 1.0]]).astype(np.float32)
    hdr = values.reshape(50, 60, 3)
    ldr = pyoracle.reference_tonemap(hdr)
    np.savez_compressed(os.path.join(GOLDEN, "tonemap.npz"), hdr=hdr, ldr=ldr)
    print("tonemap: levels used", np.unique(ldr).size)


def mesh_fixture():
    """7. config 4's geometry at FULL size (1 310 720-triangle displaced icosphere in the room)
    through the reference's OWN BVH (acceleration_bvh.h; its build takes minutes at this size):
    20 000 camera rays + the secondary rays leaving the surfaces they hit."""
    flat = scene.cornel_mesh(8)
    ref = pyoracle.Reference(flat)
    rng = np.random.default_rng(20261104)
    n = 20000
    px, py = rng.integers(0, 2048, n).astype(np.uint32), rng.integers(0, 2048, n).astype(np.uint32)
    co, cd, _ = ref.eye_rays(2048, 2048, *FILM, px, py, rng.random((n, 5), dtype=np.float32))
    cprim, ct = ref.cast(co, cd, accel=0)
    hit = cprim >= 0
    so = (co + ct[:, None] * cd)[hit].astype(np.float32)
    sd = unit(rng, int(hit.sum()))
    o, d = np.concatenate([co, so]), np.concatenate([cd, sd])
    prim, t = ref.cast(o, d, accel=0)
    # the reference's brute-force List agrees with its BVH (checked on a subset: 1.3 M tests per ray)
    sub = rng.choice(len(o), 300, replace=False)
    lprim, lt = ref.cast(o[sub], d[sub], accel=1)
    assert np.array_equal(lprim, prim[sub]) and np.array_equal(lt, t[sub], equal_nan=True)
    np.savez_compressed(os.path.join(GOLDEN, "mesh8_cast.npz"), origins=o, directions=d, prim=prim, t=t)
    print("mesh8: rays", len(o), "hit fraction", float((prim >= 0).mean()), "on the mesh", float((prim >= 18).mean()))


def light_tracing_fixture():
    """Row N2: light rays, SampleImportance, lens response and whole light-tracing images from the
    compiled reference (ref_light_rays / ref_sample_importance / ref_lens_response /
    ref_render_lt_stream in oracle/ref_harness.cc)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import light_tracing_box
    from test_oracle_pinned import lens_response_inputs
    rng = np.random.default_rng(20261105)
    flat = scene.cornel_box()
    ref = pyoracle.Reference(flat)
    out = {}
    xi = rng.random((2048, 5), dtype=np.float32)
    o, d, w = ref.light_rays(xi)
    out.update({"light_rays.xi": xi, "light_rays.origin": o, "light_rays.direction": d, "light_rays.weight": w})
    for name, kind, rgb, param in MATERIALS:
        material = capi.Material(kind, (capi.C.c_float * 3)(*rgb), param)
        normal, wo = unit(rng, 512), unit(rng, 512)
        xi8 = rng.random((512, 8), dtype=np.float32)
        dir_in, weight, draws = ref.sample_importance(material, normal, wo, xi8)
        for key, val in dict(normal=normal, wo=wo, xi=xi8, dir_in=dir_in, weight=weight, draws=draws).items():
            out[f"importance.{name}.{key}"] = val
        out[f"importance.{name}.desc"] = np.array([kind, *rgb, param], dtype=np.float64)
    pos, dirs = lens_response_inputs(rng, 4096, 4.0, 0.05)
    out["response.position"], out["response.dir_out"] = pos, dirs
    for (w_, h_) in ((64, 64), (1024, 512)):
        pixel, value = ref.lens_response(w_, h_, 0.036, 0.036 * h_ / w_, pos, dirs)
        out[f"response.{w_}x{h_}.pixel"], out[f"response.{w_}x{h_}.value"] = pixel, value
    image, stats = ref.render_lt_stream(64, 64, *FILM, 3, 64, 23)
    out["cornell.64x64.begin3.spp64.seed23.image"] = image
    out["cornell.64x64.begin3.spp64.seed23.stats"] = np.array([stats["rays"], stats["truncated"], stats["splats"]], dtype=np.uint64)
    box = pyoracle.Reference(light_tracing_box())
    for depth in (0, 3):
        image, stats = box.render_lt_stream(96, 64, 0.036, 0.024, 0, 64, 24, depth)
        out[f"box.96x64.spp64.seed24.depth{depth}.image"] = image
        out[f"box.96x64.spp64.seed24.depth{depth}.stats"] = np.array([stats["rays"], stats["truncated"], stats["splats"]], dtype=np.uint64)
    np.savez_compressed(os.path.join(GOLDEN, "light_tracing.npz"), **out)


def pinhole_fixture():
    """Row N1's pinhole lens (MakePinholeLens): camera rays, an image on the shared stream and
    the lens response, from the compiled reference."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import pinhole_box
    rng = np.random.default_rng(20261106)
    ref = pyoracle.Reference(pinhole_box())
    out = {}
    for (w, h) in ((64, 48), (1024, 512)):
        k = 2048
        px, py = rng.integers(0, w, k).astype(np.uint32), rng.integers(0, h, k).astype(np.uint32)
        xi = rng.random((k, 5), dtype=np.float32)
        origin, direction, weight = ref.eye_rays(w, h, 0.036, 0.036 * h / w, px, py, xi)
        for key, val in dict(px=px, py=py, xi=xi, origin=origin, direction=direction, weight=weight).items():
            out[f"{w}x{h}.{key}"] = val
    image, stats = ref.render_stream(48, 32, 0.036, 0.024, 2, 32, 5, 0)
    out["48x32.begin2.spp32.seed5.image"] = image
    out["48x32.begin2.spp32.seed5.rays"] = np.array([stats["rays"], stats["truncated"]], dtype=np.uint64)
    d = unit(rng, 4096)
    pos = np.tile(np.array([[0.2, 0.1, 0.9]], np.float32), (4096, 1))
    pixel, value = ref.lens_response(64, 48, 0.036, 0.027, pos, d)
    out.update({"response.position": pos, "response.dir_out": d, "response.64x48.pixel": pixel, "response.64x48.value": value})
    np.savez_compressed(os.path.join(GOLDEN, "pinhole.npz"), **out)


def bdpt_fixture():
    """Row N3: images of the reference's own GenerateLightSubpath / GenerateEyeSubpath / Combine on the
    sequential subpath streams (ref_render_bdpt_stream)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import light_tracing_box, pinhole_box
    out = {}
    ref = pyoracle.Reference(scene.cornel_box())
    for (w, spp, seed, cap) in ((48, 16, 31, 12), (32, 16, 33, 3)):
        image, stats = ref.render_bdpt_stream(w, w, *FILM, 2, spp, seed, cap)
        tag = f"cornell.{w}x{w}.begin2.spp{spp}.seed{seed}.cap{cap}"
        out[f"{tag}.image"] = image
        out[f"{tag}.stats"] = np.array([stats["light_events"], stats["eye_events"], stats["splats"]], dtype=np.uint64)
    for name, flat in (("ltbox", light_tracing_box()), ("pinhole", pinhole_box())):
        box = pyoracle.Reference(flat)
        image, stats = box.render_bdpt_stream(48, 32, 0.036, 0.024, 0, 16, 32, 10)
        tag = f"{name}.48x32.begin0.spp16.seed32.cap10"
        out[f"{tag}.image"] = image
        out[f"{tag}.stats"] = np.array([stats["light_events"], stats["eye_events"], stats["splats"]], dtype=np.uint64)
    np.savez_compressed(os.path.join(GOLDEN, "bdpt.npz"), **out)


def sphere_light_fixture():
    """9. emitting spheres (Sphere::SampleSurfacePoint in LightSet::GenerateRay): light rays and images of
    the path tracer, the light tracer and BDPT from the compiled reference on tests/conftest.py::sphere_light_box."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import sphere_light_box
    ref = pyoracle.Reference(sphere_light_box())
    rng = np.random.default_rng(20261107)
    out = {}
    xi = rng.random((4096, 5), dtype=np.float32)
    o, d, w = ref.light_rays(xi)
    out.update({"light_rays.xi": xi, "light_rays.origin": o, "light_rays.direction": d, "light_rays.weight": w})
    image, stats = ref.render_stream(48, 32, 0.036, 0.024, 1, 32, 7, 0)
    out["pt.48x32.begin1.spp32.seed7.image"] = image
    out["pt.48x32.begin1.spp32.seed7.rays"] = np.array([stats["rays"], stats["truncated"]], dtype=np.uint64)
    image, stats = ref.render_lt_stream(96, 64, 0.036, 0.024, 0, 64, 25, 0)
    out["lt.96x64.begin0.spp64.seed25.image"] = image
    out["lt.96x64.begin0.spp64.seed25.stats"] = np.array([stats["rays"], stats["truncated"], stats["splats"]], dtype=np.uint64)
    image, stats = ref.render_bdpt_stream(48, 32, 0.036, 0.024, 0, 16, 33, 10)
    out["bdpt.48x32.begin0.spp16.seed33.cap10.image"] = image
    out["bdpt.48x32.begin0.spp16.seed33.cap10.stats"] = np.array([stats["light_events"], stats["eye_events"], stats["splats"]], dtype=np.uint64)
    np.savez_compressed(os.path.join(GOLDEN, "sphere_lights.npz"), **out)
    print("sphere lights: lt splats", stats)


def converged_fixture():
    """8. the reference's own path tracer (MakeRGBPathTracing()->Render, its own non-reproducible
    mt19937_64 engines), many independent runs on the built-in Cornell box: the converged 64 x 64 image
    the GPU estimators are compared with block by block, and run means at TWO frame sizes (pixel values
    scale with 1 / (W * H), src/amber/scene/lens_thin.cc:139-148).  Heavy-tailed: the 2 cm light is found
    by BSDF sampling only, so a run mean is dominated by a few paths."""
    builtin = pyoracle.Reference(builtin_cornel_box=True)
    out = {}
    for (w, spp, runs) in ((64, 1024, 32), (128, 256, 16)):
        total = np.zeros((w, w, 3), dtype=np.float64)
        means, blocks = [], []
        seconds = 0.0
        for _ in range(runs):
            image, _, s = builtin.render(w, w, *FILM, spp, 8, algo=0)
            total += image
            means.append(float(image.mean()))
            blocks.append(box_downsample(luminance(image), w // 4))  # luminance of a 4 x 4 grid of big blocks
            seconds += s
        out[f"{w}x{w}.image"] = (total / runs).astype(np.float32)
        out[f"{w}x{w}.run_means"] = np.array(means, dtype=np.float64)
        out[f"{w}x{w}.run_block_luminance"] = np.array(blocks, dtype=np.float64)
        out[f"{w}x{w}.spp_per_run"] = np.array([spp], dtype=np.int64)
        print(f"converged {w}x{w}: {runs} runs x {spp} spp in {seconds:.0f} s; mean {np.mean(means):.4f} +- "
              f"{np.std(means, ddof=1) / np.sqrt(runs):.4f} (std of a run {np.std(means, ddof=1):.4f})")
    np.savez_compressed(os.path.join(GOLDEN, "cornell_converged.npz"), **out)


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    if len(sys.argv) > 2 and sys.argv[1] == "--only" and sys.argv[2] == "spheres":
        sphere_light_fixture()
        return
    if len(sys.argv) > 2 and sys.argv[1] == "--only" and sys.argv[2] == "converged":
        converged_fixture()
        return
    if len(sys.argv) > 2 and sys.argv[1] == "--only" and sys.argv[2] == "bdpt":
        bdpt_fixture()
        return
    if len(sys.argv) > 2 and sys.argv[1] == "--only" and sys.argv[2] == "pinhole":
        pinhole_fixture()
        return
    if len(sys.argv) > 2 and sys.argv[1] == "--only" and sys.argv[2] == "lt":
        light_tracing_fixture()
        return
    if len(sys.argv) > 2 and sys.argv[1] == "--only" and sys.argv[2] == "mesh":
        mesh_fixture()
        return
    if len(sys.argv) > 2 and sys.argv[1] == "--only" and sys.argv[2] == "tonemap":
        tonemap_fixture()  # (the statistics of section 5 are not reproducible: do not regenerate them casually)
        return
    flat = scene.cornel_box()
    ref = pyoracle.Reference(flat)
    builtin = pyoracle.Reference(builtin_cornel_box=True)
    oracle = pyoracle.Oracle(flat)
    rng = np.random.default_rng(20261104)

    # 1. closest hits on the fixed ray-set recipe (camera rays + secondary rays)
    n = 8192
    o, d = oracle.ray_set(1024, 1024, *FILM, 20261104, n)
    prim_bvh, t_bvh = ref.cast(o, d, accel=0)
    prim_list, t_list = ref.cast(o, d, accel=1)
    assert (prim_bvh == prim_list).all() and np.array_equal(t_bvh, t_list, equal_nan=True)
    np.savez_compressed(os.path.join(GOLDEN, "cornell_cast.npz"), origins=o, directions=d,
                        prim=prim_bvh, t=t_bvh, seed=20261104, n=n)

    # 2. materials: SampleLight with scripted uniforms, BSDF, Radiance
    out = {}
    m = 512
    for name, kind, rgb, param in MATERIALS:
        material = capi.Material(kind, (capi.C.c_float * 3)(*rgb), param)
        normal, wo, wi = unit(rng, m), unit(rng, m), unit(rng, m)
        xi = rng.random((m, 8), dtype=np.float32)
        dir_in, weight, draws = ref.sample_light(material, normal, wo, xi)
        bsdf, radiance = ref.bsdf(material, normal, wo, wi)
        for key, val in dict(normal=normal, wo=wo, wi=wi, xi=xi, dir_in=dir_in, weight=weight,
                             draws=draws, bsdf=bsdf, radiance=radiance).items():
            out[f"{name}.{key}"] = val
        out[f"{name}.desc"] = np.array([kind, *rgb, param], dtype=np.float64)
    np.savez_compressed(os.path.join(GOLDEN, "materials.npz"), **out)

    # 3. camera rays (thin lens, pixel-bound sensor), two resolutions
    out = {}
    for (w, h) in ((64, 64), (1024, 512)):
        k = 2048
        px, py = rng.integers(0, w, k).astype(np.uint32), rng.integers(0, h, k).astype(np.uint32)
        xi = rng.random((k, 5), dtype=np.float32)
        film = (0.036, 0.036 * h / w)
        origin, direction, weight = ref.eye_rays(w, h, *film, px, py, xi)
        origin_b, direction_b, weight_b = builtin.eye_rays(w, h, *film, px, py, xi)
        assert np.array_equal(origin, origin_b) and np.array_equal(direction, direction_b) and np.array_equal(weight, weight_b)
        tag = f"{w}x{h}"
        for key, val in dict(px=px, py=py, xi=xi, origin=origin, direction=direction, weight=weight).items():
            out[f"{tag}.{key}"] = val
    np.savez_compressed(os.path.join(GOLDEN, "eye_rays.npz"), **out)

    # 4. whole images: the reference's loop on the shared Philox stream
    out = {}
    for (w, spp, seed, depth) in ((32, 64, 7, 0), (48, 32, 11, 4)):
        image, stats = ref.render_stream(w, w, *FILM, 0, spp, seed, depth)
        image_b, stats_b = builtin.render_stream(w, w, *FILM, 0, spp, seed, depth)
        assert np.array_equal(image, image_b) and stats == stats_b
        tag = f"{w}x{w}x{spp}.seed{seed}.depth{depth}"
        out[f"{tag}.image"] = image
        out[f"{tag}.rays"] = np.array([stats["rays"], stats["truncated"]], dtype=np.uint64)
    np.savez_compressed(os.path.join(GOLDEN, "cornell_stream.npz"), **out)

    # 5. statistics of the reference's own renderer (non-reproducible engines): mean
    #    luminance over K independent PT runs, and a BDPT image for the NEE comparison
    k_runs, w, spp = 16, 64, 256
    means = []
    for _ in range(k_runs):
        image, _, _ = builtin.render(w, w, *FILM, spp, 8, algo=0)
        means.append(float(image.mean()))
    bdpt, _, seconds = builtin.render(w, w, *FILM, 1024, 8, algo=1)
    np.savez_compressed(os.path.join(GOLDEN, "cornell_bdpt_64x64.npz"), image=bdpt, spp=1024)
    with open(os.path.join(GOLDEN, "cornell_luminance.json"), "w") as f:
        json.dump({"width": w, "height": w, "spp": spp, "runs": k_runs, "pt_means": means,
                   "pt_mean": float(np.mean(means)), "pt_std_of_run": float(np.std(means, ddof=1)),
                   "bdpt_mean": float(bdpt.mean()), "bdpt_spp": 1024, "bdpt_seconds": seconds}, f, indent=1)
    tonemap_fixture()
    mesh_fixture()
    print("pt mean", np.mean(means), "+-", np.std(means, ddof=1) / np.sqrt(k_runs), "bdpt mean", bdpt.mean(),
          f"({seconds:.1f}s)")


if __name__ == "__main__":
    main()
