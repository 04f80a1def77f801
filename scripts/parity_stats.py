#!/usr/bin/env python
"""Image parity of the benchmarked estimator (NEE) against the reference estimator, on the GPU.

    python scripts/parity_stats.py [--quick] > profiles/rNN_parity_stats.json

ESTIMATOR_REFERENCE is sample for sample the reference's `algorithm_pt.cc:125-160` (pinned image for
image to the compiled reference on a shared random stream, tests/test_gpu_parity.py), so a converged
render with it IS the reference's converged image.  It finds the 2 cm light by BSDF sampling only
(relative standard deviation of one path sample ~ 100-250), so "converged" means 1e10-1e11 paths;
the GPU does that in seconds.  Printed, per frame size: mean luminance of both estimators with the
Monte-Carlo standard error from independent seeds, and the 4x4-block rRMSE of
NEE at 1024 spp against the converged reference image above / below the water line, next to the
reference image's own residual noise (from its two independent halves).
The tolerances in tests/test_image_parity.py are set from this output.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from amber_b200 import capi, cornel_box  # noqa: E402
from amber_b200.metrics import rrmse, water_line_rows  # noqa: E402
from amber_b200.render import Device, Sensor  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def render(dev, estimator, spp, seed, window=(0, 0), begin=0):
    """spp samples per pixel in calls of at most 2^20 samples (sample indices stay below 2^32)."""
    dev.film_clear()
    done = 0
    while done < spp:
        n = min(spp - done, 1 << 20)
        dev.render(begin + done, n, seed=seed, max_depth=0, estimator=estimator, window=window, wait=False)
        done += n
    dev.wait()
    image, samples = dev.resolve()
    assert samples == spp
    return image.astype(np.float64)


def converged_reference(dev, spp, seed, window=(0, 0)):
    """Two independent halves (different seeds) and their mean."""
    a = render(dev, capi.ESTIMATOR_REFERENCE, spp // 2, seed, window)
    b = render(dev, capi.ESTIMATOR_REFERENCE, spp // 2, seed + 1, window)
    return 0.5 * (a + b), a, b


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true", help="a tenth of the samples")
    args = ap.parse_args()
    scale = 0.1 if args.quick else 1.0
    out = {"max_depth": "unbounded (as the reference)", "nee_spp": 1024}
    with Device(0) as dev:
        dev.upload_scene(cornel_box())

        # ---- mean luminance, 64 x 64 and 128 x 128 (the golden's two sizes) --------------------
        golden = json.load(open(os.path.join(GOLDEN, "cornell_luminance.json")))
        conv = np.load(os.path.join(GOLDEN, "cornell_converged.npz"))
        for size in (64, 128):
            dev.set_sensor(Sensor(size, size))
            t0 = time.perf_counter()
            spp_ref = int(scale * (1 << 18) * (64 * 64) / (size * size))
            ref = np.array([render(dev, capi.ESTIMATOR_REFERENCE, spp_ref, 100 + s).mean() for s in range(8)])
            nee = np.array([render(dev, capi.ESTIMATOR_NEE, max(64, spp_ref // 16), 200 + s).mean() for s in range(8)])
            runs = conv[f"{size}x{size}.run_means"]
            out[f"mean_{size}"] = {
                "gpu_reference_estimator": {"mean": ref.mean(), "sem": ref.std(ddof=1) / np.sqrt(len(ref)), "spp_per_seed": spp_ref, "seeds": len(ref)},
                "gpu_nee": {"mean": nee.mean(), "sem": nee.std(ddof=1) / np.sqrt(len(nee)), "spp_per_seed": max(64, spp_ref // 16), "seeds": len(nee)},
                "cpu_reference_golden": {"mean": runs.mean(), "sem": runs.std(ddof=1) / np.sqrt(len(runs)), "runs": len(runs),
                                         "spp_per_run": int(conv[f"{size}x{size}.spp_per_run"][0])},
                "seconds": time.perf_counter() - t0,
            }
        out["mean_64"]["cpu_reference_golden_r01"] = {"mean": golden["pt_mean"], "sem": golden["pt_std_of_run"] / np.sqrt(golden["runs"])}

        # ---- 4x4-block rRMSE against the converged reference estimator ------------------------
        for tag, size, spp_ref, window_rows in (("64", 64, 1 << 22, None), ("256", 256, 1 << 18, None),
                                               ("1024_rows_320_352", 1024, 1 << 19, (320, 32)),
                                               ("1024_rows_800_832", 1024, 1 << 19, (800, 32))):
            spp_ref = max(1024, int(scale * spp_ref))
            dev.set_sensor(Sensor(size, size))
            window = (0, 0) if window_rows is None else (window_rows[0] * size, window_rows[1] * size)
            t0 = time.perf_counter()
            ref, half_a, half_b = converged_reference(dev, spp_ref, 300, window)
            ref_seconds = time.perf_counter() - t0
            nee = render(dev, capi.ESTIMATOR_NEE, 1024, 1, window)
            nee_b = render(dev, capi.ESTIMATOR_NEE, 1024, 2, window)
            nee_hi = render(dev, capi.ESTIMATOR_NEE, max(1024, spp_ref // 16), 3, window)
            if window_rows is None:
                above, below = water_line_rows(size)
                crop = lambda im: im  # noqa: E731
            else:
                above, below = water_line_rows(size, window_rows[0], window_rows[1])
                crop = lambda im: im[window_rows[0]: window_rows[0] + window_rows[1]]  # noqa: E731
            entry = {"size": size, "reference_spp": spp_ref, "reference_seconds": ref_seconds,
                     "paths_reference": float(spp_ref) * (window[1] or size * size)}
            for name, rows in (("above_water", above), ("below_water", below)):
                if rows.stop - rows.start <= 0:
                    continue
                entry[name] = {
                    "block_rows": [rows.start, rows.stop],
                    "rrmse_nee_1024spp_vs_reference": rrmse(crop(nee), crop(ref), rows),
                    "rrmse_nee_hi_spp_vs_reference": rrmse(crop(nee_hi), crop(ref), rows),
                    "nee_hi_spp": max(1024, spp_ref // 16),
                    # noise of the mean of two halves = half the rRMSE between them
                    "reference_own_noise": 0.5 * rrmse(crop(half_a), crop(half_b), rows),
                    "nee_1024spp_own_noise": rrmse(crop(nee), crop(nee_b), rows) / np.sqrt(2.0),
                }
            out[f"rrmse_{tag}"] = entry
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
