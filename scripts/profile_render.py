#!/usr/bin/env python
"""Short Cornell render through the C ABI, for `ncu --set full` captures (keeps the
bench's scene, frame size, estimator and pool; only the sample count is small)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from amber_b200 import capi, cornel_box, cornel_mesh  # noqa: E402
from amber_b200.render import Device, Sensor  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--spp", type=int, default=16)
ap.add_argument("--size", type=int, default=1024)
ap.add_argument("--estimator", default="nee", choices=["nee", "reference", "lt", "bdpt"])
ap.add_argument("--mesh", type=int, default=-1, help="icosphere subdivisions (config 4); -1 = built-in Cornell box")
ap.add_argument("--pool", type=int, default=0)
ap.add_argument("--lib", default=None, help="alternative libamber_cuda.so (build variants)")
ap.add_argument("--repeat", type=int, default=1)
ap.add_argument("--profile", action="store_true", help="per-kernel CUDA-event times")
args = ap.parse_args()

scene = cornel_box() if args.mesh < 0 else cornel_mesh(args.mesh)
with Device(0, lib_path=args.lib) as dev:
    dev.upload_scene(scene)
    dev.set_sensor(Sensor(args.size, args.size))
    est = {"nee": capi.ESTIMATOR_NEE, "reference": capi.ESTIMATOR_REFERENCE, "lt": capi.ESTIMATOR_LIGHT_TRACING, "bdpt": capi.ESTIMATOR_BDPT}[args.estimator]
    dev.set_profiling(args.profile)
    for _ in range(args.repeat):
        dev.film_clear()
        dev.render(0, args.spp, seed=1, max_depth=0 if args.estimator == 'bdpt' else 32, estimator=est, pool_paths=args.pool)
    s = dev.stats()
    rays = s["closest_rays"] + s["shadow_rays"]
    print(f"{rays / s['last_render_ms'] / 1e3:.1f} Mrays/s, {s['paths'] / s['last_render_ms'] / 1e3:.1f} Mpaths/s, "
          f"{s['last_render_ms']:.2f} ms, launches {s['kernel_launches']}, accel {s['accel']}")
    if args.profile:
        n = max(1, s["extend_launches"])
        print("per launch (us): " + ", ".join(f"{k} {s[k + '_ms'] / n * 1e3:.1f}" for k in ("raygen", "extend", "shade", "connect"))
              + f"; wavefront iterations {n}, closest rays/iteration {s['closest_rays'] / n:.0f}, shadow rays/iteration {s['shadow_rays'] / n:.0f}")
        if s["accel"]:
            print(f"BVH work per closest ray: {s['bvh_node_visits'] / max(1, s['closest_rays']):.2f} node steps, "
                  f"{s['bvh_triangle_tests'] / max(1, s['closest_rays']):.2f} triangle tests; per shadow ray: "
                  f"{s['bvh_node_visits_connect'] / max(1, s['shadow_rays']):.2f} / {s['bvh_triangle_tests_connect'] / max(1, s['shadow_rays']):.2f}; "
                  f"extend {s['closest_rays'] / max(s['extend_ms'], 1e-9) / 1e6:.2f} Grays/s, connect {s['shadow_rays'] / max(s['connect_ms'], 1e-9) / 1e6:.2f} Grays/s")
