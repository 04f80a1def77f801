#!/usr/bin/env python
"""A small render of every estimator (path tracing reference / NEE, light tracing, BDPT in both implementations) on both
acceleration layouts (constant-memory list, 4-wide BVH from the host builder and from the device builder), the casts
and the tonemapped resolve: the smallest program that launches every kernel of the library.  Meant to run under
compute-sanitizer (memcheck, racecheck, initcheck)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from amber_b200 import Device, Sensor, capi, cornel_box, cornel_mesh  # noqa: E402

for scene, size, builder in ((cornel_box(), 24, None), (cornel_mesh(2), 16, "host"), (cornel_mesh(2), 16, "lbvh")):
    if builder:
        os.environ["AMBER_CUDA_BVH_BUILDER"] = builder
    with Device(0) as dev:
        dev.upload_scene(scene)
        dev.set_sensor(Sensor(size, size))
        for estimator, depth, bdpt in ((capi.ESTIMATOR_REFERENCE, 8, None), (capi.ESTIMATOR_NEE, 8, None),
                                       (capi.ESTIMATOR_LIGHT_TRACING, 8, None), (capi.ESTIMATOR_BDPT, 0, "wavefront"),
                                       (capi.ESTIMATOR_BDPT, 0, "megakernel")):
            if bdpt:
                os.environ["AMBER_CUDA_BDPT"] = bdpt
            dev.film_clear()
            dev.render(0, 2, seed=3, max_depth=depth, estimator=estimator, pool_paths=2048)
            image, _ = dev.resolve()
            print("accel", dev.stats()["accel"], "builder", builder, "estimator", estimator, bdpt or "", "mean", float(image.mean()))
        rng = np.random.default_rng(5)
        o = rng.uniform(-0.9, 0.9, size=(500, 3)).astype(np.float32)
        d = rng.normal(size=(500, 3)).astype(np.float32)
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        prim, t = dev.cast(o, d)
        print("cast hits", int((prim >= 0).sum()))
        dev.resolve_ldr(flip_x=True)
print("SANITIZE_RUN_OK")
