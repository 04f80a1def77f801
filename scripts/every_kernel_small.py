#!/usr/bin/env python
"""A small render of every estimator (path tracing reference / NEE, light tracing, BDPT) on both
acceleration layouts (constant-memory list, 4-wide BVH), plus the tonemapped resolve: the smallest
program that launches every kernel of the library."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from amber_b200 import Device, Sensor, capi, cornel_box, cornel_mesh  # noqa: E402

for scene, size in ((cornel_box(), 24), (cornel_mesh(2), 16)):
    with Device(0) as dev:
        dev.upload_scene(scene)
        dev.set_sensor(Sensor(size, size))
        for estimator, depth in ((capi.ESTIMATOR_REFERENCE, 8), (capi.ESTIMATOR_NEE, 8), (capi.ESTIMATOR_LIGHT_TRACING, 8),
                                 (capi.ESTIMATOR_BDPT, 0)):
            dev.film_clear()
            dev.render(0, 2, seed=3, max_depth=depth, estimator=estimator, pool_paths=2048)
            image, _ = dev.resolve()
            print("accel", dev.stats()["accel"], "estimator", estimator, "mean", float(image.mean()))
        dev.resolve_ldr(flip_x=True)
print("SANITIZE_RUN_OK")
