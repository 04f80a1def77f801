#!/usr/bin/env python
"""Timing of the render queue: the host algorithm (batches of Context.iterate() queued ahead of the device)
against one blocking call of the same samples, for several frame sizes / batch sizes / depth caps."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from amber_b200 import Context, CudaPathTracing, Sensor, cornel_box  # noqa: E402
from amber_b200.render import Device  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--cases", default="128:64:64:0,512:256:16:32,512:256:8:16,1024:256:16:32,1024:1024:64:32")
args = ap.parse_args()
scene = cornel_box()
for case in args.cases.split(","):
    size, spp, batch, depth = (int(x) for x in case.split(":"))
    algo = CudaPathTracing(max_depth=depth, seed=5, batch_spp=batch)
    algo.render(scene, Sensor(size, size), Context(1, batch))          # contexts, upload, pool
    t0 = time.perf_counter()
    algo.render(scene, Sensor(size, size), Context(1, spp))
    queued = time.perf_counter() - t0
    stats = algo.last_stats
    dev = algo._devices[0]
    dev.film_clear()
    t0 = time.perf_counter()
    dev.render(0, spp, seed=5, max_depth=depth)
    blocking = time.perf_counter() - t0
    print(json.dumps({"size": size, "spp": spp, "batch_spp": batch, "max_depth": depth, "queued_s": queued,
                      "one_call_s": blocking, "chained_renders": stats["chained_renders"], "paths": stats["paths"],
                      "Mpaths_per_s_queued": size * size * spp / queued / 1e6}), flush=True)
    algo.close()
