"""GPU debugging aid: find (pixel, sample) pairs whose radiance differs between the device
and the CPU oracle, and print the oracle's vertex list for them."""
import collections
import sys

import numpy as np

sys.path.insert(0, ".")
from amber_b200 import Sensor, cornel_box, cornel_mesh  # noqa: E402
from amber_b200.render import Device  # noqa: E402
from oracle import pyoracle  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "cornell"
estimator = int(sys.argv[2]) if len(sys.argv) > 2 else 1
scene = cornel_box() if which == "cornell" else cornel_mesh(4)
W, DEPTH, SEED = 64, 16, 5
oracle = pyoracle.Oracle(scene)
dev = Device(0)
dev.upload_scene(scene)
dev.set_sensor(Sensor(W, W))
first = collections.Counter()
kinds = collections.Counter()
shown = 0
total_bad = 0
for sample in range(16):
    dev.film_clear()
    dev.render(sample, 1, seed=SEED, max_depth=DEPTH, estimator=estimator)
    got, _ = dev.resolve()
    want, _ = oracle.render(W, W, 0.036, 0.036, sample, 1, SEED, DEPTH, estimator, threads=16)
    floor = 1e-6 * np.abs(want).mean()
    bad = ~(np.abs(got - want) <= 2e-3 * np.abs(want) + floor).all(-1)
    total_bad += bad.sum()
    for py, px in zip(*np.nonzero(bad)):
        path = oracle.trace_path(W, W, 0.036, 0.036, int(px), int(py), sample, SEED, DEPTH, estimator)
        prims = [int(v[0]) for v in path]
        first[prims[0]] += 1
        for p in prims:
            kinds[p] += 1
        if shown < 25:
            shown += 1
            print(f"sample {sample} pixel ({px},{py}) got {got[py, px]} want {want[py, px]}")
            for v in path:
                print("    prim %3d t %.7g pos (%.6f %.6f %.6f) draws %d thr %.4g p_rr %.4f" % (v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]))
print("divergent (pixel,sample) pairs:", total_bad, "of", 16 * W * W)
print("first-hit prim of divergent paths:", sorted(first.items())[:40])
print("prims anywhere on divergent paths:", sorted(kinds.items(), key=lambda kv: -kv[1])[:30])
