import sys
import numpy as np
sys.path.insert(0, ".")
from amber_b200 import Sensor, cornel_box
from amber_b200.render import Device
from oracle import pyoracle
scene = cornel_box()
oracle = pyoracle.Oracle(scene)
W = 64
dev = Device(0); dev.upload_scene(scene); dev.set_sensor(Sensor(W, W))
for sample in range(2):
    dev.film_clear()
    dev.render(sample, 1, seed=5, max_depth=1, estimator=1)
    got, _ = dev.resolve()
    want, _ = oracle.render(W, W, 0.036, 0.036, sample, 1, 5, 1, 1, threads=16)
    n = 0
    for py in range(W):
        for px in range(W):
            path = oracle.trace_path(W, W, 0.036, 0.036, px, py, sample, 5, 1, 1)
            if int(path[0][0]) in (22, 23, 24) :
                g, w = got[py, px, 0], want[py, px, 0]
                if abs(g - w) > 2e-3 * abs(w) and n < 40:
                    n += 1
                    print("prim %d pix (%d,%d) got %.6g want %.6g ratio %.4f pos (%.4f %.4f %.4f) t %.6f" % (path[0][0], px, py, g, w, g / w if w else -1, path[0][2], path[0][3], path[0][4], path[0][1]))
    bad = ~(np.abs(got - want) <= 2e-3 * np.abs(want) + 1e-6 * want.mean()).all(-1)
    print("sample", sample, "bad pixels", bad.sum())
