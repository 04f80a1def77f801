import sys
import numpy as np
sys.path.insert(0, ".")
from amber_b200 import Sensor, cornel_box
from amber_b200.render import Device
from oracle import pyoracle
scene = cornel_box()
oracle = pyoracle.Oracle(scene)
dev = Device(0); dev.upload_scene(scene); dev.set_sensor(Sensor(64, 64))
o, d = oracle.ray_set(64, 64, 0.036, 0.036, 1, 200000)
o, d = o[:200000], d[:200000]
prim, t = oracle.cast(o, d)
sel = prim == 22
o, d, t = o[sel], d[sel], t[sel]
x = (o + (t[:, None] * d).astype(np.float32)).astype(np.float32)
light = np.array([0.0, 0.99, 0.0], np.float32)
w = light - x
w = (w / np.linalg.norm(w, axis=1, keepdims=True)).astype(np.float32)
p_o, t_o = oracle.cast(x, w)
p_g, t_g = dev.cast(x, w)
print("rays", len(x), "oracle prims", np.unique(p_o, return_counts=True))
print("gpu prims", np.unique(p_g, return_counts=True))
bad = p_o != p_g
print("mismatch", bad.sum())
for i in np.nonzero(bad)[0][:10]:
    print(x[i], w[i], "oracle", p_o[i], t_o[i], "gpu", p_g[i], t_g[i])
