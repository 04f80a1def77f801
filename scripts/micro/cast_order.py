#!/usr/bin/env python
"""How much does ray ORDER matter to the BVH traversal kernel?  Secondary rays (diffuse bounce off
the first hit) of the config-4 mesh, cast in pixel order, in random order, and sorted by
(direction octant, Morton code of the origin).  Decides whether a ray-sorting pass would pay."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from amber_b200 import cornel_mesh  # noqa: E402
from amber_b200.render import Device, Sensor  # noqa: E402


def morton(p, bits=10):
    q = np.clip(((p - p.min(0)) / (np.ptp(p, axis=0) + 1e-9) * (2 ** bits - 1)).astype(np.uint64), 0, 2 ** bits - 1)
    code = np.zeros(len(p), dtype=np.uint64)
    for b in range(bits):
        for a in range(3):
            code |= ((q[:, a] >> np.uint64(b)) & np.uint64(1)) << np.uint64(3 * b + a)
    return code


rng = np.random.default_rng(1)
with Device(0) as dev:
    dev.upload_scene(cornel_mesh(8))
    dev.set_sensor(Sensor(2048, 2048))
    n = 2_400_000
    idx = np.arange(n)
    px, py = idx % 2048, (idx // 2048) + 700          # rows through the middle of the frame, pixel order
    o, d, _ = dev.eye_rays(px, py, rng.random((n, 5), dtype=np.float32))
    prim, t = dev.cast(o, d)
    hit = prim >= 0
    o2 = (o + t[:, None] * d)[hit].astype(np.float32)
    d2 = rng.normal(size=o2.shape).astype(np.float32)
    d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
    o2 = o2 + 1e-3 * d2
    octant = (d2[:, 0] < 0).astype(np.uint64) | ((d2[:, 1] < 0).astype(np.uint64) << np.uint64(1)) | ((d2[:, 2] < 0).astype(np.uint64) << np.uint64(2))
    # shadow-like rays: from the same surface points towards the (tiny) light -- one direction field, scattered origins
    d_light = np.array([0.0, 0.99, 0.0], dtype=np.float32) - o2
    d_light /= np.linalg.norm(d_light, axis=1, keepdims=True)
    shadow_orders = {"shadow-like, random": rng.permutation(len(o2)),
                     "shadow-like, sorted by origin (morton 30 bit)": np.argsort(morton(o2), kind="stable"),
                     "shadow-like, sorted by origin (morton 15 bit)": np.argsort(morton(o2, 5), kind="stable")}
    orders = {
        "pixel order": np.arange(len(o2)),
        "random": rng.permutation(len(o2)),
        "sorted by origin (morton 30 bit)": np.argsort(morton(o2), kind="stable"),
        "sorted by octant, origin": np.argsort((octant << np.uint64(30)) | morton(o2), kind="stable"),
        "sorted by origin (morton 12 bit)": np.argsort(morton(o2, 4), kind="stable"),
    }
    for name, perm in list(orders.items()) + list(shadow_orders.items()):
        m = len(perm)
        o4 = torch.zeros((m, 4), dtype=torch.float32, device="cuda")
        d4 = torch.zeros((m, 4), dtype=torch.float32, device="cuda")
        o4[:, :3] = torch.from_numpy(o2[perm]).cuda()
        d4[:, :3] = torch.from_numpy((d_light if name.startswith("shadow") else d2)[perm]).cuda()
        h4 = torch.empty((m, 4), dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        times = [dev.cast_device(o4.data_ptr(), d4.data_ptr(), m, h4.data_ptr()) for _ in range(8)]
        best = min(times[2:])
        print(f"{name:36s} {m} rays: {best * 1e3:7.1f} us, {m / best / 1e3:7.1f} Mrays/s")
