#!/usr/bin/env python
"""Times the `extend` kernel alone (amber_cuda_cast_device) on N device-resident rays:
camera rays + their secondary rays, the recipe of the fixed intersection-parity set."""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from amber_b200 import cornel_box, cornel_mesh  # noqa: E402
from amber_b200.render import Device, Sensor  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=4_000_000)
ap.add_argument("--lib", default=None)
ap.add_argument("--mesh", type=int, default=-1)
ap.add_argument("--repeat", type=int, default=10)
args = ap.parse_args()

scene = cornel_box() if args.mesh < 0 else cornel_mesh(args.mesh)
rng = np.random.default_rng(1)
with Device(0, lib_path=args.lib) as dev:
    dev.upload_scene(scene)
    dev.set_sensor(Sensor(1024, 1024))
    n = args.n // 2
    px, py = rng.integers(0, 1024, n), rng.integers(0, 1024, n)
    o, d, _ = dev.eye_rays(px, py, rng.random((n, 5), dtype=np.float32))
    prim, t = dev.cast(o, d)
    hit = prim >= 0
    # secondary rays: from the hit points into a random hemisphere-ish direction
    o2 = (o + t[:, None] * d)[hit]
    d2 = rng.normal(size=o2.shape).astype(np.float32)
    d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
    o_all = np.concatenate([o, o2]).astype(np.float32)
    d_all = np.concatenate([d, d2]).astype(np.float32)
    perm = rng.permutation(len(o_all))
    o_all, d_all = o_all[perm], d_all[perm]
    m = len(o_all)
    o4 = torch.zeros((m, 4), dtype=torch.float32, device="cuda")
    d4 = torch.zeros((m, 4), dtype=torch.float32, device="cuda")
    o4[:, :3] = torch.from_numpy(o_all).cuda()
    d4[:, :3] = torch.from_numpy(d_all).cuda()
    h4 = torch.empty((m, 4), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    times = [dev.cast_device(o4.data_ptr(), d4.data_ptr(), m, h4.data_ptr()) for _ in range(args.repeat)]
    best = min(times[2:])
    print(f"{m} rays: best {best * 1e3:.1f} us, {m / best / 1e3:.1f} Mrays/s, hit fraction {(h4[:, 1].view(torch.int32) >= 0).float().mean().item():.3f}")
