#!/usr/bin/env python
"""bench.py -- the headline benchmark: Cornell box 1024x1024, 1024 spp, path tracing.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One step = one pass of the hot path over the whole workload (BASELINE.json configs[1]):
every pixel of the 1024x1024 frame x 1024 samples, i.e. 1.07e9 camera paths through
raygen / extend / shade / connect on one B200.  With N ranks (torchrun) every rank renders
1024 spp of its own slice of the global sample range [0, N*1024) -- weak scaling, no
data-path collective -- and one NCCL all-reduce sums the film at the end of each step.

Rank 0 prints ONE JSON line (see the contract in the task description).  `value` is device
time with the scene resident; `e2e` goes through the C ABI with host buffers every step
(scene upload H2D, render, resolved image D2H).  `--impl reference` times the UNMODIFIED
reference's CPU path tracer (oracle/_ref, all host threads) on a bounded sample of the same
workload instead.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WIDTH = HEIGHT = 1024
SPP = 1024
MAX_DEPTH = 32
SEED = 1
METRIC = "Mrays/s (closest-hit + shadow rays), Cornell box 1024x1024 unidirectional path tracing"
FILM = (0.036, 0.036)
# algorithmic bytes per ray over the whole wavefront ...
BYTES_PER_CLOSEST_RAY = 160
BYTES_PER_SHADOW_RAY = 108
# ... and per kernel, from the record layout in HBM ("Data layout"): bytes one unit of
# that kernel's work must move, (per closest ray, per shadow ray, per camera path).
KERNEL_BYTES = {
    # read q_new 4, write ray 32 + state 32
    "raygen": (0, 0, 68),
    # read ray 32, write hit 8
    "extend": (40, 0, 0),
    # read hit 8 (classify) + hit 8, ray 32, state 32; write ray 32, state 32 | shadow record 32 +
    # q_shadow 4 per shadow ray | q_new 4 per finished path
    "shade": (144, 36, 4),
    # read q_shadow 4, ray origin 16, shadow record 32; film atomic 16
    "connect": (0, 68, 0),
}
KERNEL_NAMES = {"raygen": "raygen_kernel", "extend": "extend_kernel<ConstList>", "shade": "shade_kernel",
                "connect": "connect_kernel<ConstList>"}


def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def measured_peaks() -> tuple[float, str]:
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index: int):
        self.device_index = device_index
        self.lines: list[str] = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.device_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, sm_max, reasons = [], [], set()
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                sm_max.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        loaded = [c for c in sm if c > 0.5 * max(sm_max or [1])] or sm
        return {"sm_mhz": float(np.median(loaded)) if loaded else None,
                "sm_max_mhz": max(sm_max) if sm_max else None, "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------
def reference_step(ref, spp: int, threads: int):
    """The reference's own MakeRGBPathTracing()->Render on `threads` std::threads."""
    _, casts, seconds = ref.render(WIDTH, HEIGHT, *FILM, spp, threads, algo=0, want_image=False)
    return casts, seconds


def cpu_baseline(budget_s: float = 12.0) -> dict:
    """Reference CPU PT on a bounded sample of the workload (same frame, fewer spp)."""
    from amber_b200 import cornel_box
    from oracle import pyoracle
    threads = host_threads()
    if pyoracle.reference_available():
        ref = pyoracle.Reference(cornel_box())
        casts, seconds = reference_step(ref, 4, threads)          # calibrate on four frames
        spp = int(max(4, min(256, 4 * budget_s / max(seconds, 1e-3))))
        casts, seconds = reference_step(ref, spp, threads)
        kind = "reference"
    else:  # the C port of the same algorithm
        pyoracle.build(ref=False)
        oracle = pyoracle.Oracle(cornel_box())
        spp = 2
        t0 = time.perf_counter()
        _, stats = oracle.render(WIDTH, HEIGHT, *FILM, 0, spp, SEED, MAX_DEPTH, 0, threads=min(threads, 64))
        seconds = time.perf_counter() - t0
        casts = stats["closest_rays"]
        kind = "port"
    paths = WIDTH * HEIGHT * spp
    return {"value": casts / seconds / 1e6, "unit": "Mrays/s", "cores": threads, "kind": kind,
            "sample": f"Cornell {WIDTH}x{HEIGHT}, {spp} spp ({paths} paths, {casts} rays) in {seconds:.2f} s; "
                      f"reference estimator (no shadow rays), unbounded depth",
            "samples_per_s": paths / seconds}


def run_reference(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from amber_b200 import cornel_box
    from oracle import pyoracle
    threads = host_threads()
    if not pyoracle.reference_available():
        base = cpu_baseline()
        kind, ref = "port", None
    else:
        kind, ref = "reference", pyoracle.Reference(cornel_box())
    if ref is not None:
        _, calib = reference_step(ref, 4, threads)
        budget = 150.0 / max(1, args.steps + args.warmup)          # whole run within a few minutes
        spp = int(max(4, min(256, 4 * budget / max(calib, 1e-3))))
        for _ in range(args.warmup):
            reference_step(ref, spp, threads)
        rays = seconds = 0.0
        for _ in range(args.steps):
            c, s = reference_step(ref, spp, threads)
            rays += c
            seconds += s
        value = rays / seconds / 1e6
        paths = WIDTH * HEIGHT * spp * args.steps
        sample = f"each step: Cornell {WIDTH}x{HEIGHT}, {spp} spp of the 1024-spp workload"
        samples_per_s = paths / seconds
        ms = seconds / args.steps * 1e3
    else:
        value, sample, samples_per_s, ms = base["value"], base["sample"], base["samples_per_s"], None
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"Cornell box {WIDTH}x{HEIGHT} path tracing, 1024 spp (BASELINE configs[1])",
                   "estimator": "reference (algorithm_pt.cc, no light sampling)", "threads": threads},
        "samples_per_s": samples_per_s,
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------
def run_gpu(args) -> None:
    import ctypes

    import torch

    from amber_b200 import capi, cornel_box
    from amber_b200.dist import partition_samples, reduce_device_film
    from amber_b200.render import Device, Sensor

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    distributed = world > 1
    if distributed:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    estimator = capi.ESTIMATOR_REFERENCE if args.estimator == "reference" else capi.ESTIMATOR_NEE
    scene = cornel_box()
    sensor = Sensor(WIDTH, HEIGHT)
    dev = Device(local_rank)
    dev.upload_scene(scene)
    dev.set_sensor(sensor)
    begin, count = partition_samples(SPP * world, rank, world)  # weak scaling: 1024 spp per rank
    image = np.empty((HEIGHT, WIDTH, 3), dtype=np.float32)

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    def step(e2e: bool) -> float:
        """Returns device milliseconds of this rank's step."""
        if e2e:
            dev.upload_scene(scene)      # host scene description -> device (H2D)
            dev.set_sensor(sensor)       # (clears the film)
        else:
            dev.film_clear()
        dev.render(begin, count, seed=SEED, max_depth=MAX_DEPTH, estimator=estimator, pool_paths=args.pool)
        ms = dev.stats()["last_render_ms"]
        if distributed:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            reduce_device_film(dev)
            e1.record()
            torch.cuda.synchronize()
            ms += e0.elapsed_time(e1)
        if e2e and rank == 0:
            dev.resolve(image)           # mean image -> host (D2H)
        return ms

    for _ in range(args.warmup):
        step(False)

    def timed(e2e: bool):
        barrier()
        launches0 = dev.stats()["kernel_launches"]
        clocks = ClockSampler(local_rank)
        if rank == 0:
            clocks.start()
        wall0 = time.perf_counter()
        device_ms = 0.0
        rays = paths = 0
        for _ in range(args.steps):
            device_ms += step(e2e)
            s = dev.stats()
            rays += s["closest_rays"] + s["shadow_rays"]
            paths += s["paths"]
        barrier()
        wall_ms = (time.perf_counter() - wall0) * 1e3
        launches = dev.stats()["kernel_launches"] - launches0
        clock_info = clocks.stop() if rank == 0 else None
        totals = torch.tensor([device_ms, wall_ms, float(rays), float(paths), float(launches)],
                              dtype=torch.float64, device=f"cuda:{local_rank}")
        if distributed:
            mx = totals.clone()
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            sm = totals.clone()
            dist.all_reduce(sm, op=dist.ReduceOp.SUM)
            device_ms, wall_ms = float(mx[0]), float(mx[1])
            rays, paths, launches = float(sm[2]), float(sm[3]), int(sm[4])
        return device_ms, wall_ms, rays, paths, launches, clock_info

    device_ms, wall_ms, rays, paths, launches, clocks = timed(False)
    e2e_device_ms, e2e_wall_ms, e2e_rays, e2e_paths, _, _ = timed(True)

    # per-kernel timing of the dominant kernel: one extra step with an event pair per launch
    roofline = None
    if rank == 0:
        dev.set_profiling(True)
        dev.film_clear()
        dev.render(begin, count, seed=SEED, max_depth=MAX_DEPTH, estimator=estimator, pool_paths=args.pool)
        s = dev.stats()
        dev.set_profiling(False)
        peak, peak_kind = measured_peaks()
        kernel_ms = {k: s[f"{k}_ms"] for k in ("raygen", "extend", "shade", "connect")}
        dominant = max(kernel_ms, key=kernel_ms.get)
        launches_each = max(1, s["extend_launches"])          # every kernel runs once per wavefront iteration
        per_closest, per_shadow, per_path = KERNEL_BYTES[dominant]
        bytes_per_launch = (per_closest * s["closest_rays"] + per_shadow * s["shadow_rays"]
                            + per_path * s["paths"]) / launches_each
        avg_ms = kernel_ms[dominant] / launches_each
        achieved = bytes_per_launch / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
        whole = (BYTES_PER_CLOSEST_RAY * s["closest_rays"] + BYTES_PER_SHADOW_RAY * s["shadow_rays"]) / \
                (s["last_render_ms"] * 1e-3) / 1e9
        roofline = {
            "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": None, "kernel": KERNEL_NAMES[dominant], "peak_source": peak_kind,
            "algorithmic_bytes_per_launch": bytes_per_launch,
            "bytes_model": {"per_closest_ray": per_closest, "per_shadow_ray": per_shadow, "per_camera_path": per_path},
            "closest_rays_per_launch": s["closest_rays"] / launches_each,
            "shadow_rays_per_launch": s["shadow_rays"] / launches_each,
            "avg_launch_ms": avg_ms, "launches_timed": launches_each,
            "timing": "CUDA events on the context's stream around every launch of one extra full step",
            "kernel_share_of_step": {k: v / max(1e-9, sum(kernel_ms.values())) for k, v in kernel_ms.items()},
            "whole_wavefront": {"achieved": whole, "frac": whole / peak,
                                "bytes_model": "160 B/closest ray + 108 B/shadow ray (SURVEY 8d)"},
            "note": "shade is limited by DRAM bytes and instruction issue together (ncu, profiles/r01_r01h_summary.txt: "
                    "593 MB of DRAM traffic per launch, 60 % of the issue slots; adding 93 MB of writes cost 29 us, "
                    "DESIGN.md section 6); extend is issue bound (83 %), connect DRAM bound (65 % busy)",
        }
        traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(traffic_file):
            roofline["traffic"] = json.load(open(traffic_file)).get(dominant, {}).get("dram_bytes_per_launch")

    if rank != 0:
        if distributed:
            dist.destroy_process_group()
        return

    base = cpu_baseline() if world == 1 and not args.no_cpu_baseline else None
    value = rays / (device_ms * 1e-3) / 1e6
    line = {
        "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": device_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {
            "workload": f"Cornell box {WIDTH}x{HEIGHT} path tracing, {SPP} spp per GPU (BASELINE configs[1])",
            "estimator": args.estimator, "max_depth": MAX_DEPTH, "seed": SEED, "pool_paths": args.pool or "default",
            "parallelism": f"sample-partition x{world}, one film all-reduce per step" if distributed else "single GPU",
            "l2": "working set (path pool + queues, > 300 MB) exceeds the 126 MB L2; film cleared between steps",
        },
        "samples_per_s": paths / (device_ms * 1e-3),
        "rays_per_path": rays / max(1.0, paths),
        "ms_per_step_wall": wall_ms / args.steps,
        "e2e": {"value": e2e_rays / (e2e_wall_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                "h2d_bytes_per_step": int(ctypes.sizeof(scene.primitives) + ctypes.sizeof(scene.materials)
                                          + ctypes.sizeof(scene.lens)),
                "d2h_bytes_per_step": int(image.nbytes), "ms_per_step": e2e_wall_ms / args.steps,
                "samples_per_s": e2e_paths / (e2e_wall_ms * 1e-3),
                "timing": "wall clock around upload_scene + set_sensor + render + resolve through the C ABI"},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": roofline,
    }
    if base is not None:
        line["cpu_baseline"] = base
    print(json.dumps(line), flush=True)
    if distributed:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--estimator", default="nee", choices=["nee", "reference"])
    ap.add_argument("--pool", type=int, default=0, help="resident path slots (0 = library default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
