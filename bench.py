#!/usr/bin/env python
"""bench.py -- the headline benchmark: Cornell box 1024x1024, 1024 spp, path tracing.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One step = one pass of the hot path over the whole workload (BASELINE.json configs[1]):
every pixel of the 1024x1024 frame x 1024 samples, i.e. 1.07e9 camera paths through
raygen / extend / shade / connect on one B200.  With N ranks (torchrun) every rank renders
1024 spp of its own slice of the global sample range [0, N*1024) -- weak scaling, no
data-path collective -- and ONE NCCL all-reduce sums the film at the end of each step.

Rank 0 prints ONE JSON line (see the contract in the task description).  `value` is SAMPLES
(camera paths) per second of device time with the scene resident -- the unit of work both arms
share: the GPU's estimator also traces shadow rays, the reference's does not, so a ray count is
not comparable across the arms; Mrays/s is printed next to it (`mrays_per_s`, `rays_per_path`).
`e2e` goes through the C ABI with host buffers every step (scene upload H2D, render, resolved
image D2H).  `--impl reference` times the UNMODIFIED reference's CPU path tracer (oracle/_ref,
all host threads) on a bounded sample of the same workload instead.

After the timed headline the same process measures, once each, what BASELINE.json's other configs
ask for (keys `strong`, `config4`, `config5` of the same line): the headline job split over the N
ranks (strong scaling), config 4 (1.3 M-triangle mesh, 2048^2 x 64 spp split over the ranks, with the
roofline of the BVH kernels), and at N = 8 config 5 (4096^2 x 16 384 spp split over the ranks).  At N = 1
`other_estimators` adds the throughput of light tracing and bidirectional path tracing
on the headline scene.
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
METRIC = "Msamples/s (camera paths per second; Mrays/s alongside), Cornell box 1024x1024 unidirectional path tracing"
UNIT = "Msamples/s"


def config(world: int = 1) -> dict:
    """The workload both arms are quoted on (identical in the two JSON lines)."""
    return {
        "workload": f"Cornell box {WIDTH}x{HEIGHT} path tracing, {SPP} spp (BASELINE configs[1])",
        "frame": [WIDTH, HEIGHT], "spp": SPP, "scene": "built-in Cornell box (25 objects)",
        "unit_of_work": "one sample = one camera path traced to termination",
    }
FILM = (0.036, 0.036)
# algorithmic bytes per ray over the whole wavefront ...
BYTES_PER_CLOSEST_RAY = 160
BYTES_PER_SHADOW_RAY = 108
# ... and each kernel's share of those figures: bytes per (closest ray, shadow ray, camera path).
# The records in HBM are laid out differently (8-byte hits, 32-byte state records), which
# is an implementation choice: the roofline fraction is quoted on the survey's model, and next to it on the DRAM
# bytes ncu measured for the kernel (roofline.traffic).
KERNEL_BYTES = {
    # write ray 32 + path state 16 (+ 4 for the slot id it reads); not part of 8(d)'s per-ray figure
    "raygen": (0, 0, 52),
    # read ray 32, write hit 16
    "extend": (48, 0, 0),
    # read ray 32 + hit 16 + state 16, write next ray 32 + state 16 | shadow ray 32 + contribution 16
    "shade": (112, 48, 0),
    # read 48, film add 12
    "connect": (0, 60, 0),
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
    return {"value": paths / seconds / 1e6, "unit": UNIT, "cores": threads, "kind": kind,
            "sample": f"Cornell {WIDTH}x{HEIGHT}, {spp} spp ({paths} paths, {casts} rays) in {seconds:.2f} s; "
                      f"reference estimator (no shadow rays), unbounded depth",
            "mrays_per_s": casts / seconds / 1e6}


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
        paths = WIDTH * HEIGHT * spp * args.steps
        value = paths / seconds / 1e6
        mrays = rays / seconds / 1e6
        sample = f"each step: {spp} spp of the {SPP}-spp workload (a rate: the frame and the scene are the workload's)"
        ms = seconds / args.steps * 1e3
    else:
        value, sample, mrays, ms = base["value"], base["sample"], base["mrays_per_s"], None
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config(),
        "arm": {"estimator": "reference (algorithm_pt.cc, no light sampling, unbounded depth)", "threads": threads},
        "mrays_per_s": mrays, "rays_per_path": {"closest": mrays / max(value, 1e-30), "shadow": 0.0},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------
def run_gpu(args) -> None:
    import ctypes

    import torch

    from amber_b200 import capi, cornel_box, cornel_mesh
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
    # the e2e leg's result buffer: PINNED host memory (a pageable destination is staged by the driver through
    # bounce buffers at a speed that varies several-fold from box to box: 2 ms to 29 ms for these 12.6 MB)
    image = torch.empty((HEIGHT, WIDTH, 3), dtype=torch.float32, pin_memory=True).numpy()

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    def render_and_reduce(first: int, n: int, depth: int = MAX_DEPTH) -> float:
        """This rank's samples [first, first + n) + the film all-reduce; device milliseconds."""
        ms, done = 0.0, 0
        while done < n or (n == 0 and done == 0):
            k = min(n - done, 4096)           # sample indices of one call stay below 2^32
            dev.render(first + done, k, seed=SEED, max_depth=depth, estimator=estimator, pool_paths=args.pool)
            if k:
                ms += dev.stats()["last_render_ms"]
            done += max(k, 1)
        if distributed:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            reduce_device_film(dev)
            e1.record()
            torch.cuda.synchronize()
            ms += e0.elapsed_time(e1)
        return ms

    e2e_phases = {"upload_scene": 0.0, "set_sensor": 0.0, "render_and_reduce": 0.0, "resolve": 0.0}

    def step(e2e: bool) -> float:
        """Returns device milliseconds of this rank's step."""
        t0 = time.perf_counter()
        if e2e:
            dev.upload_scene(scene)      # host scene description -> device (H2D)
            t1 = time.perf_counter()
            dev.set_sensor(sensor)       # (clears the film)
        else:
            t1 = t0
            dev.film_clear()
        t2 = time.perf_counter()
        ms = render_and_reduce(begin, count)
        t3 = time.perf_counter()
        if e2e and rank == 0:
            dev.resolve(image)           # mean image -> host (D2H)
        if e2e:
            t4 = time.perf_counter()
            for key, dt in zip(e2e_phases, (t1 - t0, t2 - t1, t3 - t2, t4 - t3)):
                e2e_phases[key] += dt * 1e3
        return ms

    def over_ranks(values, how):
        t = torch.tensor(values, dtype=torch.float64, device=f"cuda:{local_rank}")
        if distributed:
            dist.all_reduce(t, op=dist.ReduceOp.MAX if how == "max" else dist.ReduceOp.SUM)
        return [float(x) for x in t]

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
        closest = shadow = paths = truncated = 0
        for _ in range(args.steps):
            device_ms += step(e2e)
            s = dev.stats()
            closest += s["closest_rays"]
            shadow += s["shadow_rays"]
            paths += s["paths"]
            truncated += s["truncated_paths"]
        barrier()
        wall_ms = (time.perf_counter() - wall0) * 1e3
        launches = dev.stats()["kernel_launches"] - launches0
        clock_info = clocks.stop() if rank == 0 else None
        device_ms, wall_ms = over_ranks([device_ms, wall_ms], "max")
        closest, shadow, paths, truncated, launches = over_ranks([closest, shadow, paths, truncated, launches], "sum")
        return device_ms, wall_ms, closest, shadow, paths, truncated, int(launches), clock_info

    device_ms, wall_ms, closest, shadow, paths, truncated, launches, clocks = timed(False)
    e2e_device_ms, e2e_wall_ms, e2e_closest, e2e_shadow, e2e_paths, _, _, _ = timed(True)

    peak, peak_kind = measured_peaks()

    # per-kernel timing of the dominant kernel: one extra step with an event pair per launch
    roofline = None
    if rank == 0:
        dev.set_profiling(True)
        dev.film_clear()
        dev.render(begin, count, seed=SEED, max_depth=MAX_DEPTH, estimator=estimator, pool_paths=args.pool)
        s = dev.stats()
        dev.set_profiling(False)
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
            "bytes_model": {"per_closest_ray": per_closest, "per_shadow_ray": per_shadow, "per_camera_path": per_path,
                            "source": "SURVEY.md 8(d): this kernel's share of the 160 B per closest ray / 108 B per shadow ray"},
            "closest_rays_per_launch": s["closest_rays"] / launches_each,
            "shadow_rays_per_launch": s["shadow_rays"] / launches_each,
            "avg_launch_ms": avg_ms, "launches_timed": launches_each,
            "timing": "CUDA events on the context's stream around every launch of one extra full step",
            "kernel_share_of_step": {k: v / max(1e-9, sum(kernel_ms.values())) for k, v in kernel_ms.items()},
            "kernel_avg_launch_us": {k: 1e3 * v / launches_each for k, v in kernel_ms.items()},
            "whole_wavefront": {"achieved": whole, "frac": whole / peak,
                                "bytes_model": "160 B/closest ray + 108 B/shadow ray (SURVEY 8d)"},
        }
        traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(traffic_file):
            measured = json.load(open(traffic_file)).get(dominant, {})
            roofline["traffic"] = measured.get("dram_bytes_per_launch")
            if roofline["traffic"]:
                roofline["frac_from_measured_traffic"] = roofline["traffic"] / (avg_ms * 1e-3) / 1e9 / peak

    # ---- strong scaling of the headline job: the SAME 1024 spp split over the ranks -------------------
    def best_of(n_repeats, first, n, depth=MAX_DEPTH):
        best = None
        for _ in range(n_repeats):
            dev.film_clear()
            barrier()
            ms = render_and_reduce(first, n, depth)
            s = dev.stats()
            ms, = over_ranks([ms], "max")
            sums = over_ranks([s["closest_rays"], s["shadow_rays"], s["paths"], s["truncated_paths"]], "sum")
            if best is None or ms < best[0]:
                best = [ms] + sums
        return best

    def summary(ms, closest_, shadow_, paths_, truncated_):
        return {"ms": ms, "msamples_per_s": paths_ / ms / 1e3, "mrays_per_s": (closest_ + shadow_) / ms / 1e3,
                "rays_per_path": {"closest": closest_ / max(1.0, paths_), "shadow": shadow_ / max(1.0, paths_)},
                "truncated_fraction": truncated_ / max(1.0, paths_)}

    extras = {}
    if not args.headline_only:
        s_begin, s_count = partition_samples(SPP, rank, world)
        strong = summary(*best_of(2, s_begin, s_count))
        strong.update({"job": f"the headline's {SPP} spp split over {world} rank(s), film all-reduce included",
                       "spp_per_rank": s_count})
        extras["strong"] = strong

        # ---- config 4: 1.3 M-triangle mesh, 2048^2 x 64 spp split over the ranks ----------------------
        t0 = time.perf_counter()
        mesh = cornel_mesh(8)
        scene_s = time.perf_counter() - t0
        t0 = time.perf_counter()
        dev.upload_scene(mesh)
        upload_s = time.perf_counter() - t0
        dev.set_sensor(Sensor(2048, 2048))
        m_begin, m_count = partition_samples(64, rank, world)
        best_of(1, m_begin, m_count)   # warm-up (pool allocation)
        c4 = summary(*best_of(2, m_begin, m_count))
        c4_profile = None
        if rank == 0:
            dev.set_profiling(True)
            dev.film_clear()
            dev.render(m_begin, m_count, seed=SEED, max_depth=MAX_DEPTH, estimator=estimator, pool_paths=args.pool)
            p = dev.stats()
            dev.set_profiling(False)
            n_launch = max(1, p["extend_launches"])
            ext_us, con_us = 1e3 * p["extend_ms"] / n_launch, 1e3 * p["connect_ms"] / n_launch
            floor_extend = p["bvh_node_visits"] * p["bvh_node_bytes"] + p["bvh_triangle_tests"] * p["bvh_triangle_bytes"] \
                + 40 * p["closest_rays"]
            floor_connect = p["bvh_node_visits_connect"] * p["bvh_node_bytes"] + p["bvh_triangle_tests_connect"] * p["bvh_triangle_bytes"] \
                + 68 * p["shadow_rays"]
            l1_demand = floor_extend / max(p["extend_ms"], 1e-9) / 1e6      # GB/s
            c4_profile = {
                "per_launch_us": {"raygen": 1e3 * p["raygen_ms"] / n_launch, "extend": ext_us,
                                  "shade": 1e3 * p["shade_ms"] / n_launch, "connect": con_us},
                "rays_per_launch": {"closest": p["closest_rays"] / n_launch, "shadow": p["shadow_rays"] / n_launch},
                "extend_grays_per_s": p["closest_rays"] / max(p["extend_ms"], 1e-9) / 1e6,
                "connect_grays_per_s": p["shadow_rays"] / max(p["connect_ms"], 1e-9) / 1e6,
                "node_visits_per_closest_ray": p["bvh_node_visits"] / max(1, p["closest_rays"]),
                "triangle_tests_per_closest_ray": p["bvh_triangle_tests"] / max(1, p["closest_rays"]),
                "node_visits_per_shadow_ray": p["bvh_node_visits_connect"] / max(1, p["shadow_rays"]),
                "triangle_tests_per_shadow_ray": p["bvh_triangle_tests_connect"] / max(1, p["shadow_rays"]),
                "roofline": {
                    "kernel": "extend_kernel<WideBvh>", "bound": "hbm", "unit": "GB/s", "peak": peak, "peak_source": peak_kind,
                    # bytes the traversal must fetch if every node and triangle record it touches moved once
                    # per touch (the L1-level demand): rays x nodes visited x node bytes + triangles tested x
                    # record bytes + 40 B per ray (ray in, hit out).  The L2 and DRAM serve the part of it that
                    # misses L1 / L2 (ncu: profiles/).
                    "achieved": l1_demand, "frac": l1_demand / peak,
                    "algorithmic_bytes_per_launch": floor_extend / n_launch,
                    "avg_launch_ms": ext_us * 1e-3,
                    "node_bytes": p["bvh_node_bytes"], "triangle_bytes": p["bvh_triangle_bytes"],
                    "connect_achieved": floor_connect / max(p["connect_ms"], 1e-9) / 1e6,
                    "traffic": None,
                },
            }
            traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
            if os.path.exists(traffic_file):
                measured = json.load(open(traffic_file)).get("extend_bvh", {})
                roof = c4_profile["roofline"]
                roof["traffic"] = measured.get("dram_bytes_per_launch")
                if roof["traffic"]:
                    # ncu's bytes of one steady-state launch (profiles/traffic.json) over the LIVE launch time
                    roof["dram_gbs"] = roof["traffic"] / (ext_us * 1e-6) / 1e9
                    roof["dram_frac"] = roof["dram_gbs"] / peak
                if measured.get("l2_read_bytes_per_launch"):
                    roof["l2_read_bytes_per_launch"] = measured["l2_read_bytes_per_launch"]
                    roof["l2_read_gbs"] = measured["l2_read_bytes_per_launch"] / (ext_us * 1e-6) / 1e9
                roof["reading"] = ("achieved = the traversal's byte demand (nodes visited x node bytes + triangles tested x record "
                                   "bytes + 40 B per ray) over the launch time: what L1 must deliver; l2_read_gbs / dram_gbs = what "
                                   "L2 / DRAM actually moved (ncu).  The kernel is bound by dependent-load latency at ~14 of 32 "
                                   "lanes per instruction, not by any of these bandwidths (profiles/r02_mesh_summary.txt)")
        c4.update({"job": f"Cornell room + 1 310 720-triangle displaced icosphere, 2048x2048, 64 spp split over {world} rank(s) "
                          "(BASELINE configs[3])", "spp_per_rank": m_count, "primitives": int(mesh.desc.n_primitives),
                   "scene_generation_s": scene_s, "upload_and_bvh_build_s": upload_s})
        if rank == 0:
            st = dev.stats()
            c4.update({"bvh_nodes": st["bvh_nodes"], "host_build_ms": st["scene_build_ms"], "host_to_device_ms": st["scene_upload_ms"]})
            c4["profile"] = c4_profile
            if world == 1:
                # the same job with the tree built on the device (AMBER_CUDA_BVH_BUILDER=lbvh, bvh_build_gpu.cuh)
                os.environ["AMBER_CUDA_BVH_BUILDER"] = "lbvh"
                t0 = time.perf_counter()
                dev.upload_scene(mesh)
                lbvh_upload_s = time.perf_counter() - t0
                os.environ.pop("AMBER_CUDA_BVH_BUILDER")
                dev.set_sensor(Sensor(2048, 2048))
                st = dev.stats()
                lbvh = summary(*best_of(2, m_begin, m_count))
                c4["device_built_tree"] = {"upload_and_bvh_build_s": lbvh_upload_s, "scene_build_ms": st["scene_build_ms"],
                                           "host_to_device_ms": st["scene_upload_ms"], "bvh_nodes": st["bvh_nodes"],
                                           "ms": lbvh["ms"], "mrays_per_s": lbvh["mrays_per_s"],
                                           "note": "Morton-order radix tree built on the GPU: the tree is ready sooner and traces "
                                                   "slower than the host's binned-SAH tree above (the default)"}
        extras["config4"] = c4

        # ---- config 5 (N = 8 only): 4096^2 x 16 384 spp split over the ranks ------------------------------
        if world >= 8:
            dev.upload_scene(scene)
            dev.set_sensor(Sensor(4096, 4096))
            b5, n5 = partition_samples(16384, rank, world)
            t0 = time.perf_counter()
            c5 = summary(*best_of(1, b5, n5))
            barrier()
            c5_wall = time.perf_counter() - t0
            c5.update({"job": f"Cornell box 4096x4096, 16 384 spp split over {world} ranks, one 268 MB film all-reduce "
                              "(BASELINE configs[4])", "seconds": c5["ms"] / 1e3, "wall_seconds": c5_wall,
                       "gpaths_per_s": c5["msamples_per_s"] / 1e3})
            if rank == 0:
                mean_image, samples = dev.resolve()
                c5.update({"mean": float(mean_image.mean()), "samples_in_film": samples,
                           "mean_times_pixels_over_1024x1024": float(mean_image.mean()) * 16.0})
            extras["config5"] = c5

        # ---- rows N2 / N3 of SURVEY 8(f) on the headline scene (one GPU only; not part of the headline) ------
        if world == 1:
            dev.upload_scene(scene)
            dev.set_sensor(Sensor(WIDTH, HEIGHT))
            rows = {}
            for name, other, spp, depth in (("light_tracing", capi.ESTIMATOR_LIGHT_TRACING, 64, MAX_DEPTH),
                                                ("bidirectional_path_tracing", capi.ESTIMATOR_BDPT, 32, 0)):
                best_ms, best_stats = None, None
                for _ in range(2):
                    dev.film_clear()
                    dev.render(0, spp, seed=SEED, max_depth=depth, estimator=other)
                    st = dev.stats()
                    if best_ms is None or st["last_render_ms"] < best_ms:
                        best_ms, best_stats = st["last_render_ms"], st
                samples = float(spp) * WIDTH * HEIGHT
                rows[name] = {"msamples_per_s": samples / best_ms / 1e3, "ms": best_ms, "spp": spp,
                              "closest_rays_per_sample": best_stats["closest_rays"] / samples,
                              "film_splats_per_sample": best_stats["light_splats"] / samples}
            # ... and the bidirectional estimator on config 4's scene (casts through the BVH traversal kernels)
            dev.upload_scene(mesh)
            dev.set_sensor(Sensor(WIDTH, HEIGHT))
            best_ms, best_stats = None, None
            for _ in range(2):
                dev.film_clear()
                dev.render(0, 16, seed=SEED, max_depth=0, estimator=capi.ESTIMATOR_BDPT)
                st = dev.stats()
                if best_ms is None or st["last_render_ms"] < best_ms:
                    best_ms, best_stats = st["last_render_ms"], st
            samples = 16.0 * WIDTH * HEIGHT
            rows["bidirectional_path_tracing_config4_scene"] = {
                "msamples_per_s": samples / best_ms / 1e3, "ms": best_ms, "spp": 16,
                "closest_rays_per_sample": best_stats["closest_rays"] / samples,
                "film_splats_per_sample": best_stats["light_splats"] / samples,
                "note": "1 310 738 triangles, 1024x1024"}
            rows["light_tracing"]["note"] = "one sample = one light path per pixel (algorithm_lt.cc:112-122)"
            rows["bidirectional_path_tracing"]["note"] = ("one sample = one light + one eye subpath per pixel, every (s, t) connected "
                                                          "(bidirectional_path_sampling.h); wavefront implementation (DESIGN 4.10)")
            extras["other_estimators"] = rows

    if rank != 0:
        if distributed:
            dist.destroy_process_group()
        return

    base = cpu_baseline() if world == 1 and not args.no_cpu_baseline else None
    value = paths / (device_ms * 1e-3) / 1e6
    e2e_value = e2e_paths / (e2e_wall_ms * 1e-3) / 1e6
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": device_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config(world),
        "arm": {
            "estimator": args.estimator, "max_depth": MAX_DEPTH, "seed": SEED, "pool_paths": args.pool or "default",
            "spp_per_gpu": SPP,
            "parallelism": f"sample-partition x{world}, one film all-reduce per step" if distributed else "single GPU",
            "l2": "working set (path pool + queues, > 300 MB) exceeds the 126 MB L2; film cleared between steps",
        },
        "mrays_per_s": (closest + shadow) / (device_ms * 1e-3) / 1e6,
        "rays_per_path": {"closest": closest / max(1.0, paths), "shadow": shadow / max(1.0, paths)},
        "truncated_fraction": truncated / max(1.0, paths),
        "ms_per_step_wall": wall_ms / args.steps,
        "e2e": {"value": e2e_value, "unit": UNIT,
                "h2d_bytes_per_step": int(ctypes.sizeof(scene.primitives) + ctypes.sizeof(scene.materials)
                                          + ctypes.sizeof(scene.lens)),
                "d2h_bytes_per_step": int(image.nbytes), "ms_per_step": e2e_wall_ms / args.steps,
                "mrays_per_s": (e2e_closest + e2e_shadow) / (e2e_wall_ms * 1e-3) / 1e6,
                "timing": "wall clock around upload_scene + set_sensor + render + resolve through the C ABI",
                "host_ms_per_step": {k: v / args.steps for k, v in e2e_phases.items()}},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": roofline,
    }
    line.update(extras)
    if base is not None:
        line["cpu_baseline"] = base
        line["vs_cpu_baseline"] = {
            "samples_per_s": value / base["value"], "e2e_samples_per_s": e2e_value / base["value"],
            "mrays_per_s": line["mrays_per_s"] / base["mrays_per_s"],
            "note": "the GPU arm also traces shadow rays (next-event estimation); the reference's estimator has "
                    "none, so the samples/s ratio is the like-for-like one",
        }
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
    ap.add_argument("--headline-only", action="store_true", help="skip the strong / config 4 / config 5 legs")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
