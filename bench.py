#!/usr/bin/env python
"""bench.py — headline benchmark of the render hot path (BASELINE.json: Mpaths/s and ray-sphere tests/s on 1/2/4/8 B200,
as a fraction of the FP32 roofline, beside the reference's CPU renderer timed on the box's host cores).

  python bench.py --gpus N --steps K --warmup W            this repo's CUDA path (one rank per GPU under torchrun for N > 1)
  python bench.py --impl reference --steps K --warmup W    the reference's CPU implementation of the path (C++ port)

A "step" is one full render of the workload (trace + framebuffer exchange + resolve).  The default workload at EVERY N is the
north_star target T: random-spheres (486 spheres), 1920x1080, 1000 spp IN TOTAL, depth 50 — STRONG scaling: N ranks trace
1000/N samples of every pixel each (sample-range sharding) and the partial framebuffers are combined by the fused
peer-memory reduce+resolve kernel over NVLink (NCCL reduce as the fallback).  --workload C2 is BASELINE configs[1]
(1920x1080 x 100 spp; weak scaling at N > 1); the N = 1 line of the default run carries C2's ms_per_step as `c2_ms_per_step`.

`value`   : Mpaths/s, whole job, scene/camera resident on the device, no host copies in the timed region (CUDA events,
            max over ranks).
`e2e`     : the same metric through the reference-facing C-ABI call with HOST buffers — rtb200_render at N = 1,
            rtb200_multi_render (ONE call driving all N GPUs from one host thread: what a Rust host's `render` calls,
            integration/rust/src/cuda_renderer.rs) at N > 1 — every step uploads the scene (host->device) and reads the RGB8
            image back into pinned host memory (device->host).
`roofline`: FP32 ALU bound (no dense contraction; KBs of scene + one framebuffer): achieved = ray-sphere tests/s x 17 FLOP
             over the trace kernel's own launch time (events around that kernel alone, recorded inside the
            library on the launching stream); peak = the nominal FP32 FMA rate 2 x 128 x SMs x f_max (MEASURED_PEAKS.json
            holds no FP32 figure and the profiling guide states no FP32 fallback); `frac_of_k0_probe` = against the FFMA
            probe measured live on this device.  `executed` = what the kernel really issues (ncu of the same workload,
            profiles/traffic.json): the two-stage sweep runs 3 FFMA on most pairs, so the binding resource is the issue slot.
`parity`  : computed OUTSIDE the timed region from the image the timed configuration produces: (1) N > 1: the N-GPU linear
            image before quantisation against the single-GPU image of the same job (north_star bar: 1e-4 per channel on the
            mean); (2) three scanlines against the CPU oracle (mixed mode = the GPU comparand), which reproduces the device's
            summation tree (K sample sub-ranges per rank in order, ranks in order): linear within 1e-4, RGB8 equal.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time
import types

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "rust-ray-tracing_b200"

FLOP_PER_TEST = 17.0   # SURVEY §8d / DESIGN.md: 5 FMA + 7 add/mul per ray-sphere test

WORKLOADS = {
    # name: (scene generator, width, height, spp, depth)
    "T": ("random_spheres", 1920, 1080, 1000, 50),      # the north_star target; spp is the TOTAL over all GPUs
    "C3": ("random_spheres", 3840, 2160, 1000, 50),     # BASELINE configs[2]; total spp
    "C2": ("random_spheres", 1920, 1080, 100, 50),      # BASELINE configs[1]; spp per GPU
    "C4": ("ten_k_spheres", 1920, 1080, 64, 50),
    "C5": ("dielectric_heavy", 1920, 1080, 256, 50),
    "C1": ("three_spheres", 400, 225, 16, 50),
    # beyond BASELINE: 40 001 spheres, the streamed trace kernel (stage-1 data fetched per warp from global memory)
    "L40K": ("forty_k_spheres", 1920, 1080, 16, 50),
}
STRONG = {"C3", "T"}   # fixed TOTAL sample counts, split over the GPUs


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        # median of the samples under load (upper half): idle samples before/after the region are not the clock we ran at
        load = sm[len(sm) // 2:] if sm else []
        med = load[len(load) // 2] if load else None
        return {"sm_mhz": med, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}


def host_only_scenes():
    """The scene generators WITHOUT the package __init__ (which loads librtb200.so): the reference arm must not map the
    repo's library.  scenes.py / scene.py need only the plain-ctypes struct mirrors (_abi_types.py)."""
    name = "rtb200_hostonly"
    if name not in sys.modules:
        pkg = types.ModuleType(name)
        pkg.__path__ = [os.path.join(ROOT, PKG)]
        sys.modules[name] = pkg
    return importlib.import_module(name + ".scenes")


def cpu_reference_run(workload: str, steps: int, warmup: int, budget_s: float, full: bool = False):
    """Times the CPU port of the reference's LIVE path (oracle/cpu_ref.cpp: TileRenderer + render_vectorized2 +
    trace_vectorized2 + hit_packed, 4 x f64 AVX2 lanes, 128x128 tiles, all host threads) on a bounded sample of the
    workload: the SAME scene, camera, resolution and depth, with the samples per pixel cut so that one step fits
    `budget_s` seconds (a calibration pass at 4 spp sets the count).  Mpaths/s is a rate: it does not depend on the
    sample count (profiles/ keeps one full-configuration CPU step of C2 that shows it)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as oracle
    scenes = host_only_scenes()
    gen, W, H, spp, depth = WORKLOADS[workload]
    scene = getattr(scenes, gen)()
    arrays = scene.arrays()
    cores = os.cpu_count() or 1
    # camera through the oracle's own Camera::new (no GPU library involved on this arm)
    cam = oracle.camera_new(W, H, 10.0, 30.0, (16.0, 2.0, 18.5), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 0.0)
    if full:
        s_spp = spp
    else:
        cal_rows = (0, H)
        _, t4, _, _ = oracle.cpuref_render(arrays, W, H, cam, 4, depth, threads=cores, want_rgb8=False, row_range=cal_rows)   # calibrate: 4 spp
        s_spp = max(4, min(spp, int(budget_s / max(t4 / 4.0, 1e-6)) // 4 * 4))
    times, segs, tests = [], 0, 0
    for i in range(warmup + steps):
        _, sec, tests, segs = oracle.cpuref_render(arrays, W, H, cam, s_spp, depth, threads=cores, want_rgb8=True)
        if i >= warmup:
            times.append(sec)
    paths = W * H * s_spp
    sec = sum(times) / len(times)
    return {
        "mpaths_s": paths / sec * 1e-6,
        "tests_s": tests / sec,
        "ms_per_step": sec * 1e3,
        "cores": cores,
        "sample_fraction": s_spp / spp,
        "same_config": s_spp == spp,
        "sample": f"{gen} {W}x{H} (full resolution, same camera), {s_spp} of the workload's {spp} spp, depth {depth}: {paths} paths/step; C++ port of "
                  f"TileRenderer + trace_vectorized2 + hit_packed (4 x f64 AVX2 lanes, 128x128 tiles) on {cores} threads",
        "n_spheres": len(scene),
    }


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    gen, W, H, spp, depth = WORKLOADS[args.workload]
    steps = max(1, args.steps)
    total = steps + args.warmup
    budget = max(2.0, min(12.0, 120.0 / total))         # the whole arm stays within ~2 minutes + calibration
    res = cpu_reference_run(args.workload, steps, args.warmup, budget, full=args.full_reference)
    line = {
        "impl": "reference",
        "metric": "Mpaths/s", "value": res["mpaths_s"], "unit": "Mpaths/s", "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup,
        "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": "strong" if args.workload in STRONG else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {gen} {W}x{H}, {spp} spp, depth {depth} (bounded sample, see cpu_baseline.sample)",
                   "n_spheres": res["n_spheres"], "sample_fraction": res["sample_fraction"], "same_config": res["same_config"]},
        "sphere_tests_per_s": res["tests_s"],
        "cpu_baseline": {"value": res["mpaths_s"], "unit": "Mpaths/s", "cores": res["cores"], "kind": "port", "sample": res["sample"]},
        "e2e": {"value": res["mpaths_s"], "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


def oracle_strips(oracle, np, scene, cam, W, H, spp_total, depth, rows, ranks_samples, splits):
    """The CPU oracle's image of scanlines `rows`, summed with the DEVICE's summation tree: rank g traced samples
    ranks_samples[g] as `splits` contiguous sub-ranges whose FP32 partial sums were added in order (k_combine), and the
    ranks' sums were added in rank order (k_resolve_peers).  Returns (sum[len(rows),W,3] float32, rgb8[len(rows),W,3])."""
    from concurrent.futures import ThreadPoolExecutor
    arrays, cam18 = scene.arrays(), cam.vectors()

    def part(job):       # one (scanline, sample sub-range): the oracle's FP32 sum over [a, b) (single-threaded; the pool is the parallelism)
        r, a, b = job
        if b <= a:
            return np.zeros((W, 3), dtype=np.float32)
        s, _, _ = oracle.render(arrays, W, H, cam18, spp_total, depth, seed=1, f32=True, sample_range=(a, b), row_range=(r, r + 1),
                                want_rgb8=False, threads=1)
        return s[r].astype(np.float32)                                 # exact: the oracle's float sum, carried in float64

    jobs = []
    for r in rows:
        for (s0, s1) in ranks_samples:
            n = s1 - s0
            for k in range(splits):
                jobs.append((r, s0 + (k * n) // splits, s0 + ((k + 1) * n) // splits))
    with ThreadPoolExecutor(max_workers=os.cpu_count() or 1) as pool:       # ctypes releases the GIL inside the oracle
        parts = list(pool.map(part, jobs))
    out_sum = np.zeros((len(rows), W, 3), dtype=np.float32)
    j = 0
    for i, r in enumerate(rows):
        total = None
        for _ in ranks_samples:
            rank_sum = None
            for k in range(splits):
                rank_sum = parts[j] if rank_sum is None else (rank_sum + parts[j]).astype(np.float32)
                j += 1
            total = rank_sum if total is None else (total + rank_sum).astype(np.float32)
        out_sum[i] = total
    mean = (out_sum / np.float32(spp_total)).astype(np.float32)            # src/renderer.rs:161
    v = (np.sqrt(mean).astype(np.float32) * np.float32(255.999)).astype(np.float32)   # Color::to_u8_array, src/color.rs:54-64
    rgb = np.where(v >= 255.0, 255, np.where(v >= 0.0, v, 0.0)).astype(np.uint8)
    return out_sum, rgb


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the rtb200 path has no CPU fallback (use --impl reference for the CPU port)")
    torch.cuda.set_device(local_rank)
    host_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        host_group = dist.new_group(backend="gloo")     # host-side waits that must not put a spinning NCCL kernel on the idle GPUs
    rtb = importlib.import_module(PKG)
    dmod = importlib.import_module(PKG + ".distributed")

    gen, W, H, spp, depth = WORKLOADS[args.workload]
    if args.spp:
        spp = args.spp
    strong = args.workload in STRONG
    if strong:
        if spp % world:
            raise SystemExit(f"{args.workload}: {spp} spp do not split over {world} GPUs")
        spp_per_gpu, spp_total = spp // world, spp     # strong scaling: total work fixed
    else:
        spp_per_gpu, spp_total = spp, spp * world      # weak scaling: per-GPU work fixed
    scene = getattr(rtb.scenes, gen)()
    cam = rtb.Camera.default(W, H)
    renderer = rtb.CudaRenderer(device=local_rank, seed=1, rays_per_lane=args.rays, stages=args.stages)
    tracer = dmod.CudaTracer(renderer, scene)
    if "x" in args.sharding:
        gr, gs = (int(v) for v in args.sharding.split("x"))
        sharded = dmod.ShardedRenderer(tracer, mode="grid", grid=(gr, gs), exchange=args.exchange, barrier=args.barrier)
    else:
        sharded = dmod.ShardedRenderer(tracer, mode=args.sharding, exchange=args.exchange, barrier=args.barrier)
    info = renderer.device_info()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=torch.device("cuda", local_rank))
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=torch.device("cuda", local_rank))
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- FP32 roofline denominators ----------------------------------------------------------------------------
    k0_tflops, _ = renderer.fp32_peak(2000)                                                   # FFMA probe, this device, now
    nominal_tflops = 2.0 * 128 * info["sm_count"] * info["sm_clock_khz"] * 1e3 * 1e-12        # 2 x 128 lanes x SMs x f_max

    # ---- device-resident timing: K steps of trace + exchange + resolve ------------------------------------------
    for _ in range(args.warmup):
        sharded.render_device(depth, spp_total, cam)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = renderer.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for i in range(args.steps):
        sharded.render_device(depth, spp_total, cam)                     # trace + exchange + resolve
    ev1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = renderer.launch_count() - launches0
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    kstat = renderer.trace_kernel_stats(min(args.steps, 64))             # the trace kernel alone, events on its launch stream
    trace_ms = kstat["avg_ms"]
    paths_local, segs_local = renderer.read_counters()                   # counters of the last trace on this rank
    paths = sum_over_ranks(float(paths_local))
    segs = sum_over_ranks(float(segs_local))
    ms_per_step = ms_total / args.steps
    n_spheres = len(scene)
    mpaths = paths / (ms_per_step * 1e-3) * 1e-6
    tests_s = segs * n_spheres / (ms_per_step * 1e-3)
    tests_per_launch = segs_local * n_spheres
    achieved_tflops = tests_per_launch * FLOP_PER_TEST / (trace_ms * 1e-3) * 1e-12

    # ---- parity, outside the timed region ------------------------------------------------------------------------
    parity = None
    if not args.no_parity:
        sharded.want_sum = True
        img_dev = sharded.render_device(depth, spp_total, cam)          # the timed configuration once more, keeping the linear image
        torch.cuda.synchronize()
        barrier()
        splits = renderer.trace_kernel_stats(1)["sample_splits"]
        if rank == 0:
            if world > 1:
                lin_n = sharded.last_sum()
                lin_n = lin_n.cpu().numpy() if lin_n is not None else None      # None: NCCL exchange (no summed image kept)
            else:
                lin_n = None
            img_n = img_dev.cpu().numpy()
            single_img, single_lin, single_stat = renderer.render(depth, spp_total, scene, cam, want_linear=True)   # the whole job on one GPU
            parity = {"tolerance": 1e-4, "computed": "outside the timed region"}
            ok = True
            if world > 1:
                d_rgb = int(np.abs(img_n.astype(np.int32) - single_img.astype(np.int32)).max())
                parity["rgb8_max_abs_vs_single_gpu"] = d_rgb
                ok = ok and d_rgb <= 1
                if lin_n is not None:
                    d_lin = float(np.abs(lin_n[..., :3].astype(np.float64) - single_lin[..., :3].astype(np.float64)).max() / spp_total)
                    parity["linear_max_abs_vs_single_gpu"] = d_lin
                    parity["sample_counts_exact"] = bool(np.all(lin_n[..., 3] == spp_total))
                    ok = ok and d_lin <= 1e-4 and parity["sample_counts_exact"]
            else:
                lin_n = single_lin
                img_n = single_img
            # three scanlines against the CPU oracle with the device's summation tree
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import oracle_lib as oracle
            rows = [H // 8, H // 2, H - 4] if not args.parity_rows else [int(r) for r in args.parity_rows.split(",")]
            shard_ranges = [sharded.shard(spp_total, H)[0]] if world == 1 else None
            if world > 1:
                R, S = sharded.grid
                shard_ranges = [dmod.shard_range(spp_total, S, k) for k in range(S)]
            strips = []
            t_or = time.perf_counter()
            o_sum, o_rgb = oracle_strips(oracle, np, scene, cam, W, H, spp_total, depth, rows, shard_ranges, splits)
            for i, r in enumerate(rows):
                if lin_n is not None:
                    d = float(np.abs(lin_n[r, :, :3].astype(np.float64) - o_sum[i].astype(np.float64)).max() / spp_total)
                else:
                    d = None
                eq = bool(np.array_equal(img_n[r], o_rgb[i]))
                strips.append({"row": r, "linear_max_abs": d, "rgb8_equal": eq,
                               "rgb8_max_abs": int(np.abs(img_n[r].astype(np.int32) - o_rgb[i].astype(np.int32)).max())})
                ok = ok and eq and (d is None or d <= 1e-4)
            parity["oracle"] = {"mode": "mixed (FP32 path state, f64 hit test) — oracle/rt_oracle.hpp", "sample_splits": splits,
                                "strips": strips, "seconds": time.perf_counter() - t_or}
            parity["ok"] = bool(ok)
        barrier()
        sharded.want_sum = False

    # ---- end to end through the reference-facing C-ABI call with host buffers --------------------------------------
    e2e_steps = max(1, args.steps)
    h2d = (n_spheres + 31) // 32 * 32 * (32 + 16 + 16 + 16 + 12 + 4)    # the scene arena: exact f64 | f32 | filter | material | stage-1 triples | kind
    d2h = W * H * 3
    e2e_split, e2e_torchrun, e2e_call = None, None, None
    if world == 1:
        e2e_call = "rtb200_render (Renderer.render): scene upload + trace + resolve + RGB8 read-back into pinned host memory"
        out_img = renderer.pinned_empty((H, W, 3), np.uint8)   # the caller's image buffer: pinned host memory
        renderer.render(depth, spp_total, scene, cam, out=out_img)      # warm-up (allocations)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            renderer.upload_scene(scene)                  # host -> device: the step's inputs
            img, stat = renderer.render(depth, spp_total, scene, cam, out=out_img)   # device -> host: RGB8 image
        e2e_s = (time.perf_counter() - t0) / e2e_steps
        e2e_split = {"trace_ms": stat.trace_ms, "resolve_ms": stat.resolve_ms, "d2h_ms": stat.d2h_ms, "call_ms": stat.duration * 1e3}
    else:
        # (a) one process per GPU (torchrun), the public ShardedRenderer.render with a pinned host image on rank 0
        host_img = torch.empty((H, W, 3), dtype=torch.uint8, pin_memory=True) if rank == 0 else None
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            tracer.upload()
            sharded.render(depth, spp_total, cam, out=host_img)
        torch.cuda.synchronize()
        e2e_tr_s = max_over_ranks((time.perf_counter() - t0) / e2e_steps)
        e2e_torchrun = {"value": paths / e2e_tr_s * 1e-6, "unit": "Mpaths/s", "ms_per_step": e2e_tr_s * 1e3,
                        "call": "ShardedRenderer.render under torchrun (one process per GPU), pinned host image on rank 0"}
        # (b) the ONE C-ABI call a Rust host would make: rank 0 drives all N GPUs through rtb200_multi_render while the other
        #     ranks wait (their processes keep their contexts; their GPUs are otherwise idle)
        e2e_s = None
        barrier()
        if rank == 0 and torch.cuda.device_count() >= world:
            try:
                multi = rtb.MultiGpuRenderer(devices=list(range(world)), seed=1,
                                             row_groups=(1 if args.sharding == "samples" else world if args.sharding == "rows" else int(args.sharding.split("x")[0])))
                out_img = renderer.pinned_empty((H, W, 3), np.uint8)
                multi.render(depth, spp_total, scene, cam, out=out_img)      # warm-up (allocations, peer mappings)
                t0 = time.perf_counter()
                for _ in range(e2e_steps):
                    multi.upload_scene(scene)
                    img, stat = multi.render(depth, spp_total, scene, cam, out=out_img)
                e2e_s = (time.perf_counter() - t0) / e2e_steps
                e2e_split = {"trace_ms": stat.trace_ms, "resolve_ms": stat.resolve_ms, "call_ms": stat.duration * 1e3,
                             "peer_access": multi.info()["peer_access"]}
                e2e_call = (f"rtb200_multi_render: ONE call, one host thread, {world} GPUs (scene upload to every GPU + traces + peer-memory "
                            "reduce+resolve + RGB8 read-back into pinned host memory)")
                if parity is not None and not args.no_parity:
                    d_rgb = int(np.abs(img.astype(np.int32) - img_n.astype(np.int32)).max())
                    parity["one_call_rgb8_max_abs_vs_torchrun"] = d_rgb
                    parity["ok"] = bool(parity["ok"] and d_rgb == 0)     # same shards, same summation order: identical
                multi.close()
                torch.cuda.set_device(local_rank)
            except Exception as e:                      # e.g. the other GPUs are not visible to this rank
                e2e_s = None
                e2e_call = f"rtb200_multi_render unavailable here ({e}); e2e = the torchrun path"
        dist.barrier(group=host_group)                  # the other ranks wait on the HOST (gloo): their GPUs stay idle meanwhile
        barrier()
        if world > 1:
            t = torch.tensor([e2e_s if (rank == 0 and e2e_s) else 0.0], dtype=torch.float64, device=torch.device("cuda", local_rank))
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_s = float(t.item()) or e2e_tr_s
            if e2e_call is None:
                e2e_call = "ShardedRenderer.render under torchrun"
    e2e_mpaths = paths / e2e_s * 1e-6

    # ---- C2 at N = 1 beside the default workload (comparability with round 1's 44.37 ms) --------------------------
    c2_ms = None
    if world == 1 and args.workload == "T" and not args.spp:
        g2, W2, H2, spp2, d2 = WORKLOADS["C2"]
        cam2 = rtb.Camera.default(W2, H2)
        for _ in range(3):
            sharded.render_device(d2, spp2, cam2)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            sharded.render_device(d2, spp2, cam2)
        b.record()
        torch.cuda.synchronize()
        c2_ms = a.elapsed_time(b) / 5

    # ---- CPU baseline (rank 0, N=1 only) -----------------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        res = cpu_reference_run(args.workload, 1, 0, 12.0)
        cpu = {"value": res["mpaths_s"], "unit": "Mpaths/s", "cores": res["cores"], "kind": "port", "sample": res["sample"],
               "sample_fraction": res["sample_fraction"], "sphere_tests_per_s": res["tests_s"]}

    traffic = None       # DRAM bytes per launch of the dominant kernel, from the committed ncu capture of this workload
    executed = None      # what the kernel really executes (ncu of the same workload): issue slots, pipes, lanes per instruction
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f).get(args.workload)
        if t and not args.spp and world == 1:
            traffic = t.get("dram_bytes_per_launch")
            executed = {k: t[k] for k in ("kernel", "issue_slots_busy_pct", "issue_cycles_busy_pct", "fma_pipe_pct", "fma_pipe_cycles_active_pct", "alu_pipe_pct",
                                          "fp64_pipe_pct", "threads_per_instruction", "warp_instructions", "note", "source") if k in t}
    except Exception:
        traffic = None

    if rank == 0:
        exch = sharded.exchange_in_use()
        line = {
            "metric": "Mpaths/s", "value": mpaths, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": f"{args.workload}: {gen} {W}x{H}, {spp_total} spp in total ({spp_per_gpu} per GPU), depth {depth}"
                                   + (" — the north_star target, fixed total work (strong scaling)" if args.workload == "T" else ""),
                       "n_spheres": n_spheres,
                       "sharding": (("sample-range" if args.sharding == "samples" else "row-range (interleaved scanlines)" if args.sharding == "rows"
                                     else "grid " + args.sharding + " (row groups x sample groups)") + "; exchange = "
                                    + {"p2p": "fused NVLink peer-memory reduce+resolve kernel (CUDA IPC); barriers = "
                                              + ("peer-memory flag kernels, no NCCL call inside a step" if args.barrier == "flags" else "two 4-byte NCCL all-reduces"),
                                       "nccl": "NCCL reduce(sum) to rank 0 + resolve"}[exch]) if world > 1 else "single GPU",
                       "l2": "no reuse across steps: each step re-traces from the camera; inputs are KBs, the framebuffer is written once",
                       "precision": "FP32 path state; the hit test of sweep candidates in f64 (the reference's precision)",
                       "seed": 1, "rays_per_lane": args.rays, "stages": args.stages},
            "sphere_tests_per_s": tests_s, "segments_per_path": segs / max(paths, 1.0),
            "roofline": {"bound": "fp32", "achieved": achieved_tflops, "peak": nominal_tflops, "unit": "TFLOP/s",
                         "frac": achieved_tflops / nominal_tflops, "traffic": traffic,
                         "kernel": kstat["kernel"], "kernel_ms": trace_ms, "kernel_launches_timed": kstat["launches"], "flop_per_test": FLOP_PER_TEST,
                         "peak_source": "nominal FP32 FMA rate: 2 FLOP x 128 lanes x %d SMs x %.3f GHz (no FP32 figure in MEASURED_PEAKS.json)"
                                        % (info["sm_count"], info["sm_clock_khz"] * 1e-6),
                         "k0_probe_tflops": k0_tflops, "frac_of_k0_probe": achieved_tflops / k0_tflops,
                         "note": "ALGORITHMIC fraction: achieved counts the reference's 17 FLOP per ray-sphere test (SURVEY 8d); the sweep executes a "
                                 "3-FFMA conservative test on most pairs, so it can exceed what the FMA pipe issues — `executed` has the hardware's view",
                         "executed": executed},
            "e2e": {"value": e2e_mpaths, "unit": "Mpaths/s", "h2d_bytes_per_step": h2d * (world if world > 1 else 1), "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_s * 1e3, "steps": e2e_steps, "call": e2e_call, "last_call": e2e_split},
            "gpu_launches": launches,
            "clocks": clocks,
            "cpu_baseline": cpu,
            "parity": parity,
        }
        if e2e_torchrun is not None:
            line["e2e_torchrun"] = e2e_torchrun
        if c2_ms is not None:
            line["c2_ms_per_step"] = c2_ms
            line["c2_mpaths_per_s"] = WORKLOADS["C2"][1] * WORKLOADS["C2"][2] * WORKLOADS["C2"][3] / (c2_ms * 1e-3) * 1e-6
        emit(line)
    sharded.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


_json_out = None


def emit(line: dict) -> None:
    """The ONE stdout line of the contract."""
    out = _json_out if _json_out is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def _resolve_workload(name: str, host_only: bool) -> None:
    """G<half>: scene-size sweep workloads, registered on demand."""
    if name in WORKLOADS:
        return
    if name.startswith("G") and name[1:].isdigit():
        half = int(name[1:])
        scenes = host_only_scenes() if host_only else importlib.import_module(PKG + ".scenes")
        setattr(scenes, "_bench_" + name, lambda: scenes._grid_scene(half, 1000 + half, 0.80, 0.15, big=False))
        WORKLOADS[name] = ("_bench_" + name, 1920, 1080, 16, 50)
        return
    raise SystemExit(f"unknown workload {name}")


def main():
    # Libraries (NCCL prints its version banner) write to file descriptor 1: keep the real stdout for the JSON line
    # and send everything else to stderr.
    global _json_out
    sys.stdout.flush()
    _json_out = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="T", help="one of " + ", ".join(sorted(WORKLOADS)) + ", or G<half>: a (2*half)^2 + 1 sphere jittered grid "
                    "at 1920x1080 x 16 spp (scene-size sweeps).  Default T = the north_star target (1000 spp in total at every N)")
    ap.add_argument("--spp", type=int, default=0, help="override the workload's samples per pixel")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity object (development runs)")
    ap.add_argument("--parity-rows", default="", help="comma-separated scanlines for the oracle strips (default: H/8, H/2, H-4)")
    ap.add_argument("--full-reference", action="store_true", help="--impl reference: the whole configuration, no sample cut (minutes)")
    ap.add_argument("--stages", type=int, default=0, choices=[0, 1, 2, 3, 4], help="sweep variant: 0 = library default, 1 = one-stage filter, 2 = two-stage, 3 = two-stage with shared-memory path state, 4 = streamed two-stage (any scene size)")
    ap.add_argument("--sharding", default="samples", help="multi-GPU sharding: samples (default), rows, or RxS = R row groups x S sample groups (tile + sample)")
    ap.add_argument("--exchange", default="auto", choices=["auto", "p2p", "nccl"], help="multi-GPU framebuffer exchange")
    ap.add_argument("--barrier", default="flags", choices=["flags", "nccl"], help="p2p exchange: stream barriers by peer-memory flag kernels (default) or 4-byte NCCL all-reduces")
    ap.add_argument("--rays", type=int, default=0, choices=[0, 1, 2], help="trace kernel variant: rays per lane (0 = library default)")
    args = ap.parse_args()
    _resolve_workload(args.workload, host_only=(args.impl == "reference"))
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = max(args.warmup, 1)
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
