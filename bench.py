#!/usr/bin/env python
"""bench.py — headline benchmark of the render hot path (BASELINE.json: Mpaths/s and ray-sphere tests/s on B200,
as a fraction of the FP32 roofline, beside the host-CPU reference restatement).

  python bench.py --gpus N --steps K --warmup W            this repo's CUDA path (one rank per GPU under torchrun)
  python bench.py --impl reference --steps K --warmup W    the reference's CPU implementation of the path (port)

A "step" is one full render of the workload (trace + framebuffer reduce + resolve).  At N=1 the workload is BASELINE
config C2: random-spheres (~485 spheres) 1920x1080, 100 spp, depth 50.  At N>1 every rank traces 100 spp of the same
image (sample-range sharding: the job is a 1920x1080 x (100*N) spp render) and the partial framebuffers are summed
onto rank 0 with an NCCL reduce — weak scaling.

`value`  : Mpaths/s, whole job, scene/camera resident on the device, no host copies in the timed region.
`e2e`    : the same metric through the public Renderer.render call with HOST buffers: every step uploads the scene
           (host->device) and reads the RGB8 image back (device->host).
`roofline`: FP32 ALU bound (this path has no dense contraction and moves ~KBs of scene + one framebuffer):
           achieved = ray-sphere tests/s x 17 FLOP, peak = FP32 FMA throughput measured live by the K0
           probe on the same device (MEASURED_PEAKS.json holds only HBM and BF16 figures).
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

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "rust-ray-tracing_b200"

FLOP_PER_TEST = 17.0   # SURVEY §8d / DESIGN.md: 5 FMA + 7 add/mul per ray-sphere test

WORKLOADS = {
    # name: (scene generator, width, height, spp per GPU, depth)
    "C2": ("random_spheres", 1920, 1080, 100, 50),
    "C4": ("ten_k_spheres", 1920, 1080, 64, 50),
    "C5": ("dielectric_heavy", 1920, 1080, 256, 50),
    "C1": ("three_spheres", 400, 225, 16, 50),
    # beyond BASELINE: 40 001 spheres, the streamed trace kernel (stage-1 data fetched per warp from global memory)
    "L40K": ("forty_k_spheres", 1920, 1080, 16, 50),
    # fixed TOTAL sample counts, split over the GPUs (strong scaling): BASELINE configs[2] and the north_star target
    "C3": ("random_spheres", 3840, 2160, 1000, 50),
    "T": ("random_spheres", 1920, 1080, 1000, 50),
}
STRONG = {"C3", "T"}


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


def cpu_reference_run(workload: str, steps: int, warmup: int, budget_s: float = 12.0):
    """Times the CPU port of the reference's LIVE path (oracle/cpu_ref.cpp: TileRenderer + render_vectorized2 +
    trace_vectorized2 + hit_packed, 4 x f64 AVX2 lanes, 128x128 tiles, all host threads) on a bounded sample of the
    workload: same scene, camera and spp at a quarter of the resolution in each dimension (1/16 of the pixels), further
    reduced in spp if a calibration pass says one step would exceed `budget_s` seconds."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as oracle
    from importlib import import_module
    rtb_scenes = import_module(PKG + ".scenes")
    gen, W, H, spp, depth = WORKLOADS[workload]
    scene = getattr(rtb_scenes, gen)()
    arrays = scene.arrays()
    cores = os.cpu_count() or 1
    sw, sh = max(16, W // 4), max(9, H // 4)
    # camera through the oracle's own Camera::new (no GPU library involved on this arm)
    cam = oracle.camera_new(sw, sh, 10.0, 30.0, (16.0, 2.0, 18.5), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 0.0)
    _, t4, _, _ = oracle.cpuref_render(arrays, sw, sh, cam, 4, depth, threads=cores, want_rgb8=False)   # calibrate: 4 spp
    s_spp = max(4, min(spp, int(budget_s / max(t4 / 4.0, 1e-6)) // 4 * 4))
    times, segs, tests = [], 0, 0
    for i in range(warmup + steps):
        _, sec, tests, segs = oracle.cpuref_render(arrays, sw, sh, cam, s_spp, depth, threads=cores, want_rgb8=True)
        if i >= warmup:
            times.append(sec)
    paths = sw * sh * s_spp
    sec = sum(times) / len(times)
    return {
        "mpaths_s": paths / sec * 1e-6,
        "tests_s": tests / sec,
        "ms_per_step": sec * 1e3,
        "cores": cores,
        "sample": f"{gen} {sw}x{sh} (same camera, 1/16 of the pixels), {s_spp} spp, depth {depth}: {paths} paths/step; C++ port of "
                  f"TileRenderer + trace_vectorized2 + hit_packed (4 x f64 AVX2 lanes, 128x128 tiles) on {cores} threads",
        "n_spheres": len(scene),
    }


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    gen, W, H, spp, depth = WORKLOADS[args.workload]
    steps = max(1, args.steps)
    budget = 12.0 if steps * 12.0 <= 150 else max(2.0, 150.0 / (steps + args.warmup))
    res = cpu_reference_run(args.workload, steps, args.warmup, budget)
    line = {
        "impl": "reference",
        "metric": "Mpaths/s", "value": res["mpaths_s"], "unit": "Mpaths/s", "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup,
        "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": f"{args.workload}: {gen} {W}x{H}, {spp} spp, depth {depth} (bounded sample, see cpu_baseline.sample)",
                   "n_spheres": res["n_spheres"]},
        "sphere_tests_per_s": res["tests_s"],
        "cpu_baseline": {"value": res["mpaths_s"], "unit": "Mpaths/s", "cores": res["cores"], "kind": "port", "sample": res["sample"]},
        "e2e": {"value": res["mpaths_s"], "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


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
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    rtb = importlib.import_module(PKG)
    dmod = importlib.import_module(PKG + ".distributed")

    gen, W, H, spp_per_gpu, depth = WORKLOADS[args.workload]
    if args.spp:
        spp_per_gpu = args.spp
    strong = args.workload in STRONG and not args.spp
    if strong:
        if spp_per_gpu % world:
            raise SystemExit(f"{args.workload}: {spp_per_gpu} spp do not split over {world} GPUs")
        spp_per_gpu //= world                  # strong scaling: total work fixed
    spp_total = spp_per_gpu * world            # (weak scaling: per-GPU work fixed)
    scene = getattr(rtb.scenes, gen)()
    cam = rtb.Camera.default(W, H)
    renderer = rtb.CudaRenderer(device=local_rank, seed=1, rays_per_lane=args.rays, stages=args.stages)
    tracer = dmod.CudaTracer(renderer, scene)
    if "x" in args.sharding:
        gr, gs = (int(v) for v in args.sharding.split("x"))
        sharded = dmod.ShardedRenderer(tracer, mode="grid", grid=(gr, gs), exchange=args.exchange)
    else:
        sharded = dmod.ShardedRenderer(tracer, mode=args.sharding, exchange=args.exchange)
    info = renderer.device_info()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- FP32 roofline denominator, measured live on this device -----------------------------------------
    peak_tflops, _ = renderer.fp32_peak(2000)
    nominal_tflops = 2.0 * 128 * info["sm_count"] * info["sm_clock_khz"] * 1e3 * 1e-12

    # ---- device-resident timing: K steps of trace + reduce + resolve ------------------------------------------
    for _ in range(args.warmup):
        sharded.render_device(depth, spp_total, cam)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = renderer.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    trace_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    ev0.record()
    out = None
    step = [0]
    sharded.on_trace = lambda phase: trace_ev[step[0]][phase].record()   # the trace kernel alone, on its own stream (roofline)
    for i in range(args.steps):
        step[0] = i
        out = sharded.render_device(depth, spp_total, cam)               # trace + exchange + resolve
    sharded.on_trace = None
    ev1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = renderer.launch_count() - launches0
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    trace_ms = sum(a.elapsed_time(b) for a, b in trace_ev) / args.steps
    paths_local, segs_local = renderer.read_counters()       # counters of the last trace on this rank
    paths = sum_over_ranks(float(paths_local))
    segs = sum_over_ranks(float(segs_local))
    ms_per_step = ms_total / args.steps
    n_spheres = len(scene)
    mpaths = paths / (ms_per_step * 1e-3) * 1e-6
    tests_s = segs * n_spheres / (ms_per_step * 1e-3)

    # roofline of the dominant kernel (k_trace_paths) on this rank
    tests_per_launch = segs_local * n_spheres
    achieved_tflops = tests_per_launch * FLOP_PER_TEST / (trace_ms * 1e-3) * 1e-12

    # ---- end to end through Renderer.render with host buffers ------------------------------------------------
    e2e_steps = max(1, min(args.steps, 3))
    sc_abi, _keep = scene.to_abi()
    h2d = (n_spheres + 31) // 32 * 32 * (16 + 16 + 16 + 4 + 12 + 32)   # the scene arena: exact | filter | material | kind | stage-1 triples | pairs
    d2h = W * H * 3
    if world == 1:
        out_img = renderer.pinned_empty((H, W, 3), np.uint8)   # the caller's image buffer: pinned host memory
        renderer.render(depth, spp_total, scene, cam, out=out_img)      # warm-up (allocations)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            renderer.upload_scene(scene)                  # host -> device: the step's inputs
            img, stat = renderer.render(depth, spp_total, scene, cam, out=out_img)   # device -> host: RGB8 image
        e2e_s = (time.perf_counter() - t0) / e2e_steps
    else:
        host_img = torch.empty((H, W, 3), dtype=torch.uint8, pin_memory=True) if rank == 0 else None   # the caller's pinned image buffer
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            tracer.upload()
            img = sharded.render(depth, spp_total, cam, out=host_img)    # numpy image on rank 0
        torch.cuda.synchronize()
        e2e_s = (time.perf_counter() - t0) / e2e_steps
    e2e_s = max_over_ranks(e2e_s)
    e2e_mpaths = paths / e2e_s * 1e-6
    e2e_split = None
    if world == 1:
        e2e_split = {"trace_ms": stat.trace_ms, "resolve_ms": stat.resolve_ms, "d2h_ms": stat.d2h_ms, "call_ms": stat.duration * 1e3}

    # ---- CPU baseline (rank 0, N=1 only) -----------------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        res = cpu_reference_run(args.workload, 1, 0, 12.0)
        cpu = {"value": res["mpaths_s"], "unit": "Mpaths/s", "cores": res["cores"], "kind": "port", "sample": res["sample"],
               "sphere_tests_per_s": res["tests_s"]}

    traffic = None       # DRAM bytes per launch of the dominant kernel, from the committed ncu capture of this workload
    ncu_facts = None     # issue-slot / pipe utilisation of the same capture (the kernel is issue-bound, not FLOP-bound)
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f).get(args.workload)
        if t and not args.spp and world == 1:
            traffic = t["dram_bytes_per_launch"]
            ncu_facts = {k: t[k] for k in ("kernel", "issue_slots_busy_pct", "fma_pipe_pct", "alu_pipe_pct", "threads_per_instruction", "source") if k in t}
    except Exception:
        traffic = None

    if rank == 0:
        line = {
            "metric": "Mpaths/s", "value": mpaths, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": f"{args.workload}: {gen} {W}x{H}, {spp_per_gpu} spp per GPU ({spp_total} spp total), depth {depth}",
                       "n_spheres": n_spheres, "sharding": (("sample-range" if args.sharding == "samples" else "row-range" if args.sharding == "rows" else "grid " + args.sharding + " (row groups x sample groups)") + "; exchange = " + {"p2p": "fused NVLink peer-memory reduce+resolve kernel (CUDA IPC), two 4-byte all-reduces as barriers",
                                                                   "nccl": "NCCL reduce(sum) to rank 0 + resolve"}[sharded.exchange_in_use()])
                       if world > 1 else "single GPU",
                       "l2": "no reuse across steps: each step re-traces from the camera; inputs are KBs, the framebuffer is written once",
                       "seed": 1, "rays_per_lane": args.rays, "stages": args.stages},
            "sphere_tests_per_s": tests_s, "segments_per_path": segs / max(paths, 1.0),
            "roofline": {"bound": "fp32", "achieved": achieved_tflops, "peak": peak_tflops, "unit": "TFLOP/s",
                         "frac": achieved_tflops / peak_tflops, "traffic": traffic,
                         "kernel": "k_trace_paths", "kernel_ms": trace_ms, "flop_per_test": FLOP_PER_TEST,
                         "peak_source": "K0 FFMA probe, measured live on this device",
                         "nominal_peak": nominal_tflops, "frac_of_nominal": achieved_tflops / nominal_tflops,
                         "note": "achieved counts the reference's 17 FLOP per ray-sphere test (SURVEY 8d); the two-stage sweep executes fewer "
                                 "(3 FFMA per pair in stage 1), so the bound that binds is the issue slot: see ncu",
                         "ncu": ncu_facts},
            "e2e": {"value": e2e_mpaths, "unit": "Mpaths/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_s * 1e3, "last_call": e2e_split},
            "gpu_launches": launches,
            "clocks": clocks,
            "cpu_baseline": cpu,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


_json_out = None


def emit(line: dict) -> None:
    """The ONE stdout line of the contract."""
    out = _json_out if _json_out is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def _resolve_workload(name: str) -> None:
    """G<half>: scene-size sweep workloads, registered on demand."""
    if name in WORKLOADS:
        return
    if name.startswith("G") and name[1:].isdigit():
        half = int(name[1:])
        scenes = importlib.import_module(PKG + ".scenes")
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
    ap.add_argument("--workload", default="C2", help="one of " + ", ".join(sorted(WORKLOADS)) + ", or G<half>: a (2*half)^2 + 1 sphere jittered grid "
                    "at 1920x1080 x 16 spp (scene-size sweeps)")
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--stages", type=int, default=0, choices=[0, 1, 2, 3, 4], help="sweep variant: 0 = library default, 1 = one-stage filter, 2 = two-stage, 3 = two-stage with shared-memory path state, 4 = streamed two-stage (any scene size)")
    ap.add_argument("--sharding", default="samples", help="multi-GPU sharding: samples (default), rows, or RxS = R row groups x S sample groups (tile + sample)")
    ap.add_argument("--exchange", default="auto", choices=["auto", "p2p", "nccl"], help="multi-GPU framebuffer exchange")
    ap.add_argument("--rays", type=int, default=0, choices=[0, 1, 2], help="trace kernel variant: rays per lane (0 = library default)")
    args = ap.parse_args()
    _resolve_workload(args.workload)
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = max(args.warmup, 1)
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
