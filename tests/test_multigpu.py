"""Two ranks on two GPUs (NCCL): sample-range and row-range sharding with both framebuffer exchanges — the fused NVLink
peer-memory reduce+resolve kernel (CUDA IPC) and the NCCL reduce baseline — against the single-GPU render of the
whole job; four ranks on four GPUs: the 2 x 2 tile + sample grid (BASELINE configs[2]).  Skipped on a box with fewer
GPUs (run with `gpurun --gpus 2` / `--gpus 4`)."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H, SPP, DEPTH, SEED = 322, 181, 12, 50, 4      # odd sizes: pixel slices that are not multiples of 4


def _worker(rank, world, port, q, modes=(("samples", None), ("rows", None))):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        rtb = importlib.import_module("rust-ray-tracing_b200")
        dmod = importlib.import_module("rust-ray-tracing_b200.distributed")
        scene = rtb.scenes.random_spheres()
        cam = rtb.Camera.default(W, H)
        r = rtb.CudaRenderer(device=rank, seed=SEED)
        tracer = dmod.CudaTracer(r, scene)
        out, lin = {}, {}
        for mode, grid in modes:
            for exchange, barrier in (("p2p", "flags"), ("p2p", "nccl"), ("nccl", "nccl")):
                sr = dmod.ShardedRenderer(tracer, mode=mode, exchange=exchange, grid=grid, barrier=barrier)
                sr.want_sum = True                               # p2p: the summed linear image before quantisation
                a = sr.render(DEPTH, SPP, cam)
                b = sr.render(DEPTH, SPP, cam)
                assert sr.exchange_in_use() == exchange
                key = (mode if grid is None else "%dx%d" % grid, exchange if barrier == "nccl" else "p2p-flags")
                if rank == 0:
                    assert np.array_equal(a, b)                  # deterministic
                    out[key] = a
                    if exchange == "p2p":
                        lin[key] = sr.last_sum().cpu().numpy()
                else:
                    assert a is None and b is None
                sr.close()                                       # collective: mappings closed before exported buffers are freed
        if rank == 0:
            whole, whole_lin, _ = r.render(DEPTH, SPP, scene, cam, want_linear=True)
            q.put((out, whole, lin, whole_lin))
        dist.barrier()
        r.close()
    finally:
        dist.destroy_process_group()


def test_two_gpu_exchanges_match_single_gpu():
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out, whole, lin, whole_lin = q.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    # rows (interleaved scanlines): disjoint pixels, no summation -> identical; samples: two partial sums added -> float re-association only
    for ex in ("p2p-flags", "p2p", "nccl"):
        assert np.array_equal(out[("rows", ex)], whole)
        d = np.abs(out[("samples", ex)].astype(int) - whole.astype(int))
        assert d.max() <= 1 and (d > 0).mean() < 1e-3
    assert np.array_equal(out[("samples", "p2p-flags")], out[("samples", "p2p")])          # the barrier kind changes nothing
    assert np.abs(out[("samples", "p2p")].astype(int) - out[("samples", "nccl")].astype(int)).max() <= 1
    # the north_star's image bar at N > 1: the linear image BEFORE quantisation against the single-GPU one, 1e-4 on the mean
    for key, img in lin.items():
        assert np.abs(img[..., :3].astype(np.float64) - whole_lin[..., :3].astype(np.float64)).max() / SPP <= 1e-4, key
        assert np.all(img[..., 3] == SPP), key
    assert np.array_equal(lin[("rows", "p2p-flags")], whole_lin)


def test_four_gpu_tile_and_sample_grid_matches_single_gpu():
    """2 row groups x 2 sample groups, 4 x 1 and 1 x 4 on four GPUs, both exchanges."""
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 4:
        pytest.skip("needs four GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    modes = (("grid", (2, 2)), ("grid", (4, 1)), ("grid", (1, 4)))
    procs = [ctx.Process(target=_worker, args=(r, 4, port, q, modes)) for r in range(4)]
    for p in procs:
        p.start()
    out, whole, lin, whole_lin = q.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for key, img in lin.items():
        assert np.abs(img[..., :3].astype(np.float64) - whole_lin[..., :3].astype(np.float64)).max() / SPP <= 1e-4, key
    for ex in ("p2p-flags", "p2p", "nccl"):
        assert np.array_equal(out[("4x1", ex)], whole)           # disjoint rows: no summation
        for g in ("2x2", "1x4"):                                 # partial sums added: float re-association only
            d = np.abs(out[(g, ex)].astype(int) - whole.astype(int))
            assert d.max() <= 1 and (d > 0).mean() < 1e-3, (g, ex)
