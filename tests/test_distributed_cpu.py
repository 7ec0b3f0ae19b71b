"""The N>1 host path (ShardedRenderer: shard -> trace -> reduce -> resolve on rank 0) on CPU with the gloo backend,
world_size 2.  The tracer here is a test double built on the CPU oracle, so the sharding / collective logic is exercised
without a GPU; the product tracer (CudaTracer) plugs into the same interface."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H, SPP, DEPTH, SEED = 48, 27, 6, 20, 9


class OracleTracer:
    """Same interface as rust-ray-tracing_b200.distributed.CudaTracer, CPU tensors, oracle<float> arithmetic."""

    def __init__(self, oracle, scene):
        self.oracle = oracle
        self.arrays = scene.arrays()

    def new_accum(self, camera):
        return torch.zeros((camera.image_height(), camera.image_width(), 4), dtype=torch.float32)

    def trace(self, accum, max_bounces, spp_total, camera, sample_range, row_range):
        r0, r1, rs = (tuple(row_range) + (1,))[:3]
        s, _, _ = self.oracle.render(self.arrays, camera.image_width(), camera.image_height(), camera.vectors(), spp_total, max_bounces,
                                     seed=SEED, sample_range=sample_range, row_range=(r0, r1), f32=True, threads=2)
        accum[r0:r1:rs, :, :3] = torch.from_numpy(s[r0:r1:rs].astype(np.float32))     # the double traces the band, keeps its scanlines
        accum[r0:r1:rs, :, 3] = float(sample_range[1] - sample_range[0])

    def resolve(self, accum, spp_total):
        mean = accum[..., :3].numpy() / np.float32(spp_total)
        v = np.sqrt(mean) * np.float32(255.999)
        v = np.where(np.isnan(v), 0, np.clip(v, 0, 255))
        return torch.from_numpy(v.astype(np.uint8))


def _make_sharded(dmod, tracer, mode):
    """mode: "samples", "rows", or "RxS" (grid of R row groups x S sample groups)."""
    if "x" in mode:
        r, s_ = (int(v) for v in mode.split("x"))
        return dmod.ShardedRenderer(tracer, mode="grid", grid=(r, s_))
    return dmod.ShardedRenderer(tracer, mode=mode)


def _worker(rank, world, port, mode, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rtb = importlib.import_module("rust-ray-tracing_b200")
        dmod = importlib.import_module("rust-ray-tracing_b200.distributed")
        import oracle_lib
        scene = rtb.scenes.three_spheres()
        cam = rtb.Camera.default(W, H)
        sr = _make_sharded(dmod, OracleTracer(oracle_lib, scene), mode)
        assert sr.world == world and sr.rank == rank
        img = sr.render(DEPTH, SPP, cam)
        if rank == 0:
            q.put((mode, img, sr._accum.numpy().copy()))
        else:
            assert img is None
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("mode,world", [("samples", 2), ("rows", 2), ("1x2", 2), ("2x1", 2), ("2x2", 4)])
def test_sharded_render_gloo(mode, world, rtb, oracle):
    """world 2: sample-range and row-range sharding and the same two written as grids; world 4: the 2 x 2 tile + sample
    grid of BASELINE configs[2]."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, mode, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    _, img, accum = got
    scene = rtb.scenes.three_spheres()
    cam = rtb.Camera.default(W, H)
    ref_sum, ref_rgb, _ = oracle.render(scene.arrays(), W, H, cam.vectors(), SPP, DEPTH, seed=SEED, f32=True)
    assert np.all(accum[..., 3] == SPP)                                   # every pixel got all its samples exactly once
    assert np.abs(accum[..., :3] / SPP - ref_sum / SPP).max() <= 1e-4     # the north_star image tolerance
    assert np.abs(img.astype(int) - ref_rgb.astype(int)).max() <= 1       # quantised: at most one code value from re-association


def test_shard_range_partitions_exactly(rtb):
    dmod = importlib.import_module("rust-ray-tracing_b200.distributed")
    for total in (0, 1, 7, 100, 1000, 1080):
        for world in (1, 2, 3, 4, 8):
            parts = [dmod.shard_range(total, world, r) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == total
            assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        dmod.shard_range(10, 2, 2)


def test_grid_shards_cover_the_job_exactly_once(rtb):
    """Every (row, sample) of the job belongs to exactly one rank, for every factorisation of 8 ranks."""
    dmod = importlib.import_module("rust-ray-tracing_b200.distributed")

    class Stub(dmod.ShardedRenderer):
        def __init__(self, world, rank, grid):
            self.world, self.rank, self.grid = world, rank, grid
            self.row_group, self.sample_group = divmod(rank, grid[1])

    Hh, spp = 37, 11
    for R, S in [(1, 8), (2, 4), (4, 2), (8, 1)]:
        cover = np.zeros((Hh, spp), dtype=int)
        for rank in range(8):
            (s0, s1), (r0, r1, rs) = Stub(8, rank, (R, S)).shard(spp, Hh)
            cover[r0:r1:rs, s0:s1] += 1
        assert np.all(cover == 1), (R, S)
