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
        s, _, _ = self.oracle.render(self.arrays, camera.image_width(), camera.image_height(), camera.vectors(), spp_total, max_bounces,
                                     seed=SEED, sample_range=sample_range, row_range=row_range, f32=True, threads=2)
        r0, r1 = row_range
        accum[r0:r1, :, :3] = torch.from_numpy(s[r0:r1].astype(np.float32))
        accum[r0:r1, :, 3] = float(sample_range[1] - sample_range[0])

    def resolve(self, accum, spp_total):
        mean = accum[..., :3].numpy() / np.float32(spp_total)
        v = np.sqrt(mean) * np.float32(255.999)
        v = np.where(np.isnan(v), 0, np.clip(v, 0, 255))
        return torch.from_numpy(v.astype(np.uint8))


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
        sr = dmod.ShardedRenderer(OracleTracer(oracle_lib, scene), mode=mode)
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


@pytest.mark.parametrize("mode", ["samples", "rows"])
def test_sharded_render_world2_gloo(mode, rtb, oracle):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, mode, q)) for r in range(2)]
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
