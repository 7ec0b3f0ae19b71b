"""Multi-GPU rendering: one process per GPU, `torch.distributed` for the one exchange step.

The path shards naturally: paths are independent, so each rank traces a shard of the job into its own
float4 accumulation image and the partial images are summed onto rank 0 (NCCL reduce over NVLink), which resolves
to RGB8.  Primary sharding is by SAMPLE RANGE — rank g traces samples [g*spp/G, (g+1)*spp/G) of every pixel (perfect
balance; the Philox counter carries the global sample index, so the image does not depend on G beyond float
summation order).  Secondary is by ROW RANGE — disjoint rows, zero elsewhere — the analogue of the reference's tile
queue (src/renderer.rs:254-266).

torch is plumbing here: device memory, the stream, and the collective.  The tracer object does the work; the CUDA
tracer calls the C ABI (rtb200_trace_device / rtb200_resolve_device).
"""
from __future__ import annotations

from typing import Optional, Tuple


def shard_range(total: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous balanced split of [0,total) — shard sizes differ by at most one."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad world/rank")
    return (total * rank) // world, (total * (rank + 1)) // world


class CudaTracer:
    """Device-resident tracer over a CudaRenderer: buffers are torch CUDA tensors, kernels run on torch's current stream."""

    def __init__(self, renderer, scene):
        import torch
        self.torch = torch
        self.r = renderer
        self.scene = scene
        self.device = torch.device("cuda", renderer.device)
        renderer.upload_scene(scene)

    def upload(self):
        self.r.upload_scene(self.scene)

    def new_accum(self, camera):
        return self.torch.zeros((camera.image_height(), camera.image_width(), 4), dtype=self.torch.float32, device=self.device)

    def trace(self, accum, max_bounces, spp_total, camera, sample_range, row_range):
        stream = self.torch.cuda.current_stream(self.device).cuda_stream
        self.r.trace_device(max_bounces, spp_total, camera, accum.data_ptr(), stream, sample_range, row_range)

    def resolve(self, accum, spp_total):
        n = accum.shape[0] * accum.shape[1]
        out = self.torch.empty((accum.shape[0], accum.shape[1], 3), dtype=self.torch.uint8, device=self.device)
        stream = self.torch.cuda.current_stream(self.device).cuda_stream
        self.r.resolve_device(accum.data_ptr(), n, spp_total, out.data_ptr(), stream)
        return out


class ShardedRenderer:
    """Renderer::render (src/renderer.rs:38-40) across the ranks of a torch.distributed group.

    Every rank calls render(); rank `dst` gets the RGB8 image, the others get None.
    """

    def __init__(self, tracer, group=None, mode: str = "samples", dst: int = 0):
        import torch.distributed as dist
        if mode not in ("samples", "rows"):
            raise ValueError("mode must be 'samples' or 'rows'")
        self.dist = dist
        self.tracer = tracer
        self.group = group
        self.mode = mode
        self.dst = dst
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self._accum = None

    def shard(self, spp_total: int, height: int):
        if self.mode == "samples":
            return shard_range(spp_total, self.world, self.rank), (0, height)
        return (0, spp_total), shard_range(height, self.world, self.rank)

    def render_device(self, max_bounces: int, spp_total: int, camera):
        """Trace the local shard, reduce, resolve on dst.  Returns the device RGB8 tensor on dst (None elsewhere).
        Asynchronous with respect to the host (everything is queued on the current stream / NCCL stream)."""
        H = camera.image_height()
        if self._accum is None or tuple(self._accum.shape[:2]) != (H, camera.image_width()):
            self._accum = self.tracer.new_accum(camera)
        accum = self._accum
        sample_range, row_range = self.shard(spp_total, H)
        if self.mode == "rows":
            accum.zero_()     # rows outside the shard must contribute zero to the sum
        self.tracer.trace(accum, max_bounces, spp_total, camera, sample_range, row_range)
        if self.world > 1:
            self.dist.reduce(accum, dst=self.dst, op=self.dist.ReduceOp.SUM, group=self.group)
        if self.rank == self.dst:
            return self.tracer.resolve(accum, spp_total)
        return None

    def render(self, max_bounces: int, spp_total: int, camera):
        """Host-facing call: returns the [H,W,3] uint8 numpy image on dst, None elsewhere."""
        out = self.render_device(max_bounces, spp_total, camera)
        if out is None:
            return None
        return out.cpu().numpy()
