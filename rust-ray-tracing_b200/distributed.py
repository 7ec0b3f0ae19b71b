"""Multi-GPU rendering: one process per GPU, `torch.distributed` for the one exchange step.

The path shards naturally: paths are independent, so each rank traces a shard of the job into its own
float4 accumulation image and the partial images are summed onto rank 0 (NCCL reduce over NVLink), which resolves
to RGB8.  Primary sharding is by SAMPLE RANGE — rank g traces samples [g*spp/G, (g+1)*spp/G) of every pixel (perfect
balance; the Philox counter carries the global sample index, so the image does not depend on G beyond float
summation order).  Secondary is by ROWS — disjoint scanlines, zero elsewhere — the analogue of the reference's tile
queue (src/renderer.rs:254-266): row group g of R takes the INTERLEAVED scanlines g, g + R, g + 2R, ... (a static split
that is balanced where contiguous bands are not: sky rows are cheap).  mode="grid" combines the two (BASELINE
configs[2]: "tile + sample sharded"): the ranks form R row groups x S sample groups, rank g = (g // S, g % S) traces its
row group's scanlines and its sample group's samples, and the exchange sums the S partial images of each row group
("samples" is the grid 1 x G, "rows" is G x 1).

Two exchanges are implemented (ShardedRenderer(exchange=...)):

  "p2p"  (default where it works) — NVLink peer memory, no reduce pass: every rank's accumulation image is a plain
         cudaMalloc buffer exported with CUDA IPC; after the trace, rank g runs ONE kernel (rtb200_resolve_peers_device)
         that, for its slice of the pixels, loads the partial sums of ALL ranks straight from peer memory, adds them in
         rank order (deterministic), quantises, and stores the RGB8 slice directly into rank dst's image.  The two
         stream-ordered barriers a step needs (before: every trace has finished; after: every slice has landed) are
         peer-memory flag kernels too (rtb200_peer_barrier_device): a step makes NO torch.distributed / NCCL call
         (barrier="nccl" keeps the earlier form, two 4-byte all-reduces, as the baseline).
  "nccl" — reduce(SUM) of the float4 image onto dst, then the resolve kernel there (the baseline; also the fallback
         when IPC / peer access is unavailable).

torch is plumbing here: device memory, the stream, and the collective.  The tracer object does the work; the CUDA
tracer calls the C ABI (rtb200_trace_device / rtb200_resolve_device / rtb200_resolve_peers_device).
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

    def new_peer_exchange(self, dist, group, world, rank, height, width):
        return PeerExchange(self, dist, group, world, rank, height, width)

    def resolve_grid(self, accum_ptrs, row_groups, sample_groups, width, pixel_begin, pixel_end, spp_total, rgb8_ptr, sum_ptr=0):
        stream = self.torch.cuda.current_stream(self.device).cuda_stream
        self.r.resolve_grid_sum_device(accum_ptrs, row_groups, sample_groups, width, pixel_begin, pixel_end, spp_total, rgb8_ptr, sum_ptr, stream)

    def peer_barrier(self, flag_ptrs, rank, epoch):
        stream = self.torch.cuda.current_stream(self.device).cuda_stream
        self.r.peer_barrier_device(flag_ptrs, rank, epoch, stream)

    def resolve_peers(self, accum_ptrs, pixel_begin, pixel_end, spp_total, rgb8_ptr):
        stream = self.torch.cuda.current_stream(self.device).cuda_stream
        self.r.resolve_peers_device(accum_ptrs, pixel_begin, pixel_end, spp_total, rgb8_ptr, stream)

    def resolve(self, accum, spp_total):
        n = accum.shape[0] * accum.shape[1]
        out = self.torch.empty((accum.shape[0], accum.shape[1], 3), dtype=self.torch.uint8, device=self.device)
        stream = self.torch.cuda.current_stream(self.device).cuda_stream
        self.r.resolve_device(accum.data_ptr(), n, spp_total, out.data_ptr(), stream)
        return out


class DeviceBuffer:
    """A raw device allocation (rtb200_device_alloc) that torch can wrap without copying (__cuda_array_interface__)."""

    def __init__(self, ptr: int, shape, typestr: str, owner=None):
        self.ptr = int(ptr)
        self.shape = tuple(shape)
        self.typestr = typestr
        self.owner = owner

    def data_ptr(self) -> int:
        return self.ptr

    @property
    def __cuda_array_interface__(self):
        return {"shape": self.shape, "typestr": self.typestr, "data": (self.ptr, False), "version": 2, "strides": None}


class PeerExchange:
    """The NVLink peer-memory exchange state of one rank for one image size: its own accumulation image and RGB8 image
    (plain device allocations), and every peer's, opened through CUDA IPC."""

    def __init__(self, tracer, dist, group, world: int, rank: int, height: int, width: int):
        import ctypes as C
        torch = tracer.torch
        self.tracer, self.world, self.rank, self.H, self.W = tracer, world, rank, height, width
        r = tracer.r
        n_pix = height * width
        self.dist, self.group = dist, group
        self.accum = DeviceBuffer(r.device_alloc(n_pix * 16), (height, width, 4), "<f4")
        self.rgb8 = DeviceBuffer(r.device_alloc(n_pix * 3 + 16), (height, width, 3), "|u1")
        self.sum = DeviceBuffer(r.device_alloc(n_pix * 16), (height, width, 4), "<f4")       # the summed linear image (on request; dst's is used)
        self.flags = DeviceBuffer(r.device_alloc(64), (16,), "<u4")                           # peer-barrier slots (zero-initialised)
        self.epoch = 0
        own = (self.accum, self.rgb8, self.sum, self.flags)
        mine = torch.tensor([x for b in own for x in r.ipc_export(b.ptr)], dtype=torch.uint8, device=tracer.device)
        every = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(every, mine, group=group)
        self.peer_accum, self.peer_rgb8, self.peer_sum, self.peer_flags, self._opened = [], [], [], [], []
        lists = (self.peer_accum, self.peer_rgb8, self.peer_sum, self.peer_flags)
        ok = 1
        for g, h in enumerate(every):
            if g == rank:
                for lst, b in zip(lists, own):
                    lst.append(b.ptr)
                continue
            hb = bytes(h.cpu().tolist())
            for k, lst in enumerate(lists):
                try:
                    ptr = r.ipc_open(hb[64 * k: 64 * (k + 1)])
                    self._opened.append(ptr)
                    lst.append(ptr)
                except Exception:
                    ok = 0
                    lst.append(0)
        flag = torch.tensor([ok], dtype=torch.int32, device=tracer.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        self.usable = bool(flag.item())

    def barrier(self):
        """Stream-ordered barrier over peer-memory flags (no NCCL call): see rtb200_peer_barrier_device."""
        self.epoch += 1
        self.tracer.peer_barrier(self.peer_flags, self.rank, self.epoch)

    def close(self):
        """Collective: every rank closes its mappings of the peers' buffers, THEN (after a barrier and a device sync) frees
        its own exported buffers — freeing an exported allocation while an importer still has it mapped is undefined."""
        for p in self._opened:
            try:
                self.tracer.r.ipc_close(p)
            except Exception:
                pass
        self._opened = []
        try:
            self.tracer.torch.cuda.synchronize(self.tracer.device)
            if self.dist.is_initialized():
                self.dist.barrier(group=self.group)
        except Exception:
            pass
        for b in (self.accum, self.rgb8, self.sum, self.flags):
            if b is not None and b.ptr:
                self.tracer.r.device_free(b.ptr)
                b.ptr = 0


class ShardedRenderer:
    """Renderer::render (src/renderer.rs:38-40) across the ranks of a torch.distributed group.

    Every rank calls render(); rank `dst` gets the RGB8 image, the others get None.
    """

    def __init__(self, tracer, group=None, mode: str = "samples", dst: int = 0, exchange: str = "auto", grid=None, barrier: str = "flags"):
        import torch.distributed as dist
        if mode not in ("samples", "rows", "grid"):
            raise ValueError("mode must be 'samples', 'rows' or 'grid'")
        if exchange not in ("auto", "p2p", "nccl"):
            raise ValueError("exchange must be 'auto', 'p2p' or 'nccl'")
        if barrier not in ("flags", "nccl"):
            raise ValueError("barrier must be 'flags' (peer-memory flag kernel) or 'nccl' (4-byte all-reduce)")
        self.exchange = exchange
        self.barrier = barrier
        self.want_sum = False     # p2p exchange: also form the summed linear image on dst (last_sum(); costs 33 MB of peer stores at 1080p)
        self._peer = None
        self.on_trace = None      # optional callback(phase): phase 0 just before, 1 just after the trace kernel is queued (timing hooks)
        self.dist = dist
        self.tracer = tracer
        self.group = group
        self.mode = mode
        self.dst = dst
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self._accum = None
        # R row groups x S sample groups; rank = row_group * S + sample_group
        if mode == "samples":
            self.grid = (1, self.world)
        elif mode == "rows":
            self.grid = (self.world, 1)
        else:
            if grid is None:
                raise ValueError("mode='grid' needs grid=(row_groups, sample_groups)")
            self.grid = (int(grid[0]), int(grid[1]))
            if self.grid[0] < 1 or self.grid[1] < 1 or self.grid[0] * self.grid[1] != self.world:
                raise ValueError(f"grid {self.grid} does not cover {self.world} ranks")
        self.row_group, self.sample_group = divmod(self.rank, self.grid[1])

    def shard(self, spp_total: int, height: int):
        """(sample range, row range) of this rank; row range = (first, end, stride): row group g of R traces the
        interleaved scanlines g, g + R, ... (rtb200_params.row_stride)."""
        R, S = self.grid
        return shard_range(spp_total, S, self.sample_group), (min(self.row_group, height), height, R)

    # ---- the NVLink peer-memory exchange ------------------------------------------------------------------
    def _peer_exchange(self, camera):
        """The PeerExchange for this image size, or None when the exchange is "nccl", the world is 1, the tracer has
        no peer support (CPU test double), or IPC / peer access failed on some rank (then every rank falls back)."""
        if self.world == 1 or self.exchange == "nccl" or not hasattr(self.tracer, "resolve_peers"):
            return None
        H, W = camera.image_height(), camera.image_width()
        if self._peer is not None and (self._peer.H, self._peer.W) != (H, W):
            self._peer.close()
            self._peer = None
        if self._peer is None:
            self._peer = self.tracer.new_peer_exchange(self.dist, self.group, self.world, self.rank, H, W)
            if not self._peer.usable:
                if self.exchange == "p2p":
                    raise RuntimeError("exchange='p2p' requested but CUDA IPC / peer access is not available on every rank")
                self._peer.close()
                self.exchange = "nccl"
                self._peer = None
                return None
            self._flag = self.tracer.torch.zeros(1, dtype=self.tracer.torch.float32, device=self.tracer.device)
        return self._peer

    def _trace(self, accum, max_bounces, spp_total, camera, sample_range, row_range):
        if self.on_trace:
            self.on_trace(0)
        self.tracer.trace(accum, max_bounces, spp_total, camera, sample_range, row_range)
        if self.on_trace:
            self.on_trace(1)

    def exchange_in_use(self) -> str:
        return "p2p" if self._peer is not None else ("none" if self.world == 1 else "nccl")

    def _stream_barrier(self):
        """Stream-ordered barrier: the kernels queued after it start only when every rank has finished what it queued before
        its own.  Default: the peer-memory flag kernel; barrier="nccl": a 4-byte all-reduce."""
        if self.barrier == "flags" and self._peer is not None and hasattr(self.tracer, "peer_barrier"):
            self._peer.barrier()
        else:
            self.dist.all_reduce(self._flag, group=self.group)

    def last_sum(self):
        """p2p exchange with want_sum: the summed linear image [H,W,4] (device tensor) of the last render on dst, else None."""
        if self._peer is None or not self.want_sum or self.rank != self.dst:
            return None
        return self.tracer.torch.as_tensor(self._peer.sum, device=self.tracer.device)

    def close(self):
        """Collective (every rank calls it): releases the peer-exchange buffers and mappings."""
        if self._peer is not None:
            self._peer.close()
            self._peer = None
        self._accum = None

    def __del__(self):
        # not collective-safe in general; only drops what needs no peer (close() is the orderly way)
        self._accum = None

    def _render_device_p2p(self, px, max_bounces: int, spp_total: int, camera):
        H, W = px.H, px.W
        sample_range, row_range = self.shard(spp_total, H)
        self._trace(px.accum, max_bounces, spp_total, camera, sample_range, row_range)
        self._stream_barrier()                                   # every rank's partial image is complete
        # Every rank reduces + resolves one contiguous slice of the image: a pixel of image row r sums the partial images
        # of the S ranks of row group r % R (in rank order: deterministic), straight into dst's image.
        R, S = self.grid
        p0, p1 = shard_range(H * W, self.world, self.rank)
        p0, p1 = (p0 + 3) // 4 * 4 if self.rank else 0, (p1 + 3) // 4 * 4 if self.rank + 1 < self.world else H * W
        self.tracer.resolve_grid(px.peer_accum, R, S, W, min(p0, H * W), min(p1, H * W), spp_total, px.peer_rgb8[self.dst],
                                 px.peer_sum[self.dst] if self.want_sum else 0)
        self._stream_barrier()                                   # every slice has landed in dst's image
        if self.rank == self.dst:
            return self.tracer.torch.as_tensor(px.rgb8, device=self.tracer.device)
        return None

    def render_device(self, max_bounces: int, spp_total: int, camera):
        """Trace the local shard, exchange, resolve.  Returns the device RGB8 tensor on dst (None elsewhere).
        Asynchronous with respect to the host (everything is queued on the current stream / NCCL stream)."""
        px = self._peer_exchange(camera)
        if px is not None:
            return self._render_device_p2p(px, max_bounces, spp_total, camera)
        H = camera.image_height()
        if self._accum is None or tuple(self._accum.shape[:2]) != (H, camera.image_width()):
            self._accum = self.tracer.new_accum(camera)
        accum = self._accum
        sample_range, row_range = self.shard(spp_total, H)
        if self.grid[0] > 1:
            accum.zero_()     # rows outside the shard must contribute zero to the sum
        self._trace(accum, max_bounces, spp_total, camera, sample_range, row_range)
        if self.world > 1:
            self.dist.reduce(accum, dst=self.dst, op=self.dist.ReduceOp.SUM, group=self.group)
        if self.rank == self.dst:
            return self.tracer.resolve(accum, spp_total)
        return None

    def render(self, max_bounces: int, spp_total: int, camera, out=None):
        """Host-facing call: returns the [H,W,3] uint8 numpy image on dst, None elsewhere.  `out`: optional host tensor
        (e.g. torch.empty(..., pin_memory=True)) that receives the image instead of a fresh array."""
        img = self.render_device(max_bounces, spp_total, camera)
        if img is None:
            return None
        if out is not None:
            out.copy_(img, non_blocking=True)
            self.tracer.torch.cuda.current_stream(self.tracer.device).synchronize()
            return out.numpy()
        return img.cpu().numpy()
