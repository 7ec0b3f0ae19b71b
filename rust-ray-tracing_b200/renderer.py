"""Renderer — Python mirror of the reference's L2 boundary (src/renderer.rs).

`Renderer.render(max_bounces, samples_per_pixel, scene, camera) -> (image, RenderStat)` is the reference's trait
(src/renderer.rs:38-40); `CudaRenderer` is the B200 implementation that stands where `TileRenderer` stands
(:232-240, :244-387).  Every call goes through the C ABI in include/rtb200.h; nothing here computes pixels.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np

from . import _abi
from .scene import Camera, Scene


@dataclass
class RenderStat:
    """RenderStat, src/renderer.rs:11-34, plus the counters the throughput metric needs."""
    duration: float
    pixels_rendered: int
    pixels_per_second: float
    paths: int = 0
    segments: int = 0
    sphere_tests: int = 0
    trace_ms: float = 0.0
    resolve_ms: float = 0.0
    d2h_ms: float = 0.0


class Renderer:
    """trait Renderer, src/renderer.rs:38-40."""

    def render(self, max_bounces: int, samples_per_pixel: int, scene: Scene, camera: Camera) -> Tuple[np.ndarray, RenderStat]:
        raise NotImplementedError


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class CudaRenderer(Renderer):
    """The sm_100a backend behind the reference's Renderer boundary.

    `TileRenderer::new(num_threads, block_size)` (src/renderer.rs:232-240) picks threads and tile size; here the
    equivalents are the CUDA device and the Philox seed of the sample stream.
    """

    def __init__(self, device: int = 0, seed: int = 1, rays_per_lane: int = 0, stages: int = 0):
        """rays_per_lane: 0 = the library's default trace kernel, 1 / 2 = force that variant (RTB200_FLAG_RAYS1/2).
        stages: 0 = the library decides, 1 / 2 = force the one- / two-stage sweep (RTB200_FLAG_ONE_STAGE / _TWO_STAGE),
        3 = two-stage with the path state in shared memory (adds RTB200_FLAG_SHARED_STATE; matters beyond 2560 spheres),
        4 = the streamed two-stage kernel (RTB200_FLAG_STREAMED: what scenes beyond ~10 600 spheres get)."""
        self.rays_per_lane = int(rays_per_lane)
        self.stages = int(stages)
        self._ctx = C.c_void_p()
        _abi.check(_abi.lib.rtb200_create(int(device), C.byref(self._ctx)))
        self.device = int(device)
        self.seed = int(seed)
        self._scene = None        # the uploaded Scene object itself (kept alive: identity, not id(), decides re-uploads)
        self._keep = None
        self._pinned = []

    def pinned_empty(self, shape, dtype) -> np.ndarray:
        """A numpy array over page-locked host memory (rtb200_host_alloc) for `render(..., out=...)`: the device->host
        copy of the image is then a single DMA.  The memory lives until close()."""
        dt = np.dtype(dtype)
        n = int(np.prod(shape)) * dt.itemsize
        p = C.c_void_p()
        _abi.check(_abi.lib.rtb200_host_alloc(n, C.byref(p)))
        self._pinned.append(p)
        buf = (C.c_uint8 * max(n, 1)).from_address(p.value)
        return np.frombuffer(buf, dtype=dt, count=int(np.prod(shape))).reshape(shape)

    def close(self) -> None:
        for p in getattr(self, "_pinned", []):
            _abi.lib.rtb200_host_free(p)
        self._pinned = []
        if getattr(self, "_ctx", None) is not None and self._ctx:
            _abi.lib.rtb200_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- facts -------------------------------------------------------------------------------------------
    def device_info(self) -> dict:
        sm, khz, mj, mn = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
        _abi.check(_abi.lib.rtb200_device_info(self._ctx, C.byref(sm), C.byref(khz), C.byref(mj), C.byref(mn)))
        return {"sm_count": sm.value, "sm_clock_khz": khz.value, "cc": (mj.value, mn.value)}

    def launch_count(self) -> int:
        return int(_abi.lib.rtb200_launch_count(self._ctx))

    def fp32_peak(self, iters: int = 2000) -> Tuple[float, float]:
        """(TFLOP/s, ms) of the K0 FFMA probe."""
        tf, ms = C.c_double(), C.c_double()
        _abi.check(_abi.lib.rtb200_fp32_peak(self._ctx, int(iters), C.byref(tf), C.byref(ms)))
        return tf.value, ms.value

    def filter_check(self, scene: Scene, n_rays: int, per_ray: int = 8, seed: int = 1, camera_point=None) -> dict:
        """rtb200_filter_check: brute-forces n_rays * per_ray adversarial (ray, sphere) pairs of `scene` on the device —
        the sweep's two conservative FP32 filters against the exact f64 discriminant, all in the arithmetic the render
        kernels run.  false_negatives / slab_false_negatives must be 0."""
        if self._scene is not scene:
            self.upload_scene(scene)
        out = (C.c_uint64 * 9)()
        slack = C.c_double()
        cam = (C.c_double * 3)(*camera_point) if camera_point is not None else None
        _abi.check(_abi.lib.rtb200_filter_check(self._ctx, int(n_rays), int(per_ray), int(seed) & 0xFFFFFFFF, cam, out, C.byref(slack)))
        return {"pairs": int(out[0]), "exact_positive": int(out[1]), "filter_candidates": int(out[2]), "false_negatives": int(out[3]),
                "slab_candidates": [int(out[4]), int(out[5]), int(out[6])], "slab_false_negatives": int(out[7]), "min_slack": slack.value}

    def set_progress(self, callback=None, interval_ms: int = 100) -> None:
        """callback(items_claimed, items_total, elapsed_seconds) is called from render() every interval_ms while the
        trace kernel runs and once when it has finished (the reference's progress grid, src/renderer.rs:300-365, as a
        callback).  None switches it off."""
        if callback is None:
            self._progress_c = None
            _abi.check(_abi.lib.rtb200_set_progress(self._ctx, None, None, int(interval_ms)))
            return
        proto = C.CFUNCTYPE(None, C.c_void_p, C.c_uint64, C.c_uint64, C.c_double)
        self._progress_c = proto(lambda user, done, total, secs: callback(int(done), int(total), float(secs)))
        _abi.check(_abi.lib.rtb200_set_progress(self._ctx, C.cast(self._progress_c, C.c_void_p), None, int(interval_ms)))

    def set_hit_mode(self, both_roots: bool) -> None:
        """False (default): the reference's packed path — near root only (src/objects.rs:272-276).
        True: the scalar Sphere::hit rule (src/objects.rs:227-234) — the far root when the near one is out of range,
        i.e. rays inside a glass sphere hit its far wall.  Not what the reference's binary renders."""
        _abi.check(_abi.lib.rtb200_set_hit_mode(self._ctx, 1 if both_roots else 0))

    # ---- scene ---------------------------------------------------------------------------------------------
    def upload_scene(self, scene: Scene) -> None:
        sc, keep = scene.to_abi()
        _abi.check(_abi.lib.rtb200_upload_scene(self._ctx, C.byref(sc)))
        self._scene = scene
        self._keep = keep
        self.n_spheres = len(scene)

    def _params(self, max_bounces, spp, camera, sample_range, row_range) -> _abi.Params:
        s0, s1 = sample_range if sample_range is not None else (0, spp)
        r0, r1, rs = (tuple(row_range) + (1,))[:3] if row_range is not None else (0, camera.image_height(), 1)
        flags = {0: 0, 1: 1, 2: 2}[self.rays_per_lane] | {0: 0, 1: 4, 2: 8, 3: 8 | 32, 4: 64}[self.stages]
        return _abi.Params(int(max_bounces), int(spp), self.seed & 0xFFFFFFFFFFFFFFFF, int(s0), int(s1), int(r0), int(r1), flags, int(rs))

    # ---- Renderer::render -------------------------------------------------------------------------------------
    def render(self, max_bounces: int, samples_per_pixel: int, scene: Scene, camera: Camera, *, want_linear: bool = False,
               sample_range=None, row_range=None, out: Optional[np.ndarray] = None):
        """Renderer::render (src/renderer.rs:38-40).  Returns (image[H,W,3] uint8, RenderStat); with want_linear=True
        returns (image, linear[H,W,4] float32 per-pixel SUM, RenderStat).  row_range: (first, end) or (first, end, stride) —
        stride R traces every R-th scanline from `first` (interleaved row shards).  `out`: optional [H,W,3] uint8 buffer to
        receive the image (e.g. from pinned_empty) instead of a fresh array."""
        if self._scene is not scene:
            self.upload_scene(scene)
        W, H = camera.image_width(), camera.image_height()
        if out is not None:
            if out.shape != (H, W, 3) or out.dtype != np.uint8 or not out.flags.c_contiguous:
                raise ValueError("out must be a C-contiguous [H,W,3] uint8 array")
            img = out
        else:
            img = np.empty((H, W, 3), dtype=np.uint8)
        lin = np.empty((H, W, 4), dtype=np.float32) if want_linear else None
        st = _abi.Stats()
        p = self._params(max_bounces, samples_per_pixel, camera, sample_range, row_range)
        _abi.check(_abi.lib.rtb200_render(self._ctx, C.byref(camera.abi), C.byref(p), _ptr(img), _ptr(lin), C.byref(st)))
        stat = RenderStat(st.seconds, st.pixels_rendered, st.pixels_per_second, st.paths, st.segments, st.sphere_tests,
                          st.trace_ms, st.resolve_ms, st.d2h_ms)
        return (img, lin, stat) if want_linear else (img, stat)

    # ---- the sweep alone (Object::hit_packed over the scene) --------------------------------------------------
    def intersect_batch(self, scene: Scene, origins: np.ndarray, directions: np.ndarray, enabled: Optional[np.ndarray] = None):
        """Closest near-root hit per ray.  origins/directions: [n,3] float32.  Returns (index[n] int32, t[n] float32)."""
        if self._scene is not scene:
            self.upload_scene(scene)
        o = np.ascontiguousarray(np.asarray(origins, dtype=np.float32).reshape(-1, 3).T)
        d = np.ascontiguousarray(np.asarray(directions, dtype=np.float32).reshape(-1, 3).T)
        n = o.shape[1]
        if d.shape[1] != n:
            raise ValueError("origins and directions differ in length")
        en = None
        if enabled is not None:
            en = np.ascontiguousarray(np.asarray(enabled).astype(np.uint8))
            if en.shape != (n,):
                raise ValueError("enabled mask has the wrong length")
        idx = np.empty(n, dtype=np.int32)
        t = np.empty(n, dtype=np.float32)
        _abi.check(_abi.lib.rtb200_intersect_batch(self._ctx, _ptr(o[0]), _ptr(o[1]), _ptr(o[2]), _ptr(d[0]), _ptr(d[1]), _ptr(d[2]),
                                                   _ptr(en), n, _ptr(idx), _ptr(t)))
        return idx, t

    def intersect_batch_f64(self, scene: Scene, origins: np.ndarray, directions: np.ndarray, enabled: Optional[np.ndarray] = None):
        """The same routine at the reference's precision (rtb200_intersect_batch_f64): f64 rays against the scene's f64
        spheres, every pair.  origins/directions: [n,3] float64.  Returns (index[n] int32, t[n] float64)."""
        if self._scene is not scene:
            self.upload_scene(scene)
        o = np.ascontiguousarray(np.asarray(origins, dtype=np.float64).reshape(-1, 3).T)
        d = np.ascontiguousarray(np.asarray(directions, dtype=np.float64).reshape(-1, 3).T)
        n = o.shape[1]
        if d.shape[1] != n:
            raise ValueError("origins and directions differ in length")
        en = None
        if enabled is not None:
            en = np.ascontiguousarray(np.asarray(enabled).astype(np.uint8))
            if en.shape != (n,):
                raise ValueError("enabled mask has the wrong length")
        idx = np.empty(n, dtype=np.int32)
        t = np.empty(n, dtype=np.float64)
        _abi.check(_abi.lib.rtb200_intersect_batch_f64(self._ctx, _ptr(o[0]), _ptr(o[1]), _ptr(o[2]), _ptr(d[0]), _ptr(d[1]), _ptr(d[2]),
                                                       _ptr(en), n, _ptr(idx), _ptr(t)))
        return idx, t

    # ---- device-resident entry points (torch owns the memory and the stream) ----------------------------------
    def trace_device(self, max_bounces: int, samples_per_pixel: int, camera: Camera, accum_ptr: int, stream: int = 0,
                     sample_range=None, row_range=None) -> None:
        p = self._params(max_bounces, samples_per_pixel, camera, sample_range, row_range)
        _abi.check(_abi.lib.rtb200_trace_device(self._ctx, C.byref(camera.abi), C.byref(p), C.c_void_p(accum_ptr), C.c_void_p(stream)))

    def trace_device_static(self, max_bounces: int, samples_per_pixel: int, camera: Camera, accum_ptr: int, stream: int = 0,
                            sample_range=None, row_range=None) -> int:
        """rtb200_trace_device_static: the same trace step with its scratch in the library's file-scope device array — no
        allocation inside the call.  Returns the bytes of the array the call used; raises Rtb200Error (ERR_INVALID) when
        the shard's partial planes exceed RTB200_STATIC_SCRATCH_BYTES."""
        p = self._params(max_bounces, samples_per_pixel, camera, sample_range, row_range)
        used = C.c_uint64(0)
        _abi.check(_abi.lib.rtb200_trace_device_static(self._ctx, C.byref(camera.abi), C.byref(p), C.c_void_p(accum_ptr), C.c_void_p(stream),
                                                       C.byref(used)))
        return int(used.value)

    def resolve_device(self, accum_ptr: int, n_pixels: int, samples_per_pixel: int, rgb8_ptr: int, stream: int = 0) -> None:
        _abi.check(_abi.lib.rtb200_resolve_device(self._ctx, C.c_void_p(accum_ptr), int(n_pixels), int(samples_per_pixel),
                                                  C.c_void_p(rgb8_ptr), C.c_void_p(stream)))

    def intersect_batch_device(self, ptrs, n: int, enabled_ptr: int, idx_ptr: int, t_ptr: int, stream: int = 0) -> None:
        ox, oy, oz, dx, dy, dz = (C.c_void_p(p) for p in ptrs)
        _abi.check(_abi.lib.rtb200_intersect_batch_device(self._ctx, ox, oy, oz, dx, dy, dz, C.c_void_p(enabled_ptr) if enabled_ptr else None,
                                                          int(n), C.c_void_p(idx_ptr), C.c_void_p(t_ptr), C.c_void_p(stream)))

    # ---- multi-GPU exchange buffers (NVLink peer memory through CUDA IPC) -------------------------------------
    def device_alloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        _abi.check(_abi.lib.rtb200_device_alloc(self._ctx, int(nbytes), C.byref(p)))
        return int(p.value)

    def device_free(self, ptr: int) -> None:
        _abi.check(_abi.lib.rtb200_device_free(self._ctx, C.c_void_p(ptr)))

    def ipc_export(self, ptr: int) -> bytes:
        h = (C.c_uint8 * 64)()
        _abi.check(_abi.lib.rtb200_ipc_export(self._ctx, C.c_void_p(ptr), h))
        return bytes(h)

    def ipc_open(self, handle: bytes) -> int:
        h = (C.c_uint8 * 64)(*handle)
        p = C.c_void_p()
        _abi.check(_abi.lib.rtb200_ipc_open(self._ctx, h, C.byref(p)))
        return int(p.value)

    def ipc_close(self, ptr: int) -> None:
        _abi.check(_abi.lib.rtb200_ipc_close(self._ctx, C.c_void_p(ptr)))

    def resolve_peers_device(self, accum_ptrs, pixel_begin: int, pixel_end: int, samples_per_pixel: int, rgb8_ptr: int, stream: int = 0) -> None:
        """Fused reduce (sum over accum_ptrs, in order) + resolve of pixels [pixel_begin, pixel_end) into the whole-image
        RGB8 buffer at rgb8_ptr (local or peer)."""
        arr = (C.c_void_p * len(accum_ptrs))(*[C.c_void_p(p) for p in accum_ptrs])
        _abi.check(_abi.lib.rtb200_resolve_peers_device(self._ctx, arr, len(accum_ptrs), int(pixel_begin), int(pixel_end),
                                                        int(samples_per_pixel), C.c_void_p(rgb8_ptr), C.c_void_p(stream)))

    def resolve_grid_device(self, accum_ptrs, row_groups: int, sample_groups: int, image_width: int, pixel_begin: int, pixel_end: int,
                            samples_per_pixel: int, rgb8_ptr: int, stream: int = 0) -> None:
        """The same for a grid of row groups x sample groups: image row r sums the partial images
        accum_ptrs[(r % row_groups) * sample_groups + k], k < sample_groups (rtb200_resolve_grid_device)."""
        if len(accum_ptrs) != row_groups * sample_groups:
            raise ValueError("accum_ptrs must hold row_groups * sample_groups pointers")
        arr = (C.c_void_p * len(accum_ptrs))(*[C.c_void_p(p) for p in accum_ptrs])
        _abi.check(_abi.lib.rtb200_resolve_grid_device(self._ctx, arr, int(row_groups), int(sample_groups), int(image_width), int(pixel_begin),
                                                       int(pixel_end), int(samples_per_pixel), C.c_void_p(rgb8_ptr), C.c_void_p(stream)))

    def resolve_grid_sum_device(self, accum_ptrs, row_groups: int, sample_groups: int, image_width: int, pixel_begin: int, pixel_end: int,
                                samples_per_pixel: int, rgb8_ptr: int, sum_ptr: int, stream: int = 0) -> None:
        """resolve_grid_device that also stores the summed float4 pixels into the whole-image buffer at sum_ptr (0: not wanted)."""
        if len(accum_ptrs) != row_groups * sample_groups:
            raise ValueError("accum_ptrs must hold row_groups * sample_groups pointers")
        arr = (C.c_void_p * len(accum_ptrs))(*[C.c_void_p(p) for p in accum_ptrs])
        _abi.check(_abi.lib.rtb200_resolve_grid_sum_device(self._ctx, arr, int(row_groups), int(sample_groups), int(image_width), int(pixel_begin),
                                                           int(pixel_end), int(samples_per_pixel), C.c_void_p(rgb8_ptr),
                                                           C.c_void_p(sum_ptr) if sum_ptr else None, C.c_void_p(stream)))

    def peer_barrier_device(self, flag_ptrs, my_rank: int, epoch: int, stream: int = 0, timeout_s: float = 20.0) -> None:
        """rtb200_peer_barrier_device: stream-ordered barrier between the ranks through peer-memory flags."""
        arr = (C.c_void_p * len(flag_ptrs))(*[C.c_void_p(p) for p in flag_ptrs])
        _abi.check(_abi.lib.rtb200_peer_barrier_device(self._ctx, arr, len(flag_ptrs), int(my_rank), int(epoch) & 0xFFFFFFFF, float(timeout_s),
                                                       C.c_void_p(stream)))

    def trace_kernel_stats(self, n_last: int = 0) -> dict:
        """rtb200_trace_kernel_stats: the trace kernel alone — average ms of the last n_last launches (0: all recorded), the
        kernel variant of the last launch, and its K (sample sub-ranges per pixel, summed in order)."""
        ms, n, name, k = C.c_double(), C.c_int32(), C.c_char_p(), C.c_int32()
        _abi.check(_abi.lib.rtb200_trace_kernel_stats(self._ctx, int(n_last), C.byref(ms), C.byref(n), C.byref(name), C.byref(k)))
        return {"avg_ms": ms.value, "launches": n.value, "kernel": (name.value or b"").decode(), "sample_splits": k.value}

    def read_counters(self) -> Tuple[int, int]:
        p, s = C.c_uint64(), C.c_uint64()
        _abi.check(_abi.lib.rtb200_read_counters(self._ctx, C.byref(p), C.byref(s)))
        return p.value, s.value


class MultiGpuRenderer(Renderer):
    """Renderer::render over the GPUs of one box in ONE call (rtb200_multi, include/rtb200.h): the reference's `render`
    owns its worker threads (src/renderer.rs:272-296, :367-369); this one owns its devices — one host thread queues every
    shard's trace kernel, the fused reduce+resolve kernels read the partial sums over NVLink peer memory, and the image
    comes back in the caller's host buffer.  No torch, no NCCL, no other process.

    devices: device indices, one shard each (a device may repeat); None = every visible device.
    row_groups: 1 = sample-range sharding (default), len(devices) = interleaved-scanline sharding, in between = the
    row groups x sample groups grid (tile + sample)."""

    def __init__(self, devices=None, seed: int = 1, row_groups: int = 1, n_devices: Optional[int] = None):
        self._m = C.c_void_p()
        if devices is None:
            if n_devices is None:
                raise ValueError("give devices=[...] or n_devices=N")
            arr, n = None, int(n_devices)
        else:
            devices = [int(d) for d in devices]
            arr, n = (C.c_int32 * len(devices))(*devices), len(devices)
        _abi.check(_abi.lib.rtb200_multi_create(arr, n, C.byref(self._m)))
        self.n_shards = n
        self.seed = int(seed)
        self.row_groups = int(row_groups)
        self.flags = 0
        self._scene = None
        self._keep = None
        self._progress_c = None

    def close(self) -> None:
        if getattr(self, "_m", None) is not None and self._m:
            _abi.lib.rtb200_multi_destroy(self._m)
            self._m = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self) -> dict:
        n, peer = C.c_int32(), C.c_int32()
        _abi.check(_abi.lib.rtb200_multi_info(self._m, C.byref(n), C.byref(peer)))
        return {"n_shards": n.value, "peer_access": bool(peer.value)}

    def upload_scene(self, scene: Scene) -> None:
        sc, keep = scene.to_abi()
        _abi.check(_abi.lib.rtb200_multi_upload_scene(self._m, C.byref(sc)))
        self._scene, self._keep = scene, keep

    def set_hit_mode(self, both_roots: bool) -> None:
        _abi.check(_abi.lib.rtb200_multi_set_hit_mode(self._m, 1 if both_roots else 0))

    def set_progress(self, callback=None, interval_ms: int = 100) -> None:
        if callback is None:
            self._progress_c = None
            _abi.check(_abi.lib.rtb200_multi_set_progress(self._m, None, None, int(interval_ms)))
            return
        proto = C.CFUNCTYPE(None, C.c_void_p, C.c_uint64, C.c_uint64, C.c_double)
        self._progress_c = proto(lambda user, done, total, secs: callback(int(done), int(total), float(secs)))
        _abi.check(_abi.lib.rtb200_multi_set_progress(self._m, C.cast(self._progress_c, C.c_void_p), None, int(interval_ms)))

    def render(self, max_bounces: int, samples_per_pixel: int, scene: Scene, camera: Camera, *, want_linear: bool = False,
               sample_range=None, row_range=None, out: Optional[np.ndarray] = None, row_groups: Optional[int] = None):
        """Renderer::render (src/renderer.rs:38-40) on all shards.  Same returns as CudaRenderer.render; the linear image
        is the SUM over all shards as the reduce kernel formed it."""
        if self._scene is not scene:
            self.upload_scene(scene)
        W, H = camera.image_width(), camera.image_height()
        if out is not None:
            if out.shape != (H, W, 3) or out.dtype != np.uint8 or not out.flags.c_contiguous:
                raise ValueError("out must be a C-contiguous [H,W,3] uint8 array")
            img = out
        else:
            img = np.empty((H, W, 3), dtype=np.uint8)
        lin = np.empty((H, W, 4), dtype=np.float32) if want_linear else None
        s0, s1 = sample_range if sample_range is not None else (0, samples_per_pixel)
        r0, r1 = row_range if row_range is not None else (0, H)
        p = _abi.Params(int(max_bounces), int(samples_per_pixel), self.seed & 0xFFFFFFFFFFFFFFFF, int(s0), int(s1), int(r0), int(r1), int(self.flags), 1)
        st = _abi.Stats()
        _abi.check(_abi.lib.rtb200_multi_render(self._m, C.byref(camera.abi), C.byref(p), int(row_groups if row_groups is not None else self.row_groups),
                                                _ptr(img), _ptr(lin), C.byref(st)))
        stat = RenderStat(st.seconds, st.pixels_rendered, st.pixels_per_second, st.paths, st.segments, st.sphere_tests,
                          st.trace_ms, st.resolve_ms, st.d2h_ms)
        return (img, lin, stat) if want_linear else (img, stat)
