"""rtb200_trace_device_static — the trace step with its scratch in the library's file-scope `static __device__` array (no
cudaMalloc inside the call; include/rtb200.h) — against rtb200_trace_device, through the C ABI, on device buffers.

Collected last on purpose.  The test brings its own context, so it passes alone in a
fresh process and after the other GPU tests.  Tolerance: the one the image tests use (1e-4 per channel of the mean before
quantisation); the two entry points run the same kernels with the same sample splits, so the buffers are in fact
bit-identical and that is asserted too."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _trace_both(rtb, r, torch, scene, cam, spp, depth, **shard):
    W, H = cam.image_width(), cam.image_height()
    st = torch.cuda.current_stream().cuda_stream
    a = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda:0")
    b = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda:0")
    r.upload_scene(scene)
    r.trace_device(depth, spp, cam, a.data_ptr(), st, **shard)
    torch.cuda.synchronize()
    used = r.trace_device_static(depth, spp, cam, b.data_ptr(), st, **shard)
    torch.cuda.synchronize()
    return a, b, used


def _same(torch, r, a, b, spp):
    la, lb = a.cpu().numpy(), b.cpu().numpy()
    assert np.abs(la[..., :3] / spp - lb[..., :3] / spp).max() <= 1e-4
    assert np.array_equal(la, lb)                       # same kernels, same splits: bit-identical
    n = la.shape[0] * la.shape[1]
    ra = torch.zeros(n * 3 + 16, dtype=torch.uint8, device="cuda:0")
    rb = torch.zeros(n * 3 + 16, dtype=torch.uint8, device="cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    r.resolve_device(a.data_ptr(), n, spp, ra.data_ptr(), st)
    r.resolve_device(b.data_ptr(), n, spp, rb.data_ptr(), st)
    torch.cuda.synchronize()
    assert torch.equal(ra, rb)
    return la


def test_static_scratch_trace_matches_trace_device_twice_in_a_row(rtb):
    torch = pytest.importorskip("torch")
    r = rtb.CudaRenderer(device=0, seed=7)
    try:
        # call 1: BASELINE C1 (three spheres, 400x225, 16 spp): 16 sample splits -> 23 MB of partial planes in the array
        scene1 = rtb.scenes.three_spheres()
        cam1 = rtb.Camera.default(400, 225)
        a1, b1, used1 = _trace_both(rtb, r, torch, scene1, cam1, 16, 50)
        assert 256 < used1 <= 256 + rtb._abi.STATIC_SCRATCH_BYTES
        l1 = _same(torch, r, a1, b1, 16)
        assert np.all(l1[..., 3] == 16)
        # call 2, straight after, other scene / size / seed / shard: the array still holds call 1's planes and counters
        r.seed = 11
        scene2 = rtb.scenes.random_spheres()
        cam2 = rtb.Camera.default(160, 90)
        a2, b2, used2 = _trace_both(rtb, r, torch, scene2, cam2, 8, 50, sample_range=(3, 8), row_range=(10, 77))
        assert used2 >= 256
        l2 = _same(torch, r, a2, b2, 8)
        assert np.all(l2[10:77, :, 3] == 5) and np.all(l2[:10] == 0) and np.all(l2[77:] == 0)
        # call 3 repeats call 1 through the static entry point only, with no allocation visible to the driver: free device
        # memory is the same before and after (every kernel and the array's address were set up by the calls above)
        r.seed = 7
        r.upload_scene(scene1)
        c = torch.zeros_like(a1)
        torch.cuda.synchronize()
        free0, _ = torch.cuda.mem_get_info(0)
        r.trace_device_static(50, 16, cam1, c.data_ptr(), torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        free1, _ = torch.cuda.mem_get_info(0)
        assert free1 == free0, (free0, free1)
        assert torch.equal(c, a1)
    finally:
        r.close()


def test_static_scratch_rejects_oversize_shards(rtb):
    torch = pytest.importorskip("torch")
    r = rtb.CudaRenderer(device=0, seed=1)
    try:
        r.upload_scene(rtb.scenes.three_spheres())
        cam = rtb.Camera.default(1920, 1080)             # 16 splits x 2 M pixels x 16 B = 531 MB of partial planes
        buf = torch.zeros((8, 4), dtype=torch.float32, device="cuda:0")
        with pytest.raises(rtb._abi.Rtb200Error) as ei:
            r.trace_device_static(50, 100, cam, buf.data_ptr(), 0)
        assert ei.value.code == rtb._abi.ERR_INVALID
        assert "static scratch maximum" in ei.value.message and "33554432" in ei.value.message
        # a band of the same job that fits is accepted afterwards (<= 32 768 traced pixels always fit)
        full = torch.zeros((1080, 1920, 4), dtype=torch.float32, device="cuda:0")
        used = r.trace_device_static(50, 100, cam, full.data_ptr(), 0, row_range=(500, 516))
        torch.cuda.synchronize()
        assert used <= 256 + rtb._abi.STATIC_SCRATCH_BYTES
        ref = torch.zeros_like(full)
        r.trace_device(50, 100, cam, ref.data_ptr(), 0, row_range=(500, 516))
        torch.cuda.synchronize()
        assert torch.equal(full, ref)
        assert float(full[500:516, :, 3].min()) == 100.0 and float(full[:500].abs().max()) == 0.0
    finally:
        r.close()
