"""rtb200_multi: ONE synchronous render call that drives several shards — the drop-in shape of Renderer::render
(src/renderer.rs:38-40), which spawns and joins its own workers (:272-296, :367-369).  Single process, through ctypes ->
C ABI; no torch, no NCCL.

On a one-GPU box the shards share device 0 (devices=[0, 0, ...]): the sharding, the cross-stream event ordering and the
fused reduce+resolve kernel run exactly as on several GPUs, only the peer loads stay on one device — so these tests run
(and show up in the driver's launch list) wherever the GPU suite runs.  With 2 / 4 / 8 GPUs visible the same assertions
run over real NVLink peers (test_multi_real_devices)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W, H, SPP, DEPTH, SEED = 322, 181, 12, 50, 4      # odd sizes: pixel slices that are not multiples of 4


def _gpu_count():
    try:
        cudart = C.CDLL("libcudart.so")
    except OSError:
        import torch
        return torch.cuda.device_count()
    n = C.c_int(0)
    cudart.cudaGetDeviceCount(C.byref(n))
    return n.value


def _check(rtb, oracle, devices, row_groups, single_img, single_lin, single_stat, scene, cam, both_roots=False):
    m = rtb.MultiGpuRenderer(devices=devices, seed=SEED, row_groups=row_groups)
    try:
        assert m.info()["n_shards"] == len(devices)
        if both_roots:
            m.set_hit_mode(True)
        img, lin, stat = m.render(DEPTH, SPP, scene, cam, want_linear=True)
        img2, stat2 = m.render(DEPTH, SPP, scene, cam)
    finally:
        m.close()
    assert np.array_equal(img, img2)                                    # deterministic
    assert stat.paths == W * H * SPP == single_stat.paths and stat.segments == single_stat.segments
    assert np.all(lin[..., 3] == SPP)
    # the linear image BEFORE quantisation against the single-context render of the same job: 1e-4 on the mean (north_star)
    err = np.abs(lin[..., :3].astype(np.float64) - single_lin[..., :3].astype(np.float64)).max() / SPP
    assert err <= 1e-4, err
    if row_groups == len(devices):
        assert np.array_equal(lin, single_lin) and np.array_equal(img, single_img)   # row shards: no re-association at all
    else:
        assert np.abs(img.astype(int) - single_img.astype(int)).max() <= 1           # sample shards: the sum is re-associated
    return img, lin


@pytest.mark.parametrize("n_shards,row_groups", [(1, 1), (2, 1), (2, 2), (3, 1), (4, 2), (4, 4), (8, 1), (8, 2)])
def test_multi_one_call_matches_single_context(rtb, oracle, renderer, n_shards, row_groups):
    scene = rtb.scenes.random_spheres()
    cam = rtb.Camera.default(W, H)
    renderer.seed = SEED
    try:
        s_img, s_lin, s_stat = renderer.render(DEPTH, SPP, scene, cam, want_linear=True)
    finally:
        renderer.seed = 1
    img, lin = _check(rtb, oracle, [0] * n_shards, row_groups, s_img, s_lin, s_stat, scene, cam)
    # and against the oracle directly (RGB8 of three strips, linear to 1e-4)
    for r0 in (0, H // 2, H - 3):
        ref_sum, ref_rgb, _ = oracle.render(scene.arrays(), W, H, cam.vectors(), SPP, DEPTH, seed=SEED, f32=True, row_range=(r0, r0 + 3))
        assert np.abs(lin[r0:r0 + 3, :, :3].astype(np.float64) - ref_sum[r0:r0 + 3]).max() / SPP <= 1e-4
        assert np.abs(img[r0:r0 + 3].astype(int) - ref_rgb[r0:r0 + 3].astype(int)).max() <= (0 if row_groups == n_shards else 1)


def test_multi_real_devices(rtb, oracle, renderer):
    """2 / 4 / 8 visible GPUs: the same call over real peers (NVLink peer loads in k_resolve_peers, peer store of the RGB8
    slices into the root's image)."""
    n = _gpu_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs (run with gpurun --gpus 2/4/8)")
    scene = rtb.scenes.random_spheres()
    cam = rtb.Camera.default(W, H)
    renderer.seed = SEED
    try:
        s_img, s_lin, s_stat = renderer.render(DEPTH, SPP, scene, cam, want_linear=True)
    finally:
        renderer.seed = 1
    for k in (2, 4, 8):
        if k > n:
            break
        for rg in sorted({1, 2, k}):
            _check(rtb, oracle, list(range(k)), rg, s_img, s_lin, s_stat, scene, cam)
        m = rtb.MultiGpuRenderer(devices=list(range(k)))
        assert m.info()["peer_access"]
        m.close()


def test_multi_sample_subrange_rows_and_errors(rtb, renderer):
    scene = rtb.scenes.three_spheres()
    cam = rtb.Camera.default(97, 55)
    m = rtb.MultiGpuRenderer(devices=[0, 0, 0], seed=9)
    try:
        renderer.seed = 9
        a_img, a_lin, a_stat = renderer.render(20, 10, scene, cam, want_linear=True, sample_range=(3, 10), row_range=(5, 40))
        b_img, b_lin, b_stat = m.render(20, 10, scene, cam, want_linear=True, sample_range=(3, 10), row_range=(5, 40))
        assert b_stat.paths == a_stat.paths == 97 * 35 * 7 and b_stat.segments == a_stat.segments
        assert np.abs(b_lin[..., :3] - a_lin[..., :3]).max() <= 1e-4 * 10 and np.array_equal(b_lin[..., 3], a_lin[..., 3])
        assert np.all(b_lin[:5] == 0) and np.all(b_lin[40:] == 0) and np.all(b_img[:5] == 0)
        with pytest.raises(rtb.Rtb200Error):
            m.render(20, 10, scene, cam, row_groups=2)               # 2 does not divide 3 shards
        with pytest.raises(rtb.Rtb200Error):
            m.render(20, 0, scene, cam)
        img, stat = m.render(5, 2, scene, cam, row_groups=3)          # still usable
        assert img.shape == (55, 97, 3)
    finally:
        renderer.seed = 1
        m.close()
    with pytest.raises(rtb.Rtb200Error):
        rtb.MultiGpuRenderer(devices=[0, 99])                         # no such device
    with pytest.raises(rtb.Rtb200Error):
        rtb.MultiGpuRenderer(devices=[0] * 17)                        # more than RTB200_MAX_PEERS shards


def test_multi_both_roots_and_progress(rtb, oracle, renderer):
    scene = rtb.scenes.dielectric_heavy()
    cam = rtb.Camera.default(160, 90)
    renderer.seed = 2
    renderer.set_hit_mode(True)
    try:
        ref, _ = renderer.render(50, 8, scene, cam)
    finally:
        renderer.set_hit_mode(False)
        renderer.seed = 1
    m = rtb.MultiGpuRenderer(devices=[0, 0], seed=2, row_groups=2)
    calls = []
    try:
        m.set_hit_mode(True)
        m.set_progress(lambda done, total, secs: calls.append((done, total)), interval_ms=1)
        img, _ = m.render(50, 8, scene, cam)
    finally:
        m.close()
    assert np.array_equal(img, ref)                                   # row shards: identical
    assert calls and calls[-1][0] == calls[-1][1] > 0 and [c[0] for c in calls] == sorted(c[0] for c in calls)
