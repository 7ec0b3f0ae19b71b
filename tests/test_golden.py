"""The committed fixtures under tests/golden/ (written by tests/golden/make_golden.py from the CPU oracle; the reference
has no fixtures of its own for this path).

Two tags per fixture: "f64" = the reference-faithful double oracle, "mix" = the GPU comparand (FP32 path state, the
closest-hit sweep — Sphere::hit_packed + update — in f64 on the f64 scene, its root rounded once to FP32; rt_oracle.hpp).

CPU: the oracle built here reproduces them bit for bit (it cannot drift silently) and every host loads the TOML fixture
into the stored scene arrays.  GPU (-m gpu): the CUDA path, through the C ABI, reproduces the files written on another
machine — hit/miss and index EQUAL TO THE f64 ROUTINE'S, t within 1e-5 relative of the f64 routine's (BASELINE.json's bar;
it is the f64 root rounded once, 6e-8), images within 1e-4 per channel before quantisation of the mixed oracle (bit-equal up
to the order of partial sums), RGB8 identical."""
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RENDERS = ["three_spheres", "random_spheres", "dielectric_heavy", "depth3"]
GENERATOR = {"three_spheres": "three_spheres", "random_spheres": "random_spheres", "dielectric_heavy": "dielectric_heavy",
             "depth3": "random_spheres"}


def _load(name):
    return np.load(os.path.join(GOLDEN, name))


def _arrays(g):
    return tuple(g[k] for k in ("spheres", "mat_index", "kind", "hollow", "albedo", "param"))


@pytest.mark.parametrize("name", RENDERS)
def test_scene_generators_reproduce_the_stored_scenes(rtb, name):
    g = _load(f"render_{name}.npz")
    scene = getattr(rtb.scenes, GENERATOR[name])()
    for a, b in zip(scene.arrays(), _arrays(g)):
        assert np.array_equal(np.asarray(a), b)
    cam = rtb.Camera.default(int(g["W"]), int(g["H"]))
    assert np.array_equal(np.asarray(cam.vectors(), dtype=np.float64), g["cam18"])


@pytest.mark.parametrize("tag", ["mix", "f64"])
@pytest.mark.parametrize("name", RENDERS)
def test_oracle_reproduces_golden_renders(oracle, name, tag):
    g = _load(f"render_{name}.npz")
    W, H, spp, depth, seed = (int(g[k]) for k in ("W", "H", "spp", "depth", "seed"))
    s, rgb, seg = oracle.render(_arrays(g), W, H, g["cam18"], spp, depth, seed=seed, f32=(tag == "mix"))
    assert seg == int(g["segments_" + tag])
    assert np.array_equal(s, g["sum_" + tag])            # bit for bit, whatever the thread count
    assert np.array_equal(rgb, g["rgb8_" + tag])


@pytest.mark.parametrize("tag", ["mix", "f64"])
def test_oracle_reproduces_golden_intersections(oracle, tag):
    g = _load("intersect_random_spheres.npz")
    idx, t = oracle.intersect(g["spheres"], g["rays"], enabled=g["enabled"], f32=(tag == "mix"))
    assert np.array_equal(idx, g["idx_" + tag])
    assert np.array_equal(t, g["t_" + tag])
    assert np.all(idx[g["enabled"] == 0] == -1)           # a disabled lane never hits (src/objects.rs:259,277)


def _rel_vs_f64(t, g):
    hit = g["idx_f64"] >= 0
    return np.abs(np.asarray(t, dtype=np.float64)[hit] - g["t_f64"][hit]) / np.abs(g["t_f64"][hit])


def test_mixed_golden_meets_the_intersection_bar_against_f64():
    """BASELINE.json: 'a ray-batch intersection test must match the reference's packed-SIMD hit routine exactly on hit/miss
    and object index, with hit distance within 1e-5 relative error'.  The reference's routine is f64 (src/objects.rs:249-290);
    the GPU comparand runs it in f64 from FP32-valued rays, so on the stored batch (rays exactly representable in FP32):
    same index for every ray, t = the f64 root rounded once to FP32."""
    g = _load("intersect_random_spheres.npz")
    assert np.array_equal(g["idx_mix"], g["idx_f64"])
    hit = g["idx_f64"] >= 0
    assert np.array_equal(g["t_mix"][hit], g["t_f64"][hit].astype(np.float32).astype(np.float64))
    rel = _rel_vs_f64(g["t_mix"], g)
    assert rel.max() <= 1e-5 and rel.max() <= 6.0e-8


def test_pure_fp32_hit_arithmetic_misses_the_bar(oracle):
    """Why the exact test is f64: the same routine in FP32 (reference operation order) agrees on every index of the stored
    batch but cancels in the discriminant — 1000-unit ground sphere, grazing hits from 13 units away — so only ~77 % of the
    hits are inside 1e-5 and the worst is ~6e-4.  No product path runs this; the figures are kept as the measured reason."""
    g = _load("intersect_random_spheres.npz")
    idx, t = oracle.intersect(g["spheres"], g["rays"], enabled=g["enabled"], f32="pure")
    assert np.array_equal(idx, g["idx_f64"])
    rel = _rel_vs_f64(t, g)
    assert np.median(rel) < 2e-6 and 0.5 < (rel <= 1e-5).mean() < 0.9 and 1e-4 < rel.max() < 1e-3


def test_toml_fixture_loads_into_the_stored_scene(rtb):
    g = _load("render_three_spheres.npz")
    scene = rtb.Scene.from_toml(open(os.path.join(GOLDEN, "three_spheres.toml")).read())
    for a, b in zip(scene.arrays(), _arrays(g)):
        assert np.array_equal(np.asarray(a), b)


# ---- GPU -------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("stages", [0, 1, 2, 4])
@pytest.mark.parametrize("name", RENDERS)
def test_cuda_render_matches_golden(rtb, renderer, name, stages):
    g = _load(f"render_{name}.npz")
    W, H, spp, depth, seed = (int(g[k]) for k in ("W", "H", "spp", "depth", "seed"))
    scene = getattr(rtb.scenes, GENERATOR[name])()
    cam = rtb.Camera.default(W, H)
    renderer.seed = seed
    renderer.stages = stages
    try:
        img, lin, stat = renderer.render(depth, spp, scene, cam, want_linear=True)
    finally:
        renderer.stages = 0
        renderer.seed = 1
    err = np.abs(lin[..., :3].astype(np.float64) / spp - g["sum_mix"] / spp).max()
    assert err <= 1e-4, err                                # the bar; bit-identical up to the order of the partial sums
    assert stat.segments == int(g["segments_mix"])
    assert np.array_equal(img, g["rgb8_mix"])
    # and against the reference-faithful f64 oracle: the FP32 path state (hit points, normals, scattered directions) makes a
    # few silhouette samples take another branch, so this comparison is statistical
    d64 = np.abs(lin[..., :3].astype(np.float64) / spp - g["sum_f64"] / spp)
    # (measured on the four fixtures: 0 / 0.26 % / 0.04 % / 0 of the pixels beyond 1e-4, median 8e-9; the all-FP32 sweep of round 1: < 5 %)
    assert np.median(d64) < 1e-7 and (d64.max(axis=2) > 1e-4).mean() < 0.01


@pytest.mark.gpu
def test_cuda_intersect_matches_golden(rtb, renderer):
    """The north_star's intersection bar on the PRODUCTION sweep (K1 = the sweep of every render kernel), against the
    reference-faithful f64 routine: index equal for every ray, t within 1e-5 relative (it is the f64 root rounded once)."""
    g = _load("intersect_random_spheres.npz")
    scene = rtb.scenes.random_spheres()
    rays = g["rays"].astype(np.float32)
    idx, t = renderer.intersect_batch(scene, rays[:, :3], rays[:, 3:], enabled=g["enabled"].astype(np.uint8))
    assert np.array_equal(idx, g["idx_f64"])
    rel = _rel_vs_f64(t, g)
    assert rel.max() <= 1e-5, rel.max()
    hit = idx >= 0
    assert np.array_equal(t[hit], g["t_f64"][hit].astype(np.float32)) and np.all(np.isinf(t[~hit]))
    assert np.array_equal(idx, g["idx_mix"]) and np.array_equal(t[hit].astype(np.float64), g["t_mix"][hit])   # = the mixed oracle, bit for bit


@pytest.mark.gpu
def test_cuda_renders_the_toml_fixture(rtb, renderer):
    g = _load("render_three_spheres.npz")
    W, H, spp, depth, seed = (int(g[k]) for k in ("W", "H", "spp", "depth", "seed"))
    scene = rtb.Scene.from_toml(open(os.path.join(GOLDEN, "three_spheres.toml")).read())
    renderer.seed = seed
    img, stat = renderer.render(depth, spp, scene, rtb.Camera.default(W, H))
    assert np.array_equal(img, g["rgb8_mix"]) and stat.segments == int(g["segments_mix"])


@pytest.mark.gpu
def test_cuda_f64_probe_equals_the_f64_golden_bit_for_bit(rtb, renderer):
    """rtb200_intersect_batch_f64: the reference's routine at the reference's precision on the device, every pair, no
    filter.  On the stored batch it returns oracle<double>'s index and t exactly (same operation order, mul_add sites,
    IEEE sqrt/div), and it is the on-device yardstick for the production sweep (FP32 filters + the same f64 test on what
    they let through) on the same rays."""
    g = _load("intersect_random_spheres.npz")
    scene = rtb.scenes.random_spheres()
    rays = g["rays"]
    idx, t = renderer.intersect_batch_f64(scene, rays[:, :3], rays[:, 3:], enabled=g["enabled"].astype(np.uint8))
    assert np.array_equal(idx, g["idx_f64"])
    hit = idx >= 0
    assert np.array_equal(t[hit], g["t_f64"][hit]) and np.all(np.isinf(t[~hit]))
    idx32, t32 = renderer.intersect_batch(scene, rays[:, :3].astype(np.float32), rays[:, 3:].astype(np.float32),
                                          enabled=g["enabled"].astype(np.uint8))
    assert np.array_equal(idx32, idx)                                     # the filtered production sweep against the filter-free f64 kernel:
    assert np.array_equal(t32[hit], t[hit].astype(np.float32))            # same object, the same root rounded once
