#!/usr/bin/env python
"""Writes the committed golden fixtures under tests/golden/ from the CPU oracle (oracle/rt_oracle.hpp).

The reference has no tests, fixtures or images for this path, cannot be built here (no Rust toolchain) and seeds its
randomness from the OS, so these are NOT outputs of the reference: they freeze the oracle, which is pinned
by the hand-derived known-answer vectors in tests/test_oracle_kat.py and the Random123 Philox vectors.  What they buy:
  * the oracle cannot drift silently (compiler, flags, an edit to rt_oracle.hpp): tests/test_golden.py (CPU) re-runs it
    against these files bit for bit, in both modes: "f64" (reference-faithful) and "mix" (the GPU comparand: FP32 path
    state, closest-hit sweep in f64 — what the CUDA kernels compute);
  * the CUDA path is checked on the GPU box against files that were written HERE, by a different build of the oracle
    on a different machine (tests/test_golden.py, -m gpu);
  * the three-sphere scene in the reference's TOML schema (three_spheres.toml) is what every host (Python,
    native C++) must load into exactly the arrays stored in render_three_spheres.npz.

  python tests/golden/make_golden.py        (re-)writes the fixtures; run from the repo root after `make -C oracle`
"""
import importlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

RENDERS = [  # name, scene generator, W, H, spp, depth, seed
    ("three_spheres", "three_spheres", 96, 54, 8, 50, 1),
    ("random_spheres", "random_spheres", 64, 36, 4, 50, 7),
    ("dielectric_heavy", "dielectric_heavy", 64, 36, 4, 50, 3),
    ("depth3", "random_spheres", 48, 27, 4, 3, 1),
]


def main():
    rtb = importlib.import_module("rust-ray-tracing_b200")
    import oracle_lib as oracle

    for name, gen, W, H, spp, depth, seed in RENDERS:
        scene = getattr(rtb.scenes, gen)()
        cam = rtb.Camera.default(W, H)
        arrays = scene.arrays()
        out = {"W": W, "H": H, "spp": spp, "depth": depth, "seed": seed, "cam18": np.asarray(cam.vectors(), dtype=np.float64)}
        for k, a in zip(("spheres", "mat_index", "kind", "hollow", "albedo", "param"), arrays):
            out[k] = np.asarray(a)
        for tag, f32 in (("mix", True), ("f64", False)):   # mix = the GPU comparand (FP32 state, f64 hit routine), oracle_lib._mode
            s, rgb, seg = oracle.render(arrays, W, H, cam.vectors(), spp, depth, seed=seed, f32=f32, threads=1)
            out["sum_" + tag], out["rgb8_" + tag], out["segments_" + tag] = s, rgb, np.uint64(seg)
        np.savez_compressed(os.path.join(HERE, f"render_{name}.npz"), **out)
        print(name, W, H, spp, "segments", int(out["segments_mix"]), int(out["segments_f64"]))

    # ray batch against the random-spheres scene: camera-like rays, rays from inside the scene, non-unit directions,
    # a few disabled lanes
    scene = rtb.scenes.random_spheres()
    spheres = np.asarray(scene.arrays()[0], dtype=np.float64)
    rng = np.random.default_rng(20241104)
    m = 4096
    o = np.empty((m, 3)); d = np.empty((m, 3))
    o[: m // 2] = (13.0, 2.0, 3.0)
    tgt = np.stack([rng.uniform(-11, 11, m // 2), rng.uniform(0.0, 1.0, m // 2), rng.uniform(-11, 11, m // 2)], axis=1)
    d[: m // 2] = tgt - o[: m // 2]
    o[m // 2:] = np.stack([rng.uniform(-11, 11, m - m // 2), rng.uniform(0.05, 1.5, m - m // 2), rng.uniform(-11, 11, m - m // 2)], axis=1)
    d[m // 2:] = rng.normal(size=(m - m // 2, 3)) * rng.uniform(0.1, 4.0, size=(m - m // 2, 1))
    o = o.astype(np.float32).astype(np.float64)      # exactly representable in f32: both scalar types see the same rays
    d = d.astype(np.float32).astype(np.float64)
    enabled = (rng.random(m) > 0.05).astype(np.int32)
    rays = np.concatenate([o, d], axis=1)
    out = {"spheres": spheres, "rays": rays, "enabled": enabled}
    for tag, f32 in (("mix", True), ("f64", False)):   # mix = the GPU comparand (FP32 state, f64 hit routine), oracle_lib._mode
        idx, t = oracle.intersect(spheres, rays, enabled=enabled, f32=f32)
        out["idx_" + tag], out["t_" + tag] = idx, t
    np.savez_compressed(os.path.join(HERE, "intersect_random_spheres.npz"), **out)
    print("intersect", m, "hits", int((out["idx_mix"] >= 0).sum()), "mix/f64 index disagreements", int((out["idx_mix"] != out["idx_f64"]).sum()))

    with open(os.path.join(HERE, "three_spheres.toml"), "w") as f:
        f.write("# BASELINE configs[0]: the three-sphere scene (diffuse / metal / glass + ground) in the reference's scene schema\n"
                "# (Scene::from_toml, src/ray_tracing.rs:217-228; SURVEY 2.1).  Written by tests/golden/make_golden.py.\n")
        f.write(rtb.scenes.three_spheres().to_toml())


if __name__ == "__main__":
    main()
