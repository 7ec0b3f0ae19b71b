"""tests/abi_c_caller.c: a plain-C program (gcc, include/rtb200.h, no CUDA headers) that replays the call sequence of the
Rust binding's `render` (integration/rust/src/cuda_renderer.rs) — the Rust side itself cannot be compiled here.  CPU: it
builds and links against librtb200.so.  GPU: its images equal the golden fixture."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "rust-ray-tracing_b200")


def _build(tmp_path):
    exe = tmp_path / "abi_c_caller"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"), "-o", str(exe),
                    os.path.join(ROOT, "tests", "abi_c_caller.c"), "-L", PKG, "-lrtb200", f"-Wl,-rpath,{PKG}"], check=True)
    return exe


def test_c_caller_builds_as_plain_c(rtb, tmp_path):
    assert _build(tmp_path).exists()


@pytest.mark.gpu
@pytest.mark.parametrize("n_shards,row_groups", [(2, 1), (3, 3), (4, 2)])
def test_c_caller_reproduces_the_golden_image(rtb, tmp_path, n_shards, row_groups):
    exe = _build(tmp_path)
    prefix = str(tmp_path / "img")
    r = subprocess.run([str(exe), prefix, str(n_shards), str(row_groups)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    g = np.load(os.path.join(ROOT, "tests", "golden", "render_three_spheres.npz"))
    W, H = int(g["W"]), int(g["H"])
    single = np.fromfile(prefix + "_single.rgb", dtype=np.uint8).reshape(H, W, 3)
    multi = np.fromfile(prefix + "_multi.rgb", dtype=np.uint8).reshape(H, W, 3)
    assert np.array_equal(single, g["rgb8_mix"])
    seg = int(g["segments_mix"])
    assert f"single paths={W * H * int(g['spp'])} segments={seg}" in r.stdout and f"multi paths={W * H * int(g['spp'])} segments={seg}" in r.stdout
    if row_groups == n_shards:
        assert np.array_equal(multi, single)                              # row shards: nothing is re-associated
    else:
        assert np.abs(multi.astype(int) - single.astype(int)).max() <= 1  # sample shards: partial sums added in shard order
