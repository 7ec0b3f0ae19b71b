"""bench.py's contract on the arms that run without a GPU: the reference arm (CPU port of the reference's live renderer)
prints exactly one JSON line with the keys the driver reads; the GPU arm refuses to run without a device instead of
falling back to the CPU."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True, cwd=ROOT, env=e, timeout=300)


def test_reference_arm_prints_one_contract_line():
    r = _run(["--impl", "reference", "--workload", "C1", "--steps", "1", "--warmup", "0"])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "Mpaths/s" and d["unit"] == "Mpaths/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["steps"] == 1 and d["warmup"] == 0 and d["n_gpus"] == 1
    assert d["vs_baseline"] is None and d["data"] == "synthetic" and d["dtype"] == "f64"
    assert d["config"]["workload"].startswith("C1:") and 0 < d["config"]["sample_fraction"] <= 1.0
    assert d["config"]["same_config"] == (d["config"]["sample_fraction"] == 1.0)
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "sample" in cb
    assert d["e2e"] == {"value": d["value"], "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0


def test_reference_arm_does_not_map_the_product_library():
    """The reference arm builds its scenes from the host-only half of the package (scene generators + plain-ctypes struct
    mirrors): librtb200.so must not be mapped into that process (the driver records which .so files each arm loaded)."""
    code = ("import sys; sys.path.insert(0, %r); import bench; sc = bench.host_only_scenes().random_spheres(); assert len(sc.arrays()[0]) == 486; "
            "bench.cpu_reference_run('C1', 1, 0, 1.0); maps = open('/proc/self/maps').read(); "
            "assert 'librtb200' not in maps, 'librtb200.so is mapped'; assert 'libcpuref' in maps; print('ok')" % ROOT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=ROOT, timeout=300)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stderr[-2000:]


def test_reference_arm_other_ranks_exit_quietly():
    r = _run(["--impl", "reference", "--workload", "C1", "--steps", "1", "--warmup", "0", "--gpus", "2"], env={"RANK": "1", "WORLD_SIZE": "2"})
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_gpu_arm_refuses_to_run_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = _run(["--steps", "1", "--warmup", "1"])
    assert r.returncode != 0 and "no CPU fallback" in (r.stderr + r.stdout)


@pytest.mark.gpu
def test_gpu_arm_line_carries_the_contract_keys():
    """The product arm on the smallest workload (C1 in full): one JSON line with roofline, e2e, clocks, gpu_launches."""
    r = _run(["--workload", "C1", "--steps", "2", "--warmup", "3", "--no-cpu-baseline"])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert "impl" not in d and d["metric"] == "Mpaths/s" and d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 3
    assert d["value"] > 0 and d["dtype"] == "f32" and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert d["config"]["workload"].startswith("C1:") and d["config"]["n_spheres"] == 4
    rf = d["roofline"]
    assert rf["bound"] == "fp32" and rf["unit"] == "TFLOP/s" and rf["peak"] > 20 and rf["achieved"] > 0
    assert abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-9 and rf["flop_per_test"] == 17.0
    e = d["e2e"]
    assert e["value"] > 0 and e["value"] <= d["value"] * 1.05 and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] == 400 * 225 * 3
    assert d["gpu_launches"] >= 2 * 2                      # at least trace + resolve per step
    assert d["clocks"]["sm_max_mhz"] is None or d["clocks"]["sm_max_mhz"] > 0
    assert d["cpu_baseline"] is None                       # --no-cpu-baseline
    assert rf["kernel"].startswith("k_trace_paths") and rf["kernel_ms"] > 0 and rf["frac_of_k0_probe"] > 0
    par = d["parity"]                                      # computed outside the timed region: three scanlines against the CPU oracle
    assert par["ok"] is True and len(par["oracle"]["strips"]) == 3
    assert all(st["rgb8_equal"] and st["linear_max_abs"] <= 1e-4 for st in par["oracle"]["strips"])
    assert "rtb200_render" in e["call"] and e["steps"] == 2
