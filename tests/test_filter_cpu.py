"""The sweep's conservative candidate filter (rust-ray-tracing_b200/csrc/rt_filter.h) must never reject a (ray, sphere)
pair whose reference-order float discriminant (src/objects.rs:252-257) is >= 0.  oracle/filter_check compiles the
product header for the host (same explicit-fma arithmetic as the device) and brute-forces adversarial near-silhouette
pairs; the GPU parity tests then confirm the device build end to end (bit-identical images)."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "filter_check")


@pytest.fixture(scope="module")
def checker():
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "filter_check"], check=True, capture_output=True)
    return BIN


@pytest.mark.parametrize("scenario", [0, 1, 2])
def test_filter_has_no_false_negatives(checker, scenario):
    threads = min(8, os.cpu_count() or 1)
    r = subprocess.run([checker, str(scenario), "16000000", "3", str(threads)], capture_output=True, text=True)
    out = json.loads(r.stdout)
    assert out["false_negatives"] == 0 and r.returncode == 0, out
    assert out["ref_positive"] > 500000          # the sample really contains reference candidates
    assert out["min_slack"] > 0.5                 # worst observed error used less than half of the proven margin
    assert out["flagged"] < 3 * out["ref_positive"]   # and the filter is not vacuous on silhouette-aimed rays


def test_packed_slab_sign_convention_keeps_every_scalar_candidate():
    """The shared-memory kernels evaluate stage 1 as q' = fma(p, p, -R'') with R'' = R' one ulp up and take the sign bit SET as
    "candidate" (rt_kernels.cuh, SmemSlab); the scalar form is q = fma(-p, p, R') with the sign bit CLEAR.  Every candidate of the
    scalar form must be one of the packed form: p^2 <= R' (exactly) implies the exact p^2 - R'' < 0, and a negative exact value
    rounds to a negative, non-zero float.  Exact FP32 fma here = one rounding of the exact f64 value (48-bit product of floats
    minus a float of comparable exponent: representable in f64 for the ranges sampled)."""
    import numpy as np
    rng = np.random.default_rng(7)
    n = 2_000_000
    rp = np.exp(rng.uniform(np.log(1e-12), np.log(1e7), n)).astype(np.float32)          # R' over the range scenes produce
    # p near the knife edge |p| = sqrt(R'), both sides, down to the last ulps, plus exact hits of p*p == R' where they exist
    p = (np.sqrt(rp.astype(np.float64)) * (1.0 + rng.choice([-1, 1], n) * np.exp(rng.uniform(np.log(1e-9), np.log(1e-2), n)))).astype(np.float32)
    p[: n // 8] = np.sqrt(rp[: n // 8].astype(np.float64)).astype(np.float32)
    p64, rp64 = p.astype(np.float64), rp.astype(np.float64)
    q_scalar = (rp64 - p64 * p64).astype(np.float32)                                     # fma(-p, p, R'), rounded once
    rpp = np.nextafter(rp, np.float32(np.inf))
    q_packed = (p64 * p64 - rpp.astype(np.float64)).astype(np.float32)                   # fma(p, p, -R''), rounded once
    scalar_candidate = ~np.signbit(q_scalar)
    packed_candidate = np.signbit(q_packed) & (q_packed != 0)
    assert scalar_candidate.sum() > n // 4 and (~scalar_candidate).sum() > n // 4         # both sides of the edge are sampled
    assert not np.any(scalar_candidate & ~packed_candidate)
    # and the packed form is only one ulp of R' looser: the two disagree nowhere outside that sliver
    extra = packed_candidate & ~scalar_candidate
    ulp = (rpp - rp).astype(np.float64)
    assert extra.any() and np.all((p64 * p64 - rp64)[extra] <= ulp[extra])
