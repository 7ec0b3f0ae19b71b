#!/usr/bin/env python
"""Debug: start / queue-dry / exit times of the warps of one C2 trace launch (library built with -DRTB_TIMELINE;
tools/gpu_timeline.sh swaps it in).  Prints when the queue ran dry and how the warp exits spread out after that."""
import ctypes as C
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rtb = importlib.import_module("rust-ray-tracing_b200")
lib = rtb._abi.lib
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 100
scene = rtb.scenes.random_spheres()
cam = rtb.Camera.default(1920, 1080)
r = rtb.CudaRenderer(device=0, seed=1)
for _ in range(3):
    r.render(50, spp, scene, cam)
buf = (C.c_ulonglong * (8 + 4096))()
lib.rtb200_debug_timeline(1, buf)
img, st = r.render(50, spp, scene, cam)
lib.rtb200_debug_timeline(0, buf)
a = np.array(buf[:], dtype=np.uint64)
t0 = int(a[0])
rel = lambda x: (int(x) - t0) * 1e-6
print(f"spp {spp}: trace_ms {st.trace_ms:.3f}")
print(f"warp starts: {rel(a[0]):.3f} .. {rel(a[1]):.3f} ms;  queue dry seen: {rel(a[2]):.3f} .. {rel(a[3]):.3f} ms;  exits: {rel(a[4]):.3f} .. {rel(a[5]):.3f} ms")
bins = a[8:].astype(np.int64)
first_bin = (int(a[4]) >> 15) & 4095
last_bin = (int(a[5]) >> 15) & 4095
n = (last_bin - first_bin) % 4096 + 1
cum = 0
total = int(bins.sum())
print(f"{total} warps; exits per 32.8 us bin after the first exit:")
for k in range(n):
    c = int(bins[(first_bin + k) % 4096])
    cum += c
    print(f"  +{k * 0.032768:.3f} ms: {c:5d}  ({100.0 * cum / total:5.1f} % done)")
r.close()
