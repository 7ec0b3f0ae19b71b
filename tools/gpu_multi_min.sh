#!/bin/bash
# minimal N-GPU confirmation: the multi-GPU tests over real peers, then the default bench line (T, strong scaling)
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 600 python -m pytest tests/test_multigpu.py tests/test_multi_one_call.py tests/test_c_caller.py -m gpu -v > gpurun_out/pytest_mgpu_${N}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_mgpu_${N}.log
grep -E "passed|failed|rc=" gpurun_out/pytest_mgpu_${N}.log | tail -3
bash tools/gpu_bench_multi.sh T
