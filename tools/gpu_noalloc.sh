#!/bin/bash
# the static-scratch entry point: its test alone in a fresh process, then the whole GPU suite (it runs last), smoke, default bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_noalloc_alone.log
tail -3 gpurun_out/pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; tail -c 1500 gpurun_out/bench_default.json
