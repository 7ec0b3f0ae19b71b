#!/bin/bash
# streamed (per-warp cp.async) trace kernel: parity, then A/B against the tiled kernel on L40K and C4
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tiled" > gpurun_out/pytest_gpu3.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu3.log
tail -3 gpurun_out/pytest_gpu3.log
for shape in g640x1 g512x1 g320x2 g256x2; do
  RTB200_TILED_SHAPE=$shape timeout 120 python bench.py --workload L40K --stages 4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_l40k_$shape.json 2>> gpurun_out/bench3.err
  RTB200_TILED_SHAPE=$shape timeout 120 python bench.py --workload C4 --stages 4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_tiled_$shape.json 2>> gpurun_out/bench3.err
done
for f in gpurun_out/bench_l40k_g*.json gpurun_out/bench_c4_tiled_g*.json; do echo $f; python -c "import json,sys; d=json.load(open('$f')); print(d['ms_per_step'], d['value'], d['roofline']['frac'])"; done
