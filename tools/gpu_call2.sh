#!/bin/bash
# tiled trace kernel: golden + tiled parity tests, CTA-shape A/B on the 40 001-sphere scene, ncu of the chosen shape
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_golden.py tests/test_gpu_parity.py -m gpu -x -q -k "golden or tiled or beyond" > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu2.log
tail -3 gpurun_out/pytest_gpu2.log
for shape in 512x1 640x1 256x2 320x2 128x4 160x4; do
  RTB200_TILED_SHAPE=$shape timeout 120 python bench.py --workload L40K --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_l40k_$shape.json 2>> gpurun_out/bench2.err
  RTB200_TILED_SHAPE=$shape timeout 120 python bench.py --workload C4 --stages 4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_tiled_$shape.json 2>> gpurun_out/bench2.err
done
for f in gpurun_out/bench_l40k_*x*.json gpurun_out/bench_c4_tiled_*x*.json; do echo $f; python -c "import json,sys; d=json.load(open('$f')); print(d['ms_per_step'], d['value'], d['roofline']['frac'])"; done
