#!/bin/bash
# round-1 re-entry confirmation: full GPU suite, then bench lines (C2, C4 default vs tiled, L40K tiled at both CTA sizes)
mkdir -p gpurun_out
timeout 700 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; tail -c 600 gpurun_out/bench_c2.json
python bench.py --workload C4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4.json 2>> gpurun_out/bench.err
python bench.py --workload C4 --stages 4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_tiled.json 2>> gpurun_out/bench.err
RTB200_TILED_THREADS=512 python bench.py --workload C4 --stages 4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_tiled512.json 2>> gpurun_out/bench.err
python bench.py --workload L40K --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_l40k.json 2>> gpurun_out/bench.err
RTB200_TILED_THREADS=512 python bench.py --workload L40K --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_l40k_512.json 2>> gpurun_out/bench.err
RTB200_TILE_SPHERES=8192 python bench.py --workload L40K --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_l40k_tile8k.json 2>> gpurun_out/bench.err
for f in gpurun_out/bench_c4*.json gpurun_out/bench_l40k*.json; do echo $f; python -c "import json,sys; d=json.load(open('$f')); print(d['ms_per_step'], d['value'], d['roofline']['frac'])"; done
