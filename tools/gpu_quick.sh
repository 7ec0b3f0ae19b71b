#!/bin/bash
# quick one-GPU check: the parity suite, then the bench workloads named on the command line (default C2 C5 C4 G14 G25)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
WL=${@:-C2 C5 C4 G14 G25}
for w in $WL; do
  timeout 300 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${w}.json 2>> gpurun_out/bench.err
  python -c "import json,sys; d=json.load(open('gpurun_out/bench_${w}.json')); print('$w', d['config'].get('n_spheres'), round(d['ms_per_step'],3), round(d['value'],1), round(d['roofline']['frac'],4))"
done
