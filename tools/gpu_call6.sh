#!/bin/bash
# where does parameter space stop paying?  two-stage from parameter space (default) vs staged in shared memory (RTB200_PARAM_MAX=0)
mkdir -p gpurun_out
for w in G8 G11 G14 G18 G22 G25; do for pm in 2560 0; do
  RTB200_PARAM_MAX=$pm timeout 120 python bench.py --workload $w --stages 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${w}_pm$pm.json 2>> gpurun_out/bench6.err
  echo $w param_max $pm; python -c "import json,sys; d=json.load(open('gpurun_out/bench_${w}_pm$pm.json')); print(d['config']['n_spheres'], d['ms_per_step'], d['value'], d['roofline']['frac'])"
done; done
RTB200_PARAM_MAX=0 timeout 120 python bench.py --workload C2 --stages 2 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2_pm0.json 2>> gpurun_out/bench6.err
python -c "import json,sys; d=json.load(open('gpurun_out/bench_c2_pm0.json')); print('C2 pm0', d['config']['n_spheres'], d['ms_per_step'], d['value'], d['roofline']['frac'])"
