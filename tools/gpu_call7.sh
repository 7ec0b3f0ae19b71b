#!/bin/bash
# new parameter-space threshold (768 spheres): full GPU suite, smoke, default bench line, mid-size scenes, C1/C4/C5
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu7.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu7.log
tail -3 gpurun_out/pytest_gpu7.log
python __graft_entry__.py smoke > gpurun_out/smoke7.log 2>&1; tail -1 gpurun_out/smoke7.log
python bench.py > gpurun_out/bench_default7.json 2> gpurun_out/bench_default7.err; python -c "import json; d=json.load(open('gpurun_out/bench_default7.json')); print('C2', d['ms_per_step'], d['value'], d['roofline']['frac'], d['e2e']['value'], d['cpu_baseline']['value'])"
for w in G14 G18 G25 C4 C5 C1; do
  timeout 200 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${w}_7.json 2>> gpurun_out/bench7.err
  python -c "import json,sys; d=json.load(open('gpurun_out/bench_${w}_7.json')); print('$w', d['config']['n_spheres'], d['ms_per_step'], d['value'], d['roofline']['frac'])"
done
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref7.json 2>> gpurun_out/bench7.err; tail -c 300 gpurun_out/bench_ref7.json
