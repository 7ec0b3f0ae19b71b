#!/bin/bash
# one-GPU validation: full GPU suite, smoke, the default bench line (T), BASELINE and mid-size workloads, the reference arm,
# then the ncu launch list of the default bench command (gpu__time_duration of every launch)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; python -c "
import json; d=json.load(open('gpurun_out/bench_default.json'))
print('T', d['ms_per_step'], d['value'], 'frac', d['roofline']['frac'], 'e2e', d['e2e']['value'], 'cpu', d['cpu_baseline']['value'], 'c2', d.get('c2_ms_per_step'), 'parity', d['parity']['ok'], d['parity']['oracle']['seconds'])"
for w in ${WORKLOADS:-C2 G14 G18 G25 C4 C5 C1 L40K}; do
  timeout 300 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${w}.json 2>> gpurun_out/bench.err
  python -c "import json,sys; d=json.load(open('gpurun_out/bench_${w}.json')); print('$w', d['config']['n_spheres'], d['ms_per_step'], d['value'], d['roofline']['frac'], d['parity']['ok'])"
done
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err; tail -c 400 gpurun_out/bench_ref.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches_T.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity > gpurun_out/ncu_launches.log 2>&1
grep -c k_trace_paths gpurun_out/launches_T.csv
