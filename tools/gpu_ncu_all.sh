#!/bin/bash
# ncu of the trace kernel of T (the default bench command), C2, C4, C5: launch list of the default command, then one
# --set full capture per workload (after the plain command exited 0)
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain_T.log 2>&1 || { echo "plain T failed"; tail -5 gpurun_out/plain_T.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches_T.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_Ta.log 2>&1
grep -c k_trace_paths gpurun_out/launches_T.csv
for W in ${WORKLOADS:-T C2 C4 C5}; do
  CMD="python bench.py --workload $W --steps 1 --warmup 3 --no-cpu-baseline --no-parity"
  $CMD > gpurun_out/plain_$W.log 2>&1 || { echo "plain $W failed"; tail -5 gpurun_out/plain_$W.log; continue; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_trace_paths -s 3 -c 1 -o gpurun_out/prof_$W -f $CMD > gpurun_out/ncu_$W.log 2>&1
  ls -la gpurun_out/prof_$W.ncu-rep
done
