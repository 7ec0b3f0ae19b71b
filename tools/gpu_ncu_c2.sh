#!/bin/bash
# ncu of the default bench command's trace kernel (C2): launch list, then one --set full capture
mkdir -p gpurun_out
CMD="python bench.py --workload C2 --steps 1 --warmup 3 --no-cpu-baseline --no-parity"
$CMD > gpurun_out/plain_c2.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_c2.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_c2.csv $CMD > gpurun_out/ncu_c2a.log 2>&1
grep -c k_trace_paths gpurun_out/launches_c2.csv
ncu --set full --clock-control none --import-source on -k regex:k_trace_paths_v -s 3 -c 1 -o gpurun_out/prof_c2 $CMD > gpurun_out/ncu_c2b.log 2>&1
ls -la gpurun_out/prof_c2.ncu-rep
