#!/bin/bash
# ncu of the streamed kernel k_trace_paths_g on the 40 001-sphere scene at 2 spp (launch list, then one --set full capture)
mkdir -p gpurun_out
CMD="python bench.py --workload L40K --spp 2 --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain9.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain9.log; exit 1; }
tail -c 400 gpurun_out/plain9.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_l40k.csv $CMD > gpurun_out/ncu9a.log 2>&1
tail -3 gpurun_out/launches_l40k.csv
ncu --set full --clock-control none --import-source on -k regex:k_trace_paths_g -s 3 -c 1 -o gpurun_out/prof_l40k $CMD > gpurun_out/ncu9b.log 2>&1
ls -la gpurun_out/prof_l40k.ncu-rep
