#!/bin/bash
# usage: gpu_ab3.sh "<variant> ..." "<workload[@ENV=VAL][:spp]> ..."   -> gpurun_out/ab3.txt
mkdir -p gpurun_out
cp rust-ray-tracing_b200/librtb200.so /tmp/librtb200_keep.so
for v in $1; do
  cp build_ab/librtb200_$v.so rust-ray-tracing_b200/librtb200.so
  for spec in $2; do
    w=${spec%%[@:]*}; e=""; spp=""
    case $spec in *@*) e=${spec#*@}; e=${e%%:*};; esac
    case $spec in *:*) spp="--spp ${spec##*:}";; esac
    env $e timeout 300 python bench.py --workload $w $spp --steps 3 --warmup 3 --no-cpu-baseline --no-parity > gpurun_out/ab3_tmp.json 2>> gpurun_out/ab3.err
    python - "$v" "$spec" gpurun_out/ab3_tmp.json <<'PY' | tee -a gpurun_out/ab3.txt
import json, sys
v, w, f = sys.argv[1:4]
try:
    d = json.load(open(f)); print(v, w, round(d["ms_per_step"], 3), round(d["value"], 1), d["roofline"].get("kernel", "")[:34])
except Exception as ex:
    print(v, w, "FAILED", ex)
PY
  done
done
cp /tmp/librtb200_keep.so rust-ray-tracing_b200/librtb200.so
