#!/bin/bash
# A/B of library builds under build_ab/: each is copied over the product library and benched.
# usage: gpu_ab.sh <workload>[,<workload>...] <variant> [<variant> ...]   ("base" = the library as shipped)
mkdir -p gpurun_out
WL=$1; shift
cp rust-ray-tracing_b200/librtb200.so /tmp/librtb200_keep.so
for v in "$@"; do
  if [ "$v" = base ]; then cp /tmp/librtb200_keep.so rust-ray-tracing_b200/librtb200.so; else cp build_ab/librtb200_$v.so rust-ray-tracing_b200/librtb200.so; fi
  for w in ${WL//,/ }; do
    timeout 200 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ab_${w}_$v.json 2>> gpurun_out/ab.err
    python -c "import json; d=json.load(open('gpurun_out/ab_${w}_$v.json')); print('$v $w', round(d['ms_per_step'],3), round(d['value'],1))"
  done
done
cp /tmp/librtb200_keep.so rust-ray-tracing_b200/librtb200.so
