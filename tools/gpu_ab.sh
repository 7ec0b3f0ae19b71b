#!/bin/bash
# A/B of library builds under build_ab/: each is copied over the product library and benched on C2 (and T)
mkdir -p gpurun_out
cp rust-ray-tracing_b200/librtb200.so /tmp/librtb200_keep.so
for v in "$@"; do
  cp build_ab/librtb200_$v.so rust-ray-tracing_b200/librtb200.so
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ab_c2_$v.json 2>> gpurun_out/ab.err
  python -c "import json; d=json.load(open('gpurun_out/ab_c2_$v.json')); print('$v C2', round(d['ms_per_step'],3), round(d['value'],1))"
done
cp /tmp/librtb200_keep.so rust-ray-tracing_b200/librtb200.so
