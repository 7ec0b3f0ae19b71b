#!/bin/bash
# A/B of library builds under build_ab/ (tools/build_ab.sh): GPU suite on the candidates, then benches per variant.
# Lines of "variant workload[/env] ms Mpaths/s" go to gpurun_out/ab2.txt.
mkdir -p gpurun_out
cp rust-ray-tracing_b200/librtb200.so /tmp/librtb200_keep.so
use() { cp build_ab/librtb200_$1.so rust-ray-tracing_b200/librtb200.so; }
run() {  # variant workload [ENV=VAL]
  local v=$1 w=$2 e=$3
  env $e timeout 300 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline --no-parity > gpurun_out/ab2_${v}_${w}_${e%%=*}.json 2>> gpurun_out/ab2.err
  python - "$v" "$w" "$e" gpurun_out/ab2_${v}_${w}_${e%%=*}.json <<'PY' | tee -a gpurun_out/ab2.txt
import json, sys
v, w, e, f = sys.argv[1:5]
try:
    d = json.load(open(f)); print(v, w, e or "-", round(d["ms_per_step"], 3), round(d["value"], 1), d["roofline"].get("kernel", "")[:40])
except Exception as ex:
    print(v, w, e or "-", "FAILED", ex)
PY
}
for v in w1 f2; do
  use $v
  timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/ab2_pytest_$v.log 2>&1; echo "$v pytest rc=$? $(tail -1 gpurun_out/ab2_pytest_$v.log)" | tee -a gpurun_out/ab2.txt
done
use head; for w in C2 T C5 C4 G18 G25 L40K; do run head $w; done; run head C2 RTB200_PARAM_MAX=0
use np;   for w in C2 C4; do run np $w; done
use w1;   for w in C2 T C5 C4 G18 G25; do run w1 $w; done
use f2;   for w in C4 G14 G18 G25 L40K; do run f2 $w; done; for w in C2 T C5; do run f2 $w RTB200_PARAM_MAX=0; done
use head; run head C2
cp /tmp/librtb200_keep.so rust-ray-tracing_b200/librtb200.so
