#!/bin/bash
# where does the ~1.2 ms fixed cost of a C2 render come from?  spp sweep at constant samples per item (K scaled) and at constant K
mkdir -p gpurun_out
for cfg in "25 2" "50 3" "100 4" "200 5" "400 6" "25 4" "50 4" "200 4" "400 4"; do set -- $cfg
  RTB200_SPLITS_LOG2=$2 python bench.py --workload C2 --spp $1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/fc_$1_$2.json 2>> gpurun_out/fc.err
  python -c "import json; d=json.load(open('gpurun_out/fc_$1_$2.json')); print('spp $1 log2K $2', round(d['ms_per_step'],3), round(d['roofline']['kernel_ms'],3))"
done
