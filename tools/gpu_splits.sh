#!/bin/bash
# item-size rule at low sample counts: C2 geometry at 8/16/25/50 spp and the 1297-sphere grid at 16 spp, K = 2^L
mkdir -p gpurun_out
for cfg in "C2 8 0" "C2 8 1" "C2 8 2" "C2 8 3" "C2 16 1" "C2 16 2" "C2 16 3" "C2 16 4" "C2 25 2" "C2 25 3" "C2 25 4" "C2 50 3" "C2 50 4" "C2 50 5" "G18 16 1" "G18 16 2" "G18 16 3" "G18 16 4" "C4 16 1" "C4 16 2" "C4 16 3" "C4 16 4"; do set -- $cfg
  RTB200_SPLITS_LOG2=$3 python bench.py --workload $1 --spp $2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/sp_$1_$2_$3.json 2>> gpurun_out/sp.err
  python -c "import json; d=json.load(open('gpurun_out/sp_$1_$2_$3.json')); print('$1 spp $2 log2K $3', round(d['ms_per_step'],3))"
done
