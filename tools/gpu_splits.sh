#!/bin/bash
# item-count rule, K = 2^L: usage gpu_splits.sh "<workload> <spp|0> <L>" ...
mkdir -p gpurun_out
for cfg in "$@"; do set -- $cfg
  if [ "$2" = 0 ]; then S=""; else S="--spp $2"; fi
  RTB200_SPLITS_LOG2=$3 python bench.py --workload $1 $S --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/sp_$1_$2_$3.json 2>> gpurun_out/sp.err
  python -c "import json; d=json.load(open('gpurun_out/sp_$1_$2_$3.json')); print('$1 spp $2 log2K $3', round(d['ms_per_step'],3))"
done
