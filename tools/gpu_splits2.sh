#!/bin/bash
# item-count sweep for the per-GPU share of T on 8 GPUs (1920x1080 x 125 spp): RTB200_SPLITS_LOG2 = 2..5 (and the library's own choice)
mkdir -p gpurun_out
for lg in auto 2 3 4 5; do
  if [ $lg = auto ]; then unset RTB200_SPLITS_LOG2; else export RTB200_SPLITS_LOG2=$lg; fi
  python bench.py --workload T --spp 125 --steps 10 --warmup 3 --no-cpu-baseline --no-parity > gpurun_out/splits_$lg.json 2>/dev/null
  python -c "import json; d=json.load(open('gpurun_out/splits_$lg.json')); print('125spp lg=$lg', round(d['ms_per_step'],3), round(d['roofline']['kernel_ms'],3))"
done
unset RTB200_SPLITS_LOG2
for c in 16 32 64 128; do
  RTB200_CLAIM_CHUNK=$c python bench.py --workload T --spp 125 --steps 10 --warmup 3 --no-cpu-baseline --no-parity > gpurun_out/chunk_$c.json 2>/dev/null
  python -c "import json; d=json.load(open('gpurun_out/chunk_$c.json')); print('125spp chunk=$c', round(d['ms_per_step'],3))"
done
