#!/bin/bash
# N-GPU bench lines only (N = visible GPUs): T default, barrier/exchange variants, C2 weak, C3 strong
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
run() {  # name, extra args
  name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29541 + RANDOM % 200)) bench.py --gpus $N "$@" > gpurun_out/bench_${name}_${N}gpu.json 2> gpurun_out/bench_${name}_${N}gpu.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_${name}_${N}gpu.json'))
print('${name} ${N}gpu', round(d['ms_per_step'],3), round(d['value'],1), 'e2e', round(d['e2e']['ms_per_step'],3), 'torchrun-e2e', round(d.get('e2e_torchrun',{}).get('ms_per_step',0),3), 'parity', d['parity'] and d['parity']['ok'], d['parity'] and d['parity'].get('linear_max_abs_vs_single_gpu'))" || grep -m3 -A3 "Error" gpurun_out/bench_${name}_${N}gpu.err
}
for spec in "$@"; do
  case $spec in
    T) run T --steps 5 --warmup 3 ;;
    Tnb) run T_ncclbarrier --steps 5 --warmup 3 --barrier nccl --no-parity ;;
    Tnccl) run T_nccl --steps 5 --warmup 3 --exchange nccl --no-parity ;;
    C2) run C2 --workload C2 --steps 5 --warmup 3 --no-parity ;;
    C3) run C3 --workload C3 --steps 2 --warmup 3 --no-parity ;;
  esac
done
