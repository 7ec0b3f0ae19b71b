#!/bin/bash
# N-GPU validation (N = number of visible GPUs, >= 2): the multi-GPU tests (torchrun path and the one-call C-ABI path over real
# peers), the default bench line (north_star target T, strong scaling) with both barrier kinds, the weak-scaling C2 line, and
# the one-call CLI.  The pytest log goes to gpurun_out/pytest_mgpu_${N}.log (copied to profiles/ by hand).
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 900 python -m pytest tests/test_multigpu.py tests/test_multi_one_call.py tests/test_c_caller.py -m gpu -v > gpurun_out/pytest_mgpu_${N}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_mgpu_${N}.log
grep -E "passed|failed|PASSED|FAILED|SKIPPED|rc=" gpurun_out/pytest_mgpu_${N}.log | tail -30
run() {  # name, extra args
  name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29541 + RANDOM % 200)) bench.py --gpus $N "$@" > gpurun_out/bench_${name}_${N}gpu.json 2> gpurun_out/bench_${name}_${N}gpu.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_${name}_${N}gpu.json'))
print('${name} ${N}gpu', round(d['ms_per_step'],3), round(d['value'],1), 'e2e', round(d['e2e']['ms_per_step'],3), 'torchrun-e2e', round(d.get('e2e_torchrun',{}).get('ms_per_step',0),3), 'parity', d['parity'] and d['parity']['ok'], d['parity'] and d['parity'].get('linear_max_abs_vs_single_gpu'))" || tail -5 gpurun_out/bench_${name}_${N}gpu.err
}
run T --steps 5 --warmup 3
run T_ncclbarrier --steps 5 --warmup 3 --barrier nccl --no-parity
run T_nccl --steps 5 --warmup 3 --exchange nccl --no-parity
run C2 --workload C2 --steps 5 --warmup 3 --no-parity
run C3 --workload C3 --steps 2 --warmup 3 --no-parity
python -c "
import importlib,sys; sys.path.insert(0,'.'); rtb=importlib.import_module('rust-ray-tracing_b200'); open('gpurun_out/rs.toml','w').write(rtb.scenes.random_spheres().to_toml())"
rust-ray-tracing_b200/rtb200_cli gpurun_out/rs.toml -o gpurun_out/cli_${N}gpu.png --width 1920 --height 1080 --spp 1000 --gpus $N --stats-json --quiet | tail -1
rust-ray-tracing_b200/rtb200_cli gpurun_out/rs.toml -o gpurun_out/cli_1gpu.png --width 1920 --height 1080 --spp 1000 --stats-json --quiet | tail -1
