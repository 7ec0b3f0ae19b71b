#!/bin/bash
# N-GPU validation (N = number of visible GPUs, >= 2): multi-GPU tests, the weak-scaling bench line, the strong-scaling target T
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/pytest_mgpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_mgpu.log
tail -3 gpurun_out/pytest_mgpu.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err
python -c "import json; d=json.load(open('gpurun_out/bench_${N}gpu.json')); print('C2 ${N}gpu', d['ms_per_step'], d['value'], d['e2e']['value'])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus $N --workload T --steps 3 --warmup 3 > gpurun_out/bench_T_${N}gpu.json 2>> gpurun_out/bench_${N}gpu.err
python -c "import json; d=json.load(open('gpurun_out/bench_T_${N}gpu.json')); print('T ${N}gpu', d['ms_per_step'], d['value'])"
