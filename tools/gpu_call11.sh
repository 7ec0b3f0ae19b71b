#!/bin/bash
# interleaved rows: single-GPU parity (runs on GPU 0), then the 2-rank and 4-rank exchanges, then C3 at 4 GPUs per sharding
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu11.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu11.log
tail -4 gpurun_out/pytest_gpu11.log
for sh in samples 2x2 4x1; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 4 --workload C3 --sharding $sh --steps 3 --warmup 3 > gpurun_out/bench_c3_4gpu_${sh}_il.json 2>> gpurun_out/bench11.err
python -c "import json; d=json.load(open('gpurun_out/bench_c3_4gpu_${sh}_il.json')); print('C3 4gpu $sh', d['ms_per_step'], d['value'], d['e2e']['value'], d['config']['sharding'][:70])"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus 4 --workload C3 --sharding 2x2 --exchange nccl --steps 3 --warmup 3 > gpurun_out/bench_c3_4gpu_2x2_nccl.json 2>> gpurun_out/bench11.err
python -c "import json; d=json.load(open('gpurun_out/bench_c3_4gpu_2x2_nccl.json')); print('C3 4gpu 2x2 nccl', d['ms_per_step'], d['value'], d['e2e']['value'], d['config']['sharding'][:70])"
