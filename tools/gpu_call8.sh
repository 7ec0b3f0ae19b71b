#!/bin/bash
# 2-GPU confirmation: multi-GPU tests, weak-scaling bench line, reference arm under torchrun
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/pytest_mgpu8.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_mgpu8.log
tail -3 gpurun_out/pytest_mgpu8.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_2gpu8.json 2> gpurun_out/bench_2gpu8.err
python -c "import json; d=json.load(open('gpurun_out/bench_2gpu8.json')); print('2gpu', d['ms_per_step'], d['value'], d['e2e']['value'], d['config']['sharding'][:60])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload T --steps 3 --warmup 3 > gpurun_out/bench_T_2gpu8.json 2>> gpurun_out/bench_2gpu8.err
python -c "import json; d=json.load(open('gpurun_out/bench_T_2gpu8.json')); print('T 2gpu', d['ms_per_step'], d['value'], d['scaling'])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > gpurun_out/bench_ref_2gpu8.json 2>> gpurun_out/bench_2gpu8.err
head -c 200 gpurun_out/bench_ref_2gpu8.json
