#!/bin/bash
# 4-GPU confirmation of the tile + sample grid: tests, C3 as a 2x2 grid vs pure sample sharding
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/pytest_mgpu10.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_mgpu10.log
tail -3 gpurun_out/pytest_mgpu10.log
for sh in 2x2 samples 4x1; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 4 --workload C3 --sharding $sh --steps 3 --warmup 3 > gpurun_out/bench_c3_4gpu_$sh.json 2>> gpurun_out/bench10.err
python -c "import json; d=json.load(open('gpurun_out/bench_c3_4gpu_$sh.json')); print('C3 4gpu $sh', d['ms_per_step'], d['value'], d['e2e']['value'], d['config']['sharding'][:70])"
done
