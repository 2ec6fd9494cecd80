#!/bin/bash
# round 2, GPU call 2 (--gpus 2): multi-device tests, 2-rank bench (strong scaling, group e2e)
set -x
mkdir -p gpurun_out/r2
nvidia-smi -L > gpurun_out/r2/gpus_n2.txt
nvidia-smi topo -m >> gpurun_out/r2/gpus_n2.txt 2>&1
python -m pytest tests/test_gpu_group.py -x -q -m gpu 2>&1 | tail -15 > gpurun_out/r2/pytest_n2.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2/bench_n2.json 2> gpurun_out/r2/bench_n2.err
tail -5 gpurun_out/r2/pytest_n2.txt; tail -c 1500 gpurun_out/r2/bench_n2.err; tail -c 300 gpurun_out/r2/bench_n2.json
