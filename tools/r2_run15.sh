#!/bin/bash
# 8-GPU box: group tests, f4 A/B (static vs dynamic split, p2p vs nccl), strong-scaling bench at N=8 and N=2 on the new kernels
set -x
mkdir -p gpurun_out/r2
timeout 600 python -m pytest tests/test_gpu_group.py -x -q -m gpu 2>&1 | tail -5 > gpurun_out/r2/pytest_group_n8_b.txt
cat gpurun_out/r2/pytest_group_n8_b.txt
timeout 900 python tools/ab_split.py volumes:256 simple:256 mesh_hdri:128 benchmark:1024 > gpurun_out/r2/ab_split_n8.txt 2>&1
cat gpurun_out/r2/ab_split_n8.txt
for n in 8 2; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/r2/bench_n${n}_b.json 2> gpurun_out/r2/bench_n${n}_b.err
  tail -c 300 gpurun_out/r2/bench_n${n}_b.err
done
