#!/bin/bash
# round 2: 8-GPU box — group tests over 8 devices, strong-scaling bench at N=8 and N=4
set -x
mkdir -p gpurun_out/r2
nvidia-smi -L > gpurun_out/r2/gpus_n8.txt
timeout 600 python -m pytest tests/test_gpu_group.py -x -q -m gpu 2>&1 | tail -8 > gpurun_out/r2/pytest_group_n8.txt
cat gpurun_out/r2/pytest_group_n8.txt
for n in 8 4; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/r2/bench_n$n.json 2> gpurun_out/r2/bench_n$n.err
  tail -c 400 gpurun_out/r2/bench_n$n.err
done
ls -la gpurun_out/r2/
