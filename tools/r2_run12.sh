#!/bin/bash
set -x
mkdir -p gpurun_out/r2
python bench.py --steps 5 --warmup 3 > gpurun_out/r2/bench_n1_b.json 2> gpurun_out/r2/bench_n1_b.err
tail -c 300 gpurun_out/r2/bench_n1_b.err
bash tools/r2_run9.sh > gpurun_out/r2/run9b.log 2>&1
tail -3 gpurun_out/r2/run9b.log
