#!/bin/bash
set -x
mkdir -p gpurun_out/r2
timeout 600 python -m pytest tests/test_gpu_scan.py tests/test_gpu_render.py tests/test_gpu_edge.py -x -q -m gpu 2>&1 | tail -8 > gpurun_out/r2/pytest_run4.txt
cat gpurun_out/r2/pytest_run4.txt
RTR_LIBRARY=$PWD/variants/librtr_base.so timeout 300 python tools/ab.py "base" -- benchmark:1024 benchmark:128 > gpurun_out/r2/ab_scan.txt 2>&1
timeout 300 python tools/ab.py "new_const" "new_global:RTR_SCAN_KERNEL=2" -- benchmark:1024 benchmark:128 >> gpurun_out/r2/ab_scan.txt 2>&1
cat gpurun_out/r2/ab_scan.txt
