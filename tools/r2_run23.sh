#!/bin/bash
mkdir -p gpurun_out/r2
timeout 900 python -m pytest tests/test_gpu_bvh_build.py -m gpu -x -q 2>&1 | tail -25 | tee gpurun_out/r2/run23_pytest_bvh_build.txt
