#!/bin/bash
set -x
mkdir -p gpurun_out/r2
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2/pytest_noalloc_alone.txt 2>&1
tail -15 gpurun_out/r2/pytest_noalloc_alone.txt
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -15 > gpurun_out/r2/pytest_run6_full.txt
cat gpurun_out/r2/pytest_run6_full.txt
