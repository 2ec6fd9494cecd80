#!/bin/bash
set -x
mkdir -p gpurun_out/r2
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -15 > gpurun_out/r2/pytest_run8.txt
cat gpurun_out/r2/pytest_run8.txt
