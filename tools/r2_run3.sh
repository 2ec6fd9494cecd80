#!/bin/bash
set -x
mkdir -p gpurun_out/r2
python -m pytest tests/test_gpu_wavefront.py tests/test_gpu_render.py tests/test_gpu_group.py -x -q -m gpu 2>&1 | tail -8 > gpurun_out/r2/pytest_run3.txt
python tools/ab.py "off:RTR_WF_BIN=0" "oct8:RTR_WF_BIN=8" "oct24:RTR_WF_BIN=24" "oct8_b1:RTR_WF_BIN=8,RTR_WF_BIN_LAST=1" "oct8_b2:RTR_WF_BIN=8,RTR_WF_BIN_LAST=2" -- mesh_hdri:64 volumes:32 simple:128 bvh_stress:16 mesh_depth50:32 > gpurun_out/r2/ab_bin.txt 2>&1
cat gpurun_out/r2/pytest_run3.txt; cat gpurun_out/r2/ab_bin.txt
