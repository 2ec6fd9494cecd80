#!/bin/bash
# 4-wide traversal tree (RTR_WF_WIDE=1): same-box A/B against the 2-wide walk, then the GPU suite with it switched on
mkdir -p gpurun_out/r2
{
timeout 900 python tools/ab.py "bvh2" "bvh4:RTR_WF_WIDE=1" -- mesh_hdri:32 volumes:16 simple:64 mesh_depth50:16 grid_depth50:8 2>&1 | grep -v "^$"
RTR_BVH_BUILD=host timeout 600 python tools/ab.py "bvh2_hostbuild" "bvh4_hostbuild:RTR_WF_WIDE=1" -- bvh_stress:16 2>&1 | grep -v "^$"
} > gpurun_out/r2/ab_bvh4.txt 2>&1
cat gpurun_out/r2/ab_bvh4.txt
RTR_WF_WIDE=1 timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/r2/pytest_gpu_wide4.txt
