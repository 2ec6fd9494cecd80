#!/bin/bash
# merged pop site (4-wide walk) against the previous build; device-built 4-wide tree against the 2-wide walk; GPU suite
mkdir -p gpurun_out/r2
WL="mesh_hdri:32 volumes:16 simple:64 mesh_depth50:16"
{
RTR_LIBRARY=$PWD/variants/librtr_v8.so timeout 600 python tools/ab.py "before_popmerge" -- $WL 2>&1 | grep Mrays
timeout 600 python tools/ab.py "popmerge" -- $WL 2>&1 | grep Mrays
timeout 600 python tools/ab.py "bvh2_devicebuilt:RTR_WF_WIDE=0" "bvh4_devicebuilt" -- bvh_stress:16 grid_depth50:8 2>&1 | grep -v "^$"
} > gpurun_out/r2/ab_popmerge.txt 2>&1
cat gpurun_out/r2/ab_popmerge.txt
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/r2/run20_pytest.txt
