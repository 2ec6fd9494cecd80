#!/bin/bash
# leaf postponement in the 4-wide walk against the same build without it; GPU suite
mkdir -p gpurun_out/r2
WL="mesh_hdri:32 volumes:16 simple:64 mesh_depth50:16 bvh_stress:16 grid_depth50:8"
{
RTR_LIBRARY=$PWD/variants/librtr_nopostpone.so timeout 600 python tools/ab.py "no_postpone" -- $WL 2>&1 | grep Mrays
timeout 600 python tools/ab.py "postpone" "bvh2:RTR_WF_WIDE=0" -- $WL 2>&1 | grep Mrays
} > gpurun_out/r2/ab_postpone.txt 2>&1
cat gpurun_out/r2/ab_postpone.txt
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/r2/run21_pytest.txt
