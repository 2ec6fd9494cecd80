#!/bin/bash
mkdir -p gpurun_out/r2
WL="simple:256 mesh_hdri:64 volumes:64 bvh_stress:16 mesh_depth50:128 grid_depth50:16"
{
RTR_LIBRARY=$PWD/variants/librtr_n8.so timeout 600 python tools/ab.py "8_ctas_63_regs" -- $WL 2>&1 | grep Mrays
timeout 600 python tools/ab.py "9_ctas_56_regs" -- $WL 2>&1 | grep Mrays
RTR_LIBRARY=$PWD/variants/librtr_n8.so timeout 600 python tools/ab.py "8_ctas_again" -- $WL 2>&1 | grep Mrays
} >> gpurun_out/r2/ab_extend_occupancy_two_sequences.txt 2>&1
tail -18 gpurun_out/r2/ab_extend_occupancy_two_sequences.txt | awk '{print $1, $2, $7}'
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/r2/run38_pytest.txt
