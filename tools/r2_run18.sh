#!/bin/bash
# 4-wide walk: occupancy (launch bounds) and short-stack variants, each library in its own process on the same box
mkdir -p gpurun_out/r2
WL="mesh_hdri:32 volumes:16 simple:64 mesh_depth50:16"
{
timeout 600 python tools/ab.py "default_n8_v7_ss8" -- $WL 2>&1 | grep Mrays
for v in n9 n7 v6 v8 ss12 ss16; do
  RTR_LIBRARY=$PWD/variants/librtr_$v.so timeout 600 python tools/ab.py "$v" -- $WL 2>&1 | grep Mrays
done
timeout 600 python tools/ab.py "default_again" "bvh2:RTR_WF_WIDE=0" "thr12:RTR_WF_REFILL_THRESHOLD=12" "thr20:RTR_WF_REFILL_THRESHOLD=20" "thr24:RTR_WF_REFILL_THRESHOLD=24" -- $WL 2>&1 | grep Mrays
} > gpurun_out/r2/ab_bvh4_variants.txt 2>&1
cat gpurun_out/r2/ab_bvh4_variants.txt
