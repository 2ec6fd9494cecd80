#!/bin/bash
# wf_shade under two wave sequences: L2 prefetch of the next trip, occupancy 3 / 4 / 6 CTAs of 256 threads
mkdir -p gpurun_out/r2
WL="simple:256 mesh_hdri:64 volumes:64 bvh_stress:16"
{
timeout 600 python tools/ab.py "default_4cta" -- $WL 2>&1 | grep Mrays
for v in shpf sh3 sh6; do
  RTR_LIBRARY=$PWD/variants/librtr_$v.so timeout 600 python tools/ab.py "$v" -- $WL 2>&1 | grep Mrays
done
timeout 600 python tools/ab.py "default_again" -- $WL 2>&1 | grep Mrays
} > gpurun_out/r2/ab_shade_prefetch_occupancy.txt 2>&1
cat gpurun_out/r2/ab_shade_prefetch_occupancy.txt
