#!/bin/bash
mkdir -p gpurun_out/r2
: > gpurun_out/r2/ab_shade_cta.txt
for v in s128 s256_4 s256_5 s512_2 s1024_1; do
  RTR_LIBRARY=$PWD/variants/librtr_$v.so timeout 600 python tools/ab.py "$v" -- simple:64 mesh_hdri:32 volumes:16 bvh_stress:16 mesh_depth50:32 2>&1 | grep Mrays >> gpurun_out/r2/ab_shade_cta.txt
done
cat gpurun_out/r2/ab_shade_cta.txt
