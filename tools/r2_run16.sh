#!/bin/bash
mkdir -p gpurun_out/r2
for c in none 0 25 32 40 50 100; do
  if [ $c = none ]; then unset RTR_WF_CARVEOUT; else export RTR_WF_CARVEOUT=$c; fi
  timeout 600 python tools/ab.py "carveout_$c" -- mesh_hdri:32 volumes:16 bvh_stress:16 simple:64 2>&1 | grep Mrays
done > gpurun_out/r2/ab_carveout.txt
cat gpurun_out/r2/ab_carveout.txt
