#!/bin/bash
# GPU box: paths in flight per wave (RTR_WF_PATHS) on the wavefront workloads
for p in 4194304 8388608 16777216 33554432; do
  echo "== RTR_WF_PATHS=$p"
  RTR_WF_PATHS=$p tools/run_variants.sh "new" simple:64 mesh_hdri:32 bvh_stress:16 volumes:16
done 2>&1 | tee gpurun_out/sweep_paths.txt
