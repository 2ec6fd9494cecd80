#!/bin/bash
# GPU box: wf_extend occupancy variants and refill thresholds on the no-Volume instantiation
tools/run_variants.sh "new wf6 wf7 wf9 wf10" simple:128 mesh_hdri:32 bvh_stress:16 2>&1 | tee gpurun_out/sweep_wf.txt
for t in 8 12 20 24; do
  echo "== RTR_WF_REFILL_THRESHOLD=$t" | tee -a gpurun_out/sweep_wf.txt
  RTR_WF_REFILL_THRESHOLD=$t tools/run_variants.sh "new" simple:128 mesh_hdri:32 bvh_stress:16 2>&1 | tee -a gpurun_out/sweep_wf.txt
done
