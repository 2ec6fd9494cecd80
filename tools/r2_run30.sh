#!/bin/bash
# where does the two-sequence gain come from: wave size or overlap?
mkdir -p gpurun_out/r2
{
timeout 1200 python tools/ab.py "one_seq_2^24:RTR_WF_LANES=1" "one_seq_2^23:RTR_WF_LANES=1,RTR_WF_PATHS=8388608" "two_seq_2x2^23:RTR_WF_LANES=2,RTR_WF_LANES_MIN_DEPTH=0" "two_seq_2x2^24:RTR_WF_LANES=2,RTR_WF_LANES_MIN_DEPTH=0,RTR_WF_PATHS=33554432" "two_seq_2x2^22:RTR_WF_LANES=2,RTR_WF_LANES_MIN_DEPTH=0,RTR_WF_PATHS=8388608" -- volumes:64 mesh_hdri:64 simple:128 bvh_stress:16 2>&1 | grep -v "^$"
} > gpurun_out/r2/ab_wave_sequences_sizes.txt 2>&1
cat gpurun_out/r2/ab_wave_sequences_sizes.txt
