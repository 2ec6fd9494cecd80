#!/bin/bash
mkdir -p gpurun_out/r2
timeout 900 python tools/ab.py "2_seq" "3_seq:RTR_WF_LANES=3" "4_seq:RTR_WF_LANES=4" "4_seq_2^24:RTR_WF_LANES=4,RTR_WF_PATHS=16777216" -- mesh_depth50:256 grid_depth50:32 volumes:128 2>&1 | grep -v "^$" > gpurun_out/r2/ab_wave_sequences_deep.txt
cat gpurun_out/r2/ab_wave_sequences_deep.txt
