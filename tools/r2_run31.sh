#!/bin/bash
# number of wave sequences and wave size
mkdir -p gpurun_out/r2
{
timeout 1500 python tools/ab.py "1_seq:RTR_WF_LANES=1" "2_seq:RTR_WF_LANES=2" "3_seq:RTR_WF_LANES=3" "4_seq:RTR_WF_LANES=4" "2_seq_2^25:RTR_WF_LANES=2,RTR_WF_PATHS=33554432" "3_seq_2^23:RTR_WF_LANES=3,RTR_WF_PATHS=8388608" -- volumes:128 mesh_hdri:128 simple:256 bvh_stress:16 mesh_depth50:128 2>&1 | grep -v "^$"
} > gpurun_out/r2/ab_wave_sequences_count.txt 2>&1
cat gpurun_out/r2/ab_wave_sequences_count.txt
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r2/run31_pytest.txt
RTR_WF_LANES=3 timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r2/run31_pytest_3seq.txt
