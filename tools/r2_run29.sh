#!/bin/bash
# two wave sequences on two streams (RTR_WF_LANES): A/B on every wavefront workload, then the GPU suite
mkdir -p gpurun_out/r2
{
timeout 900 python tools/ab.py "one_sequence:RTR_WF_LANES=1" "two_sequences:RTR_WF_LANES=2,RTR_WF_LANES_MIN_DEPTH=0" -- mesh_depth50:64 grid_depth50:16 volumes:64 mesh_hdri:64 simple:128 bvh_stress:16 2>&1 | grep -v "^$"
} > gpurun_out/r2/ab_wave_sequences.txt 2>&1
cat gpurun_out/r2/ab_wave_sequences.txt
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/r2/run29_pytest.txt
RTR_WF_LANES_MIN_DEPTH=0 timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/r2/run29_pytest_lanes_forced.txt
