#!/bin/bash
set -x
mkdir -p gpurun_out/r2
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -12 > gpurun_out/r2/pytest_run10.txt
cat gpurun_out/r2/pytest_run10.txt
timeout 1200 python tools/ab.py "old:RTR_WF_FETCH_CHUNK=0,RTR_WF_ATOMICS=0" "gen:RTR_WF_FETCH_CHUNK=0,RTR_WF_ATOMICS=1" "gen_shade:RTR_WF_FETCH_CHUNK=0,RTR_WF_ATOMICS=3" \
  "c32:RTR_WF_FETCH_CHUNK=32,RTR_WF_ATOMICS=3" "c64:RTR_WF_FETCH_CHUNK=64,RTR_WF_ATOMICS=3" "c128:RTR_WF_FETCH_CHUNK=128,RTR_WF_ATOMICS=3" "c256:RTR_WF_FETCH_CHUNK=256,RTR_WF_ATOMICS=3" \
  "c64_noshade:RTR_WF_FETCH_CHUNK=64,RTR_WF_ATOMICS=1" \
  -- simple:64 mesh_hdri:32 volumes:16 bvh_stress:16 mesh_depth50:32 > gpurun_out/r2/ab_atomics.txt 2>&1
cat gpurun_out/r2/ab_atomics.txt
