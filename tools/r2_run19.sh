#!/bin/bash
# ncu --set full source-level capture of the 4-wide wf_extend on C4 and C5 (bounces 0..2)
mkdir -p gpurun_out/r2
python tools/profile_one.py mesh_hdri 8 > gpurun_out/r2/prof_plain.log 2>&1 || { tail -5 gpurun_out/r2/prof_plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:wf_extend -c 3 -f -o gpurun_out/r2/ncu_full_mesh_hdri_extend4 python tools/profile_one.py mesh_hdri 8 > gpurun_out/r2/ncu_full4.log 2>&1
tail -3 gpurun_out/r2/ncu_full4.log
python tools/profile_one.py volumes 8 > gpurun_out/r2/prof_plain_v.log 2>&1 || { tail -5 gpurun_out/r2/prof_plain_v.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:wf_extend -c 3 -f -o gpurun_out/r2/ncu_full_volumes_extend4 python tools/profile_one.py volumes 8 > gpurun_out/r2/ncu_full4v.log 2>&1
tail -3 gpurun_out/r2/ncu_full4v.log
ls -la gpurun_out/r2/*.ncu-rep
