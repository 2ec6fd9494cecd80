#!/bin/bash
# ncu: whole-render counters per workload (metrics list), one --set full capture of wf_extend on C4
set -x
mkdir -p gpurun_out/r2/ncu
MET=gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct
for ws in benchmark:64 simple:8 mesh_hdri:8 volumes:8 bvh_stress:4 mesh_depth50:8; do
  w=${ws%%:*}; s=${ws##*:}
  python tools/profile_one.py $w $s > gpurun_out/r2/ncu/$w.json 2> gpurun_out/r2/ncu/$w.err &&
  timeout 900 ncu --metrics $MET --clock-control none --profile-from-start off -f -o gpurun_out/r2/ncu/$w python tools/profile_one.py $w $s > gpurun_out/r2/ncu/$w.ncu.log 2>&1
  tail -2 gpurun_out/r2/ncu/$w.ncu.log
done
python tools/profile_one.py mesh_hdri 8 > /dev/null 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:wf_extend -c 3 -f -o gpurun_out/r2/ncu_full_mesh_hdri_extend python tools/profile_one.py mesh_hdri 8 > gpurun_out/r2/ncu_full.log 2>&1
tail -3 gpurun_out/r2/ncu_full.log
ls -la gpurun_out/r2/ncu gpurun_out/r2/*.ncu-rep
