#!/bin/bash
# closing N=1 measurements on the final build: suite, smoke, bench, reference arm, launch list of the bench command, whole-render counters
mkdir -p gpurun_out/r2/ncu
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r2/run40_pytest.txt
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3 | tee gpurun_out/r2/run40_smoke.txt
python bench.py --steps 5 --warmup 3 > gpurun_out/r2/run40_bench_n1.json 2> gpurun_out/r2/run40_bench_n1.err
tail -c 300 gpurun_out/r2/run40_bench_n1.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2/run40_bench_reference.json 2> gpurun_out/r2/run40_bench_reference.err
tail -c 200 gpurun_out/r2/run40_bench_reference.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2/run40_launches_bench_default.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/r2/run40_ncu_bench.log 2>&1
tail -2 gpurun_out/r2/run40_ncu_bench.log
MET=gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct
for ws in benchmark:64 simple:8 mesh_hdri:8 volumes:8 bvh_stress:4 mesh_depth50:8; do
  w=${ws%%:*}; s=${ws##*:}
  python tools/profile_one.py $w $s > gpurun_out/r2/ncu/$w.json 2> gpurun_out/r2/ncu/$w.err &&
  timeout 900 ncu --metrics $MET --clock-control none --profile-from-start off -f -o gpurun_out/r2/ncu/$w python tools/profile_one.py $w $s > gpurun_out/r2/ncu/$w.ncu.log 2>&1
  tail -1 gpurun_out/r2/ncu/$w.ncu.log
done
