#!/bin/bash
# Round-end measurement run (GPU box): tests, the default bench with both arms, every workload at full size with
# both kernels, ncu launch lists of the bench commands and full captures of the dominant kernels.
# Outputs under gpurun_out/final_*.
set -x
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -4 > gpurun_out/final_pytest.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.txt 2>&1
python bench.py > gpurun_out/final_bench_default.json 2> gpurun_out/final_bench_default.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_reference.json 2> gpurun_out/final_bench_reference.err
for w in simple mesh_hdri volumes bvh_stress benchmark_bvh; do
  python bench.py --workload $w --steps 5 --warmup 3 > gpurun_out/final_bench_$w.json 2> gpurun_out/final_bench_$w.err
done
for w in simple mesh_hdri volumes bvh_stress; do
  python bench.py --workload $w --steps 5 --warmup 3 --no-cpu --kernel 1 > gpurun_out/final_bench_${w}_mega.json 2>/dev/null
done
python bench.py --workload benchmark --steps 5 --warmup 3 --no-cpu --kernel 2 > gpurun_out/final_bench_benchmark_wf.json 2>/dev/null
RTR_SCAN_KERNEL=0 python bench.py --workload benchmark --steps 5 --warmup 3 --no-cpu > gpurun_out/final_bench_benchmark_generic.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/final_launches_benchmark.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/final_ncu_bench.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/final_launches_mesh_hdri.csv python bench.py --workload mesh_hdri --spp 16 --steps 2 --warmup 1 --no-cpu > gpurun_out/final_ncu_mesh.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_scan -s 1 -c 1 -f -o gpurun_out/final_prof_benchmark_scan python bench.py --steps 1 --warmup 1 --no-cpu --spp 128 > gpurun_out/final_prof_scan.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:wf_(extend|shade)' -s 24 -c 12 -f -o gpurun_out/final_prof_mesh_wave python bench.py --workload mesh_hdri --spp 16 --steps 1 --warmup 1 --no-cpu > gpurun_out/final_prof_mesh.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:wf_extend' -s 32 -c 3 -f -o gpurun_out/final_prof_volumes_extend python bench.py --workload volumes --spp 16 --steps 1 --warmup 1 --no-cpu > gpurun_out/final_prof_volumes.log 2>&1
tail -3 gpurun_out/final_pytest.txt
