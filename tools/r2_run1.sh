#!/bin/bash
# round 2, GPU call 1: tests, smoke, the default bench (with sub-workloads) and the CPU arm, depth-50 scenes
set -x
mkdir -p gpurun_out/r2
nvidia-smi -L > gpurun_out/r2/gpus.txt
python -m pytest tests -x -q -m gpu 2>&1 | tail -15 > gpurun_out/r2/pytest.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2/smoke.txt 2>&1
python bench.py --steps 5 --warmup 3 > gpurun_out/r2/bench_default.json 2> gpurun_out/r2/bench_default.err
for w in mesh_depth50 grid_depth50; do
  for k in 1 2; do
    python bench.py --workload $w --steps 3 --warmup 3 --no-cpu --kernel $k > gpurun_out/r2/bench_${w}_k$k.json 2> gpurun_out/r2/bench_${w}_k$k.err
  done
done
python bench.py --workload mesh_hdri --steps 3 --warmup 3 --no-cpu > gpurun_out/r2/bench_mesh_hdri.json 2> gpurun_out/r2/bench_mesh_hdri.err
python bench.py --workload volumes --steps 3 --warmup 3 --no-cpu > gpurun_out/r2/bench_volumes.json 2> gpurun_out/r2/bench_volumes.err
tail -5 gpurun_out/r2/pytest.txt; cat gpurun_out/r2/smoke.txt | tail -3; tail -c 600 gpurun_out/r2/bench_default.err
