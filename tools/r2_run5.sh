#!/bin/bash
set -x
mkdir -p gpurun_out/r2
timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | tail -8 > gpurun_out/r2/pytest_run5.txt
cat gpurun_out/r2/pytest_run5.txt
timeout 600 python - > gpurun_out/r2/trace_batches.json 2> gpurun_out/r2/trace_batches.err <<'PY'
import json, sys, argparse
sys.argv = ["bench.py", "--no-cpu"]
import bench
args = argparse.Namespace(gpus=1, steps=1, warmup=1, impl="b200", workload="benchmark", spp=0, cpu_spp=16, no_cpu=True,
                          no_workloads=True, kernel=0, scaling="strong")
b = bench.Bench(args)
print(json.dumps(bench.measure_trace_batches(b)))
PY
cat gpurun_out/r2/trace_batches.json; tail -5 gpurun_out/r2/trace_batches.err
