#!/bin/bash
# usage: tools/sweep_build.sh "<EXTRA nvcc flags 1>" "<flags 2>" ...  (on the GPU box: rebuild kernels + short benches, both kernels)
for v in "$@"; do
  rm -f raytracingrust_b200/csrc/build/kernels.o raytracingrust_b200/csrc/build/wavefront.o
  make -s -C raytracingrust_b200/csrc EXTRA="$v" 2>&1 | grep -E "error"
  grep -h "Used" raytracingrust_b200/csrc/build/ptxas_kernels.log | sed -n 3p
  grep -h "Used" raytracingrust_b200/csrc/build/ptxas_wavefront.log | sed -n 3p
  for k in 1 2; do
  for w in ${WL:-benchmark:128 simple:128 mesh_hdri:32 volumes:8}; do
    python bench.py --workload ${w%%:*} --spp ${w##*:} --steps 2 --warmup 1 --no-cpu --kernel $k 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('[$v] k=$k', '${w%%:*}', 'Mrays/s %.0f kernel_ms %.2f'%(d['value'], d['roofline']['kernel_ms']))"
  done; done
done
