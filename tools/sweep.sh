#!/bin/bash
# usage: tools/sweep.sh "<EXTRA flags variant 1>" "<variant 2>" ...   (runs on the GPU box: rebuild + short benches)
for v in "$@"; do
  rm -f raytracingrust_b200/csrc/build/kernels.o
  make -s -C raytracingrust_b200/csrc EXTRA="$v" 2>&1 | grep -E "error" 
  for w in benchmark:64 simple:64 mesh_hdri:16 volumes:8; do
    python bench.py --workload ${w%%:*} --spp ${w##*:} --steps 2 --warmup 1 --no-cpu 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v', '${w%%:*}', 'Mrays/s %.0f kernel_ms %.2f'%(d['value'], d['roofline']['kernel_ms']))"
  done
done
