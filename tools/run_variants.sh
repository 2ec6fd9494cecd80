#!/bin/bash
# usage (GPU box): tools/run_variants.sh "A B C" [workload:spp ...]
names=$1; shift
wl=${@:-"simple:128 mesh_hdri:32 volumes:8 bvh_stress:16"}
cp raytracingrust_b200/librtr_b200.so /tmp/librtr_default.so
for v in $names; do
  cp variants/librtr_$v.so raytracingrust_b200/librtr_b200.so
  for w in $wl; do
    python bench.py --workload ${w%%:*} --spp ${w##*:} --steps 3 --warmup 1 --no-cpu $BENCH_EXTRA 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v', '${w%%:*}', d['config']['kernel'], 'Mrays/s %.0f kernel_ms %.2f'%(d['value'], d['roofline']['kernel_ms']))"
  done
done
cp /tmp/librtr_default.so raytracingrust_b200/librtr_b200.so
