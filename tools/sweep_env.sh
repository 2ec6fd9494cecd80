#!/bin/bash
# usage: tools/sweep_env.sh VAR "v1 v2 ..." [workload:spp ...]   (on the GPU box: short benches per env value)
var=$1; shift; vals=$1; shift
wl=${@:-"benchmark:128 simple:128 mesh_hdri:32 volumes:8"}
for v in $vals; do
  for w in $wl; do
    env $var=$v python bench.py --workload ${w%%:*} --spp ${w##*:} --steps 2 --warmup 1 --no-cpu $BENCH_EXTRA 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$var=$v', '${w%%:*}', 'Mrays/s %.0f kernel_ms %.2f'%(d['value'], d['roofline']['kernel_ms']))"
  done
done
