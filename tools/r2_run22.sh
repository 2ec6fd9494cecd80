#!/bin/bash
# N=1 bench record with the 4-wide walk
mkdir -p gpurun_out/r2
python bench.py --steps 5 --warmup 3 > gpurun_out/r2/bench_n1_bvh4.json 2> gpurun_out/r2/bench_n1_bvh4.err
tail -c 400 gpurun_out/r2/bench_n1_bvh4.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2/bench_n1_bvh4.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d.get('clocks'))
for k,w in d['workloads'].items(): print(k, w.get('value'), (w.get('e2e') or {}).get('value'))
PY
