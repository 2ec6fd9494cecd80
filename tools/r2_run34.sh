#!/bin/bash
mkdir -p gpurun_out/r2
python bench.py --steps 5 --warmup 3 > gpurun_out/r2/run44_bench_n1.json 2> gpurun_out/r2/run44_bench_n1.err
tail -c 300 gpurun_out/r2/run44_bench_n1.err
python - <<'PY'
import json
b=json.loads(open('gpurun_out/r2/run44_bench_n1.json').read().strip().splitlines()[-1])
print('C2', round(b['value']), round(b['e2e']['value']), b['clocks']['reasons'], 'launches', b['gpu_launches'], 'roofline', b['roofline']['bound'], round(b['roofline']['frac'],3))
for k,w in b['workloads'].items():
    if k=='trace_batches': print(k, round(w['primary']['value']), round(w['secondary']['value'])); continue
    print(k, round(w['value']), round(w['e2e']['value']), round(w['ms_per_step'],1), round(w['per_rank']['render_ms_max'],1), w['roofline']['bound'])
PY
