#!/bin/bash
# closing multi-GPU record on the final build: bench.py at N = $1 (strong scaling), group tests first when N = 8
N=${1:-8}
mkdir -p gpurun_out/r2
if [ $N -ge 8 ]; then timeout 600 python -m pytest tests/test_gpu_group.py -x -q -m gpu 2>&1 | tail -3 | tee gpurun_out/r2/run42_pytest_group_n$N.txt; fi
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$N bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2/run42_bench_n${N}.json 2> gpurun_out/r2/run42_bench_n${N}.err
tail -c 200 gpurun_out/r2/run42_bench_n${N}.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r2/run42_bench_n${N}.json').read().strip().splitlines()[-1])
print('N=${N}', round(d['value']), round(d['e2e']['value']), d['clocks'].get('reasons'), d['per_rank'])
for k,w in d['workloads'].items(): print('  ', k, w.get('value') and round(w['value']), (w.get('e2e') or {}).get('value') and round(w['e2e']['value']))
PY
