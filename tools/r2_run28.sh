#!/bin/bash
# closing multi-GPU records on the 4-wide walk: group tests on all GPUs of the box, strong-scaling bench at N = all and N = 2
N=${1:-8}
mkdir -p gpurun_out/r2
nvidia-smi -L | wc -l
timeout 600 python -m pytest tests/test_gpu_group.py -x -q -m gpu 2>&1 | tail -4 | tee gpurun_out/r2/run28_pytest_group_n$N.txt
for n in $N 2; do
  [ $n -gt $N ] && continue
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/r2/run28_bench_n${n}.json 2> gpurun_out/r2/run28_bench_n${n}.err
  tail -c 200 gpurun_out/r2/run28_bench_n${n}.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/r2/run28_bench_n${n}.json').read().strip().splitlines()[-1])
print('N=${n}', round(d['value']), round(d['e2e']['value']), d['clocks'].get('reasons'))
for k,w in d['workloads'].items(): print('  ', k, w.get('value') and round(w['value']), (w.get('e2e') or {}).get('value') and round(w['e2e']['value']))
PY
done
