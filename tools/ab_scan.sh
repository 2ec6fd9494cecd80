#!/bin/bash
# GPU box: full GPU suite, then A/B of the scan kernel (default build, 7 CTAs/SM variant) and the other workloads
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -8 ) > gpurun_out/ab_pytest.txt 2>&1
cat gpurun_out/ab_pytest.txt
{
echo "== generic megakernel (RTR_SCAN_KERNEL=0)"
RTR_SCAN_KERNEL=0 python bench.py --spp 256 --steps 3 --warmup 2 --no-cpu 2>/dev/null | tail -1 | cut -c1-140
echo "== scan8 (default build)"
python bench.py --spp 256 --steps 3 --warmup 2 --no-cpu 2>/dev/null | tail -1 | cut -c1-140
for w in simple:128 bvh_stress:16 mesh_hdri:32 volumes:8; do
python bench.py --workload ${w%%:*} --spp ${w##*:} --steps 3 --warmup 2 --no-cpu 2>/dev/null | tail -1 | cut -c1-140
done
} > gpurun_out/ab_scan_bench.txt
tools/run_variants.sh "scan7" benchmark:256 >> gpurun_out/ab_scan_bench.txt 2>&1
cat gpurun_out/ab_scan_bench.txt
