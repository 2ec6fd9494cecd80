#!/bin/bash
# GPU box: scan-kernel parity tests, then A/B of the scan kernel's occupancy variants against the generic megakernel
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_scan.py -x -q 2>&1 | tail -15 > gpurun_out/ab_scan_pytest.txt
cat gpurun_out/ab_scan_pytest.txt
{
echo "== generic megakernel (RTR_SCAN_KERNEL=0)"
RTR_SCAN_KERNEL=0 python bench.py --spp 256 --steps 3 --warmup 2 --no-cpu 2>/dev/null | tail -1
echo "== scan8 (default build)"
python bench.py --spp 256 --steps 3 --warmup 2 --no-cpu 2>/dev/null | tail -1
} > gpurun_out/ab_scan_bench.txt
BENCH_EXTRA="" tools/run_variants.sh "scan7 scan6 scan5" benchmark:256 >> gpurun_out/ab_scan_bench.txt 2>&1
cat gpurun_out/ab_scan_bench.txt
