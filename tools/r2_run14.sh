#!/bin/bash
mkdir -p gpurun_out/r2
: > gpurun_out/r2/ab_scan_occupancy.txt
for v in scan7 scan6 scan5 scan8 scan7; do
  RTR_LIBRARY=$PWD/variants/librtr_$v.so timeout 600 python tools/ab.py "$v" -- benchmark:1024 benchmark:128 2>&1 | grep Mrays >> gpurun_out/r2/ab_scan_occupancy.txt
done
cat gpurun_out/r2/ab_scan_occupancy.txt
