#!/bin/bash
mkdir -p gpurun_out/r2
for rep in 1 2 3; do
  RTR_LIBRARY=$PWD/variants/librtr_n8.so timeout 600 python tools/ab.py "8_ctas_rep$rep" -- simple:256 simple:1024 2>&1 | grep Mrays
  timeout 600 python tools/ab.py "9_ctas_rep$rep" -- simple:256 simple:1024 2>&1 | grep Mrays
done | awk '{print $1, $2, $4, $7}' | tee -a gpurun_out/r2/ab_extend_occupancy_two_sequences.txt
