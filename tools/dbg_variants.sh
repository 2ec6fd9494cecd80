#!/bin/bash
for v in "$@"; do
  rm -f raytracingrust_b200/csrc/build/kernels.o
  make -s -C raytracingrust_b200/csrc EXTRA="$v" 2>&1 | grep -E "error"
  echo "== $v: $(python tools/dbg_bounds.py 2>&1 | tail -2 | tr '\n' ' ' | cut -c1-200)"
done
