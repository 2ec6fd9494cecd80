#!/bin/bash
# usage: tools/build_variants.sh NAME="-DFLAG=1 ..." ...   builds variants/librtr_NAME.so (here, no GPU needed)
# on the GPU box: tools/run_variants.sh "NAME ..." [workload:spp ...]
mkdir -p variants
for kv in "$@"; do
  name=${kv%%=*}; flags=${kv#*=}
  rm -f raytracingrust_b200/csrc/build/kernels.o raytracingrust_b200/csrc/build/scan.o raytracingrust_b200/csrc/build/wavefront.o raytracingrust_b200/csrc/build/capi.o
  make -s -C raytracingrust_b200/csrc EXTRA="$flags" 2>&1 | grep -E "error|rror:" 
  cp raytracingrust_b200/librtr_b200.so variants/librtr_$name.so
  echo "built $name ($flags)"
done
rm -f raytracingrust_b200/csrc/build/kernels.o raytracingrust_b200/csrc/build/scan.o raytracingrust_b200/csrc/build/wavefront.o raytracingrust_b200/csrc/build/capi.o
make -s -C raytracingrust_b200/csrc 2>&1 | grep -E "rror"
