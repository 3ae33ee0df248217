#!/bin/bash
# Experiment builds of the mixture kernels: only mix_kernels.cu is recompiled (bench-only instantiations), the other
# objects come from the default build.  Usage: scripts/build_mix_variants.sh NAME "FLAGS" ...
set -e
B=lmc_atomi_b200/build/_lmc_b200
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  ( nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -diag-suppress 177 \
      -DLMC_MIX_BENCH_ONLY $flags -c lmc_atomi_b200/csrc/mix_kernels.cu -o build_tmp/mix_$name.o &&
    nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build_tmp/lmc_$name.so build_tmp/mix_$name.o \
      $(ls $B/*.o | grep -v mix_kernels.o) -lcudart && echo built $name ) &
done
wait
