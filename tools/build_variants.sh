#!/bin/bash
# Tuning aid: build several variants of the library that differ in -D flags of the stage kernels.
# usage: tools/build_variants.sh name1:"-DFLAG_A" name2:"-DFLAG_A -DFLAG_B" ...   -> tools/ab_libs/<name>.so
set -e
cd "$(dirname "$0")/../sharding-the-sphere-fall-2025-jax-devlab-examples_b200/csrc"
make -j8 >/dev/null
mkdir -p ../../tools/ab_libs
ARCH="-gencode arch=compute_100a,code=sm_100a"
FLAGS="$ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -I../../include -I."
for spec in "$@"; do
  name="${spec%%:*}"; defs="${spec#*:}"
  dir=/tmp/cs_variant_$name; mkdir -p $dir
  ( nvcc $FLAGS $defs -c cs_stage.cu -o $dir/cs_stage.o ) &
  ( nvcc $FLAGS $defs -c cs_step_kernel.cu -o $dir/cs_step_kernel.o ) &
done
wait
for spec in "$@"; do
  name="${spec%%:*}"; dir=/tmp/cs_variant_$name
  objs=$(ls build/*.o | grep -v "cs_stage.o\|cs_step_kernel.o")
  nvcc $ARCH -shared -o ../../tools/ab_libs/$name.so $objs $dir/cs_stage.o $dir/cs_step_kernel.o -lcudart_static -ldl -lrt -lpthread
  echo built tools/ab_libs/$name.so
done
