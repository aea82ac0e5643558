#!/bin/bash
# Development: link an experimental library from the built objects with build.cu recompiled under extra flags.
#   tools/build_variant.sh NAME -DMTB_EXP_...   ->  mpitree_b200/lib/exp/libNAME.so  (select with MTB_LIB=<path>)
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p mpitree_b200/lib/exp
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -I include "$@" \
  -c mpitree_b200/csrc/build.cu -o mpitree_b200/lib/exp/build_$name.o
objs=$(ls mpitree_b200/lib/obj/*.o | grep -v '/build.o')
nvcc -shared -o mpitree_b200/lib/exp/lib$name.so mpitree_b200/lib/exp/build_$name.o $objs
rm mpitree_b200/lib/exp/build_$name.o
echo mpitree_b200/lib/exp/lib$name.so
