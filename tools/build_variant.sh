#!/bin/bash
# build a variant of the library with extra -D flags (A/B experiments): tools/build_variant.sh <name> "<flags>"  ->  rebvio_b200/lib_<name>.so (use with RBV_LIB)
set -e
cd "$(dirname "$0")/.."
N=$1; shift
mkdir -p /tmp/rbv_var_$N
for f in rbv_api rbv_frontend rbv_tracker rbv_fusion; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Iinclude -Xcompiler -fPIC $* -c rebvio_b200/csrc/$f.cu -o /tmp/rbv_var_$N/$f.o &
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o rebvio_b200/lib_$N.so /tmp/rbv_var_$N/*.o
echo built rebvio_b200/lib_$N.so
