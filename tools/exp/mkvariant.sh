#!/bin/bash
# tools/exp/mkvariant.sh <name> <nvcc flags...>: libme_gpu.so with me_hme.cu / me_kernels.cu rebuilt under extra flags -> tools/exp/lib_<name>.so
# (selected by ME_B200_TRACE_LIB=tools/exp/lib_<name>.so; experiments only, the files are git-ignored)
set -e
name=$1; shift
cd "$(dirname "$0")/../../svt-av1-psyex_b200"
B=build_var_$name
mkdir -p $B
for f in me_kernels me_hme; do
  /usr/local/cuda/bin/nvcc "$@" -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xptxas -v -I../include -Icsrc -c csrc/$f.cu -o $B/$f.o 2> $B/$f.log &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../tools/exp/lib_$name.so $B/me_kernels.o $B/me_hme.o build/me_dg.o build/me_engine.o build/me_derive.o -lcudart
grep -h "registers\|spill" $B/me_hme.log | head -4
