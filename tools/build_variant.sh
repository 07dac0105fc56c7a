#!/bin/bash
# Builds gpurun_variants/libmpccbf_<name>.so: the current objects with impc_thread.cu recompiled with extra flags.
# Usage: tools/build_variant.sh <name> [extra nvcc flags...]   (run `make -C mpc-cbf_b200/csrc` first)
set -e
NAME=$1; shift
cd "$(dirname "$0")/../mpc-cbf_b200/csrc"
mkdir -p ../../gpurun_variants /tmp/exp
ARCH="-gencode arch=compute_100a,code=sm_100a"
nvcc $ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xptxas -v "$@" -c -o /tmp/exp/impc_thread_$NAME.o impc_thread.cu 2> /tmp/exp/impc_thread_$NAME.log || { cat /tmp/exp/impc_thread_$NAME.log; exit 1; }
grep -A2 "impc_thread_kernelILb0ELb0ELi0ELi0" /tmp/exp/impc_thread_$NAME.log | grep -E "stack|Used" 
EXTRA=""
if echo "$@" | grep -q MPCCBF_WITH_R1; then
  [ -f /tmp/exp/impc_thread_r1.o ] || nvcc $ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -I. -I../../include -c -o /tmp/exp/impc_thread_r1.o ../../tools/r1_kernel/impc_thread_r1.cu
  EXTRA=/tmp/exp/impc_thread_r1.o
fi
OBJS=$(ls build/*.o | grep -v "build/impc_thread.o")
nvcc $ARCH -shared -o ../../gpurun_variants/libmpccbf_$NAME.so $OBJS /tmp/exp/impc_thread_$NAME.o $EXTRA -lcudart -ldl
ls -la ../../gpurun_variants/libmpccbf_$NAME.so
