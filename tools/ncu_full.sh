#!/bin/bash
# ncu --set full of one warm launch of the fused kernel for every "name:lib:core:cfg". Usage: bash tools/ncu_full.sh <tag> spec...
TAG=$1; shift
O=gpurun_out; mkdir -p $O
for spec in "$@"; do
  IFS=: read name lib core cfg <<< "$spec"
  export MPCCBF_LIB=$PWD/$lib MPCCBF_THREAD_CORE=$core MPCCBF_THREAD_CFG=$cfg
  [ "$lib" = "-" ] && unset MPCCBF_LIB
  ncu --set full --metrics smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum \
      --clock-control none --import-source on -k regex:impc_thread_kernel -s 30 -c 1 -f -o $O/${TAG}_$name \
      python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline --no-graph --device-only > $O/${TAG}_$name.log 2>&1
  tail -1 $O/${TAG}_$name.log
done
