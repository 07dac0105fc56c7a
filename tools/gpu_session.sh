#!/bin/bash
# One GPU-box session: tests, bench, occupancy variants, kernel benches, ncu launch list + full capture of the
# fused kernel, reference arm. Everything lands under gpurun_out/<tag>_*. Usage (from the repo root on the box):
#   bash tools/gpu_session.sh <tag>
TAG=${1:-r2x}
O=gpurun_out
mkdir -p $O
echo "== tests"; python -m pytest tests -m gpu -q -rs 2>&1 | tail -40 > $O/${TAG}_tests.log; tail -15 $O/${TAG}_tests.log
echo "== bench"; python bench.py --steps 48 --warmup 8 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; tail -c 600 $O/${TAG}_bench.err; cut -c1-400 $O/${TAG}_bench.json
for c in 2 3 4; do
  MPCCBF_THREAD_CTAS=$c python bench.py --steps 48 --warmup 8 --no-extras --no-cpu-baseline > $O/${TAG}_occ_$c.json 2>> $O/${TAG}_bench.err
done
echo "== kernels"; python bench_kernels.py --out $O/${TAG}_kernels.jsonl > $O/${TAG}_kernels.log 2>&1; tail -3 $O/${TAG}_kernels.log
echo "== ncu launch list"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --steps 4 --warmup 3 --no-extras --no-cpu-baseline --no-graph > $O/${TAG}_ncu_list.log 2>&1
echo "== ncu full"
ncu --set full --metrics smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum \
    --clock-control none --import-source on -k regex:impc_thread_kernel -s 30 -c 1 -f -o $O/${TAG}_fused \
    python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline --no-graph > $O/${TAG}_ncu_full.log 2>&1
tail -2 $O/${TAG}_ncu_full.log
echo "== reference arm"; python bench.py --impl reference --steps 20 --warmup 5 > $O/${TAG}_bench_ref.json 2>> $O/${TAG}_bench.err; cut -c1-300 $O/${TAG}_bench_ref.json
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > $O/${TAG}_smi.csv
echo "== done"
