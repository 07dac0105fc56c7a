#!/bin/bash
# One GPU-box session: kernel variants, tests, bench, kernel benches, ncu launch list + full capture of the fused
# kernel, reference arm. Everything lands under gpurun_out/<tag>_*. Usage (from the repo root on the box):
#   bash tools/gpu_session.sh <tag> [skip_tests]
TAG=${1:-r2x}
O=gpurun_out
mkdir -p $O
echo "== kernel variants (device leg only)"
: > $O/${TAG}_variants.jsonl
for core in 0; do for cfg in 0 2 3 4; do
  MPCCBF_THREAD_CORE=$core MPCCBF_THREAD_CFG=$cfg python bench.py --steps 48 --warmup 8 --device-only --no-extras --no-cpu-baseline >> $O/${TAG}_variants.jsonl 2>> $O/${TAG}_bench.err
done; done
python - <<PY
import json
best=None
for l in open("$O/${TAG}_variants.jsonl"):
    try: d=json.loads(l)
    except ValueError: continue
    print("core", d["variant"]["core"], "cfg", d["variant"]["cfg"], "value %.4g" % d["value"], "kernel_ms %.4f" % d["kernel_ms"])
    if best is None or d["value"] > best["value"]: best=d
open("$O/${TAG}_best.env","w").write("export MPCCBF_THREAD_CORE=%s MPCCBF_THREAD_CFG=%s\n" % (best["variant"]["core"], best["variant"]["cfg"]))
PY
cat $O/${TAG}_best.env; source $O/${TAG}_best.env
if [ -z "$2" ]; then
echo "== tests"; python -m pytest tests -m gpu -q -rs --durations=25 > $O/${TAG}_tests.log 2>&1; grep -E "^FAILED|^ERROR|passed|failed" $O/${TAG}_tests.log | tail -20
fi
echo "== bench (best variant through the environment)"; python bench.py --steps 48 --warmup 8 > $O/${TAG}_bench.json 2>> $O/${TAG}_bench.err; tail -c 600 $O/${TAG}_bench.err; cut -c1-300 $O/${TAG}_bench.json
echo "== kernels"; python bench_kernels.py --out $O/${TAG}_kernels.jsonl > $O/${TAG}_kernels.log 2>&1; tail -3 $O/${TAG}_kernels.log | cut -c1-400
echo "== ncu launch list"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --steps 4 --warmup 3 --no-extras --no-cpu-baseline --no-graph --device-only > $O/${TAG}_ncu_list.log 2>&1
echo "== ncu full"
ncu --set full --metrics smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum \
    --clock-control none --import-source on -k regex:impc_thread_kernel -s 30 -c 1 -f -o $O/${TAG}_fused \
    python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline --no-graph --device-only > $O/${TAG}_ncu_full.log 2>&1
tail -2 $O/${TAG}_ncu_full.log
echo "== ncu full, latency path (one 8-agent scenario: 32 lanes per agent)"
ncu --set full --clock-control none --import-source on -k regex:impc_reduced_kernel -s 5 -c 1 -f -o $O/${TAG}_lanes \
    python bench.py --scenarios 1 --agents 8 --steps 4 --warmup 3 --no-extras --no-cpu-baseline --no-graph --device-only > $O/${TAG}_ncu_lanes.log 2>&1
tail -2 $O/${TAG}_ncu_lanes.log
echo "== ncu full, sub-kernels (rows, rollout, filter, episode, dense)"
ncu --set full --clock-control none -k regex:"cbf_rows_kernel|rollout_kernel|cbf_filter_kernel|episode_advance_kernel|dense_qp_kernel" -c 40 -f -o $O/${TAG}_subkernels \
    python bench_kernels.py --out $O/${TAG}_kernels_ncu.jsonl > $O/${TAG}_ncu_sub.log 2>&1
tail -1 $O/${TAG}_ncu_sub.log
echo "== reference arm"; python bench.py --impl reference --steps 20 --warmup 5 > $O/${TAG}_bench_ref.json 2>> $O/${TAG}_bench.err; cut -c1-200 $O/${TAG}_bench_ref.json
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > $O/${TAG}_smi.csv
echo "== done"
