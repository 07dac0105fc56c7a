#!/bin/bash
# A/B of kernel variants on one box: for every "name:lib:core:cfg" argument a device-only bench line and three warm launches under ncu
# (instruction counts, duration). Usage: bash tools/ab_variants.sh <tag> name:lib:core:cfg ...
TAG=$1; shift
O=gpurun_out; mkdir -p $O
: > $O/${TAG}_ab.jsonl
for spec in "$@"; do
  IFS=: read name lib core cfg <<< "$spec"
  export MPCCBF_LIB=$PWD/$lib MPCCBF_THREAD_CORE=$core MPCCBF_THREAD_CFG=$cfg
  [ "$lib" = "-" ] && unset MPCCBF_LIB
  for rep in 1 2; do
    python bench.py --steps 48 --warmup 8 --device-only --no-extras --no-cpu-baseline 2>> $O/${TAG}_ab.err | python -c "
import sys, json
d = json.loads(sys.stdin.readline()); d['name'] = '$name'
print(json.dumps(d))" >> $O/${TAG}_ab.jsonl
  done
  ncu --metrics smsp__inst_executed.sum,smsp__thread_inst_executed.sum,gpu__time_duration.sum,smsp__inst_executed_op_local_ld.sum,smsp__inst_executed_op_local_st.sum,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio --clock-control none \
      -k regex:impc_thread_kernel -s 30 -c 3 --csv --log-file $O/${TAG}_ncu_$name.csv \
      python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline --no-graph --device-only > /dev/null 2>> $O/${TAG}_ab.err
done
python - <<PY
import json, csv, glob
for l in open("$O/${TAG}_ab.jsonl"):
    d = json.loads(l); print("%-10s value %.4g kernel_ms %.4f solved %s" % (d["name"], d["value"], d["kernel_ms"], d.get("solved_value")))
for f in sorted(glob.glob("$O/${TAG}_ncu_*.csv")):
    rows = [r for r in csv.reader(open(f)) if len(r) > 10]
    agg = {}
    for r in rows[1:]:
        agg.setdefault(r[-3], []).append(float(r[-1].replace(",", "")))
    short = {"gpu__time_duration.sum": "ns", "smsp__inst_executed.sum": "winst", "smsp__thread_inst_executed.sum": "tinst",
             "smsp__inst_executed_op_local_ld.sum": "ldl", "smsp__inst_executed_op_local_st.sum": "stl"}
    print("%-14s" % f.split("_ncu_")[1][:-4], " ".join("%s %.4g" % (short.get(k, k.split("stalled_")[-1].split("_per_")[0]), sum(v) / len(v)) for k, v in agg.items()))
PY
