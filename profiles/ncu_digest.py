#!/usr/bin/env python
"""Digest of an .ncu-rep (read in the build container): headline raw metrics of each captured launch
and the share of executed warp instructions / stall samples per source line.
usage: python profiles/ncu_digest.py gpurun_out/x.ncu-rep [top_lines] [--json profiles/rNx_fused_kernel.json --agents A]
With --json the per-launch DRAM traffic and executed FP64 flops of the LAST captured launch are written for bench.py
(roofline.traffic / roofline.executed), stamped with the current commit."""
import collections
import csv
import io
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__waves_per_multiprocessor", "launch__shared_mem_per_block_dynamic",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sass__inst_executed_local_loads",
        "sass__inst_executed_local_stores", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__cycles_elapsed.max",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum", "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum",
        "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum", "smsp__sass_thread_inst_executed_ops_dadd_dmul_dfma_pred_on.sum"]


def ncu(rep, *args):
    return subprocess.run(["ncu", "-i", rep, *args], capture_output=True, text=True).stdout


def to_bytes(val, unit):
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    return float(val.replace(",", "")) * mult.get(unit, 1)


def write_json(path, hdr, units, row, agents):
    import datetime
    import json
    def get(name):
        return (row[hdr.index(name)], units[hdr.index(name)]) if name in hdr else (None, None)
    rd, wr = get("dram__bytes_read.sum"), get("dram__bytes_write.sum")
    out = {"when": datetime.datetime.utcnow().isoformat() + "Z",
           "commit": subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip(),
           "kernel": row[hdr.index("Kernel Name")].split("(")[0], "agents_per_launch": agents}
    if rd[0] is not None:
        out["dram_bytes_per_launch"] = to_bytes(*rd) + to_bytes(*wr)
    fl = 0.0
    for name, w in (("dfma", 2.0), ("dmul", 1.0), ("dadd", 1.0)):
        v, _ = get(f"smsp__sass_thread_inst_executed_op_{name}_pred_on.sum")
        if v is not None:
            out[f"thread_inst_{name}"] = float(v.replace(",", ""))
            fl += w * out[f"thread_inst_{name}"]
    if fl:
        out["fp64_flops_executed_per_launch"] = fl
    d, du = get("gpu__time_duration.sum")
    if d is not None:
        out["duration_us_under_ncu"] = float(d.replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(du, 1.0)
    json.dump(out, open(path, "w"), indent=1)
    print("wrote", path, out)


def main():
    rep = sys.argv[1]
    argv = sys.argv[2:]
    jpath, agents = None, None
    if "--json" in argv:
        i = argv.index("--json")
        jpath = argv[i + 1]
        del argv[i:i + 2]
    if "--agents" in argv:
        i = argv.index("--agents")
        agents = int(argv[i + 1])
        del argv[i:i + 2]
    top = int(argv[0]) if argv else 30
    rows = list(csv.reader(io.StringIO(ncu(rep, "--page", "raw", "--csv"))))
    hdr, units, data = rows[0], rows[1], rows[2:]
    if jpath:
        write_json(jpath, hdr, units, data[-1], agents)
    print("kernel:", [r[hdr.index("Kernel Name")].split("(")[0] for r in data])
    for w in WANT:
        if w in hdr:
            i = hdr.index(w)
            print(f"{w} [{units[i]}]: " + " | ".join(r[i] for r in data))
    rows = list(csv.reader(io.StringIO(ncu(rep, "--page", "source", "--csv", "--print-source", "cuda,sass"))))
    cur, agg, src = None, collections.defaultdict(lambda: [0, 0, 0]), {}
    iS = iI = iT = None
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif r[0] == "Line No":
            iS, iI, iT = r.index("# Samples"), r.index("Instructions Executed"), r.index("Thread Instructions Executed")
        elif iS is not None:
            try:
                k = (cur, int(r[0]))
                agg[k][0] += int(r[iS]); agg[k][1] += int(r[iI]); agg[k][2] += int(r[iT])
                src[k] = r[1].strip()[:90]
            except (ValueError, IndexError):
                pass
    ts = sum(v[0] for v in agg.values()) or 1
    ti = sum(v[1] for v in agg.values()) or 1
    print(f"-- source lines: {ts} samples, {ti} warp instructions (all captured launches)")
    byfile = collections.defaultdict(lambda: [0, 0, 0])
    for k, v in agg.items():
        for j in range(3):
            byfile[k[0]][j] += v[j]
    for f, v in byfile.items():
        print(f"   {f}: samples {100*v[0]/ts:.1f}%  inst {100*v[1]/ti:.1f}%  threads/inst {v[2]/max(v[1],1):.1f}")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{k[0][:16]:16s} {k[1]:4d} samp {100*v[0]/ts:4.1f}% inst {100*v[1]/ti:4.1f}% thr/inst {v[2]/max(v[1],1):4.1f}  {src[k]}")


if __name__ == "__main__":
    main()
