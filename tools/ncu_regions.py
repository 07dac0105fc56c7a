#!/usr/bin/env python
"""Executed warp / thread instructions and stall samples of an .ncu-rep (captured with --import-source on) aggregated by
the FUNCTION the source line belongs to (function starts are found in the current source files, so run it at the commit
the report was taken at). usage: python tools/ncu_regions.py gpurun_out/x.ncu-rep"""
import bisect, collections, csv, io, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FILES = ["mpc-cbf_b200/csrc/impc_thread.cu", "mpc-cbf_b200/csrc/gi_solver.h"]
PAT = re.compile(r"^\s*(?:template.*>\s*)?(?:static\s+)?(?:MPCCBF_HD|__host__ __device__|__global__|__device__ __forceinline__|static inline)\b.*?(\w+)\s*\(")


def regions(path):
    out = [(1, "top")]
    for n, line in enumerate(open(os.path.join(ROOT, path)), 1):
        m = PAT.match(line)
        if m and not line.strip().endswith(";"):
            out.append((n, m.group(1)))
    # split gi_run into its phases by comment markers
    if path.endswith("gi_solver.h"):
        for n, line in enumerate(open(os.path.join(ROOT, path)), 1):
            for key, name in (("// ---- Q R = active normals", "gi_run:MGS"), ("// r = R^-1 d1", "gi_run:steplen"),
                              ("const double t = fmin(fmin(t1, t2), t3);", "gi_run:update")):
                if key in line and n < 330:
                    out.append((n, name))
    return sorted(out)


def main():
    rep = sys.argv[1]
    regs = {os.path.basename(f): regions(f) for f in FILES}
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    cur, agg, cols = None, collections.defaultdict(lambda: [0, 0, 0, 0]), None
    num = lambda x: int(x) if x.strip().isdigit() else 0
    for r in csv.reader(io.StringIO(out)):
        if not r:
            continue
        if r[0] == "File Path":
            cur = os.path.basename(r[1])
        elif r[0] == "Line No":
            cols = (r.index("# Samples"), r.index("Instructions Executed"), r.index("Thread Instructions Executed"),
                    r.index("stall_no_inst") if "stall_no_inst" in r else None)
        elif cols and r[0].strip().isdigit():
            ln, name = int(r[0]), cur
            if cur in regs:
                starts = [s for s, _ in regs[cur]]
                name = cur.split(".")[0][:6] + ":" + regs[cur][bisect.bisect_right(starts, ln) - 1][1]
            a = agg[name]
            a[0] += num(r[cols[0]]); a[1] += num(r[cols[1]]); a[2] += num(r[cols[2]])
            if cols[3] is not None:
                a[3] += num(r[cols[3]])
    ts, ti, tt = (sum(v[i] for v in agg.values()) or 1 for i in range(3))
    print(f"warp inst {ti}  thread inst {tt}  samples {ts}")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        if v[1] * 1000 > ti:
            print("%-34s inst %5.1f%%  thread %5.1f%%  thr/inst %5.1f  samples %5.1f%%  no_inst %4d" % (k, 100 * v[1] / ti, 100 * v[2] / tt, v[2] / max(v[1], 1), 100 * v[0] / ts, v[3]))


if __name__ == "__main__":
    main()
