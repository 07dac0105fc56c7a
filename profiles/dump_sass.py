#!/usr/bin/env python
"""SASS of every kernel of libmpccbf.so, one gzip file per kernel under profiles/<tag>_sass/, plus an index with
instruction counts and the mnemonics that matter for the roofline argument (DFMA/DMUL/DADD, LDL/STL local-memory
traffic, LDS/STS). usage: python profiles/dump_sass.py r2"""
import collections
import gzip
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r2"
    so = os.path.join(ROOT, "mpc-cbf_b200", "libmpccbf.so")
    out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
    commit = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True, cwd=ROOT).stdout.strip()
    dirty = subprocess.run(["git", "status", "--porcelain", "--", "mpc-cbf_b200/csrc", "include"], capture_output=True, text=True, cwd=ROOT).stdout.strip()
    d = os.path.join(ROOT, "profiles", f"{tag}_sass")
    os.makedirs(d, exist_ok=True)
    kernels = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = []
        elif cur is not None:
            kernels[cur].append(line)
    rows = []
    for name, lines in kernels.items():
        dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
        short = dem.replace("(anonymous namespace)::", "").replace("mpccbf::", "")
        short = re.sub(r"^void ", "", re.sub(r"\(.*", "", short))
        short = re.sub(r"[^A-Za-z0-9_<>,]", "_", short)
        ins = [l for l in lines if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", l)]
        def cnt(pat):
            return sum(1 for l in ins if re.search(pat, l))
        with gzip.open(os.path.join(d, short + ".sass.gz"), "wt") as f:
            f.write(f"// {dem}\n// libmpccbf.so at commit {commit}{' (+ uncommitted changes)' if dirty else ''}\n")
            f.write("\n".join(lines))
        rows.append((short, len(ins), cnt(r"\bDFMA\b"), cnt(r"\bDMUL\b"), cnt(r"\bDADD\b"), cnt(r"\bLDL"), cnt(r"\bSTL"),
                     cnt(r"\bLDS"), cnt(r"\bSTS"), cnt(r"\bDMMA|\bHMMA|UTCMMA|tcgen05"), cnt(r"\bMUFU")))
    with open(os.path.join(d, "INDEX.md"), "w") as f:
        f.write(f"# SASS of every kernel in libmpccbf.so at commit {commit}{' (+ uncommitted changes)' if dirty else ''}\n\n")
        f.write("Static instruction counts (cuobjdump -sass). No kernel is GEMM-shaped: no tensor-core instruction anywhere.\n\n")
        f.write("| kernel | instructions | DFMA | DMUL | DADD | LDL | STL | LDS | STS | MMA | MUFU |\n|---|---|---|---|---|---|---|---|---|---|---|\n")
        for r in rows:
            f.write("| " + " | ".join(str(x) for x in r) + " |\n")
    print(open(os.path.join(d, "INDEX.md")).read())


if __name__ == "__main__":
    main()
