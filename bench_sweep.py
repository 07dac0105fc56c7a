#!/usr/bin/env python
"""Sweep of the bench over BASELINE.json's other configurations on ONE GPU plus
the CBF-only filter and the episode driver. Each line is a plain `bench.py` run (same timing rules)
with a different --scenarios / --agents / --k-hor; results go to a .jsonl file and a markdown table.

    python bench_sweep.py --out profiles/r1f_sweep            # on the GPU box
"""
import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))

# (label, scenarios, agents, k_hor)
CONFIGS = [
    ("cfg1 2-agent swap shape, K=16, 1 scenario", 1, 2, 16),
    ("cfg2 8-agent circle, K=20, 1 scenario", 1, 8, 20),
    ("cfg3 4096 x 16, K=20", 4096, 16, 20),
    ("cfg4 256-agent dense, K=20, 16 scenarios", 16, 256, 20),
    ("cfg4 256-agent dense, K=20, 256 scenarios", 256, 256, 20),
]
for K in (10, 20, 40):
    for N in (2, 4, 8, 16, 32, 64):
        CONFIGS.append((f"cfg5 K={K} N={N}, 2^20 QPs/step", (1 << 19) // N, N, K))


def run_bench(S, N, K, steps, warmup):
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--scenarios", str(S), "--agents", str(N), "--k-hor", str(K),
           "--steps", str(steps), "--warmup", str(warmup), "--no-cpu-baseline", "--no-extras", "--snapshots", "4"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    if r.returncode != 0:
        return {"error": (r.stderr or r.stdout)[-400:]}
    return json.loads(r.stdout.strip().splitlines()[-1])


def filter_and_episode(out):
    """CBF-only filter (one warp per scenario) and episode driver throughput, CUDA events."""
    import torch
    sys.path.insert(0, ROOT)
    from __graft_entry__ import load, load_oracle
    M, O = load(), load_oracle()
    rows = []
    for N, S in ((8, 8192), (16, 4096), (32, 2048)):
        cfg = M.default_config(num_agents=N, d_min=0.5, d_max=3.0, max_scenarios=S)
        h = M.Handle(cfg)
        rng = np.random.default_rng(N)
        st = np.zeros((S, N, 6))
        st[:, :, :2] = rng.uniform(-0.25 * N ** 0.5 - 1.5, 0.25 * N ** 0.5 + 1.5, size=(S, N, 2))
        st[:, :, 3:5] = rng.uniform(-0.8, 0.8, size=(S, N, 2))
        du = rng.uniform(-2, 2, size=(S, N, 3))
        d_st, d_du = torch.tensor(st, device="cuda"), torch.tensor(du, device="cuda")
        u = torch.zeros((S, N, 3), dtype=torch.float64, device="cuda")
        status = torch.zeros((S, N), dtype=torch.int32, device="cuda")
        l2 = torch.zeros(S, dtype=torch.float64, device="cuda")
        for _ in range(3):
            h.cbf_filter_batch(S, d_st, d_du, u, status, l2)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20
        e0.record()
        for _ in range(reps):
            h.cbf_filter_batch(S, d_st, d_du, u, status, l2)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        oc = O.base_config(num_agents=N, d_min=0.5, d_max=3.0)
        Sc = max(1, min(S, 4096 // N))
        t0 = time.perf_counter()
        O.cbf_filter_batch(oc, st[:Sc], du[:Sc], prec=0)
        cpu_qps = Sc * N / (time.perf_counter() - t0)
        rows.append({"kernel": "cbf_filter_kernel", "agents": N, "scenarios": S, "ms": ms, "qps_per_s": S * N / (ms * 1e-3),
                     "connected_frac": float((l2.cpu().numpy() > 0.1).mean()), "optimal_frac": float((status.cpu().numpy() == 0).mean()),
                     "cpu_port_qps_per_s_1thread": cpu_qps})
        h.close()
    # episode driver: 1024 x 16 agents, 50 control steps
    S, N, steps = 1024, 16, 50
    cfg = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, max_scenarios=S)
    h = M.Handle(cfg)
    import bench
    st, tg = bench.make_swarm(S, N, 7)
    d_tg = torch.tensor(tg, device="cuda")
    succ = torch.zeros(S, dtype=torch.int32, device="cuda")
    for rep in range(2):
        d_st = torch.tensor(st, device="cuda")
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        h.run_episode(S, d_st, d_tg, steps, 1.0, 0.3, succ)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    rows.append({"kernel": "run_episode (impc_thread + episode_advance per step)", "agents": N, "scenarios": S, "control_steps": steps,
                 "ms": ms, "qps_per_s": S * N * 2 * steps / (ms * 1e-3), "ms_per_control_step": ms / steps,
                 "success_frac": float(succ.cpu().numpy().mean())})
    h.close()
    with open(out + "_filter_episode.json", "w") as f:
        json.dump(rows, f, indent=1)
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep"))
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--quick", action="store_true", help="cfg1..cfg4 only")
    a = ap.parse_args()
    os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
    lines = []
    with open(a.out + ".jsonl", "w") as f:
        for label, S, N, K in (CONFIGS[:5] if a.quick else CONFIGS):
            d = run_bench(S, N, K, a.steps, a.warmup)
            d["label"] = label
            f.write(json.dumps(d) + "\n")
            f.flush()
            lines.append(d)
            print(label, "->", d.get("value", d.get("error")), flush=True)
    with open(a.out + ".md", "w") as f:
        f.write("| config | scenarios x agents, K | QPs/s (device) | QPs/s (e2e, host buffers) | p50 step ms | p99 step ms | mean GI iters/QP | FP64 frac |\n|---|---|---|---|---|---|---|---|\n")
        for d in lines:
            if "error" in d:
                f.write(f"| {d['label']} | error: {d['error'][-120:]} | | | | | | |\n")
                continue
            c, r = d["config"], d["roofline"]
            f.write(f"| {d['label']} | {c['scenarios_total']} x {c['agents']}, {c['k_hor']} | {d['value']:.4g} | "
                    f"{d['e2e']['value']:.4g} | {d['p50_step_ms']:.4f} | {d['p99_step_ms']:.4f} | "
                    f"{r['mean_active_set_iterations_per_qp']:.2f} | {r['frac']:.3f} |\n")
    rows = filter_and_episode(a.out)
    for r in rows:
        print(json.dumps(r))


if __name__ == "__main__":
    main()
