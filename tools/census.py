#!/usr/bin/env python
"""Active-set step census of the bench workload on the GPU: steps per agent split by phase (stage A + yaw measured with the
safety rows switched off, IMPC iteration 0 / 1), and per warp (32 consecutive agents) the maximum against the mean —
what lock-step execution pays for the spread. usage: python tools/census.py [scenarios]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import importlib
import bench
M = importlib.import_module("mpc-cbf_b200")
import torch

S, N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096, 16
st, tg = bench.make_swarm(S, N, 20261104)
cfg = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, max_scenarios=S)
cfg0 = M.default_config(k_hor=20, num_pieces=4, d_min=0.01, num_agents=N, max_scenarios=S)
h, h0 = M.Handle(cfg), M.Handle(cfg0)
cur, d_tg = torch.tensor(st, device="cuda"), torch.tensor(tg, device="cuda")
ctrl = torch.zeros((S, N, h.n), dtype=torch.float64, device="cuda")
status = torch.zeros((S, N, 2), dtype=torch.int32, device="cuda")
ncv = torch.zeros((S, N), dtype=torch.int32, device="cuda")
for step in range(41):
    h.step_batch(S, cur, d_tg, ctrl, status, ncv)
    torch.cuda.synchronize()
    if step % 5 == 0:
        gi = h.last_iteration_counts(S * N * 2).reshape(S * N, 2).astype(np.float64)
        c2, s2, n2 = torch.zeros_like(ctrl), torch.zeros_like(status), torch.zeros_like(ncv)
        h0.step_batch(S, cur, d_tg, c2, s2, n2)
        torch.cuda.synchronize()
        ga = h0.last_iteration_counts(S * N * 2).reshape(S * N, 2).astype(np.float64)
        tot = gi.sum(1)
        w = tot.reshape(-1, 32)
        a = ga[:, 0].reshape(-1, 32)
        i0 = np.maximum(gi[:, 0] - ga[:, 0], 0).reshape(-1, 32)
        i1 = gi[:, 1].reshape(-1, 32)
        print(f"step {step:2d}: per agent stageA+yaw {ga[:,0].mean():.2f}  iter0 {gi[:,0].mean()-ga[:,0].mean():.2f}  iter1 {gi[:,1].mean():.2f}  total {tot.mean():.2f} | "
              f"per warp: mean-of-max total {w.max(1).mean():.2f}  sum of per-phase max {a.max(1).mean()+i0.max(1).mean()+i1.max(1).mean():.2f} "
              f"(A {a.max(1).mean():.2f} i0 {i0.max(1).mean():.2f} i1 {i1.max(1).mean():.2f})  zero-step agents {np.mean(tot==0):.2f}  "
              f"status!=0 {np.mean(status.cpu().numpy()[:,:,0]!=0):.3f}", flush=True)
    h.rollout(S, cur, ctrl, ncv, cur)

# tail of the distribution at the last snapshot: the kernel cannot finish before its slowest agent
st_np = status.cpu().numpy().reshape(-1, 2)
for name, sel in (("all", np.ones(len(tot), bool)), ("both iterations OPTIMAL", (st_np[:, 0] == 0) & (st_np[:, 1] == 0)), ("some iteration not OPTIMAL", (st_np[:, 0] != 0) | (st_np[:, 1] != 0))):
    t = tot[sel]
    if len(t):
        print(f"steps per agent, {name} ({len(t)}): mean {t.mean():.2f} p50 {np.percentile(t,50):.0f} p90 {np.percentile(t,90):.0f} p99 {np.percentile(t,99):.0f} p99.9 {np.percentile(t,99.9):.0f} max {t.max():.0f}")
print("it0 max", gi[:, 0].max(), "it1 max", gi[:, 1].max(), "stageA+yaw max", ga[:, 0].max())
