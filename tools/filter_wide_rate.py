#!/usr/bin/env python
"""Rate of the CBF-only filter for swarms of more than 32 agents (one CTA per scenario). usage: python tools/filter_wide_rate.py"""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
M = importlib.import_module("mpc-cbf_b200")
import torch
for N, S, spread in ((32, 2048, 3.5), (48, 1024, 4.5), (64, 592, 5.0), (128, 296, 7.0), (256, 148, 10.0)):
    rng = np.random.default_rng(N)
    st = np.zeros((S, N, 6)); st[:, :, :2] = rng.uniform(-spread, spread, (S, N, 2)); st[:, :, 3:5] = rng.uniform(-0.8, 0.8, (S, N, 2))
    du = rng.uniform(-2, 2, (S, N, 3))
    h = M.Handle(M.default_config(num_agents=N, d_min=0.5, d_max=3.0, max_scenarios=S, v_min=[-1, -1, -1], v_max=[1, 1, 1]))
    d_st, d_du = torch.tensor(st, device="cuda"), torch.tensor(du, device="cuda")
    u = torch.zeros((S, N, 3), dtype=torch.float64, device="cuda"); stat = torch.zeros((S, N), dtype=torch.int32, device="cuda")
    l2 = torch.zeros(S, dtype=torch.float64, device="cuda")
    for _ in range(2):
        h.cbf_filter_batch(S, d_st, d_du, u, stat, l2)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        h.cbf_filter_batch(S, d_st, d_du, u, stat, l2)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"N={N} S={S}: {ms:.3f} ms per call, {S * N / ms * 1e3:.3g} filter-QPs/s, optimal {float((stat == 0).float().mean()):.2f}, connected {float((l2 > 0.1).float().mean()):.2f}", flush=True)
    h.close()
