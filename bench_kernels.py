#!/usr/bin/env python
"""Per-kernel measurement of the sub-kernels beside the fused one: K1 safety-row
construction, K3 rollout, K2d dense ADMM QP, K-filter, K-episode. CUDA events on the launching stream,
warm-up first, inputs larger than L2 where the kernel streams. One JSON line per kernel with the
algorithmic bytes / flops per launch and the fraction of the measured roofline.

    python bench_kernels.py [--out profiles/r1g_kernels.jsonl]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def timed(fn, reps, torch):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    import torch
    from __graft_entry__ import load, load_oracle
    M, O = load(), load_oracle()
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    peaks = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm = float(json.load(open(peaks))["hbm_gbs"]) if os.path.exists(peaks) else 6650.0
    lines = []

    def emit(d):
        lines.append(d)
        print(json.dumps(d), flush=True)

    # ---- K1: all-pairs safety rows, cfg4 shape (256 agents), 512 scenarios: 1.07 GB of rows > L2 -------------
    S, N = 512, 256
    cfg = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, max_scenarios=S)
    h = M.Handle(cfg)
    fp64_peak = h.measure_fp64_tflops()
    rng = np.random.default_rng(1)
    st = torch.tensor(rng.uniform(-8, 8, size=(S, N, 6)), device="cuda")
    rows = torch.empty((S, N, N - 1, 4), dtype=torch.float64, device="cuda")
    ms = timed(lambda: h.cbf_rows(S, st, rows), a.reps, torch)
    by = S * N * (N - 1) * 32 + S * N * 48
    emit({"kernel": "cbf_rows_kernel (K1)", "shape": f"{S} x {N} agents, {S * N * (N - 1)} rows", "ms": ms, "bound": "hbm",
          "algorithmic_bytes": by, "achieved_GBs": by / ms / 1e6, "peak_GBs": hbm, "frac": by / ms / 1e6 / hbm,
          "flops": 45.0 * S * N * (N - 1), "fp64_frac": 45.0 * S * N * (N - 1) / ms / 1e9 / fp64_peak})
    del rows
    h.close()

    # ---- K3: rollout, 2^20 agents (n = 48): 400 MB of control points > L2 -----------------------------------------
    S, N = 65536, 16
    cfg = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, max_scenarios=S)
    h = M.Handle(cfg)
    st = torch.tensor(rng.uniform(-8, 8, size=(S, N, 6)), device="cuda")
    cp = torch.tensor(rng.uniform(-8, 8, size=(S, N, h.n)), device="cuda")
    nc = torch.ones((S, N), dtype=torch.int32, device="cuda")
    nxt = torch.empty_like(st)
    ms = timed(lambda: h.rollout(S, st, cp, nc, nxt), a.reps, torch)
    # the kernel needs piece 0 only: 3 axes x c control points of the n
    by = S * N * (3 * cfg.num_ctrl_pts * 8 + 4 + 48 + 48)
    emit({"kernel": "rollout_kernel (K3)", "shape": f"{S * N} agents", "ms": ms, "bound": "hbm", "algorithmic_bytes": by,
          "achieved_GBs": by / ms / 1e6, "peak_GBs": hbm, "frac": by / ms / 1e6 / hbm,
          "note": "reads the 96 B of piece 0 out of each agent's 384 B curve: 32-byte sectors make the fetched bytes ~1.3x the algorithmic ones"})
    del cp, nxt
    h.close()

    # ---- K2d: dense one-CTA-per-QP ADMM on assembled MPC-CBF QPs (n = 48, m = 177 / 192) ---------------------------
    from conftest import swarm
    N = 16
    oc = O.base_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N)
    stn, tgn = swarm(4, N, seed=3)
    stn[:, :, :2] *= 0.45
    tgn[:, :, :2] *= 0.45
    qps = [O.build_qp(oc, stn[s], ag, np.tile(tgn[s, ag], 20), it=0) for s in range(4) for ag in range(N)]
    reps_b = 32
    P = np.stack([q["P"] for q in qps] * reps_b); q = np.stack([q_["q"] for q_ in qps] * reps_b)
    A = np.stack([q_["A"] for q_ in qps] * reps_b); l = np.stack([q_["l"] for q_ in qps] * reps_b)
    u = np.stack([q_["u"] for q_ in qps] * reps_b)
    cfg = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, max_scenarios=1)
    h = M.Handle(cfg)
    import time
    h.solve_qp_dense(P[:64], q[:64], A[:64], l[:64], u[:64])
    t0 = time.perf_counter()
    r = h.solve_qp_dense(P, q, A, l, u)
    dt = time.perf_counter() - t0
    n, m = q.shape[1], l.shape[1]
    it_mean = float(r["iters"].mean())
    fl = float(np.sum(n ** 3 / 3 + m * n * n + r["iters"] * (4.0 * m * n + 2.0 * n * n)))
    emit({"kernel": "dense_qp_kernel (K2d)", "shape": f"{P.shape[0]} QPs, n = {n}, m = {m}", "ms_host_call": dt * 1e3,
          "qps_per_s_host_call": P.shape[0] / dt, "mean_admm_iterations": it_mean, "optimal_frac": float((r["status"] == 0).mean()),
          "bound": "fp64", "algorithmic_flops": fl, "achieved_TFLOPs_incl_copies": fl / dt / 1e12, "peak_TFLOPs": fp64_peak,
          "note": "host-pointer entry: the time includes H2D of P, A (42 MB) and D2H; see the ncu duration for the kernel alone"})
    h.close()

    # ---- slack-mode controller step (capped multipliers inside the fused thread-per-agent kernel) ----------------
    import bench
    from parity import dense_scenarios
    N = 16
    for label, S, maker in (("ring swarms of bench.py", 4096, lambda S: bench.make_swarm(S, N, 7)),
                            ("packed grids near d_min (half infeasible without slack)", 4096,
                             lambda S: dense_scenarios(S, N, N, 0.8, spread=1.15, vmax=1.0))):
        stn, tgn = maker(S)
        d_st, d_tg = torch.tensor(stn, device="cuda"), torch.tensor(tgn, device="cuda")
        for slack in (1, 0):
            cfg = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, max_scenarios=S, slack_mode=slack,
                                   slack_cost=50000.0, slack_decay=0.1)
            h = M.Handle(cfg)
            cp = torch.zeros((S, N, h.n), dtype=torch.float64, device="cuda")
            stat = torch.zeros((S, N, 2), dtype=torch.int32, device="cuda")
            ms = timed(lambda: h.step_batch(S, d_st, d_tg, cp, stat), 5, torch)
            its = h.last_iteration_counts(S * N * 2)
            sc = np.bincount(stat.cpu().numpy().ravel(), minlength=8)
            emit({"kernel": "impc_thread_kernel<SOFT=%d> controller step, slack_mode = %d" % (slack, slack), "agents": S * N,
                  "shape": f"{S} x {N} agents, {label}",
                  "ms": ms, "qps_per_s": S * N * 2 / (ms * 1e-3),
                  "active_set_steps_p50_p90_p99": [float(v) for v in np.percentile(its, [50, 90, 99])],
                  "status_counts": {str(i): int(v) for i, v in enumerate(sc) if v}})
            h.close()

    if a.out:
        with open(a.out, "w") as f:
            for d in lines:
                f.write(json.dumps(d) + "\n")


if __name__ == "__main__":
    main()
