#!/usr/bin/env python
"""bench.py — MPC-CBF QPs solved/sec (BASELINE.json metric) on N B200s of one node.

Workload (BASELINE.json configs[2]): independent 16-agent swarm scenarios,
horizon K=20 (4 cubic pieces, n=48), d_min 0.8, 2 IMPC iterations; one "step" = one control step
(ConnectivityIMPCCBF::optimize for every agent of every scenario = S*N*impc_iter QPs).
Scenarios are sharded across ranks with no collective: --scenarios is PER GPU
("scaling": "weak").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

`value`   : QPs/s with inputs resident in HBM (device-pointer C-ABI call, CUDA-event timed).
`e2e`     : same metric through mpccbf_step_batch_host (pinned HOST buffers, H2D + D2H inside).
`roofline`: fused kernel, algorithmic HBM bytes / measured kernel time vs MEASURED_PEAKS.json.
`cpu_baseline`: the oracle port (oracle/) on the host cores over a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
METRIC = "mpc_cbf_qps_per_sec"
# dram__bytes_read.sum + dram__bytes_write.sum of one impc_thread_kernel launch of this workload
# (ncu --set full, profiles/r1g_impc_thread_ncu_summary.md: 11.3 MB read + 67.9 MB written; the part
# above the 31.7 MB algorithmic bytes is local memory — the stage-A snapshot and register spills —
# written back from L1/L2); None for other shapes
TRAFFIC_BYTES_PER_LAUNCH = 79.1e6
UNIT = "QPs/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scenarios", type=int, default=4096, help="scenarios per GPU")
    ap.add_argument("--agents", type=int, default=16)
    ap.add_argument("--k-hor", type=int, default=20)
    ap.add_argument("--snapshots", type=int, default=8, help="distinct closed-loop state sets rotated through")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--seed", type=int, default=20261104)
    return ap.parse_args()


def workload_name(a):
    return (f"{a.scenarios} independent {a.agents}-agent swarm scenarios per GPU, horizon {a.k_hor}, "
            f"pairwise safety HOCBF rows, impc_iter 2 (BASELINE configs[2])")


def make_swarm(S, N, seed):
    """circle of radius U(3,6), random rotation, antipodal goals, |v0| <= 0.5."""
    rng = np.random.default_rng(seed)
    R = rng.uniform(3.0, 6.0, size=(S, 1))
    phi = rng.uniform(0, 2 * np.pi, size=(S, 1))
    ang = phi + 2 * np.pi * np.arange(N)[None, :] / N
    st = np.zeros((S, N, 6))
    st[:, :, 0], st[:, :, 1] = R * np.cos(ang), R * np.sin(ang)
    st[:, :, 3:5] = rng.uniform(-0.5, 0.5, size=(S, N, 2))
    tg = np.zeros((S, N, 3))
    tg[:, :, :2] = -st[:, :, :2]
    return st, tg


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i] == "Active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def dist_setup(args):
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        torch.cuda.set_device(local)
        dist.init_process_group(backend="nccl", rank=rank, world_size=world,
                                device_id=torch.device("cuda", local))
    return world, rank, local


def rollout_rows(O, ocfg):
    """position / velocity basis rows at t = h (piece 0) for the noise-free state advance."""
    n, c = ocfg.num_pieces * 3 * ocfg.num_ctrl_pts, ocfg.num_ctrl_pts
    b0 = O.bernstein_basis(c - 1, ocfg.piece_max_param, ocfg.h, 0)
    b1 = O.bernstein_basis(c - 1, ocfg.piece_max_param, ocfg.h, 1)
    R = np.zeros((6, n))
    for ax in range(3):
        R[ax, ax * c:(ax + 1) * c] = b0
        R[3 + ax, ax * c:(ax + 1) * c] = b1
    return R


def oracle_cfg(O, a):
    p = int(np.ceil((a.k_hor - 1) * 0.1 / 0.5 - 1e-9))
    return O.base_config(k_hor=a.k_hor, num_pieces=max(p, 1), d_min=0.8, num_agents=a.agents)


def cpu_closed_loop_snapshots(O, ocfg, st, tg, n_snap, stride, threads):
    """Closed-loop snapshots of a (small) scenario sample computed by the CPU port."""
    R = rollout_rows(O, ocfg)
    snaps = []
    st = st.copy()
    for i in range(n_snap * stride):
        if i % stride == 0:
            snaps.append(st.copy())
        if i == n_snap * stride - 1:
            break
        r = O.step_batch(ocfg, st, tg, threads=threads)
        ok = r["n_curves"] > 0
        nxt = np.einsum("cn,san->sac", R, r["ctrl_pts"])
        st[ok] = nxt[ok]
        st[~ok, 3:] = 0.0
    return snaps


def run_reference(args):
    """Reference arm: the reference's CPU algorithm for the path (oracle port: CPLEX / Eigen /
    GiNaC are absent here) on all host threads, bounded sample per step."""
    if int(os.environ.get("RANK", "0")) != 0:   # torchrun: rank 0 alone runs the CPU arm
        return
    from __graft_entry__ import load_oracle
    O = load_oracle()
    ocfg = oracle_cfg(O, args)
    threads = O.hardware_threads()
    S_full, N = args.scenarios, args.agents
    st, tg = make_swarm(S_full, N, args.seed)
    # bounded sample: calibrate so that steps+warmup take about 2 * cpu_seconds in total
    S0 = min(S_full, 4 * threads)
    t0 = time.perf_counter()
    O.step_batch(ocfg, st[:S0], tg[:S0], threads=threads)
    per_scn = (time.perf_counter() - t0) / S0
    total = max(1, args.steps + args.warmup)
    S = int(max(threads, min(S_full, (2 * args.cpu_seconds / total) / max(per_scn, 1e-9))))
    snaps = cpu_closed_loop_snapshots(O, ocfg, st[:S], tg[:S], args.snapshots, 5, threads)
    for i in range(args.warmup):
        O.step_batch(ocfg, snaps[i % len(snaps)], tg[:S], threads=threads)
    times = []
    for i in range(args.steps):
        t0 = time.perf_counter()
        O.step_batch(ocfg, snaps[i % len(snaps)], tg[:S], threads=threads)
        times.append(time.perf_counter() - t0)
    elapsed = sum(times)
    qps = S * N * ocfg.impc_iter * args.steps / elapsed
    sample = f"{S} of {S_full} scenarios per step ({S * N * ocfg.impc_iter} QPs), closed-loop snapshots"
    line = {"impl": "reference", "metric": METRIC, "value": qps, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * elapsed / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args), "agents": N, "k_hor": args.k_hor,
                       "note": "CPU restatement of the reference path (not CPLEX: unavailable on this box)"},
            "cpu_baseline": {"value": qps, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": qps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "p50_step_ms": 1e3 * float(np.median(times))}
    print(json.dumps(line), flush=True)


def run_b200(args):
    import torch
    from __graft_entry__ import load
    world, rank, local = dist_setup(args)
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    M = load()
    S, N, K = args.scenarios, args.agents, args.k_hor
    p = int(np.ceil((K - 1) * 0.1 / 0.5 - 1e-9))
    cfg = M.default_config(k_hor=K, num_pieces=max(p, 1), d_min=0.8, num_agents=N, max_scenarios=S, device=local)
    h = M.Handle(cfg)
    n, iters = h.n, cfg.impc_iter
    qps_per_step = S * N * iters

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            torch.distributed.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        return float(t.item())

    # ---- synthetic closed-loop snapshots, generated on the device by the product itself ---------
    st_h, tg_h = make_swarm(S, N, args.seed + 7919 * rank)
    d_tg = torch.tensor(tg_h, device=dev)
    cur = torch.tensor(st_h, device=dev)
    R = max(2, args.snapshots)
    sets = []
    for r in range(R):
        sets.append(dict(states=torch.empty_like(cur), ctrl=torch.zeros((S, N, n), dtype=torch.float64, device=dev),
                         status=torch.zeros((S, N, iters), dtype=torch.int32, device=dev),
                         n_curves=torch.zeros((S, N), dtype=torch.int32, device=dev),
                         objective=torch.zeros((S, N), dtype=torch.float64, device=dev),
                         nxt=torch.empty_like(cur)))
    stream = torch.cuda.Stream(device=dev)
    sp = stream.cuda_stream
    stride = 5
    with torch.cuda.stream(stream):
        for i in range(R * stride):
            b = sets[i // stride]
            if i % stride == 0:
                b["states"].copy_(cur)
            h.step_batch(S, cur, d_tg, b["ctrl"], b["status"], b["n_curves"], b["objective"], stream=sp)
            h.rollout(S, cur, b["ctrl"], b["n_curves"], cur, stream=sp)
    stream.synchronize()
    set_bytes = sum(t.numel() * t.element_size() for t in sets[0].values())

    def step(i):
        b = sets[i % R]
        h.step_batch(S, b["states"], d_tg, b["ctrl"], b["status"], b["n_curves"], b["objective"], stream=sp)
        h.rollout(S, b["states"], b["ctrl"], b["n_curves"], b["nxt"], stream=sp)

    # ---- device-resident timing ------------------------------------------------------------
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    with torch.cuda.stream(stream):
        for i in range(args.warmup):
            step(i)
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    l0, _ = h.counters()
    with torch.cuda.stream(stream):
        for i in range(args.steps):
            b = sets[(i + args.warmup) % R]
            ev[i][0].record(stream)
            h.step_batch(S, b["states"], d_tg, b["ctrl"], b["status"], b["n_curves"], b["objective"], stream=sp)
            ev[i][1].record(stream)
            h.rollout(S, b["states"], b["ctrl"], b["n_curves"], b["nxt"], stream=sp)
            ev[i][2].record(stream)
    barrier()
    l1, _ = h.counters()
    total_ms = max_over_ranks(ev[0][0].elapsed_time(ev[-1][2]))
    step_ms = [e[0].elapsed_time(e[2]) for e in ev]
    kern_ms = float(np.mean([e[0].elapsed_time(e[1]) for e in ev]))
    value = world * qps_per_step * args.steps / (total_ms * 1e-3)

    # ---- end to end: pinned host buffers through mpccbf_step_batch_host ---------------------------
    hs = [b["states"].cpu().pin_memory() for b in sets]
    htg = torch.tensor(tg_h).pin_memory()
    ho = dict(ctrl_pts=torch.zeros((S, N, n), dtype=torch.float64).pin_memory(),
              status=torch.zeros((S, N, iters), dtype=torch.int32).pin_memory(),
              n_curves=torch.zeros((S, N), dtype=torch.int32).pin_memory(),
              objective=torch.zeros((S, N), dtype=torch.float64).pin_memory())
    ho_np = {k: v.numpy() for k, v in ho.items()}
    hs_np = [x.numpy() for x in hs]
    htg_np = htg.numpy()
    for i in range(args.warmup):
        h.step_batch_host(hs_np[i % R], htg_np, out=ho_np)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        h.step_batch_host(hs_np[i % R], htg_np, out=ho_np)
    torch.cuda.synchronize()
    e2e_serial_s = max_over_ranks(time.perf_counter() - t0)
    # the same calls pipelined two deep over two handles (mpccbf_step_batch_host_async / mpccbf_synchronize):
    # the PCIe phase of batch i overlaps the solve of batch i + 1; every batch still crosses PCIe both ways
    h2 = M.Handle(cfg)
    ho2 = {k: torch.zeros_like(v).pin_memory() for k, v in ho.items()}
    ho2_np = {k: v.numpy() for k, v in ho2.items()}
    hh, oo = (h, h2), (ho_np, ho2_np)
    for i in range(max(args.warmup, 4)):  # warm the asynchronous (staged-copy) path itself on both handles
        hh[i & 1].step_batch_host_async(hs_np[i % R], htg_np, oo[i & 1])
        hh[i & 1].synchronize()
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        k = i & 1
        if i >= 2:
            hh[k].synchronize()          # batch i - 2 is complete, its output buffers are free again
        hh[k].step_batch_host_async(hs_np[i % R], htg_np, oo[k])
    hh[0].synchronize()
    hh[1].synchronize()
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    h2.close()
    clk = clocks.stop() if rank == 0 else None
    e2e = world * qps_per_step * args.steps / e2e_s
    e2e_serial = world * qps_per_step * args.steps / e2e_serial_s
    h2d = S * N * (6 + 3) * 8
    d2h = S * N * (n * 8 + iters * 4 + 4 + 8)
    # SolveStatus histogram over all R snapshot sets of the device-resident leg (OPTIMAL 0, INFEASIBLE 3,
    # SKIPPED 7 = IMPC iteration after a failed one)
    status_np = np.concatenate([b["status"].cpu().numpy().ravel() for b in sets])
    counts = {str(k): int(v) for k, v in zip(*np.unique(status_np, return_counts=True))}
    fp64_peak = h.measure_fp64_tflops()
    # algorithmic FP64 flops per launch: SURVEY 8d reduced-space formula with the
    # MEASURED active-set iteration count I of every QP
    #   F_qp = r^3/3 + m_cbf*8r + m_cbf*r^2 + I*(4*m*r + 2*r^2),  r = 6,  m = 6K + m_cbf
    # (averaged over the R snapshot sets the timed loop cycles through, so the figure does not depend on
    # which snapshot --steps happens to end on)
    r_dof = 6.0
    flops = 0.0
    mean_iters = 0.0
    for b_set in sets:
        h.step_batch(S, b_set["states"], d_tg, b_set["ctrl"], b_set["status"], b_set["n_curves"], b_set["objective"], stream=sp)
        stream.synchronize()
        gi = h.last_iteration_counts(S * N * iters).reshape(S * N, iters).astype(np.float64)
        for it_idx in range(iters):
            m_cbf = (N - 1) * (1 if it_idx == 0 else cfg.cbf_horizon)
            m_rows = 6 * K + m_cbf
            flops += float(np.sum(r_dof ** 3 / 3 + m_cbf * 8 * r_dof + m_cbf * r_dof ** 2
                                  + gi[:, it_idx] * (4 * m_rows * r_dof + 2 * r_dof ** 2))) / len(sets)
        mean_iters += float(gi.mean()) / len(sets)
    if rank != 0:
        h.close()
        return

    # ---- roofline of the fused kernel (algorithmic HBM bytes per launch) ----------------
    alg_bytes = S * N * (6 * 8 + 3 * 8 + n * 8 + iters * 4 + 4 + 8 + iters * 4)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    else:
        peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
    achieved = alg_bytes / (kern_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": TRAFFIC_BYTES_PER_LAUNCH if (S, N, K) == (4096, 16, 20) else None, "kernel": "impc_thread_kernel", "kernel_ms": kern_ms,
                "algorithmic_bytes_per_launch": alg_bytes, "peak_source": peak_src,
                "note": "FP64-issue/latency bound by design (484 B per agent-step); the fp64 entry is the "
                        "meaningful fraction, see DESIGN.md section 4",
                "fp64": {"bound": "fp64", "achieved": flops / (kern_ms * 1e-3) / 1e12, "peak": fp64_peak,
                         "unit": "TFLOP/s", "frac": flops / (kern_ms * 1e-3) / 1e12 / fp64_peak,
                         "algorithmic_flops_per_launch": flops, "mean_active_set_iterations_per_qp": mean_iters,
                         "peak_source": "register-resident DFMA loop measured in this run (mpccbf_measure_fp64_tflops)"}}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        from __graft_entry__ import load_oracle
        O = load_oracle()
        ocfg = oracle_cfg(O, args)
        threads = O.hardware_threads()
        S0 = min(S, 4 * threads)
        t0 = time.perf_counter()
        O.step_batch(ocfg, hs_np[0][:S0], tg_h[:S0], threads=threads)
        per = (time.perf_counter() - t0) / S0
        Sc = int(max(threads, min(S, args.cpu_seconds / R / max(per, 1e-9))))
        t0 = time.perf_counter()
        for r in range(R):
            O.step_batch(ocfg, hs_np[r][:Sc], tg_h[:Sc], threads=threads)
        dt = time.perf_counter() - t0
        cpu = {"value": Sc * N * iters * R / dt, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"first {Sc} of {S} scenarios of each of the {R} snapshots ({Sc * N * iters * R} QPs, {dt:.1f} s)"}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args), "agents": N, "k_hor": K, "n": n, "impc_iter": iters,
                       "cbf_horizon": cfg.cbf_horizon, "qps_per_step_per_gpu": qps_per_step,
                       "l2": f"rotating {R} closed-loop snapshot sets, {R * set_bytes / 2**20:.0f} MiB > 126 MB L2",
                       "status_counts_all_snapshots": counts, "solver": "reduced"},
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": 1e3 * e2e_s / args.steps, "pipeline_depth": 2,
                    "serial_value": e2e_serial, "serial_ms_per_step": 1e3 * e2e_serial_s / args.steps,
                    "note": "mpccbf_step_batch_host[_async] on pinned host buffers; value = two batches in flight "
                            "over two handles, serial_value = one blocking call per batch"},
            "gpu_launches": int(l1 - l0), "p50_step_ms": float(np.median(step_ms)),
            "p99_step_ms": float(np.percentile(step_ms, 99)), "roofline": roofline,
            "cpu_baseline": cpu, "clocks": clk}
    print(json.dumps(line), flush=True)
    h.close()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
