#!/usr/bin/env python
"""bench.py — MPC-CBF QPs solved/sec (BASELINE.json metric) on N B200s of one node.

Workload (BASELINE.json configs[2]): 4096 independent 16-agent swarm scenarios IN TOTAL,
horizon K=20 (4 cubic pieces, n=48), d_min 0.8, 2 IMPC iterations; one "step" = one control step
(ConnectivityIMPCCBF::optimize for every agent of every scenario = S*N*impc_iter QPs). Scenarios are sharded
contiguously across the ranks, 4096/N per GPU, with no collective: "scaling": "strong". The weak
variant (4096 scenarios per GPU) and the one path with a collective (BASELINE configs[3]: 256-agent swarms
whose agents are split across the ranks, one NCCL all-gather per control step) are carried as sub-records.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

`value`   : QPs/s with inputs resident in HBM (device-pointer C-ABI call, CUDA-event timed, max over ranks).
`e2e`     : same metric through the host-pointer C-ABI calls on pinned HOST buffers, H2D + D2H inside the
            timed region (compact result form: 6 doubles per agent instead of n control points).
`roofline`: fused kernel; FP64 flops (SURVEY 8d formula with measured iteration counts) over the measured DFMA
            peak, the HBM figure nested.
`cpu_baseline`: the oracle port (oracle/) on the host cores over a bounded sample.
"""
import argparse
import glob
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
UNIT = "QPs/s"
SNAP_STRIDE = 5     # control steps between two closed-loop snapshots (both arms)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=48)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scenarios", type=int, default=4096, help="scenarios in TOTAL (sharded over the GPUs)")
    ap.add_argument("--agents", type=int, default=16)
    ap.add_argument("--k-hor", type=int, default=20)
    ap.add_argument("--snapshots", type=int, default=8, help="distinct closed-loop state sets rotated through")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the weak-scaling and sharded-cfg4 sub-records")
    ap.add_argument("--device-only", action="store_true", help="skip the end-to-end legs (kernel variant comparisons)")
    ap.add_argument("--device-streams", type=int, default=1, choices=[1, 2],
                    help="device leg: 1 = every step after the previous one on one stream (default); 2 = consecutive "
                         "(independent) steps alternate between two streams / handles, joined once per pass over the snapshot "
                         "sets (measured slower: 0.197 vs 0.182 ms per step)")
    ap.add_argument("--e2e-depth", type=int, default=3, help="batches in flight in the end-to-end leg (one handle each)")
    ap.add_argument("--no-graph", action="store_true", help="launch every timed step from Python instead of CUDA graphs")
    ap.add_argument("--seed", type=int, default=20261104)
    return ap.parse_args()


def config_of(a, n=None, iters=2, cbf_horizon=2):
    """The `config` object: identical for both arms (it names the workload, not the run)."""
    p = max(int(np.ceil((a.k_hor - 1) * 0.1 / 0.5 - 1e-9)), 1)
    return {"workload": (f"{a.scenarios} independent {a.agents}-agent swarm scenarios in total, sharded over the GPUs, "
                         f"horizon {a.k_hor}, pairwise safety HOCBF rows, impc_iter 2 (BASELINE configs[2])"),
            "scenarios_total": a.scenarios, "agents": a.agents, "k_hor": a.k_hor, "n": n if n is not None else p * 12,
            "impc_iter": iters, "cbf_horizon": cbf_horizon, "qps_per_step": a.scenarios * a.agents * iters,
            "snapshots": a.snapshots, "snapshot_stride": SNAP_STRIDE, "seed": a.seed, "solver": "reduced",
            "l2": "closed-loop snapshot sets rotated between timed steps (inputs + outputs of all sets > 126 MB L2 at N = 1)"}


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


def grid_swarm(S, N, seed):
    """sqrt(N) x sqrt(N) grid, pitch 1.25 d_min = 1.0 with jitter, goals mirrored through the centroid."""
    rng = np.random.default_rng(seed)
    side = int(round(np.sqrt(N)))
    gx, gy = np.meshgrid(np.arange(side), np.arange(side), indexing="ij")
    st = np.zeros((S, N, 6))
    st[:, :, :2] = np.stack([gx.ravel(), gy.ravel()], -1)[None] * 1.0 + rng.uniform(-0.05, 0.05, size=(S, N, 2))
    tg = np.zeros((S, N, 3))
    tg[:, :, :2] = 2 * st[:, :, :2].mean(1, keepdims=True) - st[:, :, :2]
    return st, tg


def scenario_range(total, rank, world):
    base, extra = divmod(total, world)
    return rank * base + min(rank, extra), base + (1 if rank < extra else 0)


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


def bind_to_gpu_numa_node(local, world):
    """Pins this rank to the CPUs of its GPU's NUMA node (its share of them when several ranks sit on one node) so
    that pinned host buffers and the launching thread are local to the GPU's PCIe root. Returns a description."""
    try:
        bus = subprocess.run(["nvidia-smi", f"--id={local}", "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                             capture_output=True, text=True, timeout=20).stdout.strip().lower()
        dom, rest = bus.split(":", 1)
        dev = f"/sys/bus/pci/devices/{dom[-4:]}:{rest}"
        node = int(open(dev + "/numa_node").read())
        cpus = []
        for part in open(dev + "/local_cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus += list(range(int(lo), int(hi or lo) + 1))
        allowed = sorted(set(cpus) & os.sched_getaffinity(0))
        if not allowed:
            return {"numa_node": node, "bound": False}
        # ranks whose GPUs share this node split its CPUs
        peers = []
        for r in range(world):
            b = subprocess.run(["nvidia-smi", f"--id={r}", "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                               capture_output=True, text=True, timeout=20).stdout.strip().lower()
            d2, r2 = b.split(":", 1)
            try:
                if int(open(f"/sys/bus/pci/devices/{d2[-4:]}:{r2}/numa_node").read()) == node:
                    peers.append(r)
            except OSError:
                pass
        if local in peers and len(peers) > 1:
            k, share = peers.index(local), max(1, len(allowed) // len(peers))
            allowed = allowed[k * share:(k + 1) * share] or allowed
        os.sched_setaffinity(0, allowed)
        return {"numa_node": node, "bound": True, "cpus": len(allowed)}
    except Exception as e:  # sysfs / nvidia-smi not available: run unbound
        return {"bound": False, "why": str(e)[:80]}


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
    """Closed-loop snapshots of a scenario sample computed by the CPU port: the same swarm, stride and state advance
    (curve(h) where optimize succeeded, else stay with zero velocity) as the device-generated snapshots of the GPU arm."""
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
    """Reference arm: the reference's CPU algorithm for the path (oracle port: CPLEX / Eigen / GiNaC are absent
    here) on all host threads. Same workload object as the GPU arm (same swarm, seed, snapshot
    stride); every step solves a bounded sample of it (the first S' scenarios of a snapshot)."""
    if int(os.environ.get("RANK", "0")) != 0:   # torchrun: rank 0 alone runs the CPU arm
        return
    from __graft_entry__ import load_oracle
    O = load_oracle()
    ocfg = oracle_cfg(O, args)
    threads = O.hardware_threads()
    S_full, N = args.scenarios, args.agents
    st, tg = make_swarm(S_full, N, args.seed)
    # bounded sample: calibrate so that snapshot generation + warm-up + steps take about 3 * cpu_seconds in total
    S0 = min(S_full, 4 * threads)
    t0 = time.perf_counter()
    O.step_batch(ocfg, st[:S0], tg[:S0], threads=threads)
    per_scn = (time.perf_counter() - t0) / S0
    total = max(1, args.steps + args.warmup + args.snapshots * SNAP_STRIDE)
    S = int(max(threads, min(S_full, (3 * args.cpu_seconds / total) / max(per_scn, 1e-9))))
    snaps = cpu_closed_loop_snapshots(O, ocfg, st[:S], tg[:S], args.snapshots, SNAP_STRIDE, threads)
    for i in range(args.warmup):
        O.step_batch(ocfg, snaps[i % len(snaps)], tg[:S], threads=threads)
    times = []
    for i in range(args.steps):
        t0 = time.perf_counter()
        O.step_batch(ocfg, snaps[(i + args.warmup) % len(snaps)], tg[:S], threads=threads)
        times.append(time.perf_counter() - t0)
    elapsed = sum(times)
    qps = S * N * ocfg.impc_iter * args.steps / elapsed
    sample = (f"first {S} of {S_full} scenarios of each closed-loop snapshot per step ({S * N * ocfg.impc_iter} QPs per step, "
              f"{elapsed:.1f} s timed)")
    line = {"impl": "reference", "metric": METRIC, "value": qps, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * elapsed / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_of(args, iters=ocfg.impc_iter, cbf_horizon=ocfg.cbf_horizon),
            "cpu_baseline": {"value": qps, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": qps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "p50_step_ms": 1e3 * float(np.median(times)),
            "note": "CPU restatement of the reference path (not CPLEX: unavailable on this box)"}
    print(json.dumps(line), flush=True)


def profile_record():
    """Newest profiles/r2*_fused_kernel.json (written by profiles/ncu_digest.py --json from an `ncu --set full`
    capture of this workload): DRAM traffic and executed FP64 instruction counts per launch, with the commit."""
    best = None
    for p in glob.glob(os.path.join(ROOT, "profiles", "*_fused_kernel.json")):
        try:
            d = json.load(open(p))
            if best is None or d.get("when", "") > best.get("when", ""):
                best = d
                best["file"] = os.path.relpath(p, ROOT)
        except (OSError, ValueError):
            pass
    return best


class DeviceLeg:
    """Device-resident workload of one rank: closed-loop snapshot sets generated by the product itself, the timed
    step (controller step + state advance), per-launch kernel time and iteration counts."""

    def __init__(self, torch, M, args, S, scen_begin, local, total_for_swarm):
        self.torch, self.M, self.S, self.N, self.K = torch, M, S, args.agents, args.k_hor
        dev = torch.device("cuda", local)
        self.dev = dev
        p = int(np.ceil((self.K - 1) * 0.1 / 0.5 - 1e-9))
        self.cfg = M.default_config(k_hor=self.K, num_pieces=max(p, 1), d_min=0.8, num_agents=self.N, max_scenarios=max(S, 1),
                                    device=local)
        self.h = M.Handle(self.cfg)
        self.n, self.iters = self.h.n, self.cfg.impc_iter
        N, n, iters = self.N, self.n, self.iters
        st_all, tg_all = make_swarm(total_for_swarm, N, args.seed)
        self.st_h = np.ascontiguousarray(st_all[scen_begin:scen_begin + S])
        self.tg_h = np.ascontiguousarray(tg_all[scen_begin:scen_begin + S])
        self.d_tg = torch.tensor(self.tg_h, device=dev)
        cur = torch.tensor(self.st_h, device=dev)
        self.R = max(2, args.snapshots)
        self.sets = []
        for _ in range(self.R):
            self.sets.append(dict(states=torch.empty_like(cur), ctrl=torch.zeros((S, N, n), dtype=torch.float64, device=dev),
                                  status=torch.zeros((S, N, iters), dtype=torch.int32, device=dev),
                                  n_curves=torch.zeros((S, N), dtype=torch.int32, device=dev),
                                  objective=torch.zeros((S, N), dtype=torch.float64, device=dev),
                                  nxt=torch.empty_like(cur)))
        self.stream = torch.cuda.Stream(device=dev)
        self.sp = self.stream.cuda_stream
        # second lane for the timed loop: consecutive steps are independent batches, so odd steps run on a second stream
        # through a second handle (its own workspaces) and the kernels of neighbouring steps fill each other's tails
        self.lanes = 1 if args.device_streams <= 1 else 2
        self.h2 = M.Handle(self.cfg) if self.lanes == 2 else None
        self.stream2 = torch.cuda.Stream(device=dev) if self.lanes == 2 else None
        h, sp = self.h, self.sp
        with torch.cuda.stream(self.stream):
            for i in range(self.R * SNAP_STRIDE):
                b = self.sets[i // SNAP_STRIDE]
                if i % SNAP_STRIDE == 0:
                    b["states"].copy_(cur)
                h.step_batch(S, cur, self.d_tg, b["ctrl"], b["status"], b["n_curves"], b["objective"], stream=sp)
                h.rollout(S, cur, b["ctrl"], b["n_curves"], cur, stream=sp)
        self.stream.synchronize()
        self.set_bytes = sum(t.numel() * t.element_size() for t in self.sets[0].values())
        self.graph = None

    def step(self, i, lane=0):
        b = self.sets[i % self.R]
        h, sp = (self.h, self.sp) if lane == 0 else (self.h2, self.stream2.cuda_stream)
        h.step_batch(self.S, b["states"], self.d_tg, b["ctrl"], b["status"], b["n_curves"], b["objective"], stream=sp)
        h.rollout(self.S, b["states"], b["ctrl"], b["n_curves"], b["nxt"], stream=sp)

    def one_pass(self):
        """The R steps of a pass: on one stream, or alternating between two that fork from and join the main stream."""
        torch = self.torch
        if self.lanes == 1:
            for i in range(self.R):
                self.step(i)
            return
        self.stream2.wait_stream(self.stream)
        for i in range(self.R):
            if i % 2 == 0:
                self.step(i, 0)
            else:
                with torch.cuda.stream(self.stream2):
                    self.step(i, 1)
        self.stream.wait_stream(self.stream2)

    def capture(self):
        """One CUDA graph = one pass over the R snapshot sets (R steps): the timed loop replays it, so that the
        launch path costs one graph launch per R steps instead of 2 R Python -> C-ABI calls."""
        torch = self.torch
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(self.stream):
            self.one_pass()
        self.stream.synchronize()
        with torch.cuda.graph(g, stream=self.stream):
            self.one_pass()
        self.graph = g

    def run_steps(self, first, count):
        """Enqueues `count` steps starting at snapshot index `first` (graph replays where whole passes fit)."""
        i, end = first, first + count
        with self.torch.cuda.stream(self.stream):
            while i < end:
                if self.graph is not None and i % self.R == 0 and end - i >= self.R:
                    self.graph.replay()
                    i += self.R
                elif self.graph is None and self.lanes == 2 and i % self.R == 0 and end - i >= self.R:
                    self.one_pass()
                    i += self.R
                else:
                    self.step(i)
                    i += 1

    def close(self):
        self.graph = None
        self.h.close()
        if self.h2 is not None:
            self.h2.close()


def time_device(torch, leg, args, barrier, max_over_ranks):
    """W warm-up steps, then exactly K timed steps between two events on the launching stream."""
    use_graph = not args.no_graph
    if use_graph:
        try:
            leg.capture()
        except Exception as e:  # capture unavailable: fall back to per-step launches, and say so
            leg.graph = None
            use_graph = False
            sys.stderr.write(f"bench: CUDA graph capture failed ({e}); launching every step from Python\n")
    leg.run_steps(0, args.warmup)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0, _ = leg.h.counters()
    first = ((args.warmup + leg.R - 1) // leg.R) * leg.R   # start the timed region on a pass boundary
    with torch.cuda.stream(leg.stream):
        e0.record(leg.stream)
    leg.run_steps(first, args.steps)
    with torch.cuda.stream(leg.stream):
        e1.record(leg.stream)
    barrier()
    l1, _ = leg.h.counters()
    total_ms = max_over_ranks(e0.elapsed_time(e1))
    launches = 2 * args.steps   # controller kernel + state-advance kernel per step (graph replays launch the same kernels)
    return total_ms, launches, use_graph, int(l1 - l0)


def step_latency(torch, leg, count):
    """Per-step device latency and the fused kernel's own duration, launched step by step with events."""
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(count)]
    h, sp = leg.h, leg.sp
    with torch.cuda.stream(leg.stream):
        for i in range(count):
            b = leg.sets[i % leg.R]
            ev[i][0].record(leg.stream)
            h.step_batch(leg.S, b["states"], leg.d_tg, b["ctrl"], b["status"], b["n_curves"], b["objective"], stream=sp)
            ev[i][1].record(leg.stream)
            h.rollout(leg.S, b["states"], b["ctrl"], b["n_curves"], b["nxt"], stream=sp)
            ev[i][2].record(leg.stream)
    leg.stream.synchronize()
    step_ms = [e[0].elapsed_time(e[2]) for e in ev]
    kern_ms = [e[0].elapsed_time(e[1]) for e in ev]
    return step_ms, float(np.mean(kern_ms[2:] if count > 4 else kern_ms))


def time_e2e(torch, M, leg, args, barrier, max_over_ranks, compact):
    """The host-pointer C-ABI calls on pinned HOST buffers, --e2e-depth batches in flight (one handle each): every step's
    inputs cross PCIe host -> device and its result device -> host inside the timed region."""
    S, N, n, iters, R = leg.S, leg.N, leg.n, leg.iters, leg.R
    hs = [b["states"].cpu().pin_memory() for b in leg.sets]
    htg = torch.tensor(leg.tg_h).pin_memory()
    key = "v_free" if compact else "ctrl_pts"

    def outs():
        o = {key: torch.zeros((S, N, 6 if compact else n), dtype=torch.float64).pin_memory(),
             "status": torch.zeros((S, N, iters), dtype=torch.int32).pin_memory(),
             "n_curves": torch.zeros((S, N), dtype=torch.int32).pin_memory(),
             "objective": torch.zeros((S, N), dtype=torch.float64).pin_memory()}
        return o, {k: v.numpy() for k, v in o.items()}
    D = max(2, args.e2e_depth)                       # batches in flight, one handle (stream + workspaces) each
    keeps, oo = zip(*[outs() for _ in range(D)])
    o0 = oo[0]
    hs_np, htg_np = [x.numpy() for x in hs], htg.numpy()
    extra = [M.Handle(leg.cfg) for _ in range(D - 1)]
    hh = (leg.h,) + tuple(extra)
    call = (lambda hd, st, o: hd.step_batch_host_compact_async(st, htg_np, o)) if compact else \
           (lambda hd, st, o: hd.step_batch_host_async(st, htg_np, o))
    for i in range(max(args.warmup, 2 * D)):
        call(hh[i % D], hs_np[i % R], oo[i % D])
        hh[i % D].synchronize()
    # the K-step loop is run three times (the whole loop lasts ~8 ms at K = 48: one host-side hiccup is several steps'
    # worth); the best repetition is reported, all are listed, and so is the median step-to-step interval
    reps, gaps = [], []
    for rep in range(3):
        barrier()
        t0 = time.perf_counter()
        last = t0
        for i in range(args.steps):
            k = i % D
            if i >= D:
                hh[k].synchronize()          # batch i - D is complete, its output buffers are free again
            call(hh[k], hs_np[(i + args.warmup) % R], oo[k])
            now = time.perf_counter()
            gaps.append(now - last)
            last = now
        for hd in hh:
            hd.synchronize()
        torch.cuda.synchronize()
        reps.append(max_over_ranks(time.perf_counter() - t0))
    secs = min(reps)
    barrier()
    # one blocking call per batch, for the latency of a single host call
    t0 = time.perf_counter()
    for i in range(args.steps):
        call(leg.h, hs_np[i % R], o0)
        leg.h.synchronize()
    serial = max_over_ranks(time.perf_counter() - t0)
    barrier()
    for hd in extra:
        hd.close()
    del keeps
    h2d = S * N * (6 + 3) * 8
    d2h = S * N * ((6 if compact else n) * 8 + iters * 4 + 4 + 8)
    diag = {"repetitions_s": reps, "p50_step_interval_ms": 1e3 * float(np.median(gaps)), "max_step_interval_ms": 1e3 * float(np.max(gaps))}
    return secs, serial, h2d, d2h, diag


def sharded_cfg4(torch, M, args, world, rank, local, barrier, max_over_ranks, swarms, steps):
    """BASELINE configs[3]: 256-agent dense swarms, all-pairs safety rows, agents split across the ranks, one NCCL
    all-gather of the [S][N/G][6] new states per control step (mpc-cbf_b200/sharded.py)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("mpc_cbf_b200_sharded", os.path.join(ROOT, "mpc-cbf_b200", "sharded.py"))
    sh = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sh)
    N = 256
    dev = torch.device("cuda", local)
    cfg = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, max_scenarios=swarms, device=local)
    st, tg = grid_swarm(swarms, N, 20261104)
    stepper = sh.AgentShardedStepper(M, cfg, swarms, rank, world, dev)
    states = torch.tensor(st, device=dev)
    tg_local = torch.tensor(np.ascontiguousarray(tg[:, stepper.begin:stepper.begin + stepper.cnt]), device=dev)
    for _ in range(5):
        stepper.step(states, tg_local)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        stepper.step(states, tg_local)
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    stepper.close()
    return {"swarms": swarms, "agents": N, "agents_per_gpu": N // world, "steps": steps, "ms_per_step": ms / steps,
            "value": swarms * N * 2 * steps / (ms * 1e-3), "unit": UNIT,
            "collective": "all_gather_into_tensor (NCCL), 1 per control step" if world > 1 else "none (1 rank)",
            "allgather_bytes_per_step": swarms * N * 48 if world > 1 else 0, "nranks": world}


def run_b200(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    affinity_all = os.sched_getaffinity(0)
    affinity = bind_to_gpu_numa_node(local, world)
    import torch
    from __graft_entry__ import load
    world, rank, local = dist_setup(args)
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    M = load()
    N, K = args.agents, args.k_hor

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

    # ---- the stated workload: 4096 scenarios in total, this rank's contiguous shard ------------------------------
    begin, S = scenario_range(args.scenarios, rank, world)
    leg = DeviceLeg(torch, M, args, S, begin, local, args.scenarios)
    n, iters = leg.n, leg.iters
    qps_total = args.scenarios * N * iters
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    total_ms, launches, used_graph, lib_launches = time_device(torch, leg, args, barrier, max_over_ranks)
    value = qps_total * args.steps / (total_ms * 1e-3)
    one_stream = None
    if leg.lanes == 2:   # the same K steps strictly one after the other on one stream, for comparison
        leg.lanes, leg.graph = 1, None
        s_ms, _, _, _ = time_device(torch, leg, args, barrier, max_over_ranks)
        leg.lanes, leg.graph = 2, None
        one_stream = {"value": qps_total * args.steps / (s_ms * 1e-3), "ms_per_step": s_ms / args.steps}
    step_ms, kern_ms = step_latency(torch, leg, max(16, min(args.steps, 64)))
    # ---- end to end ----------------------------------------------------------------------------------------------
    if args.device_only:
        clk = clocks.stop() if rank == 0 else None
        if rank == 0:
            print(json.dumps({"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                              "ms_per_step": total_ms / args.steps, "kernel_ms": kern_ms, "p50_step_ms": float(np.median(step_ms)),
                              "one_stream": one_stream,
                              "variant": {"core": os.environ.get("MPCCBF_THREAD_CORE"), "cfg": os.environ.get("MPCCBF_THREAD_CFG")},
                              "e2e": None, "clocks": clk, "note": "--device-only: kernel variant comparison, not a bench line"}), flush=True)
        leg.close()
        return
    e2e_s, e2e_serial_s, h2d, d2h, e2e_diag = time_e2e(torch, M, leg, args, barrier, max_over_ranks, compact=True)
    full_s, full_serial_s, _, d2h_full, full_diag = time_e2e(torch, M, leg, args, barrier, max_over_ranks, compact=False)
    clk = clocks.stop() if rank == 0 else None
    # status histogram over all snapshot sets of this rank (OPTIMAL 0, INFEASIBLE 3, SKIPPED 7 = IMPC iteration
    # after a failed one: never executed)
    status_np = np.concatenate([b["status"].cpu().numpy().ravel() for b in leg.sets])
    counts = {str(k): int(v) for k, v in zip(*np.unique(status_np, return_counts=True))}
    solved_frac = 1.0 - counts.get("7", 0) / max(1, status_np.size)
    fp64_peak = leg.h.measure_fp64_tflops()
    # algorithmic FP64 flops per launch: SURVEY 8d reduced-space formula with the MEASURED active-set iteration
    # count I of every QP,  F_qp = r^3/3 + m_cbf*8r + m_cbf*r^2 + I*(4*m*r + 2*r^2),  r = 6,  m = 6K + m_cbf
    r_dof, flops, mean_iters = 6.0, 0.0, 0.0
    for b_set in leg.sets:
        leg.h.step_batch(S, b_set["states"], leg.d_tg, b_set["ctrl"], b_set["status"], b_set["n_curves"], b_set["objective"],
                         stream=leg.sp)
        leg.stream.synchronize()
        gi = leg.h.last_iteration_counts(S * N * iters).reshape(S * N, iters).astype(np.float64)
        for it_idx in range(iters):
            m_cbf = (N - 1) * (1 if it_idx == 0 else leg.cfg.cbf_horizon)
            m_rows = 6 * K + m_cbf
            flops += float(np.sum(r_dof ** 3 / 3 + m_cbf * 8 * r_dof + m_cbf * r_dof ** 2
                                  + gi[:, it_idx] * (4 * m_rows * r_dof + 2 * r_dof ** 2))) / len(leg.sets)
        mean_iters += float(gi.mean()) / len(leg.sets)
    set_bytes, R, leg_lanes = leg.set_bytes, leg.R, leg.lanes
    snaps_h = [b["states"].cpu().numpy() for b in leg.sets] if (rank == 0 and not args.no_cpu_baseline) else None
    tg_h = leg.tg_h
    leg.close()

    # ---- sub-records: weak scaling and the agent-sharded path with its collective --------------------------------
    weak = None
    cfg4 = None
    if not args.no_extras:
        if world > 1:
            wleg = DeviceLeg(torch, M, args, args.scenarios, 0, local, args.scenarios)
            w_ms, _, _, _ = time_device(torch, wleg, args, barrier, max_over_ranks)
            weak = {"scaling": "weak", "scenarios_per_gpu": args.scenarios, "ms_per_step": w_ms / args.steps,
                    "value": world * qps_total * args.steps / (w_ms * 1e-3), "unit": UNIT}
            wleg.close()
        cfg4 = {"one_swarm": sharded_cfg4(torch, M, args, world, rank, local, barrier, max_over_ranks, 1, 40),
                "batch_of_256_swarms": sharded_cfg4(torch, M, args, world, rank, local, barrier, max_over_ranks, 256, 10)}
    if rank != 0:
        return

    # ---- roofline of the fused kernel -----------------------------------------------------------------------------
    alg_bytes = S * N * (6 * 8 + 3 * 8 + n * 8 + iters * 4 + 4 + 8 + iters * 4)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        hbm_peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    else:
        hbm_peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
    prof = profile_record()
    same_shape = prof is not None and prof.get("agents_per_launch") == S * N and world == 1
    achieved = flops / (kern_ms * 1e-3) / 1e12
    roofline = {"bound": "fp64", "achieved": achieved, "peak": fp64_peak, "unit": "TFLOP/s", "frac": achieved / fp64_peak,
                "traffic": prof.get("dram_bytes_per_launch") if same_shape else None,
                "kernel": "impc_thread_kernel", "kernel_ms": kern_ms, "agents_per_launch": S * N,
                "algorithmic_flops_per_launch": flops, "mean_active_set_iterations_per_qp": mean_iters,
                "peak_source": "register-resident DFMA loop measured in this run (mpccbf_measure_fp64_tflops); tensor cores are "
                               "not used: nothing on this path is GEMM-shaped",
                "formula": "SURVEY 8d reduced form F_qp = r^3/3 + 8 r m_cbf + r^2 m_cbf + I (4 m r + 2 r^2), r = 6, measured I",
                "hbm": {"bound": "hbm", "achieved": alg_bytes / (kern_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": alg_bytes / (kern_ms * 1e-3) / 1e9 / hbm_peak, "algorithmic_bytes_per_launch": alg_bytes,
                        "peak_source": peak_src, "note": "484 B per agent-step: not the binding roof of this kernel"}}
    if same_shape and prof.get("fp64_flops_executed_per_launch"):
        ex = float(prof["fp64_flops_executed_per_launch"])
        roofline["executed"] = {"fp64_flops_per_launch": ex, "achieved": ex / (kern_ms * 1e-3) / 1e12,
                                "frac": ex / (kern_ms * 1e-3) / 1e12 / fp64_peak,
                                "source": f"{prof['file']} (ncu smsp__sass_thread_inst_executed_op_d*_pred_on, commit {prof.get('commit')})"}

    cpu = None
    if not args.no_cpu_baseline:
        os.sched_setaffinity(0, affinity_all)   # the CPU baseline may use every host core
        from __graft_entry__ import load_oracle
        O = load_oracle()
        ocfg = oracle_cfg(O, args)
        threads = O.hardware_threads()
        S0 = min(S, 4 * threads)
        t0 = time.perf_counter()
        O.step_batch(ocfg, snaps_h[0][:S0], tg_h[:S0], threads=threads)
        per = (time.perf_counter() - t0) / S0
        Sc = int(max(min(threads, S), min(S, args.cpu_seconds / R / max(per, 1e-9))))
        t0 = time.perf_counter()
        for r in range(R):
            O.step_batch(ocfg, snaps_h[r][:Sc], tg_h[:Sc], threads=threads)
        dt = time.perf_counter() - t0
        cpu = {"value": Sc * N * iters * R / dt, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"first {Sc} of rank 0's {S} scenarios of each of the {R} snapshots ({Sc * N * iters * R} QPs, {dt:.1f} s)"}

    e2e = qps_total * args.steps / e2e_s
    cfgd = config_of(args, n=n, iters=iters, cbf_horizon=leg.cfg.cbf_horizon)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": cfgd,
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d * world if world > 1 else h2d,
                    "d2h_bytes_per_step": d2h * world if world > 1 else d2h,
                    "ms_per_step": 1e3 * e2e_s / args.steps, "pipeline_depth": max(2, args.e2e_depth), "result_form": "compact (v_free, 6 doubles per agent)",
                    "serial_value": qps_total * args.steps / e2e_serial_s, "serial_ms_per_step": 1e3 * e2e_serial_s / args.steps,
                    "full_ctrl_pts_value": qps_total * args.steps / full_s, "full_ctrl_pts_d2h_bytes_per_step": d2h_full * world,
                    "full_ctrl_pts_serial_value": qps_total * args.steps / full_serial_s,
                    "timing": e2e_diag, "full_ctrl_pts_timing": full_diag,
                    "note": "mpccbf_step_batch_host_compact_async on pinned host buffers, pipeline_depth batches in flight (one handle each); "
                            "bytes are summed over the ranks; serial_value = one blocking call per batch; full_ctrl_pts_* = the "
                            "same through mpccbf_step_batch_host_async returning all n control points per agent"},
            "gpu_launches": launches, "launched_with_cuda_graphs": used_graph,
            "p50_step_ms": float(np.median(step_ms)), "p99_step_ms": float(np.percentile(step_ms, 99)),
            "solved_value": value * solved_frac,
            "device_streams": leg_lanes, "one_stream": one_stream,
            "details": {"scenarios_this_rank": S, "qps_per_step_per_gpu": S * N * iters,
                        "status_counts_all_snapshots_rank0": counts,
                        "note_value": "value counts every (agent, IMPC iteration) slot like the reference's loop does; "
                                      "solved_value excludes the SKIPPED slots (iteration after a failed one). The K timed "
                                      "steps are independent batches: with device_streams = 2 consecutive steps alternate "
                                      "between two streams (joined once per pass over the snapshot sets) so that their "
                                      "kernels fill each other's tails; one_stream = the same K steps strictly in sequence; "
                                      "roofline.kernel_ms is one launch alone",
                        "snapshot_bytes_rank0": R * set_bytes, "affinity": affinity},
            "roofline": roofline, "cpu_baseline": cpu, "clocks": clk}
    if weak is not None:
        line["weak"] = weak
    if cfg4 is not None:
        line["sharded_cfg4"] = cfg4
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse_args()
    # Only the JSON line goes to stdout: libraries that print there (NCCL's version banner when a communicator is
    # created) are sent to stderr by pointing fd 1 at fd 2 and keeping the real stdout for print().
    sys.stdout.flush()
    _real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = _real_stdout
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
