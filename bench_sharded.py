#!/usr/bin/env python
"""BASELINE configs[3]: ONE dense swarm of 256 agents (all-pairs safety rows) whose agents are split
across the GPUs of one box; each control step every rank solves its own agents against the full state
snapshot, rolls them forward, and a single NCCL all-gather of the next states rebuilds the snapshot
(mpc-cbf_b200/sharded.py). Launch:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 \
        --master-port 29533 bench_sharded.py [--scenarios 1] [--agents 256] [--steps 100]

Timing: CUDA events around K closed-loop control steps after W warm-up steps, barrier on both sides,
max over ranks; rank 0 prints one JSON line. With G = 1 it is the unsharded closed loop.
"""
import argparse
import importlib.util
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def grid_swarm(S, N, seed):
    """sqrt(N) x sqrt(N) grid, pitch 1.25 d_min = 1.0 with jitter, goals mirrored through the centroid."""
    rng = np.random.default_rng(seed)
    side = int(round(np.sqrt(N)))
    assert side * side == N, "cfg4 wants a square number of agents"
    gx, gy = np.meshgrid(np.arange(side), np.arange(side), indexing="ij")
    base = np.stack([gx.ravel(), gy.ravel()], -1) * 1.0
    st = np.zeros((S, N, 6))
    st[:, :, :2] = base[None] + rng.uniform(-0.05, 0.05, size=(S, N, 2))
    c = st[:, :, :2].mean(1, keepdims=True)
    tg = np.zeros((S, N, 3))
    tg[:, :, :2] = 2 * c - st[:, :, :2]
    return st, tg


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scenarios", type=int, default=1)
    ap.add_argument("--agents", type=int, default=256)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    from __graft_entry__ import load
    M = load()
    spec = importlib.util.spec_from_file_location("mpc_cbf_b200_sharded", os.path.join(ROOT, "mpc-cbf_b200", "sharded.py"))
    sh = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sh)
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29533")
        dist.init_process_group(backend="nccl", rank=rank, world_size=world, device_id=dev)
    S, N = a.scenarios, a.agents
    cfg = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, max_scenarios=S, device=local)
    st, tg = grid_swarm(S, N, 20261104)
    stepper = sh.AgentShardedStepper(M, cfg, S, rank, world, dev)
    states = torch.tensor(st, device=dev)
    tg_local = torch.tensor(tg[:, stepper.begin:stepper.begin + stepper.cnt], device=dev).contiguous()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(a.warmup):
        stepper.step(states, tg_local)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        stepper.step(states, tg_local)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    nfail = int((stepper.status != 0).sum().item())
    if rank == 0:
        print(json.dumps({"metric": "mpc_cbf_qps_per_sec", "config": {"workload": f"{S} x {N}-agent dense swarm, all-pairs safety rows, K=20 "
                          f"(BASELINE configs[3]), agents sharded {N // world} per GPU", "scenarios": S, "agents": N},
                          "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms / a.steps,
                          "value": S * N * 2 * a.steps / (ms * 1e-3), "unit": "QPs/s", "scaling": "strong",
                          "allgather_bytes_per_step": S * N * 48 if world > 1 else 0,
                          "collective": "all_gather_into_tensor (NCCL)" if world > 1 else "none",
                          "non_optimal_qps_last_step_rank0": nfail}), flush=True)
    stepper.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
