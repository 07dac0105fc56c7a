"""Host-side multi-GPU logic (mpc-cbf_b200/sharded.py) on CPU: world_size-2 `gloo` process groups.

What runs here is the sharding arithmetic and the ONE collective of the agent-sharded path (the
all-gather of next states); the compute kernels need a B200 and are covered by the -m gpu tests
(tests/test_gpu_sharded.py runs the same stepper on the device)."""
import importlib.util
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load_sharded():
    spec = importlib.util.spec_from_file_location("mpc_cbf_b200_sharded", os.path.join(ROOT, "mpc-cbf_b200", "sharded.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, S, N, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sh = _load_sharded()
        begin, cnt = sh.agent_range(N, rank, world)
        rng = np.random.default_rng(7)
        full = torch.tensor(rng.normal(size=(S, N, 6)))           # same on every rank
        # every rank "advances" only its own agents (stand-in for step_agents + rollout_agents)
        local_next = full[:, begin:begin + cnt, :] * 2.0 + rank
        out = torch.zeros_like(full)
        sh.gather_states(local_next.contiguous(), out, world)
        expect = full * 2.0
        for r in range(world):
            b, c = sh.agent_range(N, r, world)
            expect[:, b:b + c, :] += r
        ok = bool(torch.equal(out, expect))
        # scenario sharding: ranks' shards tile [0, S) without a collective
        sb, sc = sh.scenario_range(S * 5 + 3, rank, world)
        spans = [None] * world
        dist.all_gather_object(spans, (sb, sc))
        q.put((rank, ok, spans))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("S,N", [(1, 256), (3, 16)])
def test_agent_sharded_all_gather_world2(S, N):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, S, N, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ok, spans in res:
        assert ok, f"rank {rank}: gathered snapshot differs"
        total = S * 5 + 3
        assert spans[0][0] == 0 and spans[-1][0] + spans[-1][1] == total
        assert all(spans[i][0] + spans[i][1] == spans[i + 1][0] for i in range(world - 1))
        assert max(c for _, c in spans) - min(c for _, c in spans) <= 1


def test_range_helpers():
    sh = _load_sharded()
    assert sh.scenario_range(4096, 3, 8) == (1536, 512)
    assert [sh.scenario_range(10, r, 4) for r in range(4)] == [(0, 3), (3, 3), (6, 2), (8, 2)]
    assert sh.agent_range(256, 5, 8) == (160, 32)
    with pytest.raises(ValueError):
        sh.agent_range(10, 0, 4)
    # world 1: gather is a plain copy
    a = torch.arange(12, dtype=torch.float64).view(1, 2, 6)
    out = torch.zeros_like(a)
    sh.gather_states(a, out, 1)
    assert torch.equal(a, out)
