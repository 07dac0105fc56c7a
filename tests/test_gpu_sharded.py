"""Agent-sharded control step on the device (BASELINE configs[3] shape: one dense swarm whose agents
are split across ranks). On the single test GPU the shards run one after the other through the
same C-ABI calls a rank makes (mpccbf_step_agents + mpccbf_rollout_agents); the union must equal
the unsharded step, and the closed loop must agree with the CPU oracle."""
import importlib.util
import os

import numpy as np
import pytest

from conftest import ROOT, north_star_cfg
from parity import check_step, dense_scenarios
from test_gpu_parity import ocfg_of

pytestmark = pytest.mark.gpu


def _sharded():
    spec = importlib.util.spec_from_file_location("mpc_cbf_b200_sharded", os.path.join(ROOT, "mpc-cbf_b200", "sharded.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.parametrize("N,G,S", [(64, 4, 2), (256, 8, 1)])
def test_agent_shards_equal_unsharded_step(M, O, N, G, S):
    import torch
    dev = torch.device("cuda", 0)
    sh = _sharded()
    cfg = north_star_cfg(M, num_agents=N, max_scenarios=S)
    st, tg = dense_scenarios(S, N, 21, 0.8, spread=1.25, vmax=0.8)   # 16x16 grid at pitch 1.0 for N = 256
    h = M.Handle(cfg)
    full = h.step_batch_host(st, tg)
    d_st = torch.tensor(st, device=dev)
    n, it = h.n, cfg.impc_iter
    nxt_full = torch.zeros((S, N, 6), dtype=torch.float64, device=dev)
    d_ctrl_full = torch.tensor(full["ctrl_pts"], device=dev)
    d_nc_full = torch.tensor(full["n_curves"], device=dev)
    h.rollout(S, d_st, d_ctrl_full, d_nc_full, nxt_full)
    torch.cuda.synchronize()
    gathered = torch.zeros_like(nxt_full)
    for r in range(G):
        b, c = sh.agent_range(N, r, G)
        ctrl = torch.zeros((S, c, n), dtype=torch.float64, device=dev)
        status = torch.zeros((S, c, it), dtype=torch.int32, device=dev)
        ncv = torch.zeros((S, c), dtype=torch.int32, device=dev)
        obj = torch.zeros((S, c), dtype=torch.float64, device=dev)
        d_tg = torch.tensor(np.ascontiguousarray(tg[:, b:b + c]), device=dev)
        h.step_agents(S, b, c, d_st, d_tg, ctrl, status, ncv, obj)
        loc = d_st[:, b:b + c, :].contiguous()
        nxt = torch.zeros((S, c, 6), dtype=torch.float64, device=dev)
        h.rollout_agents(S, c, loc, ctrl, ncv, nxt)
        torch.cuda.synchronize()
        assert np.array_equal(status.cpu().numpy(), full["status"][:, b:b + c])
        assert np.array_equal(ctrl.cpu().numpy(), full["ctrl_pts"][:, b:b + c])   # same kernel, same inputs: bit-equal
        gathered[:, b:b + c, :] = nxt
    assert torch.equal(gathered, nxt_full)
    # and the unsharded step itself is right (oracle) — a bounded sample of agents for the big swarm
    oc = ocfg_of(O, cfg)
    if N <= 64:
        check_step(O, oc, O.get_table(oc, "U_basis"), st, tg, full, check_viol_max=8)
    else:
        sel = np.arange(0, N, 16)
        ref = {k: np.zeros_like(v[:, sel]) for k, v in full.items()}
        for j, a in enumerate(sel):
            r1 = O.optimize(oc, st[0], int(a), np.tile(tg[0, a], oc.k_hor))
            ref["status"][0, j] = r1["status"]
            ref["n_curves"][0, j] = r1["n_curves"]
            if r1["n_curves"] > 0:
                ref["ctrl_pts"][0, j] = r1["curves"][-1]
        assert np.array_equal(ref["status"], full["status"][:, sel])
        ok = ref["n_curves"] > 0
        assert np.abs(ref["ctrl_pts"] - full["ctrl_pts"][:, sel])[ok].max() < 1e-5
    h.close()


def driver_follow(O, oc, cur, res, prev, et, nsub=10):
    """One control step of the reference driver's state advance (MPCCBFFormationControl_example.cpp:161-221):
    a new curve is adopted where optimize succeeded; otherwise the agent keeps following its previous curve from
    where it was, the parameter clamped to max_parameter; an agent that never had a curve stays put, velocity 0."""
    S, N = cur.shape[:2]
    maxp = oc.num_pieces * oc.piece_max_param
    for s in range(S):
        for a in range(N):
            if res["n_curves"][s, a] > 0:
                prev[s][a], et[s, a] = res["ctrl_pts"][s, a].copy(), 0.0
            if prev[s][a] is None:
                cur[s, a, 3:] = 0.0
                continue
            t = min(et[s, a] + oc.Ts * nsub, maxp)
            cur[s, a, :3] = O.eval_curve(oc, prev[s][a], t, 0)
            cur[s, a, 3:] = O.eval_curve(oc, prev[s][a], t, 1)
            et[s, a] = t


def test_stepper_world1_closed_loop(M, O):
    import torch
    dev = torch.device("cuda", 0)
    sh = _sharded()
    N, S = 32, 2
    cfg = north_star_cfg(M, num_agents=N, max_scenarios=S)
    st, tg = dense_scenarios(S, N, 4, 0.8, spread=1.15, vmax=1.2)   # packed: some optimize calls fail along the way
    stepper = sh.AgentShardedStepper(M, cfg, S, 0, 1, dev)
    d_states = torch.tensor(st, device=dev)
    d_tg = torch.tensor(tg, device=dev)
    oc = ocfg_of(O, cfg)
    cur = st.copy()
    prev = [[None] * N for _ in range(S)]
    et = np.zeros((S, N))
    failed = 0
    for step in range(8):
        stepper.step(d_states, d_tg)
        torch.cuda.synchronize()
        res = O.step_batch(oc, cur, tg, threads=O.hardware_threads(), prec=1)
        failed += int((res["n_curves"] == 0).sum())
        driver_follow(O, oc, cur, res, prev, et)
        assert np.abs(d_states.cpu().numpy() - cur).max() < 1e-6, step
    assert failed > 0   # the fallback (keep following the previous curve) was exercised
    stepper.close()


def test_sharded_loop_equals_the_device_episode(M):
    # the agent-sharded loop (here: two shards run one after the other on one GPU) retraces mpccbf_run_episode
    import torch
    dev = torch.device("cuda", 0)
    N, S, steps = 16, 3, 12
    cfg = north_star_cfg(M, num_agents=N, max_scenarios=S)
    st, tg = dense_scenarios(S, N, 9, 0.8, spread=1.2, vmax=1.0)
    h = M.Handle(cfg)
    d_ep = torch.tensor(st, device=dev)
    d_tg = torch.tensor(tg, device=dev)
    h.run_episode(S, d_ep, d_tg, steps, 1.0, 0.2)
    torch.cuda.synchronize()
    h.set_kernel_variant(h.auto_kernel_variant(S * N))
    cur = torch.tensor(st, device=dev)
    h.follow_reset(S)
    half = N // 2
    ctrl = torch.zeros((S, half, h.n), dtype=torch.float64, device=dev)
    status = torch.zeros((S, half, 2), dtype=torch.int32, device=dev)
    ncv = torch.zeros((S, half), dtype=torch.int32, device=dev)
    tgs = [d_tg[:, :half].contiguous(), d_tg[:, half:].contiguous()]
    for _ in range(steps):
        snap = cur.clone()   # both shards solve on the same snapshot (Jacobi), like two ranks would
        for r in range(2):
            h.step_agents(S, r * half, half, snap, tgs[r], ctrl, status, ncv, None)
            h.follow_step(S, r * half, half, cur, ctrl, ncv)
    torch.cuda.synchronize()
    assert torch.equal(cur, d_ep)
    h.close()


def test_multi_handle_single_device_and_two_devices(M, O):
    # mpccbf_multi_*: one process, one thread + stream per device, NCCL all-gather behind the C ABI. G = 1 always
    # runs; G = 2 when the box has two GPUs (gpurun --gpus 2). Scenario-sharded step == single-handle step bit for
    # bit; agent-sharded closed loop == G = 1 closed loop bit for bit.
    import torch
    N, S, steps = 16, 6, 10
    cfg = north_star_cfg(M, num_agents=N, max_scenarios=S)
    st, tg = dense_scenarios(S, N, 13, 0.8, spread=1.2, vmax=1.0)
    h = M.Handle(cfg)
    ref = h.step_batch_host(st, tg)
    h.close()
    finals = {}
    for G in (1, 2):
        if G > torch.cuda.device_count():
            continue
        m = M.Multi(cfg, G)
        assert m.num_devices == G and m.comm_ranks == G
        got = m.step_batch_host(st, tg)
        for k in ref:
            assert np.array_equal(got[k], ref[k]), (G, k)
        comp = m.step_batch_host(st, tg, compact=True)
        assert np.array_equal(comp["status"], ref["status"])
        finals[G], ms = m.run_sharded(st, tg, steps)
        assert ms > 0
        m.close()
    assert np.abs(finals[1] - st).max() > 0.05   # the swarm moved
    if 2 in finals:
        assert np.array_equal(finals[1], finals[2])
