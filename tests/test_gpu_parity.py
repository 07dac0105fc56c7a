"""GPU parity tests (run with -m gpu on the B200 box): the CUDA path, called through the C ABI
of libmpccbf.so, against the CPU oracle on the same seeded inputs."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, north_star_cfg, swarm
from parity import check_step, dense_scenarios

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device (no fallback)"
    return torch


def ocfg_of(O, cfg):
    oc = O.base_config()
    for name, _ in O.Config._fields_:
        v = getattr(cfg, name)
        if hasattr(v, "__len__"):
            getattr(oc, name)[:] = list(v)
        else:
            setattr(oc, name, v)
    return oc


def test_two_agent_swap_shipped_config(M, O, torch_cuda):
    # BASELINE config[0]: 2-agent head-on swap, base_config.json, one control step + closed loop
    cfg, so, sf = M.config_from_json(os.path.join(GOLDEN, "circle2_config.json"))
    cfg.max_scenarios = 1
    h = M.Handle(cfg)
    oc = ocfg_of(O, cfg)
    U = O.get_table(oc, "U_basis")
    st = np.zeros((1, 2, 6)); st[0, :, :3] = so
    tg = sf[None].copy()
    for step in range(30):
        got = h.step_batch_host(st, tg)
        check_step(O, oc, U, st, tg, got, check_viol_max=2)
        # advance (noise-free rollout, MPCCBFFormationControl_example.cpp:191-205)
        for a in range(2):
            if got["n_curves"][0, a] > 0:
                x = got["ctrl_pts"][0, a]
                st[0, a, :3] = O.eval_curve(oc, x, cfg.h, 0)
                st[0, a, 3:] = O.eval_curve(oc, x, cfg.h, 1)
            else:  # MPCCBFFormationControl_example.cpp:208-221: stay in place
                st[0, a, 3:] = 0.0
    assert abs(h.step_batch_host(np.array([[[2, 2, 0, 0, 0, 0], [-2, -2, 0, 0, 0, 0]]], float),
                                 tg)["objective"][0, 0] + 1378.9905) < 1e-3
    h.close()


def test_eight_agent_circle_k20(M, O, torch_cuda):
    # BASELINE config[1]: 8-agent circle swap, horizon 20, closed loop through the swap
    cfg, so, sf = M.config_from_json(os.path.join(GOLDEN, "circle8_config.json"))
    cfg = cfg.copy(k_hor=20, num_pieces=4, d_min=0.8, max_scenarios=1)
    h = M.Handle(cfg)
    oc = ocfg_of(O, cfg)
    U = O.get_table(oc, "U_basis")
    N = cfg.num_agents
    st = np.zeros((1, N, 6)); st[0, :, :3] = so
    tg = sf[None].copy()
    active = 0
    for step in range(60):
        got = h.step_batch_host(st, tg)
        r = check_step(O, oc, U, st, tg, got, check_viol_max=4)
        for a in range(N):
            if got["n_curves"][0, a] > 0:
                x = got["ctrl_pts"][0, a]
                st[0, a, :3] = O.eval_curve(oc, x, cfg.h, 0)
                st[0, a, 3:] = O.eval_curve(oc, x, cfg.h, 1)
            else:
                st[0, a, 3:] = 0.0
    d = np.linalg.norm(st[0, :, None, :2] - st[0, None, :, :2], axis=-1) + 10 * np.eye(N)
    assert d.min() > cfg.d_min - 1e-6
    h.close()


@pytest.mark.parametrize("K,p,N,S", [(20, 4, 16, 64), (16, 3, 2, 128), (10, 2, 8, 64), (40, 8, 4, 32), (20, 4, 64, 4)])
def test_free_space_swarm(M, O, torch_cuda, K, p, N, S, monkeypatch):
    monkeypatch.setenv("MPCCBF_LANES", "0")
    cfg = north_star_cfg(M, k_hor=K, num_pieces=p, num_agents=N, max_scenarios=S)
    st, tg = swarm(S, N, seed=K * 1000 + N)
    h = M.Handle(cfg)
    got = h.step_batch_host(st, tg)
    oc = ocfg_of(O, cfg)
    check_step(O, oc, O.get_table(oc, "U_basis"), st, tg, got, check_viol_max=8)
    h.close()


@pytest.mark.parametrize("N,seed,lanes", [(4, 11, 0), (16, 12, 0), (33, 13, 0), (16, 14, 8), (40, 15, 0)])
def test_dense_active_cbf_and_infeasible(M, O, torch_cuda, N, seed, lanes, monkeypatch):
    monkeypatch.setenv("MPCCBF_LANES", str(lanes))  # 0 = thread-per-agent kernel, 8 = lane groups
    S = 32
    cfg = north_star_cfg(M, num_agents=N, max_scenarios=S)
    st, tg = dense_scenarios(S, N, seed, 0.8, spread=1.2, vmax=1.5)
    st[0, 0, 3] = 2.5   # |v0| > v_max  -> INFEASIBLE at iter 0, SKIPPED after
    st[1, 1, :2] = st[1, 0, :2] + [0.3, 0.0]   # inside d_min
    st[1, 1, 3:5] = [-1.0, 0.0]; st[1, 0, 3:5] = [1.0, 0.0]   # closing
    h = M.Handle(cfg)
    got = h.step_batch_host(st, tg)
    oc = ocfg_of(O, cfg)
    r = check_step(O, oc, O.get_table(oc, "U_basis"), st, tg, got, check_viol_max=24)
    assert got["status"][0, 0, 0] == M.STATUS_INFEASIBLE and got["status"][0, 0, 1] == M.STATUS_SKIPPED
    h.close()


@pytest.mark.parametrize("lanes", [0, 1, 8, 32])
def test_lane_widths_agree(M, O, torch_cuda, lanes, monkeypatch):
    monkeypatch.setenv("MPCCBF_LANES", str(lanes))
    cfg = north_star_cfg(M, num_agents=9, max_scenarios=16)
    st, tg = dense_scenarios(16, 9, 5, 0.8, spread=1.25, vmax=1.0)
    h = M.Handle(cfg)
    got = h.step_batch_host(st, tg)
    oc = ocfg_of(O, cfg)
    check_step(O, oc, O.get_table(oc, "U_basis"), st, tg, got, check_viol_max=4)
    h.close()


def test_edge_cases(M, O, torch_cuda):
    # empty batch, single agent (no neighbours), batch larger than capacity
    cfg = north_star_cfg(M, num_agents=1, max_scenarios=8)
    h = M.Handle(cfg)
    out = h.step_batch_host(np.zeros((0, 1, 6)), np.zeros((0, 1, 3)))
    assert out["ctrl_pts"].shape == (0, 1, h.n)
    st, tg = swarm(8, 1, seed=3)
    got = h.step_batch_host(st, tg)
    oc = ocfg_of(O, cfg)
    check_step(O, oc, O.get_table(oc, "U_basis"), st, tg, got, check_viol_max=2)
    with pytest.raises(M.MpcCbfError):
        h.step_batch_host(np.zeros((9, 1, 6)), np.zeros((9, 1, 3)))
    h.close()


def test_one_sided_limits_and_single_agent(M, O, torch_cuda):
    # (i) a velocity limit of 1e30 ("no bound" for CPLEX) disables the key scan of the thread kernel
    # (it needs finite two-sided bounds): the step falls back to the lane-group kernel and must still
    # agree with the oracle on a batch large enough for the throughput path; (ii) N = 1: no neighbours
    S, N = 320, 16
    cfg = north_star_cfg(M, num_agents=N, max_scenarios=S, v_max=[2.0, 1e30, 2.6])
    h = M.Handle(cfg)
    st, tg = dense_scenarios(S, N, 21, 0.8, spread=2.5, vmax=1.5)
    got = h.step_batch_host(st, tg)
    oc = ocfg_of(O, cfg)
    r = check_step(O, oc, O.get_table(oc, "U_basis"), st[:24], tg[:24], {k: v[:24] for k, v in got.items()}, check_viol_max=8)
    assert r["n_infeasible"] < 24 * N
    h.close()
    cfg = north_star_cfg(M, num_agents=1, max_scenarios=8)
    h = M.Handle(cfg)
    st, tg = swarm(8, 1, seed=4)
    got = h.step_batch_host(st, tg)
    oc = ocfg_of(O, cfg)
    check_step(O, oc, O.get_table(oc, "U_basis"), st, tg, got)
    assert (got["status"] == 0).all()
    h.close()


@pytest.mark.parametrize("S,N", [(3, 16), (600, 16), (2, 200)])
def test_compact_result_form_is_the_same_answer(M, O, S, N):
    # mpccbf_step_batch_host_compact returns the 6 whitened free dof per agent instead of the n control
    # points; expanding them on the host reproduces mpccbf_step_batch_host bit for bit (both kernels: the
    # lane-group kernel for small batches, the thread-per-agent kernel beyond).
    cfg = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, max_scenarios=S)
    h = M.Handle(cfg)
    st, tg = swarm(S, N, seed=S + N)
    st[:, :, :2] *= 0.5 * N / 16
    tg[:, :, :2] *= 0.5 * N / 16
    st[0, 0, 3] = 2.5  # one infeasible agent: zeros in both forms
    full = h.step_batch_host(st, tg)
    comp = h.step_batch_host_compact(st, tg)
    for k in ("status", "n_curves", "objective"):
        assert np.array_equal(full[k], comp[k]), k
    assert np.array_equal(h.expand_ctrl_pts(st, comp["v_free"], comp["n_curves"]).reshape(S, N, -1), full["ctrl_pts"])
    assert (comp["v_free"][0, 0] == 0).all() and full["n_curves"][0, 0] == 0
    # the driver's state advance straight from the compact record
    oc = O.base_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N)
    pos = h.eval_compact(st, comp["v_free"], cfg.h, 0).reshape(S, N, 3)
    a = int(np.argmax(full["n_curves"][S - 1] > 0))
    assert np.abs(pos[S - 1, a] - O.eval_curve(oc, full["ctrl_pts"][S - 1, a], cfg.h, 0)).max() <= 1e-12
    h.close()


@pytest.mark.parametrize("extra,N,S", [
    (dict(continuity_degree=2), 8, 6),
    (dict(continuity_degree=1), 6, 4),
    (dict(num_ctrl_pts=5, continuity_degree=3), 6, 4),
    (dict(num_ctrl_pts=5, continuity_degree=2, k_hor=16, num_pieces=3), 4, 8),
])
def test_general_configurations_on_the_device(M, O, extra, N, S):
    # non-default continuity degree / control-point count: impc_general_kernel (r free dof per axis) through the same
    # controller entries, closed loop of 3 steps against the oracle; the compact form is refused for them
    base = dict(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N)
    base.update(extra)
    cfg, oc = M.default_config(max_scenarios=S, **base), O.base_config(**base)
    h = M.Handle(cfg)
    assert h.solver == M.SOLVER_REDUCED
    st, tg = dense_scenarios(S, N, 5, 0.8, spread=1.3, vmax=1.0)
    st[0, 0, 3] = 2.5
    U = O.get_table(oc, "U_basis")
    for _ in range(2):
        got = h.step_batch_host(st, tg)
        check_step(O, oc, U, st, tg, got, check_viol_max=2)
        for s in range(S):
            for a in range(N):
                if got["n_curves"][s, a] > 0:
                    st[s, a, :3] = O.eval_curve(oc, got["ctrl_pts"][s, a], cfg.h, 0)
                    st[s, a, 3:] = O.eval_curve(oc, got["ctrl_pts"][s, a], cfg.h, 1)
        st[0, 0, 3] = 2.5
    r = h.optimize_one(st[1], 0, np.tile(tg[1, 0], cfg.k_hor))
    ro = O.optimize(oc, st[1], 0, np.tile(tg[1, 0], cfg.k_hor))
    assert list(r["status"]) == list(ro["status"]) and np.abs(r["curves"] - ro["curves"]).max() <= 1e-5
    with pytest.raises(M.MpcCbfError):
        h.step_batch_host_compact(st, tg)
    h.close()
