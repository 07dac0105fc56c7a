"""CPU-only: the fused kernel's per-agent device function, compiled for the host
(tests/emul/libmpccbf_emul.so, -DMPCCBF_HOST_EMUL), against the oracle. This checks the
active-set logic of the CUDA source in a container without a GPU; the real kernel is checked by
tests/test_gpu_parity.py (-m gpu)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, swarm
from parity import check_step, dense_scenarios


@pytest.fixture(scope="module")
def emul(M):
    d = os.path.join(ROOT, "tests", "emul")
    subprocess.run(["make", "-C", d], check=True, stdout=subprocess.DEVNULL)
    L = M.lib(os.path.join(d, "libmpccbf_emul.so"))
    L.mpccbf_emul_step.argtypes = [C.c_void_p] + [C.c_int32] * 4 + [C.c_void_p] * 9
    return L


def run_emul(M, L, cfg, states, targets, compact=True):
    cfg = cfg.copy(device=-1)
    h = M.Handle(cfg, library=L)
    assert h.solver == M.SOLVER_REDUCED
    S, N = states.shape[:2]
    it = cfg.impc_iter
    out = dict(ctrl_pts=np.zeros((S, N, h.n)), status=np.zeros((S, N, it), np.int32),
               n_curves=np.zeros((S, N), np.int32), objective=np.zeros((S, N)), v_free=np.zeros((S, N, 6)))
    iters = np.zeros((S, N, it), np.int32)
    st, tg = np.ascontiguousarray(states), np.ascontiguousarray(targets)
    rc = L.mpccbf_emul_step(h._h, S, 0, N, 0, st.ctypes.data, tg.ctypes.data, out["ctrl_pts"].ctypes.data, None,
                            out["status"].ctypes.data, out["n_curves"].ctypes.data, out["objective"].ctypes.data,
                            iters.ctypes.data, out["v_free"].ctypes.data if compact else None)
    assert rc == 0
    out["iters"] = iters
    # compact result form: the host-side expansion reproduces the kernel's control points bit for bit, and
    # evaluating the curve from the compact record equals evaluating the expanded curve
    if compact:
        assert np.array_equal(h.expand_ctrl_pts(st, out["v_free"], out["n_curves"]).reshape(S, N, -1), out["ctrl_pts"])
    h.close()
    return out


def test_two_agent_swap(M, O, emul):
    cfg = M.default_config()
    st = np.array([[[2, 2, 0, 0, 0, 0], [-2, -2, 0, 0, 0, 0]]], float)
    tg = np.array([[[-2, -2, 0], [2, 2, 0]]], float)
    got = run_emul(M, emul, cfg, st, tg)
    oc = O.base_config()
    r = check_step(O, oc, O.get_table(oc, "U_basis"), st, tg, got)
    assert r["n_infeasible"] == 0
    assert abs(got["objective"][0, 0] + 1378.9905) < 1e-3


@pytest.mark.parametrize("K,p,N,S", [(16, 3, 4, 24), (20, 4, 16, 6), (10, 2, 8, 8)])
def test_free_space_swarm(M, O, emul, K, p, N, S):
    cfg = M.default_config(k_hor=K, num_pieces=p, d_min=0.8, num_agents=N)
    st, tg = swarm(S, N, seed=K * 100 + N)
    got = run_emul(M, emul, cfg, st, tg)
    oc = O.base_config(k_hor=K, num_pieces=p, d_min=0.8, num_agents=N)
    check_step(O, oc, O.get_table(oc, "U_basis"), st, tg, got, check_viol_max=8)


@pytest.mark.parametrize("N,seed", [(4, 1), (9, 2), (16, 3)])
def test_dense_active_cbf_and_infeasible(M, O, emul, N, seed):
    cfg = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N)
    st, tg = dense_scenarios(12, N, seed, 0.8, spread=1.2, vmax=1.5)
    st[0, 0, 3] = 2.5  # |v0| > v_max: the velocity row at k = 0 contradicts the initial-velocity equality
    got = run_emul(M, emul, cfg, st, tg)
    oc = O.base_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N)
    r = check_step(O, oc, O.get_table(oc, "U_basis"), st, tg, got, check_viol_max=16)
    assert got["status"][0, 0, 0] == M.STATUS_INFEASIBLE and got["status"][0, 0, 1] == M.STATUS_SKIPPED
    assert got["iters"].max() > 4  # CBF rows were active somewhere


def test_coincident_agents_zero_normal_row(M, O, emul):
    # two agents at the same xy position: a = (0, 0), b < 0 -> a zero row that no x satisfies.
    cfg = M.default_config(num_agents=3)
    st = np.zeros((1, 3, 6)); st[0, 2, :2] = [5.0, 5.0]
    tg = np.zeros((1, 3, 3)); tg[0, :, 0] = 1.0
    got = run_emul(M, emul, cfg, st, tg)
    ref = O.step_batch(O.base_config(num_agents=3), st, tg, prec=1)
    assert np.array_equal(got["status"], ref["status"])
    assert list(got["status"][0, 0]) == [M.STATUS_INFEASIBLE, M.STATUS_SKIPPED]
    assert list(got["status"][0, 2]) == [M.STATUS_OPTIMAL, M.STATUS_OPTIMAL]


@pytest.mark.parametrize("K,p", [(20, 4), (16, 3), (40, 8)])
def test_bound_rows_active_everywhere(M, O, emul, K, p, monkeypatch):
    # velocities near the limits and far targets: velocity / acceleration rows bind at interior
    # horizon samples, which exercises the extremum-located row scan (DevTables::poly) against the
    # oracle, and the same run with the exhaustive scan must give the same answer
    rng = np.random.default_rng(K)
    S, N = 40, 4
    st = np.zeros((S, N, 6))
    st[:, :, :2] = rng.uniform(-20, 20, size=(S, N, 2)) + 50 * np.arange(N)[None, :, None]
    st[:, :, 2] = rng.uniform(-1, 1, size=(S, N))
    st[:, :, 3:5] = rng.uniform(-1.95, 1.95, size=(S, N, 2))
    st[:, :, 5] = rng.uniform(-2.5, 2.5, size=(S, N))
    tg = np.zeros((S, N, 3))
    tg[:, :, :2] = st[:, :, :2] + rng.uniform(-6, 6, size=(S, N, 2))
    tg[:, :, 2] = rng.uniform(-3, 3, size=(S, N))
    cfg = M.default_config(k_hor=K, num_pieces=p, d_min=0.8, num_agents=N)
    got = run_emul(M, emul, cfg, st, tg)
    oc = O.base_config(k_hor=K, num_pieces=p, d_min=0.8, num_agents=N)
    check_step(O, oc, O.get_table(oc, "U_basis"), st, tg, got, check_viol_max=4)
    assert got["iters"].max() >= 3
    monkeypatch.setenv("MPCCBF_FULL_SCAN", "1")
    full = run_emul(M, emul, cfg, st, tg)
    assert np.array_equal(full["status"], got["status"])
    assert np.abs(full["ctrl_pts"] - got["ctrl_pts"]).max() < 1e-9


@pytest.mark.parametrize("extra,N,S,spread", [
    (dict(), 16, 24, 1.15),
    (dict(cbf_horizon=3), 16, 12, 1.1),
    (dict(cbf_horizon=4, slack_cost=5.0, slack_decay=0.7), 9, 16, 1.05),
    (dict(slack_cost=100.0, slack_decay=0.5), 16, 12, 1.05),
    (dict(slack_cost=0.05, slack_decay=0.9, impc_iter=4), 12, 12, 1.0),
])
def test_slack_mode_capped_multipliers(M, O, emul, extra, N, S, spread):
    # cbf_params.slack_mode: the capped-multiplier active-set iteration of gi_solver.h against the oracle's
    # full-space QP with slack variables (ConnectivityIMPCCBF.cpp:76-100, MPCCBFQPGeneratorBase.cpp:28-38,80-135)
    base = dict(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, slack_mode=1, slack_cost=50000.0, slack_decay=0.1)
    base.update(extra)
    cfg, oc = M.default_config(**base), O.base_config(**base)
    st, tg = dense_scenarios(S, N, 40 + N, 0.8, spread=spread, vmax=1.0)
    st[0, 1, 3] = 2.6  # the bound rows stay hard: |v0| > v_max is still infeasible
    got = run_emul(M, emul, cfg, st, tg)
    check_step(O, oc, O.get_table(oc, "U_basis"), st, tg, got, check_viol_max=0)
    assert list(got["status"][0, 1]) == [M.STATUS_INFEASIBLE] + [M.STATUS_SKIPPED] * (cfg.impc_iter - 1)
    assert (got["status"] == M.STATUS_OPTIMAL).mean() > 0.95


def test_compact_record_evaluates_like_the_expanded_curve(M, O, emul):
    # mpccbf_eval_compact == SingleParameterPiecewiseCurve::eval of the expanded control points, at the driver's
    # sub-step times and across piece boundaries (lower_bound piece lookup, clamped local parameter)
    cfg = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=8)
    oc = O.base_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=8)
    st, tg = swarm(3, 8, seed=5)
    got = run_emul(M, emul, cfg, st, tg)
    h = M.Handle(cfg.copy(device=-1), library=emul)
    for t in (0.0, 0.01, 0.1, 0.5, 0.5000001, 0.99, 1.0, 1.7, 2.0):
        for d in (0, 1, 2):
            val = h.eval_compact(st, got["v_free"], t, d).reshape(3, 8, 3)
            for s in range(3):
                for a in range(8):
                    assert np.abs(val[s, a] - O.eval_curve(oc, got["ctrl_pts"][s, a], t, d)).max() <= 1e-9
    with pytest.raises(M.MpcCbfError):
        h.eval_compact(st, got["v_free"], 2.1)
    h.close()


@pytest.mark.parametrize("N,S,seed,spread,dmax", [(8, 12, 3, 1.6, 3.0), (16, 8, 5, 1.5, 2.5), (2, 24, 4, 2.5, 1.5), (4, 16, 7, 2.45, 1.9)])
def test_mpc_lifted_connectivity_and_clf_rows(M, O, emul, N, S, seed, spread, dmax):
    # cfg.conn_mode = 1: the block the reference ships commented out (ConnectivityIMPCCBF.cpp:143-157,177-192):
    # lambda2 > 0.1 -> connectivity HOCBF row per CBF horizon step (first two cases), else distance-CLF rows per
    # neighbour and step (last two: lambda2 = 0) — ConnectivityMPCCBFQPOperations.cpp:45-86,111-173
    base = dict(k_hor=20, num_pieces=4, d_min=0.8, d_max=dmax, num_agents=N, conn_mode=1)
    cfg, oc = M.default_config(**base), O.base_config(**base)
    rng = np.random.default_rng(seed)
    st, tg = dense_scenarios(S, N, seed, 0.8, spread=spread, vmax=0.8)
    tg[:, :, :2] = st[:, :, :2] + rng.uniform(-6, 6, size=(S, N, 2))   # agents want to disperse: the rows bind
    got = run_emul(M, emul, cfg, st, tg)
    check_step(O, oc, O.get_table(oc, "U_basis"), st, tg, got, check_viol_max=0)
    off = run_emul(M, emul, cfg.copy(conn_mode=0), st, tg)
    assert (np.abs(got["ctrl_pts"] - off["ctrl_pts"]).max(-1) > 1e-6).sum() + (got["status"] != off["status"]).sum() > 0 or N == 8
    l2 = np.array([O.lambda2(st[s], dmax)[0] for s in range(S)])
    assert (l2 > 0.1).all() if dmax >= 2.5 else (l2 <= 0.1).all()


@pytest.mark.parametrize("extra,N,S", [
    (dict(continuity_degree=2), 8, 8),                                        # C^2 cubics: 5 free dof per axis
    (dict(continuity_degree=1), 6, 6),                                        # C^1 cubics: 8
    (dict(num_ctrl_pts=5, continuity_degree=3), 6, 6),                        # quartic pieces
    (dict(num_ctrl_pts=5, continuity_degree=2, k_hor=16, num_pieces=3), 4, 8),
])
def test_general_configurations(M, O, emul, extra, N, S):
    # bezier_continuity_upto_degree / num_control_points other than the shipped cubic C^3 (parsing.hpp:20-28): the
    # general reduced kernel (impc_general.cu, r free dof per axis) against the oracle's full-space QP
    base = dict(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N)
    base.update(extra)
    cfg, oc = M.default_config(**base), O.base_config(**base)
    st, tg = dense_scenarios(S, N, 5, 0.8, spread=1.3, vmax=1.0)
    st[0, 0, 3] = 2.5
    got = run_emul(M, emul, cfg, st, tg, compact=False)
    check_step(O, oc, O.get_table(oc, "U_basis"), st, tg, got, check_viol_max=4)
    assert list(got["status"][0, 0]) == [M.STATUS_INFEASIBLE, M.STATUS_SKIPPED]
    assert (got["status"] == M.STATUS_OPTIMAL).any()
