"""GPU: BASELINE.json's full-size configurations through size-independent properties (the oracle cannot
finish 131 072 QPs in seconds): feasibility of every returned curve against the reference's own rows,
bit-exact determinism and scenario-permutation equivariance, agreement of the two kernel variants,
translation invariance, and oracle parity on a random sample of the batch."""
import numpy as np
import pytest

import bench
from parity import TOL_CBF_VIOL, TOL_CONTROL

pytestmark = pytest.mark.gpu


def feasibility(O, oc, cfg, st, tg, got):
    """max violation of: initial-state + continuity equalities, acceleration / velocity boxes at every
    horizon sample, and the k = 0 safety rows of the final iterate (its predicted state at t_0 is the
    current state, ConnectivityIMPCCBF.cpp:160-175) — for every agent that returned a curve."""
    S, N = st.shape[:2]
    K = cfg.k_hor
    U, V = O.get_table(oc, "U_basis"), O.get_table(oc, "V_basis")   # the oracle's tables, not the product's
    qp = O.build_qp(oc, st[0], 0, np.tile(tg[0, 0], K))
    E = qp["A"][np.abs(qp["u"] - qp["l"]) < 1e-12]                  # initial-state rows, then continuity rows
    ok = got["n_curves"] > 0
    x = got["ctrl_pts"]
    acc = np.einsum("kn,san->sak", U, x).reshape(S, N, K, 3)
    vel = np.einsum("kn,san->sak", V, x).reshape(S, N, K, 3)
    amin, amax = np.array(cfg.a_min), np.array(cfg.a_max)
    vmin, vmax = np.array(cfg.v_min), np.array(cfg.v_max)
    box = max(np.max((amin - acc)[ok]), np.max((acc - amax)[ok]), np.max((vmin - vel)[ok]), np.max((vel - vmax)[ok]))
    rhs = np.zeros((S, N, E.shape[0]))
    rhs[:, :, :6] = st
    eq = np.abs(np.einsum("en,san->sae", E, x) - rhs)[ok].max()
    # safety HOCBF rows at k = 0 (closed form of ConnectivityCBF.cpp:152-198)
    p, v = st[:, :, None, :2], st[:, :, None, 3:5]
    q, w = st[:, None, :, :2], st[:, None, :, 3:5]
    d, dv = p - q, v - w
    hh = (d ** 2).sum(-1) - cfg.d_min ** 2
    lfh = 2 * (d * dv).sum(-1)
    b = 2 * (dv ** 2).sum(-1) + 3 * cfg.cbf_gamma * hh ** 2 * 2 * (d * v).sum(-1) + cfg.cbf_gamma * (lfh + cfg.cbf_gamma * hh ** 3) ** 3
    row = 2 * (d * acc[:, :, None, 0, :2]).sum(-1) + b          # a.u_0 + b >= 0
    row[:, np.arange(N), np.arange(N)] = 0.0
    cbf = np.max((-row)[ok])
    return dict(box=box, eq=eq, cbf=cbf, n_ok=int(ok.sum()))


def tight_swarm(S, N, seed):
    st, tg = bench.make_swarm(S, N, seed)
    st[:, :, :2] *= 0.8      # neighbours at 0.94 .. 1.9 m with d_min = 0.8: safety rows bind
    tg[:, :, :2] *= 0.8
    return st, tg


def test_north_star_full_size_4096x16(M, O, monkeypatch):
    S, N = 4096, 16
    cfg = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, max_scenarios=S)
    h = M.Handle(cfg)
    st, tg = tight_swarm(S, N, 20261104)
    got = h.step_batch_host(st, tg)
    again = h.step_batch_host(st, tg)
    for k in got:
        assert np.array_equal(got[k], again[k]), f"{k} is not deterministic"
    oc = O.base_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N)
    f = feasibility(O, oc, cfg, st, tg, got)
    assert f["n_ok"] >= 0.9 * S * N
    assert f["box"] <= 1e-9 and f["eq"] <= 1e-9 and f["cbf"] <= TOL_CBF_VIOL, f
    # scenario permutation: every scenario is solved independently of its place in the batch
    perm = np.random.default_rng(3).permutation(S)
    gp = h.step_batch_host(st[perm], tg[perm])
    for k in got:
        assert np.array_equal(gp[k], got[k][perm]), f"{k} depends on the scenario's position in the batch"
    # translation of the whole scene: same accelerations
    sh = np.array([100.0, -50.0])
    st2, tg2 = st.copy(), tg.copy()
    st2[:, :, :2] += sh
    tg2[:, :, :2] += sh
    g2 = h.step_batch_host(st2, tg2)
    assert np.array_equal(g2["status"], got["status"])
    U = O.get_table(oc, "U_basis")
    du = np.abs(np.einsum("kn,san->sak", U, g2["ctrl_pts"] - got["ctrl_pts"]))[got["n_curves"] > 0].max()
    assert du <= 1e-7, du
    # oracle on a random sample of the batch
    rng = np.random.default_rng(5)
    worst = 0.0
    for s, a in zip(rng.integers(0, S, 48), rng.integers(0, N, 48)):
        r = O.optimize(oc, st[s], int(a), np.tile(tg[s, a], cfg.k_hor))
        assert list(r["status"]) == list(got["status"][s, a])
        if r["n_curves"]:
            worst = max(worst, np.abs(r["curves"][-1] - got["ctrl_pts"][s, a]).max())
    assert worst <= TOL_CONTROL
    h.close()
    # the lane-group kernel on the same batch
    monkeypatch.setenv("MPCCBF_LANES", "32")
    h32 = M.Handle(cfg)
    g32 = h32.step_batch_host(st, tg)
    assert np.array_equal(g32["status"], got["status"])
    assert np.abs(g32["ctrl_pts"] - got["ctrl_pts"]).max() <= 1e-9
    h32.close()
    print(f"4096 x 16: {f}, oracle sample error {worst:.2e}, statuses {np.bincount(got['status'].ravel(), minlength=8)}")


def test_dense_swarm_full_size_256_agents(M, O):
    # BASELINE configs[3]: 16 x 16 grid at 1.25 d_min, every agent carries rows for all 255 others
    S, N = 8, 256
    cfg = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, max_scenarios=S)
    h = M.Handle(cfg)
    rng = np.random.default_rng(20261104)
    gx, gy = np.meshgrid(np.arange(16), np.arange(16), indexing="ij")
    st = np.zeros((S, N, 6))
    st[:, :, :2] = np.stack([gx.ravel(), gy.ravel()], -1)[None] * 1.0 + rng.uniform(-0.05, 0.05, size=(S, N, 2))
    st[:, :, 3:5] = rng.uniform(-0.3, 0.3, size=(S, N, 2))
    tg = np.zeros((S, N, 3))
    tg[:, :, :2] = 2 * st[:, :, :2].mean(1, keepdims=True) - st[:, :, :2]
    got = h.step_batch_host(st, tg)
    again = h.step_batch_host(st, tg)
    assert all(np.array_equal(got[k], again[k]) for k in got)
    oc = O.base_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N)
    f = feasibility(O, oc, cfg, st, tg, got)
    assert f["n_ok"] >= 0.5 * S * N
    assert f["box"] <= 1e-9 and f["eq"] <= 1e-9 and f["cbf"] <= TOL_CBF_VIOL, f
    worst = 0.0
    for s, a in ((0, 0), (3, 119), (7, 255), (5, 136)):
        r = O.optimize(oc, st[s], a, np.tile(tg[s, a], cfg.k_hor))
        assert list(r["status"]) == list(got["status"][s, a])
        if r["n_curves"]:
            worst = max(worst, np.abs(r["curves"][-1] - got["ctrl_pts"][s, a]).max())
    assert worst <= TOL_CONTROL
    h.close()
    print(f"8 x 256: {f}, oracle sample error {worst:.2e}")
