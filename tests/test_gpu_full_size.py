"""GPU: BASELINE.json's full-size configurations. Every agent of the batch is compared with the oracle
(all host threads; double precision for the bulk, long double for the few agents that differ), on tight
swarms and on the bench's own closed-loop snapshots, plus size-independent properties: feasibility of
every returned curve against the reference's own rows, bit-exact determinism and scenario-permutation
equivariance, agreement of the two kernel variants, translation invariance."""
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


def oracle_all(O, oc, st, tg, got, label, feas_tol=1e-9, max_recheck_frac=2e-3):
    """Every agent of the batch against the oracle: statuses identical, control points within TOL_CONTROL,
    objective within 1e-6 relative. Bulk in double on all host threads; agents that differ there are solved
    again in long double (the double run of the oracle has its own rounding at feasibility thresholds)."""
    S, N = st.shape[:2]
    ref = O.step_batch(oc, st, tg, threads=O.hardware_threads(), prec=0, feas_tol=feas_tol)
    err = np.abs(got["ctrl_pts"] - ref["ctrl_pts"]).max(-1)
    rel = np.abs(got["objective"] - ref["objective"]) / np.maximum(1.0, np.abs(ref["objective"]))
    bad = (got["status"] != ref["status"]).any(-1) | (err > TOL_CONTROL) | (rel > 1e-6)
    redo = np.argwhere(bad)
    assert len(redo) <= max(4, max_recheck_frac * S * N), f"{label}: {len(redo)} agents differ from the double-precision oracle"
    K = oc.k_hor
    for s, a in redo:
        r = O.optimize(oc, st[s], int(a), np.tile(tg[s, a], K), feas_tol=feas_tol, prec=1)
        assert list(r["status"]) == list(got["status"][s, a]), f"{label}: status of scenario {s} agent {a}: {got['status'][s, a]} vs oracle {r['status']}"
        if r["n_curves"]:
            assert np.abs(r["curves"][-1] - got["ctrl_pts"][s, a]).max() <= TOL_CONTROL
            o = r["objective"][r["n_curves"] - 1]
            assert abs(o - got["objective"][s, a]) <= 1e-6 * max(1.0, abs(o))
    hist_p = np.bincount(got["status"].ravel(), minlength=8)
    hist_o = np.bincount(ref["status"].ravel(), minlength=8)
    ok = ~bad & (ref["n_curves"] > 0)
    print(f"{label}: {S * N * got['status'].shape[-1]} QPs, status histogram product {hist_p.tolist()} oracle(double) {hist_o.tolist()}, "
          f"{len(redo)} agents re-checked in long double, max control-point error {err[ok].max() if ok.any() else 0.0:.2e}")
    return ref


def closed_loop_snapshot(M, h, st, tg, steps):
    """The bench's workload: the swarm stepped `steps` control steps by the product itself (bench.py)."""
    import torch
    S, N = st.shape[:2]
    cur, d_tg = torch.tensor(st, device="cuda"), torch.tensor(tg, device="cuda")
    ctrl = torch.zeros((S, N, h.n), dtype=torch.float64, device="cuda")
    status = torch.zeros((S, N, 2), dtype=torch.int32, device="cuda")
    ncv = torch.zeros((S, N), dtype=torch.int32, device="cuda")
    for _ in range(steps):
        h.step_batch(S, cur, d_tg, ctrl, status, ncv)
        h.rollout(S, cur, ctrl, ncv, cur)
    torch.cuda.synchronize()
    return cur.cpu().numpy()


def inject_infeasible(st, frac, d_min, seed):
    """`frac` of the agents get |v0| > v_max (the velocity row at k = 0 then contradicts the
    initial-velocity equality), another `frac` are put inside d_min of a neighbour with closing velocity."""
    rng = np.random.default_rng(seed)
    S, N = st.shape[:2]
    st = st.copy()
    n = max(1, int(frac * S * N))
    idx = rng.choice(S * N, size=2 * n, replace=False)
    for g in idx[:n]:
        st[g // N, g % N, 3] = 2.0 + rng.uniform(0.05, 1.0)
    for g in idx[n:]:
        s, a = g // N, g % N
        b = (a + 1) % N
        dirn = rng.normal(size=2)
        dirn /= np.linalg.norm(dirn)
        st[s, a, :2] = st[s, b, :2] + dirn * d_min * rng.uniform(0.6, 0.95)
        st[s, a, 3:5] = -dirn * rng.uniform(0.5, 1.5)   # closing
    return st


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
    # the oracle on EVERY agent of the batch
    oracle_all(O, oc, st, tg, got, "4096 x 16 tight swarm")
    # ... and on the bench's own workload: closed-loop snapshots (safety rows active, some QPs infeasible)
    # with 1 % infeasible QPs of each kind injected on top
    st0, tg0 = bench.make_swarm(S, N, 20261104)
    for steps in (15, 30):
        snap = closed_loop_snapshot(M, h, st0, tg0, steps)
        snap = inject_infeasible(snap, 0.01, cfg.d_min, steps)
        g = h.step_batch_host(snap, tg0)
        oracle_all(O, oc, snap, tg0, g, f"4096 x 16 closed-loop snapshot after {steps} steps + 1 % injected")
        assert (g["status"] == M.STATUS_INFEASIBLE).sum() >= 0.01 * S * N
    h.close()
    # the lane-group kernel on the same batch
    monkeypatch.setenv("MPCCBF_LANES", "32")
    h32 = M.Handle(cfg)
    g32 = h32.step_batch_host(st, tg)
    assert np.array_equal(g32["status"], got["status"])
    assert np.abs(g32["ctrl_pts"] - got["ctrl_pts"]).max() <= 1e-9
    h32.close()
    print(f"4096 x 16: {f}, statuses {np.bincount(got['status'].ravel(), minlength=8)}")


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
    oracle_all(O, oc, st, tg, got, "8 x 256 dense swarm")        # every one of the 2048 agents (m = 672 rows each)
    snap = closed_loop_snapshot(M, h, st, tg, 10)
    oracle_all(O, oc, snap, tg, h.step_batch_host(snap, tg), "8 x 256 dense swarm, closed-loop snapshot after 10 steps")
    h.close()
    print(f"8 x 256: {f}")


@pytest.mark.parametrize("K,N,S", [(40, 64, 2), (10, 64, 4), (40, 32, 4)])
def test_sweep_corners_against_the_oracle(M, O, K, N, S):
    # BASELINE configs[4] corners that bench_sweep.py times: horizon 40 (8 pieces, n = 96) and 64-agent swarms
    p = int(np.ceil((K - 1) * 0.1 / 0.5 - 1e-9))
    cfg = M.default_config(k_hor=K, num_pieces=p, d_min=0.8, num_agents=N, max_scenarios=S)
    oc = O.base_config(k_hor=K, num_pieces=p, d_min=0.8, num_agents=N)
    h = M.Handle(cfg)
    st, tg = bench.make_swarm(S, N, 100 * K + N)
    st[:, :, :2] *= N / 16 * 0.55   # ring spacing ~ 1.2 .. 2.3 d_min: safety rows bind
    tg[:, :, :2] *= N / 16 * 0.55
    st = inject_infeasible(st, 0.01, cfg.d_min, K + N)
    got = h.step_batch_host(st, tg)
    oracle_all(O, oc, st, tg, got, f"sweep corner K={K} N={N}")
    snap = closed_loop_snapshot(M, h, st, tg, 8)
    oracle_all(O, oc, snap[:1], tg[:1], h.step_batch_host(snap[:1], tg[:1]), f"sweep corner K={K} N={N}, closed loop")
    h.close()


def test_cplex_default_feasibility_tolerance(M, O):
    # CPLEX's default feasibility tolerance is 1e-6 (qpcpp/src/solvers/CPLEX.cpp:114-126 sets none); the
    # product defaults to 1e-9. Agent 0 of every scenario gets |v0| = v_max + delta, |delta| < 1e-6: the velocity
    # row at k = 0 then contradicts the initial-velocity equality by delta. With feas_tol = 1e-9 the QP is INFEASIBLE
    # for delta > 2e-9, with feas_tol = 1e-6 (threshold 1e-6 * max(1, |v_max|) = 2e-6) it is OPTIMAL - in product and
    # oracle alike. Every other agent must agree with the oracle at both tolerances; for the
    # marginal agent itself only the status is compared at 1e-6 (a QP solved to a 1e-6 row tolerance next to
    # infeasibility has no minimiser that is unique to 1e-5).
    S, N = 256, 16
    kw = dict(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N)
    st, tg = tight_swarm(S, N, 77)
    delta = np.where(np.arange(S) % 2 == 0, 1.0, -1.0) * np.linspace(2e-8, 8e-7, S)   # clear of both thresholds
    st[:, 0, 3] = 2.0 + delta
    res = {}
    for tol in (1e-6, 1e-9):
        h = M.Handle(M.default_config(max_scenarios=S, feas_tol=tol, **kw))
        got = h.step_batch_host(st, tg)
        ref = O.step_batch(O.base_config(feas_tol=tol, **kw), st, tg, threads=O.hardware_threads(), prec=1, feas_tol=tol)
        assert np.array_equal(got["status"], ref["status"]), tol
        ok = ref["n_curves"] > 0
        if tol == 1e-6:
            ok[:, 0] = False
        assert np.abs(got["ctrl_pts"] - ref["ctrl_pts"])[ok].max() <= TOL_CONTROL
        res[tol] = got["status"][:, 0, 0].copy()
        h.close()
    pos = delta > 0
    assert (res[1e-9][pos] == M.STATUS_INFEASIBLE).all()            # v_max exceeded by more than 1e-9 * v_max
    assert (res[1e-6][pos] == M.STATUS_OPTIMAL).mean() >= 0.5       # ... but by less than 1e-6 * v_max (the rest: safety rows of the tight swarm)
    assert np.array_equal(res[1e-6][~pos], res[1e-9][~pos])         # below v_max the tolerance decides nothing
    assert (res[1e-9][~pos] == M.STATUS_OPTIMAL).mean() >= 0.5
