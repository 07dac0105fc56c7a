"""GPU parity of the dense one-CTA-per-QP ADMM kernel (mpccbf_solve_qp_dense, the S-qp seam that
replaces qpcpp::CPLEXSolver::solve) against the exact CPU oracle on the QPs the reference's
generator assembles (oracle build_qp follows ConnectivityIMPCCBF.cpp:104-197 row for row)."""
import numpy as np
import pytest

from conftest import north_star_cfg, swarm
from parity import TOL_CBF_VIOL, TOL_CONTROL, TOL_OBJ_REL, dense_scenarios

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def handle(M):
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device (no fallback)"
    h = M.Handle(north_star_cfg(M, max_scenarios=1))
    yield h
    h.close()


def assembled_qps(O, oc, st, tg, agents, iters=(0, 1), slack=False):
    """QPs of the given (scenario, agent) pairs for IMPC iterations `iters`, stacked."""
    K = oc.k_hor
    out = []
    for s, a in agents:
        ref = np.tile(tg[s, a], K)
        r0 = O.optimize(oc, st[s], a, ref) if 1 in iters else None
        for it in iters:
            pred = None
            if it > 0:
                if r0["n_curves"] < 1:
                    continue
                x0 = r0["curves"][0]
                pred = np.array([np.concatenate([O.eval_curve(oc, x0, k * oc.h, 0), O.eval_curve(oc, x0, k * oc.h, 1)])
                                 for k in range(oc.cbf_horizon)])
            sw = None
            if slack:
                N = st.shape[1]
                sw = oc.slack_cost * oc.slack_decay ** np.arange(N - 1)
            out.append(O.build_qp(oc, st[s], a, ref, it=it, pred_states=pred, slack_weights=sw))
    return out


def check_against_oracle(O, h, qps, U=None, bounds=False):
    """Solves the stacked QPs (same shape) in ONE batched call and checks the north-star tolerances."""
    P = np.stack([q["P"] for q in qps]); q = np.stack([q_["q"] for q_ in qps])
    A = np.stack([q_["A"] for q_ in qps]); l = np.stack([q_["l"] for q_ in qps]); u = np.stack([q_["u"] for q_ in qps])
    lb = np.stack([q_["lb"] for q_ in qps]) if bounds else None
    ub = np.stack([q_["ub"] for q_ in qps]) if bounds else None
    got = h.solve_qp_dense(P, q, A, l, u, lb, ub)
    worst = dict(du=0.0, obj=0.0, viol=0.0)
    for b, qp in enumerate(qps):
        ex = O.solve_qp(qp["P"], qp["q"], qp["A"], qp["l"], qp["u"], qp["lb"] if bounds else None,
                        qp["ub"] if bounds else None)
        assert got["status"][b] == ex["status"], (b, got["status"][b], ex["status"], got["iters"][b])
        if ex["status"] != 0:
            assert np.all(got["x"][b] == 0.0)
            continue
        x = got["x"][b]
        n0 = U.shape[1] if U is not None else x.size
        du = np.abs(U @ (x[:n0] - ex["x"][:n0])).max() if U is not None else np.abs(x - ex["x"]).max()
        assert du <= TOL_CONTROL, (b, du)
        assert np.abs(x - ex["x"]).max() <= TOL_CONTROL
        rel = abs(got["objective"][b] - ex["objective"]) / max(1.0, abs(ex["objective"]))
        assert rel <= TOL_OBJ_REL, (b, rel)
        Ax = qp["A"] @ x
        fin_u, fin_l = qp["u"] < 1e20, qp["l"] > -1e20
        viol = max(np.max((Ax - qp["u"])[fin_u], initial=0.0), np.max((qp["l"] - Ax)[fin_l], initial=0.0))
        assert viol <= TOL_CBF_VIOL, (b, viol)
        worst = dict(du=max(worst["du"], du), obj=max(worst["obj"], rel), viol=max(worst["viol"], viol))
    return got, worst


def test_toy_qp_of_the_reference_cplex_test(handle):
    # lib/qpcpp/tests/CPLEXTest.cpp:28-56: min x^2 + y^2 s.t. x + y >= 1  ->  OPTIMAL, (0.5, 0.5) +- 1e-6
    P = np.array([[[1.0, 0.0], [0.0, 1.0]]]); q = np.zeros((1, 2))
    A = np.array([[[1.0, 1.0]]]); l = np.array([[1.0]]); u = np.array([[np.finfo(float).max]])
    r = handle.solve_qp_dense(P, q, A, l, u)
    assert r["status"][0] == 0
    assert np.abs(r["x"][0] - 0.5).max() < 1e-6 and abs(r["objective"][0] - 0.5) < 1e-6


def test_unbounded_and_infeasible_status(handle):
    # min -x0 + x1^2 with no rows bounding x0 from above: UNBOUNDED (CPLEX.cpp:161-163)
    P = np.array([[[0.0, 0.0], [0.0, 1.0]]]); q = np.array([[-1.0, 0.0]])
    A = np.array([[[1.0, 0.0]]]); l = np.array([[0.0]]); u = np.array([[1e30]])
    r = handle.solve_qp_dense(P, q, A, l, u)
    assert r["status"][0] == 2, r
    # x0 >= 1 and x0 <= 0: INFEASIBLE
    A = np.array([[[1.0, 0.0], [1.0, 0.0]]]); l = np.array([[1.0, -1e30]]); u = np.array([[1e30, 0.0]])
    P = np.array([[[1.0, 0.0], [0.0, 1.0]]])
    r = handle.solve_qp_dense(P, q, A, l, u)
    assert r["status"][0] == 3, r


@pytest.mark.parametrize("K,p,N,dense", [(20, 4, 16, False), (20, 4, 16, True), (16, 3, 2, False), (10, 2, 8, True)])
def test_assembled_mpc_cbf_qps(M, O, handle, K, p, N, dense):
    oc = O.base_config(k_hor=K, num_pieces=p, d_min=0.8, num_agents=N)
    if dense:
        st, tg = dense_scenarios(3, N, 12 + K, 0.8, spread=1.2, vmax=1.5)
        st[0, 0, 3] = 2.5  # |v0| > v_max: infeasible
    else:
        st, tg = swarm(3, N, seed=K)
    agents = [(s, a) for s in range(3) for a in range(min(N, 6))]
    U = O.get_table(oc, "U_basis")
    for it in (0, 1):  # shapes differ between iterations (N-1 vs H(N-1) safety rows)
        qps = assembled_qps(O, oc, st, tg, agents, iters=(it,))
        got, worst = check_against_oracle(O, handle, qps, U)
        print(f"K={K} N={N} dense={dense} it={it}: {len(qps)} QPs, m={qps[0]['l'].size}, iters med "
              f"{np.median(got['iters'])} max {got['iters'].max()}, worst {worst}")


def test_slack_mode_qps_with_variable_bounds(M, O, handle):
    # slack mode: N-1 extra variables with bounds [0, inf) and a linear cost
    # (MPCCBFQPGeneratorBase.cpp:28-38,121-135)
    N = 6
    oc = O.base_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, slack_mode=1)
    st, tg = dense_scenarios(2, N, 5, 0.8, spread=1.15, vmax=1.0)
    qps = assembled_qps(O, oc, st, tg, [(s, a) for s in range(2) for a in range(N)], iters=(0,), slack=True)
    assert qps[0]["q"].size == 48 + N - 1
    U = O.get_table(oc, "U_basis")
    check_against_oracle(O, handle, qps, U, bounds=True)


def test_dense_argument_errors(M, handle):
    with pytest.raises(M.MpcCbfError):
        handle.solve_qp_dense(np.zeros((1, 200, 200)), np.zeros((1, 200)), np.zeros((1, 1, 200)), np.zeros((1, 1)),
                              np.zeros((1, 1)))
