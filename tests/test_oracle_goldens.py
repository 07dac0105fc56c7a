"""Pins the CPU oracle against every known answer the reference's own tests hold for the path
 and against the REAL reference code compiled into oracle/_ref. CPU only."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN


@pytest.fixture(scope="module")
def gold():
    with open(os.path.join(GOLDEN, "reference_goldens.json")) as f:
        return json.load(f)


def test_safety_cbf_goldens(O, gold):
    # cbf/tests/TestInitSafetyCBF.cpp:50-143 (EXPECT_NEAR 1e-6; the closed form does far better)
    g = gold["safety_cbf"]
    for case in g["cases"]:
        for prec in (0, 1):
            a, b = O.safety_cbf_row(case["ego"], case["nbr"], g["dmin"], g["gamma"], prec=prec)
            assert np.allclose(a, case["Ac"], atol=1e-12, rtol=0), case["name"]
            assert abs(b - case["Bc"]) <= 1e-12 * max(1.0, abs(case["Bc"])), (case["name"], b)
    # on-boundary case is EXPECT_EQ(Bc, 0.0)
    a, b = O.safety_cbf_row([0] * 6, [0.8, 0, 0, 0, 0, 0], 0.8)
    assert b == 0.0


def test_connectivity_cbf_goldens(O, gold):
    # cbf/tests/TestInitConnectivity.cpp:103-153: EXPECT_NEAR(Ac, 1e-6), EXPECT_DOUBLE_EQ(Bc) (4 ulp)
    g = gold["connectivity_cbf"]
    for case in g["cases"]:
        for prec in (0, 1):
            Ac, Bc, l2 = O.conn_cbf_row(case["states"], case["self"], g["dmax"], g["gamma"], prec=prec)
            assert np.allclose(Ac, case["Ac"], atol=g["tol_Ac"], rtol=0), case["name"]
            assert abs(Bc - case["Bc"]) <= 8 * np.spacing(abs(case["Bc"])), (case["name"], Bc)
            assert l2 > 0
    # lambda2 of the oracle's Jacobi solver against LAPACK on random swarms
    rng = np.random.default_rng(3)
    for N in (2, 5, 16):
        st = np.zeros((N, 6))
        st[:, :2] = rng.uniform(-2.5, 2.5, size=(N, 2))
        l2, fv = O.lambda2(st, 3.0)
        d2 = ((st[:, None, :2] - st[None, :, :2]) ** 2).sum(-1)
        A = np.where((d2 <= 9.0) & ~np.eye(N, dtype=bool), np.exp((9.0 - d2) ** 2 / (81.0 / np.log(2))) - 1, 0.0)
        w, V = np.linalg.eigh(np.diag(A.sum(1)) - A)
        assert abs(l2 - w[1]) < 1e-11 * max(1.0, w[-1])
        assert abs(abs(fv @ V[:, 1]) - 1.0) < 1e-9 and abs(np.linalg.norm(fv) - 1.0) < 1e-12


def test_cbf_filter_oracle_kkt(O):
    # ConnectivityControl::optimize restated: the filtered input satisfies every row it was built
    # from, equals the desired input when nothing binds, and is the projection otherwise
    cfg = O.base_config(d_min=0.8, d_max=3.0, num_agents=4, v_min=[-1, -1, -1], v_max=[1, 1, 1])
    st = np.zeros((2, 4, 6))
    st[0, :, :2] = [[0, 0], [1.5, 0.2], [0.4, 1.6], [1.9, 1.8]]
    st[1, :, :2] = [[0, 0], [0.9, 0], [0, 1.0], [5.0, 5.0]]      # agent 3 out of range: lambda2 = 0 -> CLF rows
    st[:, :, 3:5] = 0.1
    du = np.zeros((2, 4, 3))
    du[:, :, 0] = 0.3
    r = O.cbf_filter_batch(cfg, st, du)
    assert r["lambda2"][0] > 0.1 and r["lambda2"][1] < 1e-9
    for s in range(2):
        for a in range(4):
            if r["status"][s, a] != 0:
                continue
            u = r["u"][s, a]
            for j in range(4):
                if j != a:
                    Ac, Bc = O.safety_cbf_row(st[s, a], st[s, j], 0.8)
                    assert Ac @ u + Bc >= -1e-8 * max(1.0, abs(Bc))
            assert np.all(u >= -(st[s, a, 3:] + 1.0) - 1e-9) and np.all(u <= 1.0 - st[s, a, 3:] + 1e-9)
    assert (r["status"][0] == 0).all()


def test_double_integrator_one_step(O):
    # model/tests/DoubleIntegratorXYYawTest.cpp:19-47
    nxt = O.double_integrator_apply(0.1, [1.0, 2.0, 0.5, 0.1, 0.2, 0.3], [0.5, 0.6, 0.1])
    exp = [1.0 + 0.1 * 0.1 + 0.5 * 0.5 * 0.01, 2.0 + 0.2 * 0.1 + 0.5 * 0.6 * 0.01,
           0.5 + 0.3 * 0.1 + 0.5 * 0.1 * 0.01, 0.1 + 0.5 * 0.1, 0.2 + 0.6 * 0.1, 0.3 + 0.1 * 0.1]
    assert np.allclose(nxt, exp, atol=1e-10, rtol=0)


def test_toy_qp(O, gold):
    # qpcpp/tests/CPLEXTest.cpp:28-56: min x^2+y^2 s.t. x+y>=1 -> OPTIMAL, (0.5, 0.5) +-1e-6
    g = gold["toy_qp"]
    r = O.solve_qp(g["P"], g["q"], g["A"], g["l"], g["u"])
    assert r["status"] == 0
    assert np.allclose(r["x"], g["x"], atol=g["tol"], rtol=0)
    assert abs(r["objective"] - 0.5) < 1e-12


def test_combinatorics_against_real_reference(O):
    # math/tests/CombinatoricsTest.cpp + the reference's own Combinatorics.cpp (oracle/_ref)
    L = O.lib()
    assert L.oracle_fac(0) == 1 and L.oracle_fac(5) == 120 and L.oracle_fac(20) == 2432902008176640000
    assert L.oracle_comb(5, 2) == 10 and L.oracle_comb(10, 0) == 1 and L.oracle_perm(5, 2) == 20
    R = O.ref_lib()
    if R is None:
        pytest.skip("oracle/_ref not built (reference tree absent)")
    import ctypes as C
    threw = C.c_int()
    for n in range(0, 21):
        assert R.ref_fac(n, C.byref(threw)) == L.oracle_fac(n) and threw.value == 0
        for k in range(0, n + 1):
            assert R.ref_comb(n, k) == L.oracle_comb(n, k)
            assert R.ref_perm(n, k) == L.oracle_perm(n, k)
    R.ref_fac(21, C.byref(threw))
    assert threw.value == 1  # fac(21) throws in the reference


def test_bernstein_closed_forms(O):
    # SURVEY App. A closed forms of splines/src/detail/BezierOperations.cpp:11-50
    T = 0.5
    for t in (0.0, 0.1, 0.30000000000000004, 0.5):
        s = t / T
        b0 = [(1 - s) ** 3, 3 * s * (1 - s) ** 2, 3 * s * s * (1 - s), s ** 3]
        b1 = np.array([-(1 - s) ** 2, (1 - s) * (1 - 3 * s), s * (2 - 3 * s), s * s]) * 3 / T
        b2 = np.array([1 - s, 3 * s - 2, 1 - 3 * s, s]) * 6 / T ** 2
        b3 = np.array([-1, 3, -3, 1]) * 6 / T ** 3
        for d, e in enumerate((b0, b1, b2, b3)):
            assert np.allclose(O.bernstein_basis(3, T, t, d), e, rtol=1e-13, atol=1e-13)
    with pytest.raises(RuntimeError):
        O.bernstein_basis(3, T, 0.6, 0)


def test_structure_known_answers(O):
    # rank(E) = 30 @K16, 6 free dof, reduced-Hessian eigenvalues
    cfg = O.base_config()
    st = np.array([[2, 2, 0, 0, 0, 0], [-2, -2, 0, 0, 0, 0]], float)
    qp = O.build_qp(cfg, st, 0, np.tile([-2, -2, 0], cfg.k_hor))
    ne = qp["n_eq"]
    assert ne == 30 and qp["n_cbf"] == 1 and qp["A"].shape == (30 + 1 + 96, 36)
    E = qp["A"][:ne]
    assert np.linalg.matrix_rank(E) == 30
    _, _, vt = np.linalg.svd(E)
    Z = vt[30:].T
    w = np.linalg.eigvalsh(Z.T @ qp["P"] @ Z)
    assert np.allclose(sorted(w), [26.26] * 3 + [986.3] * 3, rtol=2e-3)
    assert np.linalg.matrix_rank(qp["P"], tol=1e-6) == 27
    # App. C worked example: CBF row and objective
    row = qp["A"][ne]
    U = O.get_table(cfg, "U_basis")
    assert np.allclose(row, -(8 * U[0] + 8 * U[1]), atol=1e-12)
    assert qp["u"][ne] == 6611534970880000.0
    r = O.optimize(cfg, st, 0, np.tile([-2, -2, 0], cfg.k_hor))
    assert r["success"] and abs(r["objective"][0] + 1378.9905) < 1e-3
    assert np.allclose(U[:3] @ r["curves"][0], [-270 / 91, -270 / 91, 0], atol=1e-9)
    assert r["kkt"].max() < 1e-8


def test_connectivity_goldens_lifted_into_the_mpc_qp(O, gold):
    # The MPC-lifted connectivity row (ConnectivityMPCCBFQPOperations::connectivityConstraint, :45-70): with the
    # reference's own golden (Ac, Bc) the row of the assembled QP must be -(Ac' U_0) with upper bound Bc, inserted
    # after the safety rows (ConnectivityIMPCCBF.cpp:143-151, commented out there; cfg.conn_mode = 1 here).
    g = gold["connectivity_cbf"]
    for case in g["cases"]:
        st = np.array(case["states"], float)
        N = st.shape[0]
        oc = O.base_config(k_hor=20, num_pieces=4, d_min=0.8, d_max=g["dmax"], num_agents=N, conn_mode=1)
        l2, _ = O.lambda2(st, g["dmax"])
        ref = np.zeros(3 * oc.k_hor)
        qp = O.build_qp(oc, st, case["self"], ref, it=0)
        off = O.build_qp(O.base_config(k_hor=20, num_pieces=4, d_min=0.8, d_max=g["dmax"], num_agents=N), st, case["self"], ref, it=0)
        U = O.get_table(oc, "U_basis")
        if l2 <= 0.1:
            # weakly connected golden ("Misc2", lambda2 = 0.03): the controller takes the CLF branch — one row per
            # neighbour, -(a' U_0) x <= -Bc as ConnectivityMPCCBFQPOperations::clfConstraint (:72-86) writes it
            assert qp["n_cbf"] == off["n_cbf"] + (N - 1)
            others = [j for j in range(N) if j != case["self"]]
            for i, j in enumerate(others):
                a, bc = O.clf_row(st[case["self"]], st[j])
                r = qp["n_eq"] + (N - 1) + i
                assert np.abs(qp["A"][r] + a @ U[0:3]).max() <= 1e-12 * np.abs(U[0:3]).sum(0).max()
                assert abs(qp["u"][r] + bc) <= 1e-12 * max(1.0, abs(bc))
            continue
        assert qp["n_cbf"] == off["n_cbf"] + 1 and qp["A"].shape[0] == off["A"].shape[0] + 1
        r = qp["n_eq"] + (N - 1)              # first row after the safety rows
        assert np.abs(qp["A"][r] + np.array(case["Ac"]) @ U[0:3]).max() <= g["tol_Ac"] * np.abs(U[0:3]).sum(0).max()
        assert abs(qp["u"][r] - case["Bc"]) <= 8 * np.spacing(abs(case["Bc"])) and qp["l"][r] < -1e300
        # iter >= 1: one row per CBF horizon step at the predicted ego state; with pred = current state the k = 0
        # row has the same (Ac, Bc) and sits on U_0, the k = 1 row on U_1
        pred = np.tile(st[case["self"]], (oc.cbf_horizon, 1))
        qp1 = O.build_qp(oc, st, case["self"], ref, it=1, pred_states=pred)
        r1 = qp1["n_eq"] + (N - 1) * oc.cbf_horizon
        for k in range(oc.cbf_horizon):
            assert np.abs(qp1["A"][r1 + k] + np.array(case["Ac"]) @ U[3 * k:3 * k + 3]).max() <= g["tol_Ac"] * np.abs(U[3 * k:3 * k + 3]).sum(0).max()
            assert abs(qp1["u"][r1 + k] - case["Bc"]) <= 8 * np.spacing(abs(case["Bc"]))
