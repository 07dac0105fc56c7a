"""GPU parity of the CBF-only safety / connectivity filter (cbf::ConnectivityControl::optimize,
SURVEY 8f rows 1-2) against the CPU oracle, through the C ABI."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def random_swarms(S, N, seed, spread, vmax=0.8):
    rng = np.random.default_rng(seed)
    st = np.zeros((S, N, 6))
    st[:, :, :2] = rng.uniform(-spread, spread, size=(S, N, 2))
    st[:, :, 2] = rng.uniform(-1, 1, size=(S, N))
    st[:, :, 3:5] = rng.uniform(-vmax, vmax, size=(S, N, 2))
    st[:, :, 5] = rng.uniform(-0.5, 0.5, size=(S, N))
    du = rng.uniform(-2, 2, size=(S, N, 3))
    return st, du


def test_connectivity_rows_goldens_and_oracle(M, O):
    import torch
    with open(os.path.join(GOLDEN, "reference_goldens.json")) as f:
        g = json.load(f)["connectivity_cbf"]
    # the reference's own known answers (cbf/tests/TestInitConnectivity.cpp:103-153)
    cfg = M.default_config(num_agents=3, d_min=0.8, d_max=g["dmax"], max_scenarios=len(g["cases"]))
    h = M.Handle(cfg)
    st = torch.tensor(np.array([c["states"] for c in g["cases"]], float), device="cuda")
    rows = torch.zeros((len(g["cases"]), 3, 4), dtype=torch.float64, device="cuda")
    l2 = torch.zeros(len(g["cases"]), dtype=torch.float64, device="cuda")
    h.conn_rows(len(g["cases"]), st, rows, l2)
    torch.cuda.synchronize()
    r = rows.cpu().numpy()
    for i, c in enumerate(g["cases"]):
        assert np.allclose(r[i, c["self"], :3], c["Ac"], atol=g["tol_Ac"], rtol=0), c["name"]
        assert abs(r[i, c["self"], 3] - c["Bc"]) < 1e-12 * max(1.0, abs(c["Bc"])), c["name"]
    h.close()
    # random swarms against the oracle (long double Jacobi)
    for N, S, spread in ((2, 16, 1.2), (7, 32, 2.0), (16, 64, 3.0), (32, 8, 4.0)):
        cfg = M.default_config(num_agents=N, d_min=0.5, d_max=3.0, max_scenarios=S)
        h = M.Handle(cfg)
        stn, _ = random_swarms(S, N, 100 + N, spread)
        rows = torch.zeros((S, N, 4), dtype=torch.float64, device="cuda")
        l2 = torch.zeros(S, dtype=torch.float64, device="cuda")
        h.conn_rows(S, torch.tensor(stn, device="cuda"), rows, l2)
        torch.cuda.synchronize()
        r, l2 = rows.cpu().numpy(), l2.cpu().numpy()
        for s in range(S):
            for a in range(N):
                Ac, Bc, lam = O.conn_cbf_row(stn[s], a, 3.0, 5.0, prec=1)
                assert abs(l2[s] - lam) < 1e-10 * max(1.0, lam)
                scale = max(1.0, np.abs(Ac).max(), abs(Bc))
                assert np.abs(r[s, a, :3] - Ac).max() < 1e-7 * scale and abs(r[s, a, 3] - Bc) < 1e-7 * scale, (N, s, a)
        h.close()


@pytest.mark.parametrize("N,S,spread", [(2, 64, 1.0), (4, 128, 1.2), (8, 128, 1.8), (16, 256, 2.5), (32, 16, 3.5),
                                        (33, 6, 3.5), (48, 4, 4.5), (65, 4, 5.0)])   # > 32: one CTA per scenario (oracle: seconds)
def test_filter_parity(M, O, N, S, spread):
    vlim = [1.0, 1.0, 1.0]
    cfg = M.default_config(num_agents=N, d_min=0.5, d_max=3.0, max_scenarios=S, v_min=[-v for v in vlim], v_max=vlim)
    h = M.Handle(cfg)
    st, du = random_swarms(S, N, 7 * N, spread)
    st[0, 0, 3] = 1.4          # outside the velocity box: u_x must brake
    if N > 2:
        st[1, 1, :2] = st[1, 0, :2] + [0.3, 0.0]   # inside d_min: strongly negative safety bound
    got = h.cbf_filter_batch_host(st, du)
    oc = O.base_config(num_agents=N, d_min=0.5, d_max=3.0, v_min=[-v for v in vlim], v_max=vlim)
    ref = O.cbf_filter_batch(oc, st, du, prec=1)
    assert np.allclose(got["lambda2"], ref["lambda2"], atol=1e-10, rtol=1e-10)
    # lambda2 within rounding of the 0.1 switch could legitimately flip the row set: exclude
    clear = np.abs(ref["lambda2"] - 0.1) > 1e-8
    assert clear.mean() > 0.9
    assert np.array_equal(got["status"][clear], ref["status"][clear])
    ok = (ref["status"] == 0) & clear[:, None]
    assert ok.mean() > (0.3 if N <= 32 else 0.05)   # large random swarms: most agents' QPs are infeasible (statuses compared above)
    assert np.abs(got["u"] - ref["u"])[ok].max() < 1e-5        # north_star tolerance on controls
    assert np.abs(got["u"] - ref["u"])[ok].max() < 1e-7        # what the kernel actually achieves
    assert np.all(got["u"][~(got["status"] == 0)] == 0.0)
    assert (ref["lambda2"] > 0.1).any() or N == 2
    h.close()


def test_filter_unsupported_and_errors(M):
    cfg = M.default_config(num_agents=257, d_min=0.5, d_max=3.0, max_scenarios=1)
    h = M.Handle(cfg)
    with pytest.raises(M.MpcCbfError) as e:
        h.cbf_filter_batch_host(np.zeros((1, 257, 6)), np.zeros((1, 257, 3)))
    assert e.value.code == M.ERR_UNSUPPORTED
    h.close()


@pytest.mark.parametrize("N,S,spread,cost,decay", [(2, 64, 0.8, 1000.0, 0.2), (4, 128, 1.0, 1000.0, 0.2), (8, 128, 1.6, 1000.0, 0.2),
                                                    (8, 128, 1.2, 10.0, 0.7), (16, 128, 2.2, 10.0, 0.8), (34, 1, 3.0, 10.0, 0.9)])
def test_filter_slack_mode_parity(M, O, N, S, spread, cost, decay):
    # ConnectivityControl.cpp:30-45: slack s_i on safety row i and CLF row i (shared), the last slack on the connectivity
    # row, linear cost slack_cost * slack_decay^i; velocity rows stay hard. Tight swarms: without slack many QPs are
    # infeasible, with slack only the velocity rows can make one infeasible.
    vlim = [1.0, 1.0, 1.0]
    kw = dict(num_agents=N, d_min=0.5, d_max=3.0, v_min=[-v for v in vlim], v_max=vlim)
    st, du = random_swarms(S, N, 11 * N + 1, spread)
    if N > 2 and S > 1:
        st[1, 1, :2] = st[1, 0, :2] + [0.3, 0.0]   # inside d_min
    st[0, 0, 3] = 1.4                               # outside the velocity box (hard rows): still solvable by braking
    h0 = M.Handle(M.default_config(max_scenarios=S, **kw))
    hard = h0.cbf_filter_batch_host(st, du)
    h0.close()
    h = M.Handle(M.default_config(max_scenarios=S, slack_mode=1, slack_cost=cost, slack_decay=decay, **kw))
    got = h.cbf_filter_batch_host(st, du)
    h.close()
    ref = O.cbf_filter_batch(O.base_config(slack_mode=1, slack_cost=cost, slack_decay=decay, **kw), st, du, prec=1)
    clear = np.abs(ref["lambda2"] - 0.1) > 1e-8
    assert np.array_equal(got["status"][clear], ref["status"][clear])
    ok = (ref["status"] == 0) & clear[:, None]
    assert ok.mean() > 0.9
    err = np.abs(got["u"] - ref["u"])[ok].max()
    assert err < 1e-5, err
    # slack did something: QPs that are infeasible with hard rows are solved, and solved ones can differ
    assert ((hard["status"] != 0) & (got["status"] == 0)).any() or N == 2 or S == 1
    print(f"filter slack N={N}: max |u - oracle| {err:.2e}, hard infeasible {np.mean(hard['status'] != 0):.2f} -> slack infeasible {np.mean(got['status'] != 0):.3f}")
