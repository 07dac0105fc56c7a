"""GPU parity of the slack-mode controller path (cbf_params.slack_mode = true,
ConnectivityIMPCCBF.cpp:76-100, MPCCBFQPGeneratorBase.cpp:28-38,80-135) against the oracle's
full-space restatement: same status, control points and objective through mpccbf_step_batch_host."""
import numpy as np
import pytest

from conftest import swarm
from parity import TOL_CONTROL, check_step, dense_scenarios

pytestmark = pytest.mark.gpu


def cfgs(M, O, N, S, **kw):
    base = dict(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, slack_mode=1, slack_cost=50000.0, slack_decay=0.1)
    base.update(kw)
    return M.default_config(max_scenarios=S, **base), O.base_config(**base)


@pytest.mark.parametrize("N,S,seed", [(2, 6, 1), (5, 6, 2), (8, 4, 3), (16, 3, 4)])
def test_slack_step_matches_oracle(M, O, N, S, seed):
    # packed swarms: without slack about half of these QPs are infeasible; with slack every one is
    # solved and the far neighbours' rows (weights down to 5e4 * 0.1^(N-2)) are effectively soft.
    cfg, oc = cfgs(M, O, N, S)
    h = M.Handle(cfg)
    assert h.solver == M.SOLVER_REDUCED  # capped multipliers inside the fused active-set kernel
    st, tg = dense_scenarios(S, N, seed, 0.8, spread=1.15, vmax=1.0)
    got = h.step_batch_host(st, tg)
    U = O.get_table(oc, "U_basis")
    res = check_step(O, oc, U, st, tg, got, check_viol_max=0)
    assert (got["n_curves"] == 2).all()  # slack makes every packed case feasible
    assert h.slack_fallback_count() == 0
    print(f"slack N={N}: {S * N} agents, control-point error {res['err_x']:.2e}")
    h.close()


def test_slack_packed_grid_every_status_matches_the_oracle(M, O):
    # the shape on which round 1's ADMM tier returned UNKNOWN for 0.4 % of the QPs and took 63 % of the
    # agents (profiles/r1i_kernels.jsonl): 256 x 16 agents packed at 1.15 d_min. Every status must equal
    # the oracle's and every control point must agree.
    N, S = 16, 256
    cfg, oc = cfgs(M, O, N, S)
    h = M.Handle(cfg)
    st, tg = dense_scenarios(S, N, 7, 0.8, spread=1.15, vmax=1.0)
    got = h.step_batch_host(st, tg)
    ref = O.step_batch(oc, st, tg, threads=O.hardware_threads(), prec=1)
    assert np.array_equal(got["status"], ref["status"])
    assert (got["status"] == M.STATUS_OPTIMAL).all()
    assert np.abs(got["ctrl_pts"] - ref["ctrl_pts"]).max() <= TOL_CONTROL
    rel = np.abs(got["objective"] - ref["objective"]) / np.maximum(1.0, np.abs(ref["objective"]))
    assert rel.max() <= 1e-6
    assert h.slack_fallback_count() == 0
    h.close()


@pytest.mark.parametrize("extra,N,S,spread", [
    (dict(cbf_horizon=3), 16, 24, 1.1),                                   # groups of three rows
    (dict(cbf_horizon=4, slack_cost=5.0, slack_decay=0.7), 9, 32, 1.05),  # weights of the order of the tracking cost
    (dict(slack_cost=100.0, slack_decay=0.5), 16, 64, 1.05),
    (dict(slack_cost=0.05, slack_decay=0.9, impc_iter=4), 12, 64, 1.0),   # nearly free slack, four IMPC iterations
    (dict(slack_cost=30.0, slack_decay=0.6), 12, 32, 0.95),
])
def test_slack_weights_and_horizons(M, O, extra, N, S, spread):
    cfg, oc = cfgs(M, O, N, S, **extra)
    h = M.Handle(cfg)
    st, tg = dense_scenarios(S, N, 31 + N, 0.8, spread=spread, vmax=1.0)
    got = h.step_batch_host(st, tg)
    check_step(O, oc, O.get_table(oc, "U_basis"), st, tg, got, check_viol_max=0)
    h.close()


def test_slack_equals_hard_rows_when_no_row_binds_beyond_its_weight(M, O):
    # well separated ring: the slack optimum is the hard-constrained one (s = 0)
    N, S = 8, 4
    cfg, oc = cfgs(M, O, N, S)
    hard = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, max_scenarios=S)
    st, tg = swarm(S, N, seed=11)
    hs, hh = M.Handle(cfg), M.Handle(hard)
    a, b = hs.step_batch_host(st, tg), hh.step_batch_host(st, tg)
    assert np.array_equal(a["status"], b["status"]) and (a["n_curves"] == 2).all()
    assert np.abs(a["ctrl_pts"] - b["ctrl_pts"]).max() <= TOL_CONTROL
    hs.close()
    hh.close()


def test_slack_infeasible_bound_rows_and_closed_loop(M, O):
    # slack softens only the safety rows: an initial speed beyond v_max still reports INFEASIBLE and
    # the later iterate SKIPPED (same bookkeeping as the hard path); then a short closed loop
    N, S = 4, 2
    cfg, oc = cfgs(M, O, N, S)
    h = M.Handle(cfg)
    st, tg = dense_scenarios(S, N, 5, 0.8, spread=1.4, vmax=0.5)
    st[0, 1, 3] = 2.6  # v_max = 2
    U = O.get_table(oc, "U_basis")
    got = h.step_batch_host(st, tg)
    check_step(O, oc, U, st, tg, got, check_viol_max=0)
    assert list(got["status"][0, 1]) == [M.STATUS_INFEASIBLE, M.STATUS_SKIPPED] and got["n_curves"][0, 1] == 0
    st[0, 1, 3] = 0.3
    for _ in range(5):
        got = h.step_batch_host(st, tg)
        check_step(O, oc, U, st, tg, got, check_viol_max=0)
        for s in range(S):
            for a in range(N):
                if got["n_curves"][s, a] > 0:
                    st[s, a, :3] = O.eval_curve(oc, got["ctrl_pts"][s, a], cfg.h, 0)
                    st[s, a, 3:] = O.eval_curve(oc, got["ctrl_pts"][s, a], cfg.h, 1)
    h.close()


def test_slack_large_swarm_has_no_neighbour_limit(M, O):
    # round 1's ADMM tier held at most 122 neighbours; the capped active-set kernel has no such limit. 130 agents:
    # a packed cluster of 16 plus 114 agents far away. The far agents' rows can never bind and the cluster
    # members keep their distance ranks 0..14 among themselves, so the cluster's QPs are those of the 16-agent
    # swarm alone — which the oracle solves (its slack-variable QP does not converge for ~30+ neighbours: the
    # proximal loop of oracle/qp_solve.hpp runs out of iterations on weights below 1e-25).
    N, n0 = 130, 16
    cfg, _ = cfgs(M, O, N, 1)
    _, oc16 = cfgs(M, O, n0, 1)
    st = np.zeros((1, N, 6))
    tg = np.zeros((1, N, 3))
    st[:, :n0], tg[:, :n0] = dense_scenarios(1, n0, 5, 0.8, spread=1.1, vmax=0.8)
    st[0, n0:, 0] = 200.0 + 10.0 * np.arange(N - n0)
    st[0, n0:, 1] = -150.0
    tg[0, n0:, :2] = st[0, n0:, :2] + 1.0
    h = M.Handle(cfg)
    got = h.step_batch_host(st, tg)
    ref = O.step_batch(oc16, np.ascontiguousarray(st[:, :n0]), np.ascontiguousarray(tg[:, :n0]), threads=O.hardware_threads(), prec=1)
    assert (got["n_curves"] == 2).all()
    assert np.array_equal(got["status"][:, :n0], ref["status"])
    assert np.abs(got["ctrl_pts"][:, :n0] - ref["ctrl_pts"]).max() <= TOL_CONTROL
    rel = np.abs(got["objective"][:, :n0] - ref["objective"]) / np.maximum(1.0, np.abs(ref["objective"]))
    assert rel.max() <= 1e-6
    h.close()


def test_slack_episode_on_the_device_follows_the_hard_episode_when_no_slack_is_used(M, O):
    # the episode driver goes through the same controller step: on separated swarms the slack-mode
    # closed loop must retrace the hard-constrained one wherever no optimize call failed in either.
    import torch
    N, S, steps = 4, 6, 12
    cfg, _ = cfgs(M, O, N, S)
    hard = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, max_scenarios=S)
    st, tg = swarm(S, N, seed=21)
    out = []
    for c in (cfg, hard):
        h = M.Handle(c)
        d_st, d_tg = torch.tensor(st, device="cuda"), torch.tensor(tg, device="cuda")
        failed = torch.zeros(S, dtype=torch.int32, device="cuda")
        h.run_episode(S, d_st, d_tg, steps, 0.3, 0.3, failed_calls=failed)
        torch.cuda.synchronize()
        out.append((d_st.cpu().numpy(), failed.cpu().numpy()))
        h.close()
    clean = (out[0][1] == 0) & (out[1][1] == 0)
    assert clean.sum() >= S - 1, (out[0][1], out[1][1])
    assert np.abs(out[0][0] - out[1][0])[clean].max() <= 1e-6


def test_slack_ranks_in_a_symmetric_ring_match_the_oracle(M, O):
    # ring formations: the two ring neighbours of an agent are equidistant up to rounding, and their order
    # decides which of them gets slack_cost and which slack_cost * decay. Device and oracle both follow the
    # reference's double arithmetic (separately rounded squares and sum, exact ties by index).
    import bench
    N, S = 16, 48
    cfg, oc = cfgs(M, O, N, S)
    st, tg = bench.make_swarm(S, N, 20261104)
    st[:, :, :2] *= 0.8
    tg[:, :, :2] *= 0.8
    h = M.Handle(cfg)
    got = h.step_batch_host(st, tg)
    ref = O.step_batch(oc, st, tg, threads=O.hardware_threads(), prec=1)
    assert np.array_equal(got["status"], ref["status"])
    ok = ref["n_curves"] > 0
    assert np.abs(got["ctrl_pts"] - ref["ctrl_pts"])[ok].max() <= TOL_CONTROL
    h.close()
