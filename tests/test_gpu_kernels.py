"""GPU tests of the sub-kernels and the remaining C-ABI entry points (-m gpu)."""
import numpy as np
import pytest

from conftest import north_star_cfg, swarm
from parity import dense_scenarios
from test_gpu_parity import ocfg_of

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch():
    import torch
    assert torch.cuda.is_available()
    return torch


def test_cbf_rows_goldens_and_oracle(M, O, torch):
    # reference goldens (cbf/tests/TestInitSafetyCBF.cpp:50-143) through the CUDA kernel
    cfg = M.default_config(d_min=0.8, num_agents=2, max_scenarios=4)
    h = M.Handle(cfg)
    st = np.zeros((4, 2, 6))
    st[0, 1, 0] = 1.0
    st[1, 0, 3:5] = 100.0; st[1, 1, 0] = 1.0
    st[2, 1, 0] = 0.8
    st[3, 1, 0] = 0.5
    d_st = torch.tensor(st, device="cuda")
    rows = torch.zeros((4, 2, 1, 4), dtype=torch.float64, device="cuda")
    h.cbf_rows(4, d_st, rows)
    torch.cuda.synchronize()
    r = rows.cpu().numpy()
    exp_a = [-2.0, -2.0, -1.6, -1.0]
    exp_b = [0.06347497291775989, -39820583.995200224, 0.0, -0.13045522572422458]
    for s in range(4):
        assert abs(r[s, 0, 0, 0] - exp_a[s]) < 1e-15 and r[s, 0, 0, 1] == 0 and r[s, 0, 0, 2] == 0
        assert abs(r[s, 0, 0, 3] - exp_b[s]) <= 1e-12 * max(1, abs(exp_b[s]))
    assert r[2, 0, 0, 3] == 0.0
    h.close()
    # random pairs vs oracle, relative tolerance (|b| spans 1e-2 .. 1e16)
    for N, S in ((16, 200), (33, 17), (256, 3)):
        cfg = M.default_config(d_min=0.8, num_agents=N, max_scenarios=S)
        h = M.Handle(cfg)
        st, _ = dense_scenarios(S, N, N, 0.8, spread=1.3, vmax=2.0)
        rows = torch.zeros((S, N, N - 1, 4), dtype=torch.float64, device="cuda")
        h.cbf_rows(S, torch.tensor(st, device="cuda"), rows)
        r = rows.cpu().numpy()
        rng = np.random.default_rng(N)
        for _ in range(300):
            s, e, j = rng.integers(S), rng.integers(N), rng.integers(N - 1)
            o = j + (1 if j >= e else 0)
            a, b = O.safety_cbf_row(st[s, e], st[s, o], 0.8, 5.0, prec=1)
            assert np.allclose(r[s, e, j, :3], a, rtol=1e-14, atol=1e-14)
            assert abs(r[s, e, j, 3] - b) <= 1e-11 * max(1.0, abs(b)), (r[s, e, j, 3], b)
        h.close()


def test_optimize_one_matches_oracle_optimize(M, O, torch):
    # S-ctl drop-in: every IMPC iterate, general per-step reference
    cfg = north_star_cfg(M, num_agents=6, max_scenarios=1)
    h = M.Handle(cfg)
    oc = ocfg_of(O, cfg)
    st, tg = dense_scenarios(5, 6, 21, 0.8, spread=1.3, vmax=1.0)
    rng = np.random.default_rng(0)
    for s in range(5):
        for a in (0, 3, 5):
            ref = np.tile(tg[s, a], cfg.k_hor) + rng.normal(0, 0.05, 3 * cfg.k_hor)
            got = h.optimize_one(st[s], a, ref)
            exp = O.optimize(oc, st[s], a, ref)
            assert np.array_equal(got["status"][:max(1, exp["n_curves"])], exp["status"][:max(1, exp["n_curves"])])
            assert got["n_curves"] == exp["n_curves"] and got["success"] == exp["success"]
            if exp["n_curves"]:
                assert np.abs(got["curves"] - exp["curves"]).max() < 1e-5
    h.close()


def test_device_pointer_step_rollout_and_agent_shards(M, O, torch):
    S, N = 32, 16
    cfg = north_star_cfg(M, num_agents=N, max_scenarios=S)
    h = M.Handle(cfg)
    oc = ocfg_of(O, cfg)
    st, tg = swarm(S, N, seed=9)
    d_st, d_tg = torch.tensor(st, device="cuda"), torch.tensor(tg, device="cuda")
    cp = torch.zeros((S, N, h.n), dtype=torch.float64, device="cuda")
    status = torch.zeros((S, N, 2), dtype=torch.int32, device="cuda")
    nc = torch.zeros((S, N), dtype=torch.int32, device="cuda")
    obj = torch.zeros((S, N), dtype=torch.float64, device="cuda")
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        h.step_batch(S, d_st, d_tg, cp, status, nc, obj, stream=stream.cuda_stream)
    stream.synchronize()
    host = h.step_batch_host(st, tg)
    assert np.array_equal(cp.cpu().numpy(), host["ctrl_pts"])  # same kernel, same bits
    assert np.array_equal(status.cpu().numpy(), host["status"])
    # pinned output buffers: control points are stored by the kernel straight into host memory
    pinned = dict(ctrl_pts=torch.full((S, N, h.n), np.nan, dtype=torch.float64).pin_memory(),
                  status=torch.zeros((S, N, 2), dtype=torch.int32).pin_memory(),
                  n_curves=torch.zeros((S, N), dtype=torch.int32).pin_memory(),
                  objective=torch.zeros((S, N), dtype=torch.float64).pin_memory())
    h.step_batch_host(st, tg, out={k: v.numpy() for k, v in pinned.items()})
    for k in pinned:
        assert np.array_equal(pinned[k].numpy(), host[k]), k
    # ... and pinned inputs are read by the kernel in place
    pst, ptg = torch.tensor(st).pin_memory(), torch.tensor(tg).pin_memory()
    pinned["ctrl_pts"].fill_(float("nan"))
    h.step_batch_host(pst.numpy(), ptg.numpy(), out={k: v.numpy() for k, v in pinned.items()})
    for k in pinned:
        assert np.array_equal(pinned[k].numpy(), host[k]), k
    # split form: enqueue on two handles, then synchronise both (what bench.py's pipelined e2e leg does)
    h_b = M.Handle(cfg)
    pinned_b = {k: torch.zeros_like(v).pin_memory() for k, v in pinned.items()}
    pinned["ctrl_pts"].fill_(float("nan"))
    h.step_batch_host_async(pst.numpy(), ptg.numpy(), {k: v.numpy() for k, v in pinned.items()})
    h_b.step_batch_host_async(pst.numpy(), ptg.numpy(), {k: v.numpy() for k, v in pinned_b.items()})
    h.synchronize()
    h_b.synchronize()
    for k in pinned:
        assert np.array_equal(pinned[k].numpy(), host[k]) and np.array_equal(pinned_b[k].numpy(), host[k]), k
    h_b.close()
    # agent-sharded launch: two halves reproduce the whole
    for begin, cnt in ((0, 8), (8, 8), (5, 3)):
        cp2 = torch.zeros((S, cnt, h.n), dtype=torch.float64, device="cuda")
        st2 = torch.zeros((S, cnt, 2), dtype=torch.int32, device="cuda")
        h.step_agents(S, begin, cnt, d_st, d_tg[:, begin:begin + cnt].contiguous(), cp2, st2)
        torch.cuda.synchronize()
        assert np.array_equal(cp2.cpu().numpy(), host["ctrl_pts"][:, begin:begin + cnt])
    # rollout kernel vs oracle curve evaluation at t = h
    nxt = torch.zeros_like(d_st)
    nc[0, 0] = 0
    h.rollout(S, d_st, cp, nc, nxt)
    torch.cuda.synchronize()
    nx = nxt.cpu().numpy()
    assert np.array_equal(nx[0, 0, :3], st[0, 0, :3]) and np.all(nx[0, 0, 3:] == 0)
    for s, a in ((0, 1), (3, 7), (31, 15)):
        x = host["ctrl_pts"][s, a]
        exp = np.concatenate([O.eval_curve(oc, x, cfg.h, 0), O.eval_curve(oc, x, cfg.h, 1)])
        assert np.allclose(nx[s, a], exp, atol=1e-12)
    launches, qps = h.counters()
    assert launches >= 5 and qps >= 2 * S * N * 2
    assert h.measure_fp64_tflops() > 1.0
    h.close()
