"""GPU: mpccbf_solve_qp_dense_noalloc (the allocation-free variant of mpccbf_solve_qp_dense: scratch in a fixed
file-scope device array) against the existing entry point on the same QPs, two calls in a row, through the C ABI.
This file is meant to be collected LAST (its name sorts after every other test file)."""
import numpy as np
import pytest

from parity import TOL_CONTROL, TOL_OBJ_REL

pytestmark = pytest.mark.gpu


def random_qps(batch, n, m, seed):
    """Strictly convex QPs in the reference's convention min x'Px + q'x, l <= Ax <= u; a few rows one-sided."""
    rng = np.random.default_rng(seed)
    G = rng.standard_normal((batch, n, n))
    P = G @ G.transpose(0, 2, 1) / n + 0.5 * np.eye(n)
    q = rng.standard_normal((batch, n))
    A = rng.standard_normal((batch, m, n))
    c = rng.uniform(-0.5, 0.5, size=(batch, m))
    l, u = c - rng.uniform(0.1, 1.0, size=(batch, m)), c + rng.uniform(0.1, 1.0, size=(batch, m))
    u[:, ::5] = np.finfo(float).max          # "no bound" (|bound| >= 1e20)
    l[:, 3::7] = -1e30
    return P, q, A, l, u


def run_noalloc(M, cfg, P, q, A, l, u, stream=None):
    import torch
    dev = [torch.tensor(a, device="cuda") for a in (P, q, A, l, u)]
    batch, n = q.shape
    x = torch.full((batch, n), float("nan"), dtype=torch.float64, device="cuda")
    obj = torch.full((batch,), float("nan"), dtype=torch.float64, device="cuda")
    status = torch.full((batch,), -1, dtype=torch.int32, device="cuda")
    iters = torch.full((batch,), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    M.solve_qp_dense_noalloc(cfg, *dev, x, obj, status, iters, stream=stream)
    torch.cuda.synchronize()
    return dict(x=x.cpu().numpy(), objective=obj.cpu().numpy(), status=status.cpu().numpy(), iters=iters.cpu().numpy())


def assert_same(got, ref):
    assert np.array_equal(got["status"], ref["status"])
    assert (ref["status"] == 0).any()
    assert np.abs(got["x"] - ref["x"]).max() <= TOL_CONTROL
    rel = np.abs(got["objective"] - ref["objective"]) / np.maximum(1.0, np.abs(ref["objective"]))
    assert rel.max() <= TOL_OBJ_REL
    assert np.array_equal(got["iters"], ref["iters"])     # same kernel on the same inputs


def test_noalloc_dense_qp_matches_the_allocating_entry_on_two_calls_in_a_row(M):
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device (no fallback)"
    cfg = M.default_config(max_scenarios=1)
    h = M.Handle(cfg)
    # two different shapes back to back: the second call finds the first call's contents in the shared scratch.
    # (24, 48, 140): A stays in shared memory; (6, 96, 260): A goes to the scratch array as well
    for batch, n, m, seed in ((24, 48, 140, 1), (6, 96, 260, 2), (24, 48, 140, 3)):
        P, q, A, l, u = random_qps(batch, n, m, seed)
        ref = h.solve_qp_dense(P, q, A, l, u)
        assert_same(run_noalloc(M, cfg, P, q, A, l, u), ref)
        assert_same(run_noalloc(M, cfg, P, q, A, l, u), ref)
    # on a side stream too
    s = torch.cuda.Stream()
    P, q, A, l, u = random_qps(8, 30, 40, 4)
    assert_same(run_noalloc(M, cfg, P, q, A, l, u, stream=s.cuda_stream), h.solve_qp_dense(P, q, A, l, u))
    h.close()


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch(M):
    import torch
    n = 128
    batch = M.NOALLOC_SCRATCH_BYTES // 8 // (n * n) + 1      # one QP more than the array holds
    z = torch.zeros(1, device="cuda", dtype=torch.float64)
    q = torch.zeros((batch, n), device="cuda", dtype=torch.float64)
    i = torch.zeros(batch, device="cuda", dtype=torch.int32)
    with pytest.raises(M.MpcCbfError) as e:
        M.solve_qp_dense_noalloc(None, z, q, None, None, None, q, z, i, i)
    assert e.value.code == M.ERR_UNSUPPORTED and "MPCCBF_NOALLOC_SCRATCH_BYTES" in str(e.value)
