"""Parity criteria of BASELINE.json's north_star, applied to a product result vs the oracle:
identical feasible/infeasible status, controls within 1e-5 absolute, objective within 1e-6
relative, CBF constraint violation <= 1e-7."""
import numpy as np

TOL_CONTROL = 1e-5
TOL_OBJ_REL = 1e-6
TOL_CBF_VIOL = 1e-7


def dense_scenarios(S, N, seed, dmin, spread=1.35, vmax=1.0):
    """Agents packed close to d_min with random velocities: CBF rows active / some infeasible."""
    rng = np.random.default_rng(seed)
    side = int(np.ceil(np.sqrt(N)))
    gx, gy = np.meshgrid(np.arange(side), np.arange(side))
    grid = np.stack([gx.ravel(), gy.ravel()], 1)[:N] * dmin * spread
    states = np.zeros((S, N, 6))
    states[:, :, :2] = grid[None] + rng.uniform(-0.12, 0.12, size=(S, N, 2)) * dmin
    states[:, :, 2] = rng.uniform(-0.3, 0.3, size=(S, N))
    states[:, :, 3:5] = rng.uniform(-vmax, vmax, size=(S, N, 2))
    states[:, :, 5] = rng.uniform(-0.2, 0.2, size=(S, N))
    centre = grid.mean(0)
    targets = np.zeros((S, N, 3))
    targets[:, :, :2] = 2 * centre[None, None] - states[:, :, :2]
    return states, targets


def cbf_violation(O, ocfg, U, states, self_idx, curves):
    """max over IMPC iterates of the safety-row violation  -(a.u_k + b)  (scaled like CPLEX's
    row feasibility: absolute on the row as inserted)."""
    worst = 0.0
    N = states.shape[0]
    K = ocfg.k_hor
    for it, x in enumerate(curves):
        if it == 0:
            ego = [states[self_idx]]
        else:
            prev = curves[it - 1]
            ego = [np.concatenate([O.eval_curve(ocfg, prev, k * ocfg.h, 0), O.eval_curve(ocfg, prev, k * ocfg.h, 1)])
                   for k in range(ocfg.cbf_horizon)]
        for j in range(N):
            if j == self_idx:
                continue
            for k, e in enumerate(ego):
                a, b = O.safety_cbf_row(e, states[j], ocfg.d_min, ocfg.cbf_gamma)
                if abs(b) >= 1e20:
                    continue
                u = U[3 * k:3 * k + 3] @ x
                worst = max(worst, -(a @ u + b))
    return worst


def check_step(O, ocfg, U, states, targets, got, ref=None, check_viol_max=64):
    """got/ref: dicts with ctrl_pts [S][N][n], status [S][N][it], n_curves [S][N], objective [S][N]."""
    S, N = states.shape[:2]
    if ref is None:
        ref = O.step_batch(ocfg, states, targets, threads=O.hardware_threads(), prec=1)
    assert np.array_equal(got["status"], ref["status"]), \
        f"status mismatch at {np.argwhere(got['status'] != ref['status'])[:5]}"
    assert np.array_equal(got["n_curves"], ref["n_curves"])
    ok = ref["n_curves"] > 0
    K = ocfg.k_hor
    # controls = accelerations at every horizon sample, u_k = U_k x  (what the QP decides)
    ug = np.einsum("kn,san->sak", U, got["ctrl_pts"])
    ur = np.einsum("kn,san->sak", U, ref["ctrl_pts"])
    err_u = np.abs(ug - ur)[ok].max() if ok.any() else 0.0
    err_x = np.abs(got["ctrl_pts"] - ref["ctrl_pts"])[ok].max() if ok.any() else 0.0
    assert err_u <= TOL_CONTROL, f"controls differ by {err_u}"
    assert err_x <= TOL_CONTROL, f"control points differ by {err_x}"
    if ok.any():
        rel = np.abs(got["objective"] - ref["objective"])[ok] / np.maximum(1.0, np.abs(ref["objective"][ok]))
        assert rel.max() <= TOL_OBJ_REL, f"objective differs by {rel.max()} rel"
    assert np.all(got["ctrl_pts"][~ok] == 0.0)
    # CBF violation of the product's own iterate (last curve only is returned by step_batch)
    worst, checked = 0.0, 0
    full = ref["n_curves"] == ocfg.impc_iter
    for s, a in np.argwhere(full):
        if checked >= check_viol_max:
            break
        r1 = O.optimize(ocfg, states[s], int(a), np.tile(targets[s, a], K))
        curves = [r1["curves"][i] for i in range(ocfg.impc_iter - 1)] + [got["ctrl_pts"][s, a]]
        worst = max(worst, cbf_violation(O, ocfg, U, states[s], int(a), curves))
        checked += 1
    assert worst <= TOL_CBF_VIOL, f"CBF violation {worst}"
    return dict(err_u=err_u, err_x=err_x, cbf_violation=worst, n_infeasible=int((~full).sum()))
