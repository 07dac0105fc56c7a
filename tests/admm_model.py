"""numpy model of the dense one-CTA-per-QP ADMM kernel (mpc-cbf_b200/csrc/dense_qp.cu).

TEST INFRASTRUCTURE: the same algorithm, step for step, as the CUDA kernel, used to choose its
parameters and as a readable specification; the product never imports it. Conventions follow
OSQP (Stellato et al., 2020): min 1/2 x'Px + q'x, l <= Ax <= u.
"""
import numpy as np

INF = 1e20


def ruiz(P, q, A, l, u, iters=15):
    n, m = P.shape[0], A.shape[0]
    D, E, c = np.ones(n), np.ones(m), 1.0
    P, q, A = P.copy(), q.copy(), A.copy()
    for _ in range(iters):
        cn = np.maximum(np.abs(P).max(0), np.abs(A).max(0) if m else 0.0)
        cn = np.where(cn < 1e-4, 1.0, np.minimum(cn, 1e4))
        d = 1.0 / np.sqrt(cn)
        rn = np.abs(A).max(1) if m else np.zeros(0)
        rn = np.where(rn < 1e-4, 1.0, np.minimum(rn, 1e4))
        e = 1.0 / np.sqrt(rn)
        P = d[:, None] * P * d[None, :]
        A = e[:, None] * A * d[None, :]
        q = d * q
        D *= d
        E *= e
        pm = np.abs(P).max(0).mean()
        qn = np.abs(q).max()
        s = max(pm, qn)
        s = 1.0 if s < 1e-4 else min(s, 1e4)
        g = 1.0 / s
        P *= g
        q *= g
        c *= g
    return P, q, A, D, E, c


def solve(P_ref, q_ref, A, l, u, eps=1e-7, rho0=0.1, sigma=1e-6, alpha=1.6, max_iter=4000,
          check_every=10, polish=True, verbose=False, early=0.0, early_every=50):
    """P_ref in the reference convention x'Px+q'x. Returns x, status, iters, info."""
    n = q_ref.size
    m = l.size
    P0 = P_ref + P_ref.T
    l = np.where(l <= -INF, -np.inf, l)
    u = np.where(u >= INF, np.inf, u)
    P, q, As, D, E, c = ruiz(P0, q_ref, A, l, u)
    ls, us = E * l, E * u
    eq = np.isfinite(us) & np.isfinite(ls) & ((us - ls) <= 1e-12 * np.maximum(1.0, np.abs(us)))
    free = np.isinf(ls) & np.isinf(us)

    def rho_vec(rho):
        r = np.full(m, rho)
        r[eq] = 1e3 * rho
        r[free] = 1e-6
        return r

    rho = rho0
    rv = rho_vec(rho)
    x, z, y = np.zeros(n), np.zeros(m), np.zeros(m)

    def factor(rv):
        M = P + sigma * np.eye(n) + As.T @ (rv[:, None] * As)
        return np.linalg.cholesky(M)

    L = factor(rv)
    status = "max_iter"
    it = 0
    nfact = 1
    for it in range(1, max_iter + 1):
        rhs = sigma * x - q + As.T @ (rv * z - y)
        xt = np.linalg.solve(L.T, np.linalg.solve(L, rhs))
        zt = As @ xt
        xn = alpha * xt + (1 - alpha) * x
        zr = alpha * zt + (1 - alpha) * z
        zn = np.clip(zr + y / rv, ls, us)
        yn = y + rv * (zr - zn)
        dx, dy = xn - x, yn - y
        x, z, y = xn, zn, yn
        if it % check_every == 0:
            Ax = As @ x
            Px = P @ x
            Aty = As.T @ y
            rp = np.abs((Ax - z) / E).max() if m else 0.0
            rd = np.abs((Px + q + Aty) / D).max() / c
            np_ = max(np.abs(Ax / E).max(), np.abs(z / E).max()) if m else 0.0
            nd = max(np.abs(Px / D).max(), np.abs(Aty / D).max(), np.abs(q / D).max()) / c
            if verbose:
                print(it, rp, rd, rho)
            if rp <= eps + eps * np_ and rd <= eps + eps * nd:
                status = "solved"
                break
            if early > 0 and it % early_every == 0 and rp <= early * (1 + np_) and rd <= early * (1 + nd):
                xp, ok, rounds = polish_active_set(P, q, As, ls, us, x, z, y, eq)
                if ok:
                    return D * xp, "solved", dict(iters=it, nfact=nfact, rho=rho, polished=True, polish_rounds=rounds, early=True)
            # primal infeasibility certificate
            dyn = np.abs(E * dy).max()
            if dyn > 1e-30:
                Atdy = np.abs((As.T @ dy) / D).max()
                sup = (np.where(np.isinf(us), 0, us) * np.maximum(dy, 0)).sum() + \
                      (np.where(np.isinf(ls), 0, ls) * np.minimum(dy, 0)).sum()
                bad = (np.isinf(us) & (dy > 1e-5 * dyn / np.maximum(E, 1e-30).max())).any() or \
                      (np.isinf(ls) & (dy < -1e-5 * dyn)).any()
                if Atdy <= 1e-5 * dyn and sup < -1e-5 * dyn and not bad:
                    status = "primal_infeasible"
                    break
            dxn = np.abs(D * dx).max()
            if dxn > 1e-30:
                Pdx = np.abs((P @ dx) / D).max()
                qdx = q @ dx
                Adx = (As @ dx) / E
                ok = (np.where(np.isinf(us), True, Adx <= 1e-5 * dxn) &
                      np.where(np.isinf(ls), True, Adx >= -1e-5 * dxn)).all()
                if Pdx <= 1e-5 * dxn * c and qdx < -1e-5 * dxn * c and ok:
                    status = "dual_infeasible"
                    break
            # adaptive rho
            if it % (check_every * 5) == 0:
                a = rp / max(np_, 1e-30)
                b = rd / max(nd, 1e-30)
                new = rho * np.sqrt(a / max(b, 1e-30))
                new = min(max(new, 1e-6), 1e6)
                if new > 5 * rho or new < 0.2 * rho:
                    rho = new
                    rv = rho_vec(rho)
                    L = factor(rv)
                    nfact += 1
    info = dict(iters=it, nfact=nfact, rho=rho)
    if status == "solved" and polish:
        xp, ok, rounds = polish_active_set(P, q, As, ls, us, x, z, y, eq)
        info["polished"] = ok
        info["polish_rounds"] = rounds
        if ok:
            x = xp
    xo = D * x
    return xo, status, info


def polish_active_set(P, q, A, l, u, x, z, y, eq, delta=1e-7, rho_p=1e4, rounds=12, inner=12):
    """Active-set refinement in the scaled space: equality-constrained solves by the proximal
    method of multipliers on the guessed working set, repaired by adding violated rows and
    dropping rows whose multiplier has the wrong sign."""
    n, m = q.size, l.size
    lo = (z - l < -y) & ~np.isinf(l)
    up = (u - z < y) & ~np.isinf(u)
    lo |= eq
    up &= ~eq
    xk = x.copy()
    scale = 1.0
    for rnd in range(1, rounds + 1):
        W = np.where(lo | up)[0]
        b = np.where(lo[W], l[W], u[W])
        Aw = A[W]
        M = P + delta * np.eye(n) + rho_p * (Aw.T @ Aw)
        try:
            L = np.linalg.cholesky(M)
        except np.linalg.LinAlgError:
            return x, False, rnd
        lam = np.where(lo[W], np.minimum(y[W], 0), np.maximum(y[W], 0))
        lam[eq[W]] = y[W][eq[W]]
        for _ in range(inner):
            # Newton step on the augmented Lagrangian in residual-correction form: the right-hand
            # side is the (small) gradient, so the solve's rounding error scales with the step
            r = Aw @ xk - b
            g = P @ xk + q + Aw.T @ (lam + rho_p * r)
            xk = xk - np.linalg.solve(L.T, np.linalg.solve(L, g))
            r = Aw @ xk - b
            lam = lam + rho_p * r
            if np.abs(r).max(initial=0.0) < 1e-13 * max(1.0, np.abs(b).max(initial=0.0)):
                grad = np.abs(P @ xk + q + Aw.T @ lam).max()
                if grad < 1e-12 * max(1.0, np.abs(q).max()):
                    break
        Ax = A @ xk
        tolp = 1e-10 * np.maximum(1.0, np.abs(Ax))
        viol_lo = (l - Ax > tolp) & ~lo
        viol_up = (Ax - u > tolp) & ~up
        lamfull = np.zeros(m)
        lamfull[W] = lam
        ltol = 1e-10 * max(1.0, np.abs(lam).max(initial=0.0))
        bad_lo = lo & ~eq & (lamfull > ltol)
        bad_up = up & (lamfull < -ltol)
        if not (viol_lo.any() or viol_up.any() or bad_lo.any() or bad_up.any()):
            return xk, True, rnd
        lo = (lo | viol_lo) & ~bad_lo
        up = (up | viol_up) & ~bad_up
        y = lamfull
    return x, False, rounds
