// TEST INFRASTRUCTURE — NOT PRODUCT CODE (see oracle/README.md).
//
// Exact dense convex-QP solver standing in for the reference's only solver backend,
// qpcpp::CPLEXSolver::solve (qpcpp/src/solvers/CPLEX.cpp:35-178; IBM CPLEX 22.1.1, proprietary,
// not present here). CPLEX's published contract for this call: minimise the convex quadratic
//      sum_{i<=j} c_ij x_i x_j + sum_i c_i x_i (+ const)      (CPLEX.cpp:83-107)
// subject to  min_r <= a_r.x <= max_r  (CPLEX.cpp:69-80) and variable bounds (CPLEX.cpp:57-66),
// |bound| >= 1e20 meaning "no bound", returning Optimal / Infeasible / Unbounded
// (CPLEX.cpp:143-174). The minimiser of the MPC-CBF QPs is unique, so any
// exact method reproduces CPLEX's answer to within CPLEX's own 1e-6 tolerances.
//
// Method (textbook, not CPLEX's primal simplex-QP):
//   1. equality rows (min == max) are eliminated by a column-pivoted Householder QR
//      null-space parametrisation x = x_p + Z w;
//   2. the reduced strictly convex QP is solved by the Goldfarb-Idnani dual active-set method
//      (Math. Programming 27 (1983) 1-33), which terminates finitely with the exact active set
//      or a proof of infeasibility;
//   3. a reduced Hessian that is only PSD (slack mode: linear-cost slack variables) is handled by
//      an outer proximal-point loop around step 2;
//   4. KKT residuals (reduced stationarity, primal/dual feasibility, complementarity) are
//      returned as the certificate the tests assert on.
#pragma once
#include <limits>

#include "linalg.hpp"

namespace oracle {

// qpcpp/include/qpcpp/solvers/Solver.h:13-21 — same order, same integer codes.
enum SolveStatus : int {
    OPTIMAL = 0,
    FEASIBLE = 1,
    UNBOUNDED = 2,
    INFEASIBLE = 3,
    ERROR = 4,
    UNKNOWN = 5,
    INFEASIBLEORUNBOUNDED = 6
};

constexpr double kCplexInf = 1e20;  // CPLEX treats |bound| >= 1e20 as infinite

template <typename R>
struct DenseQP {
    int n = 0;
    Mat<R> P;        // n x n, objective x'Px + q'x (reference convention, Problem.cpp:89-117)
    Vec<R> q;        // n
    Mat<R> A;        // m x n
    Vec<R> l, u;     // m
    Vec<R> lb, ub;   // n (variable bounds)
    int m() const { return A.rows; }
};

template <typename R>
struct QPResult {
    int status = UNKNOWN;
    Vec<R> x;
    R objective = 0;       // x'Px + q'x
    Vec<R> y;              // per-row multiplier (>=0: upper side active, <=0: lower side)
    int iterations = 0;    // active-set changes
    int n_active = 0;
    R res_stationarity = 0, res_primal = 0, res_dual = 0, res_comp = 0;
    int rank_eq = 0, n_free = 0;
};

namespace detail {

// Goldfarb-Idnani on  min 1/2 w'Gw + g'w  s.t.  C_i.w >= d_i ,  G = L L' given by its Cholesky
// factor. Returns OPTIMAL or INFEASIBLE. `u` (size mi) receives the multipliers.
template <typename R>
int goldfarb_idnani(const Mat<R>& L, const Vec<R>& g, const Mat<R>& C, const Vec<R>& d,
                    R feas_tol, Vec<R>& w, Vec<R>& u_out, int& iters) {
    const int r = L.rows, mi = C.rows;
    const R eps = std::numeric_limits<R>::epsilon();
    // J = L^-T (columns j: solve L' J_j = e_j)
    Mat<R> J(r, r);
    for (int j = 0; j < r; ++j) {
        Vec<R> e(r, R(0));
        e[j] = 1;
        solve_lower_T(L, e);
        for (int i = 0; i < r; ++i) J(i, j) = e[i];
    }
    // unconstrained minimiser
    w.assign(r, R(0));
    {
        Vec<R> t(r);
        for (int i = 0; i < r; ++i) t[i] = -g[i];
        solve_lower(L, t);
        solve_lower_T(L, t);
        w = t;
    }
    Vec<R> cnorm(mi);
    for (int i = 0; i < mi; ++i) cnorm[i] = std::sqrt(dot(C.row(i), C.row(i), r));

    std::vector<int> act;            // active constraint ids, in order of R's columns
    std::vector<char> is_act(mi, 0);
    Mat<R> Rm(r, r);                 // upper triangular, q x q used
    Vec<R> u;                        // multipliers of active set (size q)
    iters = 0;
    const int max_iter = 50 * (r + mi) + 100;

    auto drop = [&](int lpos) {
        const int q = static_cast<int>(act.size());
        is_act[act[lpos]] = 0;
        act.erase(act.begin() + lpos);
        u.erase(u.begin() + lpos);
        // remove column lpos from R, restore triangular form with Givens on rows (lpos, lpos+1..)
        for (int c = lpos; c < q - 1; ++c)
            for (int i = 0; i < q; ++i) Rm(i, c) = Rm(i, c + 1);
        for (int i = 0; i < q; ++i) Rm(i, q - 1) = 0;
        for (int k = lpos; k < q - 1; ++k) {
            const R a = Rm(k, k), b = Rm(k + 1, k);
            const R h = std::hypot(a, b);
            if (h == 0) continue;
            const R cs = a / h, sn = b / h;
            for (int c = k; c < q - 1; ++c) {
                const R t1 = Rm(k, c), t2 = Rm(k + 1, c);
                Rm(k, c) = cs * t1 + sn * t2;
                Rm(k + 1, c) = -sn * t1 + cs * t2;
            }
            for (int i = 0; i < r; ++i) {  // rotate columns k, k+1 of J
                const R t1 = J(i, k), t2 = J(i, k + 1);
                J(i, k) = cs * t1 + sn * t2;
                J(i, k + 1) = -sn * t1 + cs * t2;
            }
        }
    };

    while (true) {
        // Step 1: most violated constraint (normalised)
        int p = -1;
        R worst = 0;
        for (int i = 0; i < mi; ++i) {
            if (is_act[i]) continue;
            const R s = dot(C.row(i), w.data(), r) - d[i];
            const R tol = feas_tol * std::max(R(1), std::fabs(d[i]));
            if (cnorm[i] == 0) {  // 0.w >= d_i: violated for every w (CPLEX presolve: infeasible row)
                if (s < -tol) return INFEASIBLE;
                continue;
            }
            if (s < -tol) {
                const R v = -s / std::max(cnorm[i], R(1e-300));
                if (v > worst) { worst = v; p = i; }
            }
        }
        if (p < 0) break;  // optimal
        R up = 0;          // multiplier of the constraint being added
        const R* np = C.row(p);
        while (true) {
            if (++iters > max_iter) return ERROR;
            const int q = static_cast<int>(act.size());
            // d = J' n+
            Vec<R> dv(r);
            for (int j = 0; j < r; ++j) {
                R s = 0;
                for (int i = 0; i < r; ++i) s += J(i, j) * np[i];
                dv[j] = s;
            }
            // z = J2 d2
            Vec<R> z(r, R(0));
            for (int j = q; j < r; ++j)
                for (int i = 0; i < r; ++i) z[i] += J(i, j) * dv[j];
            // rr = R^-1 d1
            Vec<R> rr(q);
            for (int i = q - 1; i >= 0; --i) {
                R s = dv[i];
                for (int k = i + 1; k < q; ++k) s -= Rm(i, k) * rr[k];
                rr[i] = s / Rm(i, i);
            }
            const R ztn = dot(z.data(), np, r);
            R d2n = 0;
            for (int j = q; j < r; ++j) d2n += dv[j] * dv[j];
            // z'n+ = |d2|^2 ; n+ is (numerically) in the span of the active normals when
            // |d2| is at round-off level relative to |J'n+|.
            const R dall = dot(dv.data(), dv.data(), r);
            const R zt = R(1e4) * eps;
            const bool z_zero = !(d2n > zt * zt * dall) || !(ztn > 0);
            R t1 = std::numeric_limits<R>::infinity();
            int lpos = -1;
            for (int j = 0; j < q; ++j)
                if (rr[j] > 0) {
                    const R t = u[j] / rr[j];
                    if (t < t1) { t1 = t; lpos = j; }
                }
            const R sp = dot(np, w.data(), r) - d[p];
            const R t2 = z_zero ? std::numeric_limits<R>::infinity() : (-sp / ztn);
            const R t = std::min(t1, t2);
            if (!(t < std::numeric_limits<R>::infinity())) {
                u_out.assign(mi, R(0));
                return INFEASIBLE;
            }
            if (z_zero) {  // dual step only
                for (int j = 0; j < q; ++j) u[j] -= t * rr[j];
                up += t;
                drop(lpos);
                continue;
            }
            for (int i = 0; i < r; ++i) w[i] += t * z[i];
            for (int j = 0; j < q; ++j) u[j] -= t * rr[j];
            up += t;
            if (t2 <= t1) {  // full step: add p
                // rotate d2 into a multiple of e_q, updating J2
                for (int k = r - 1; k > q; --k) {
                    const R a = dv[k - 1], b = dv[k];
                    const R h = std::hypot(a, b);
                    if (h == 0) continue;
                    const R cs = a / h, sn = b / h;
                    dv[k - 1] = h;
                    dv[k] = 0;
                    for (int i = 0; i < r; ++i) {
                        const R t1j = J(i, k - 1), t2j = J(i, k);
                        J(i, k - 1) = cs * t1j + sn * t2j;
                        J(i, k) = -sn * t1j + cs * t2j;
                    }
                }
                for (int i = 0; i <= q; ++i) Rm(i, q) = dv[i];
                act.push_back(p);
                is_act[p] = 1;
                u.push_back(up);
                break;
            }
            drop(lpos);  // partial step
        }
    }
    u_out.assign(mi, R(0));
    for (size_t j = 0; j < act.size(); ++j) u_out[act[j]] = u[j];
    return OPTIMAL;
}

}  // namespace detail

// Solve the dense QP. feas_tol: absolute row tolerance (scaled by max(1,|bound|)).
template <typename R>
QPResult<R> solve_dense_qp(const DenseQP<R>& qp, R feas_tol) {
    QPResult<R> res;
    const int n = qp.n, m = qp.m();
    const R INF = R(kCplexInf);
    res.x.assign(n, R(0));
    res.y.assign(m, R(0));
    // symmetrised Hessian of x'Px : H = P + P'
    Mat<R> H(n, n);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) H(i, j) = qp.P(i, j) + qp.P(j, i);

    // ---- split rows into equalities and one-sided inequalities  c.x >= d
    std::vector<int> eq_rows;
    struct Ineq { int row; int sign; R d; bool is_bound; };  // sign*a.x >= d
    std::vector<Ineq> ineq;
    for (int i = 0; i < m; ++i) {
        const bool lo = qp.l[i] > -INF, hi = qp.u[i] < INF;
        if (lo && hi && qp.l[i] == qp.u[i]) { eq_rows.push_back(i); continue; }
        if (lo && hi && qp.l[i] > qp.u[i]) { res.status = INFEASIBLE; return res; }
        if (lo) ineq.push_back({i, +1, qp.l[i], false});
        if (hi) ineq.push_back({i, -1, -qp.u[i], false});
    }
    std::vector<int> fixed_vars;
    for (int j = 0; j < n; ++j) {
        const R lo = qp.lb.empty() ? -INF : qp.lb[j];
        const R hi = qp.ub.empty() ? INF : qp.ub[j];
        const bool blo = lo > -INF, bhi = hi < INF;
        if (blo && bhi && lo == hi) { fixed_vars.push_back(j); continue; }
        if (blo && bhi && lo > hi) { res.status = INFEASIBLE; return res; }
        if (blo) ineq.push_back({j, +1, lo, true});
        if (bhi) ineq.push_back({j, -1, -hi, true});
    }
    const int me = static_cast<int>(eq_rows.size() + fixed_vars.size());

    // ---- null space of the equality block via QR of E' (n x me)
    Mat<R> Q = Mat<R>::identity(n);
    int rank = 0;
    Vec<R> xp(n, R(0));
    if (me > 0) {
        Mat<R> Et(n, me);
        Vec<R> e(me);
        int c = 0;
        for (int i : eq_rows) {
            for (int j = 0; j < n; ++j) Et(j, c) = qp.A(i, j);
            e[c++] = qp.l[i];
        }
        for (int j : fixed_vars) {
            Et(j, c) = 1;
            e[c++] = qp.lb[j];
        }
        std::vector<int> perm;
        Mat<R> Rf = Et;
        rank = qr_colpiv(Rf, Q, perm, R(1e3) * std::numeric_limits<R>::epsilon());
        // E' Pi = Q R  ->  (Pi' E) x = R' Q' x = Pi' e ; y1 = first `rank` comps of Q'x
        Vec<R> y1(rank);
        for (int i = 0; i < rank; ++i) {
            R s = e[perm[i]];
            for (int k = 0; k < i; ++k) s -= Rf(k, i) * y1[k];
            y1[i] = s / Rf(i, i);
        }
        for (int j = 0; j < n; ++j) {
            R s = 0;
            for (int i = 0; i < rank; ++i) s += Q(j, i) * y1[i];
            xp[j] = s;
        }
        // consistency of dependent equality rows
        for (int cidx = rank; cidx < me; ++cidx) {
            R s = 0;
            for (int k = 0; k < rank; ++k) s += Rf(k, cidx) * y1[k];
            const R rhs = e[perm[cidx]];
            if (std::fabs(s - rhs) > feas_tol * std::max(R(1), std::fabs(rhs))) {
                res.status = INFEASIBLE;
                return res;
            }
        }
    }
    const int r = n - rank;
    res.rank_eq = rank;
    res.n_free = r;
    // Z = Q[:, rank:]
    Mat<R> Z(n, r);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < r; ++j) Z(i, j) = Q(i, rank + j);

    const int mi = static_cast<int>(ineq.size());
    // reduced data
    Mat<R> HZ = matmul(H, Z);
    Mat<R> G = matmul(transpose(Z), HZ);  // r x r (Hessian of the 1/2 w'Gw form)
    for (int i = 0; i < r; ++i)
        for (int j = i + 1; j < r; ++j) G(i, j) = G(j, i) = R(0.5) * (G(i, j) + G(j, i));
    Vec<R> Hxp = matvec(H, xp);
    Vec<R> gfull(n);
    for (int i = 0; i < n; ++i) gfull[i] = Hxp[i] + qp.q[i];
    Vec<R> g = matTvec(Z, gfull);
    Mat<R> C(mi, r);
    Vec<R> d(mi);
    for (int k = 0; k < mi; ++k) {
        const Ineq& c = ineq[k];
        R shift = 0;
        if (c.is_bound) {
            for (int j = 0; j < r; ++j) C(k, j) = R(c.sign) * Z(c.row, j);
            shift = R(c.sign) * xp[c.row];
        } else {
            const R* a = qp.A.row(c.row);
            for (int j = 0; j < r; ++j) {
                R s = 0;
                for (int i = 0; i < n; ++i) s += a[i] * Z(i, j);
                C(k, j) = R(c.sign) * s;
            }
            shift = R(c.sign) * dot(a, xp.data(), n);
        }
        d[k] = c.d - shift;
    }

    Vec<R> w(r, R(0)), u(mi, R(0));
    int status = OPTIMAL, iters = 0;
    if (r == 0) {
        for (int k = 0; k < mi; ++k)
            if (d[k] > feas_tol * std::max(R(1), std::fabs(ineq[k].d))) status = INFEASIBLE;
    } else {
        R tr = 0;
        for (int i = 0; i < r; ++i) tr += G(i, i);
        const R scale = std::max(tr / R(r), R(1e-300));
        Mat<R> L = G;
        const R chol_tol = R(1e4) * std::numeric_limits<R>::epsilon() * scale;
        // tolerance scaled per original bound magnitude: pass through d-based scaling
        if (cholesky_lower(L, chol_tol)) {
            status = detail::goldfarb_idnani(L, g, C, d, feas_tol, w, u, iters);
        } else {
            // proximal-point outer loop:  w+ = argmin 1/2 w'Gw + g'w + rho/2 |w - wk|^2
            const R rho = R(1e-4) * scale;
            Mat<R> Gr = G;
            for (int i = 0; i < r; ++i) Gr(i, i) += rho;
            Mat<R> Lr = Gr;
            if (!cholesky_lower(Lr, chol_tol)) { res.status = ERROR; return res; }
            Vec<R> wk(r, R(0));
            status = ERROR;
            for (int it = 0; it < 20000; ++it) {
                Vec<R> gk(r);
                for (int i = 0; i < r; ++i) gk[i] = g[i] - rho * wk[i];
                int it_in = 0;
                const int st = detail::goldfarb_idnani(Lr, gk, C, d, feas_tol, w, u, it_in);
                iters += it_in;
                if (st != OPTIMAL) { status = st; break; }
                R dw = 0, wn = 0;
                for (int i = 0; i < r; ++i) {
                    dw = std::max(dw, std::fabs(w[i] - wk[i]));
                    wn = std::max(wn, std::fabs(w[i]));
                }
                wk = w;
                if (wn > R(1e15)) { status = UNBOUNDED; break; }
                // (absolute floor 1e-14: the long-double run of a degenerate slack QP otherwise creeps
                // for thousands of outer iterations below any tolerance the tests look at)
                if (dw <= std::max(R(1e3) * std::numeric_limits<R>::epsilon(), R(1e-14)) * std::max(R(1), wn)) {
                    status = OPTIMAL;
                    break;
                }
            }
        }
    }
    res.iterations = iters;
    res.status = status;
    if (status != OPTIMAL) return res;

    // ---- recover x, objective, multipliers, certificate
    for (int i = 0; i < n; ++i) {
        R s = xp[i];
        for (int j = 0; j < r; ++j) s += Z(i, j) * w[j];
        res.x[i] = s;
    }
    Vec<R> Px = matvec(qp.P, res.x);
    res.objective = dot(res.x.data(), Px.data(), n) + dot(qp.q.data(), res.x.data(), n);
    // gradient of the Lagrangian restricted to the null space
    Vec<R> grad = matvec(H, res.x);
    for (int i = 0; i < n; ++i) grad[i] += qp.q[i];
    R comp = 0, dual = 0, pri = 0;
    for (int k = 0; k < mi; ++k) {
        const Ineq& c = ineq[k];
        const R uk = u[k];
        if (uk != R(0)) ++res.n_active;
        dual = std::max(dual, -uk);
        R ax;
        if (c.is_bound) {
            ax = res.x[c.row];
            grad[c.row] -= uk * R(c.sign);
        } else {
            const R* a = qp.A.row(c.row);
            ax = dot(a, res.x.data(), n);
            for (int i = 0; i < n; ++i) grad[i] -= uk * R(c.sign) * a[i];
            res.y[c.row] += (c.sign < 0 ? uk : -uk);
        }
        const R slack = R(c.sign) * ax - c.d;
        const R sc = std::max(R(1), std::fabs(c.d));
        pri = std::max(pri, -slack / sc);
        comp = std::max(comp, std::fabs(uk * slack) / sc);
    }
    for (int i : eq_rows) {
        const R ax = dot(qp.A.row(i), res.x.data(), n);
        pri = std::max(pri, std::fabs(ax - qp.l[i]) / std::max(R(1), std::fabs(qp.l[i])));
    }
    Vec<R> rg = matTvec(Z, grad);
    res.res_stationarity = norm_inf(rg);
    res.res_primal = pri;
    res.res_dual = dual;
    res.res_comp = comp;
    return res;
}

}  // namespace oracle
