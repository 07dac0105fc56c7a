// TEST INFRASTRUCTURE — NOT PRODUCT CODE (see oracle/README.md).
//
// Reference-faithful CPU restatement of the MPC-CBF QP build + IMPC chain: every QP is rebuilt
// from scratch (dense GEMMs for the tracking cost, per-piece derivative Gram blocks, the
// <=100*eps drop rule, rows in the reference's insertion order) exactly as
// mpc_cbf/src/controller/ConnectivityIMPCCBF.cpp:47-215 drives it — no caching, no warm start.
// Paths below are relative to /root/reference/workspace/lib.
#pragma once
#include <array>
#include <numeric>

#include "../include/mpccbf.h"  // mpccbf_config only (plain data)
#include "bezier_model.hpp"
#include "connectivity.hpp"
#include "qp_solve.hpp"

namespace oracle {

// cbf/src/detail/ConnectivityCBF.cpp:152-198 evaluated in closed form (the reference builds the
// same expression symbolically with GiNaC and substitutes ego 0,1,3,4 / neighbour 0,1,3,4:
// cbf/include/cbf/Helpers.hpp:10-44). alpha(s) = gamma*s^3 (ConnectivityCBF.cpp:19-21,91),
// gamma = 5 (:62). Returns a (3) and b with the meaning a.u + b >= 0.
template <typename R>
void safety_cbf_row(const R ego[6], const R nbr[6], R dmin, R gamma, R a[3], R& b) {
    const R dx = ego[0] - nbr[0], dy = ego[1] - nbr[1];
    const R dvx = ego[3] - nbr[3], dvy = ego[4] - nbr[4];
    const R h = dx * dx + dy * dy - dmin * dmin;              // :161
    const R Lfh = 2 * (dx * dvx + dy * dvy);                  // :164
    const R Lf2h = 2 * (dvx * dvx + dvy * dvy);               // :168
    // grad(alpha(h)).f with f = A*state: only d/dpx * vx + d/dpy * vy survive (:171-184)
    const R Lf_alpha = 3 * gamma * h * h * (2 * dx * ego[3] + 2 * dy * ego[4]);
    const R alpha_h = gamma * h * h * h;
    const R psi1 = Lfh + alpha_h;                             // :194
    a[0] = 2 * dx;                                            // :189-191
    a[1] = 2 * dy;
    a[2] = 0;
    b = Lf2h + Lf_alpha + gamma * psi1 * psi1 * psi1;         // :195
}

// Constant tables of PiecewiseBezierMPCQPOperations (mpc/src/optimization/
// PiecewiseBezierMPCQPOperations.cpp:9-38) + ConnectivityIMPCCBF ctor (ConnectivityIMPCCBF.cpp:9-44).
template <typename R>
struct Controller {
    mpccbf_config cfg;
    int K, p, c, n, deg;
    R T;
    std::vector<R> cum;     // cumulative_max_parameters_ (:166-176)
    Vec<R> h_samples;       // LinSpaced(K, 0, (K-1)h)
    Mat<R> U;               // U_basis_ [3K, n]
    Mat<R> A0pos, A0vel, Lpos, Lvel;
    DoubleIntegratorXYYaw<R> model;

    explicit Controller(const mpccbf_config& cf)
        : cfg(cf), K(cf.k_hor), p(cf.num_pieces), c(cf.num_ctrl_pts), n(cf.num_pieces * 3 * cf.num_ctrl_pts),
          deg(cf.num_ctrl_pts - 1), T(R(cf.piece_max_param)), model(R(cf.h)) {
        for (int i = 0; i < p; ++i) cum.push_back(cum.empty() ? T : cum.back() + T);
        model.get_A0(K, A0pos, A0vel);
        model.get_lambda(K, Lpos, Lvel);
        // sample times are produced in double by the reference (Eigen LinSpaced on doubles);
        // keep that rounding even when R is long double, or (K-1)*h overshoots the last knot.
        {
            const Vec<double> hs = linspaced<double>(K, 0.0, static_cast<double>(K - 1) * cfg.h);
            h_samples.assign(hs.begin(), hs.end());
        }
        U = eval_sampling_basis(2);
    }

    // PiecewiseBezierMPCQPOperations.cpp:190-223 — lower_bound piece lookup, clamped parameter.
    void piece_and_param(R t, int& piece, R& s) const {
        if (t < 0 || t > cum.back())
            throw std::runtime_error("pieceIndexAndParameter: parameter is out of range");
        piece = static_cast<int>(std::lower_bound(cum.begin(), cum.end(), t) - cum.begin());
        if (piece >= p) throw std::runtime_error("piece_idx is out of range");
        const R local = (piece == 0) ? t : t - cum[piece - 1];
        s = std::min(std::max(local, R(0)), T);
    }

    // PiecewiseBezierMPCQPOperations.cpp:42-60 (+ BezierQPOperations.cpp:51-69 evalBasisRow)
    Mat<R> eval_sampling_basis(uint64_t d) const {
        Mat<R> out(3 * K, n);
        for (int k = 0; k < K; ++k) {
            int piece;
            R s;
            piece_and_param(h_samples[k], piece, s);
            const Vec<R> basis = bernstein_basis<R>(deg, T, s, d);
            for (int ax = 0; ax < 3; ++ax)
                for (int j = 0; j < c; ++j) out(3 * k + ax, piece * 3 * c + ax * c + j) = basis[j];
        }
        return out;
    }

    // ---- QP assembly (what the generator writes into qpcpp::Problem) -------------------------
    struct Build {
        DenseQP<R> qp;
        int n_eq = 0, n_cbf = 0;
        int cbf_row0 = 0;  // index of the first CBF row
    };

    // PiecewiseBezierMPCQPGenerator.cpp:283-321 / 350-395 — accumulate with the drop rule.
    static void add_cost(DenseQP<R>& qp, int off, const Mat<R>& Pq, const Vec<R>& lin) {
        const R epsilon = std::numeric_limits<R>::epsilon() * R(100.0);
        // NB: epsilon follows the scalar type like the reference's template does; for the
        // long-double oracle the rule only drops exact structural zeros, which is equivalent.
        for (int i = 0; i < Pq.rows; ++i) {
            if (!(std::fabs(lin[i]) <= epsilon)) qp.q[off + i] += lin[i];
            for (int j = 0; j < Pq.cols; ++j)
                if (!(std::fabs(Pq(i, j)) <= epsilon)) qp.P(off + i, off + j) += Pq(i, j);
        }
    }

    // PiecewiseBezierMPCQPOperations.cpp:64-91 — dense tracking cost, recomputed per call.
    void position_error_cost(const R x0[6], const Vec<R>& ref, Mat<R>& quad, Vec<R>& lin) const {
        Mat<R> Q(3 * K, 3 * K);
        for (int i = 3 * (K - cfg.spd_f); i < 3 * K; ++i) Q(i, i) = R(cfg.w_pos_err);
        const Mat<R> Phi = matmul(Lpos, U);
        quad = matmul(matmul(transpose(Phi), Q), Phi);
        Vec<R> x0v(x0, x0 + 6);
        const Vec<R> a0x = matvec(A0pos, x0v);
        Vec<R> coef(3 * K);
        for (int i = 0; i < 3 * K; ++i) coef[i] = R(2.0) * a0x[i] * Q(i, i) + R(-2.0) * ref[i] * Q(i, i);
        lin = matTvec(Phi, coef);
    }

    Build build_qp(const R* states, int N, int self, const Vec<R>& ref, int iter,
                   const std::vector<std::array<R, 6>>& pred_states,
                   const Vec<R>& slack_weights) const {
        Build b;
        const bool slack = cfg.slack_mode != 0;
        const int nn = N - 1;
        const int nv = n + (slack ? nn : 0);
        DenseQP<R>& qp = b.qp;
        qp.n = nv;
        qp.P = Mat<R>(nv, nv);
        qp.q.assign(nv, R(0));
        qp.lb.assign(nv, R(std::numeric_limits<double>::lowest()));  // Problem.h default bounds
        qp.ub.assign(nv, R(std::numeric_limits<double>::max()));
        if (slack)  // MPCCBFQPGeneratorBase.cpp:28-38 — slack >= 0
            for (int j = 0; j < nn; ++j) qp.lb[n + j] = 0;
        const R* x0 = states + 6 * self;

        // cost: ConnectivityIMPCCBF.cpp:108-120
        {
            Mat<R> quad;
            Vec<R> lin;
            position_error_cost(x0, ref, quad, lin);
            add_cost(qp, 0, quad, lin);
        }
        for (int d = 1; d <= cfg.continuity_degree; ++d) {
            const Mat<R> blk = integrated_squared_derivative_block<R>(c, T, d, R(cfg.w_u_eff));
            Mat<R> piece_cost(3 * c, 3 * c);
            for (int ax = 0; ax < 3; ++ax)
                for (int i = 0; i < c; ++i)
                    for (int j = 0; j < c; ++j) piece_cost(ax * c + i, ax * c + j) = blk(i, j);
            Vec<R> zero(3 * c, R(0));
            for (int piece = 0; piece < p; ++piece) add_cost(qp, piece * 3 * c, piece_cost, zero);
        }
        if (slack)  // MPCCBFQPGeneratorBase.cpp:121-135 — linear cost only
            for (int j = 0; j < nn; ++j) qp.q[n + j] += slack_weights[j];

        // rows
        std::vector<Vec<R>> rows;
        Vec<R> lo, hi;
        auto push = [&](const Vec<R>& r, R l, R u) {
            rows.push_back(r);
            lo.push_back(l);
            hi.push_back(u);
        };
        // ConnectivityIMPCCBF.cpp:123-124 -> BezierQPOperations.cpp:149-166
        for (int d = 0; d <= 1; ++d) {
            int piece;
            R s;
            piece_and_param(R(0), piece, s);
            const Vec<R> basis = bernstein_basis<R>(deg, T, s, d);
            for (int ax = 0; ax < 3; ++ax) {
                Vec<R> r(nv, R(0));
                for (int j = 0; j < c; ++j) r[piece * 3 * c + ax * c + j] = basis[j];
                const R tgt = x0[3 * d + ax];
                push(r, tgt, tgt);
            }
        }
        // ConnectivityIMPCCBF.cpp:126-132 -> PiecewiseBezierMPCQPGenerator.cpp:182-226
        for (int piece = 0; piece + 1 < p; ++piece)
            for (int d = 0; d <= cfg.continuity_degree; ++d) {
                const Vec<R> first = bernstein_basis<R>(deg, T, T, d);
                const Vec<R> second = bernstein_basis<R>(deg, T, R(0), d);
                for (int ax = 0; ax < 3; ++ax) {
                    Vec<R> r(nv, R(0));
                    for (int j = 0; j < c; ++j) {
                        r[piece * 3 * c + ax * c + j] = first[j];
                        r[(piece + 1) * 3 * c + ax * c + j] = -second[j];
                    }
                    push(r, R(0), R(0));
                }
            }
        b.n_eq = static_cast<int>(rows.size());
        b.cbf_row0 = b.n_eq;
        // CBF rows: ConnectivityIMPCCBF.cpp:135-141 (iter 0) / :158-175 (iter >= 1)
        const R NEG = R(std::numeric_limits<double>::lowest());
        auto cbf_row = [&](const R ego[6], const R nbr[6], int k, int nbr_idx) {
            R a[3], bb;
            safety_cbf_row<R>(ego, nbr, R(cfg.d_min), R(cfg.cbf_gamma), a, bb);
            // ConnectivityMPCCBFQPOperations.cpp:37-41 / :102-106:  -(A_k' U)  <= b
            Vec<R> r(nv, R(0));
            for (int j = 0; j < n; ++j) {
                R s = 0;
                for (int ax = 0; ax < 3; ++ax) s += a[ax] * U(3 * k + ax, j);
                r[j] = R(-1.0) * s;
            }
            if (slack) r[n + nbr_idx] = R(-1.0);  // ConnectivityMPCCBFQPGenerator.cpp:33-44
            push(r, NEG, bb);
        };
        int j = 0;
        for (int other = 0; other < N; ++other) {
            if (other == self) continue;
            const R* nb = states + 6 * other;
            if (iter == 0) {
                cbf_row(x0, nb, 0, j);
            } else {
                for (size_t k = 0; k < pred_states.size(); ++k)
                    cbf_row(pred_states[k].data(), nb, static_cast<int>(k), j);
            }
            ++j;
        }
        // The block the reference ships commented out (ConnectivityIMPCCBF.cpp:143-157 iter 0, :177-192 iter >= 1),
        // enabled by cfg.conn_mode: lambda2 > 0.1 -> MPC-lifted connectivity HOCBF row(s)
        // (ConnectivityMPCCBFQPOperations.cpp:45-70 / :111-150: lambda2, the eigenvector and h from the CURRENT
        // states, the ego row of robot_states replaced by pred_states[k]), else one distance-CLF row per
        // neighbour and step (:72-86 / :152-173, inserted as written there: -(a' U_k) x <= -Bc).
        if (cfg.conn_mode && !slack) {
            auto lifted = [&](const R a[3], R ub, int k) {
                Vec<R> r(nv, R(0));
                for (int jj = 0; jj < n; ++jj) {
                    R sacc = 0;
                    for (int ax = 0; ax < 3; ++ax) sacc += a[ax] * U(3 * k + ax, jj);
                    r[jj] = R(-1.0) * sacc;
                }
                push(r, NEG, ub);
            };
            Vec<R> fv;
            const R l2 = lambda2<R>(states, N, R(cfg.d_max), fv);
            const int hit = (iter == 0) ? 1 : static_cast<int>(pred_states.size());
            if (l2 > R(0.1)) {
                std::vector<R> upd(states, states + 6 * N);
                for (int k = 0; k < hit; ++k) {
                    if (iter > 0)
                        for (int cc = 0; cc < 6; ++cc) upd[6 * self + cc] = pred_states[k][cc];
                    R a[3], bb;
                    conn_cbf_row<R>(upd.data(), N, self, fv, l2, R(cfg.d_max), R(cfg.cbf_gamma), R(0.1), a, bb);
                    lifted(a, bb, k);
                }
            } else {
                for (int other = 0; other < N; ++other) {
                    if (other == self) continue;
                    for (int k = 0; k < hit; ++k) {
                        R a[3], bc;
                        clf_row<R>(iter == 0 ? x0 : pred_states[k].data(), states + 6 * other, a, bc);
                        lifted(a, R(-1.0) * bc, k);
                    }
                }
            }
        }
        b.n_cbf = static_cast<int>(rows.size()) - b.n_eq;
        // ConnectivityIMPCCBF.cpp:196-197 -> PiecewiseBezierMPCQPGenerator.cpp:148-164
        for (int d = 2; d >= 1; --d) {
            const double* LB = (d == 2) ? cfg.a_min : cfg.v_min;
            const double* UB = (d == 2) ? cfg.a_max : cfg.v_max;
            for (int k = 0; k < K; ++k) {
                int piece;
                R s;
                piece_and_param(h_samples[k], piece, s);
                const Vec<R> basis = bernstein_basis<R>(deg, T, s, d);
                for (int ax = 0; ax < 3; ++ax) {
                    Vec<R> r(nv, R(0));
                    for (int jj = 0; jj < c; ++jj) r[piece * 3 * c + ax * c + jj] = basis[jj];
                    push(r, R(LB[ax]), R(UB[ax]));
                }
            }
        }
        const int m = static_cast<int>(rows.size());
        qp.A = Mat<R>(m, nv);
        qp.l = lo;
        qp.u = hi;
        for (int i = 0; i < m; ++i) std::copy(rows[i].begin(), rows[i].end(), qp.A.row(i));
        return b;
    }

    // splines/src/curves/SingleParameterPiecewiseCurve.cpp:94-127 + Bezier.cpp:61-75
    void eval_curve(const Vec<R>& x, R t, uint64_t d, R out[3]) const {
        if (t < 0 || t > cum.back())
            throw std::invalid_argument("SingleParameterPiecewiseCurve::eval: parameter is out of range.");
        const int piece = static_cast<int>(std::lower_bound(cum.begin(), cum.end(), t) - cum.begin());
        const R s = (piece == 0) ? t : std::min(T, t - cum[piece - 1]);
        const Vec<R> basis = bernstein_basis<R>(deg, T, s, d);
        for (int ax = 0; ax < 3; ++ax) {
            R acc = 0;
            for (int j = 0; j < c; ++j) acc += x[piece * 3 * c + ax * c + j] * basis[j];
            out[ax] = acc;
        }
    }

    struct OptimizeResult {
        bool success = true;
        std::vector<Vec<R>> curves;       // control points per successful iterate
        std::vector<int> status;          // per iteration (SKIPPED=7 after failure)
        std::vector<QPResult<R>> qps;     // per attempted iteration
        std::vector<int> m_rows;
    };

    // ConnectivityIMPCCBF.cpp:47-215
    OptimizeResult optimize(const R* states, int N, int self, const Vec<R>& ref, R feas_tol) const {
        OptimizeResult out;
        out.status.assign(cfg.impc_iter, MPCCBF_STATUS_SKIPPED);
        const int nn = N - 1;
        const R* x0 = states + 6 * self;
        Vec<R> slack_w(nn, R(0));
        if (cfg.slack_mode) {  // :76-100
            std::vector<size_t> idx(nn);
            std::iota(idx.begin(), idx.end(), 0);
            Vec<R> dist(nn);
            int j = 0;
            for (int o = 0; o < N; ++o) {
                if (o == self) continue;
                // the reference computes these distances in double (std::vector<T>, T = double) whatever the
                // precision of this oracle run, and near-ties (symmetric formations) decide the ranks: plain
                // double products and sum, no contraction into an fma
                const double ddx = static_cast<double>(states[6 * o]) - static_cast<double>(x0[0]);
                const double ddy = static_cast<double>(states[6 * o + 1]) - static_cast<double>(x0[1]);
                volatile double xx = ddx * ddx, yy = ddy * ddy;
                dist[j++] = R(std::sqrt(xx + yy));
            }
            std::stable_sort(idx.begin(), idx.end(), [&](size_t a, size_t b) { return dist[a] < dist[b]; });  // exact ties: index order
            for (int i = 0; i < nn; ++i)
                slack_w[idx[i]] = R(cfg.slack_cost) * std::pow(R(cfg.slack_decay), R(i));
        }
        for (int iter = 0; iter < cfg.impc_iter; ++iter) {
            std::vector<std::array<R, 6>> pred;
            if (iter > 0) {  // :158-168
                for (int k = 0; k < cfg.cbf_horizon; ++k) {
                    std::array<R, 6> ps;
                    eval_curve(out.curves.back(), h_samples[k], 0, ps.data());
                    eval_curve(out.curves.back(), h_samples[k], 1, ps.data() + 3);
                    pred.push_back(ps);
                }
            }
            Build b = build_qp(states, N, self, ref, iter, pred, slack_w);
            QPResult<R> r = solve_dense_qp<R>(b.qp, feas_tol);
            out.status[iter] = r.status;
            out.m_rows.push_back(b.qp.m());
            if (r.status == OPTIMAL) {
                Vec<R> x(r.x.begin(), r.x.begin() + n);
                out.curves.push_back(x);
                out.qps.push_back(std::move(r));
            } else {
                out.qps.push_back(std::move(r));
                out.success = false;
                break;
            }
        }
        return out;
    }
};

}  // namespace oracle
