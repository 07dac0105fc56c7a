// Dual active-set (Goldfarb-Idnani, least-distance form) on a D-dimensional projection problem
//      min 1/2 |v - v0|^2   s.t.   n_i . v >= rhs_i
// with the whole factorisation (orthogonal J, triangular R of the active normals) in registers.
// Used by the thread-per-agent fused kernel with D = 4 (x/y axes, coupled by the CBF rows) and
// D = 2 (yaw axis, decoupled). The state is resumable: gi_run can be called again after more rows
// became visible to the row provider (the iterate of a sub-problem is a valid dual-feasible
// starting point for the larger problem).
//
// Replaces qpcpp::CPLEXSolver::solve (qpcpp/src/solvers/CPLEX.cpp:35-178) for the reduced QP.

#include <cmath>

#include "mpccbf.h"

namespace mpccbf {
namespace r1 {

#ifndef MPCCBF_HD
#define MPCCBF_HD __host__ __device__ __forceinline__
#endif

template <int D>
struct GIState {
    double J[D][D];
    double Rm[D][D];  // upper triangular, leading q x q block used
    double u[D];      // multipliers of the active rows
    double v[D];      // current iterate
    int q;            // number of active rows
};

template <int D>
MPCCBF_HD void gi_init(GIState<D>& g, const double* v0) {
#pragma unroll
    for (int i = 0; i < D; ++i) {
#pragma unroll
        for (int j = 0; j < D; ++j) {
            g.J[i][j] = (i == j) ? 1.0 : 0.0;
            g.Rm[i][j] = 0.0;
        }
        g.u[i] = 0.0;
        g.v[i] = v0[i];
    }
    g.q = 0;
}

// Rows must provide:
//   void scan(const double (&v)[D], double& best, int& best_id) const;   best_id < 0: feasible;
//        best = +inf flags a violated row that does not depend on v (infeasible)
//   void normal(int id, double (&np)[D], double& rhs) const;             row  np.v >= rhs
// Returns a SolveStatus code; n_iter is incremented per active-set step.
template <int D, class Rows>
__host__ __device__ int gi_run(GIState<D>& g, const Rows& rows, int max_iter, int& n_iter) {
    const double kInfD = HUGE_VAL;
    for (;;) {
        double best = 0.0;
        int best_id = -1;
        rows.scan(g.v, best, best_id);
        if (best_id < 0) return MPCCBF_STATUS_OPTIMAL;
        if (!(best < kInfD)) return MPCCBF_STATUS_INFEASIBLE;
        double np[D], rhs;
        rows.normal(best_id, np, rhs);

        double up_mult = 0.0;
        for (;;) {
            if (++n_iter > max_iter) return MPCCBF_STATUS_ERROR;
            // d = J' n+
            double d[D];
#pragma unroll
            for (int j = 0; j < D; ++j) {
                double s = 0.0;
#pragma unroll
                for (int i = 0; i < D; ++i) s = fma(g.J[i][j], np[i], s);
                d[j] = s;
            }
            // z = J2 d2, |d2|^2, |d|^2
            double z[D];
#pragma unroll
            for (int i = 0; i < D; ++i) z[i] = 0.0;
            double ztn = 0.0, dall = 0.0;
#pragma unroll
            for (int j = 0; j < D; ++j) {
                dall = fma(d[j], d[j], dall);
                if (j >= g.q) {
                    ztn = fma(d[j], d[j], ztn);
#pragma unroll
                    for (int i = 0; i < D; ++i) z[i] = fma(g.J[i][j], d[j], z[i]);
                }
            }
            // r = R^-1 d1
            double rr[D];
#pragma unroll
            for (int i = D - 1; i >= 0; --i) {
                double s = d[i];
#pragma unroll
                for (int k = i + 1; k < D; ++k)
                    if (k < g.q) s = fma(-g.Rm[i][k], rr[k], s);
                rr[i] = (i < g.q) ? s / g.Rm[i][i] : 0.0;
            }
            const double zt = 1e4 * 2.220446049250313e-16;
            const bool z_zero = !(ztn > zt * zt * dall);
            double t1 = kInfD;
            int lpos = -1;
#pragma unroll
            for (int j = 0; j < D; ++j)
                if (j < g.q && rr[j] > 0.0) {
                    const double t = g.u[j] / rr[j];
                    if (t < t1) {
                        t1 = t;
                        lpos = j;
                    }
                }
            double sp = -rhs;
#pragma unroll
            for (int i = 0; i < D; ++i) sp = fma(np[i], g.v[i], sp);
            const double t2 = z_zero ? kInfD : (-sp / ztn);
            const double t = fmin(t1, t2);
            if (!(t < kInfD)) return MPCCBF_STATUS_INFEASIBLE;
            if (!z_zero) {
#pragma unroll
                for (int i = 0; i < D; ++i) g.v[i] = fma(t, z[i], g.v[i]);
            }
#pragma unroll
            for (int j = 0; j < D; ++j)
                if (j < g.q) g.u[j] = fma(-t, rr[j], g.u[j]);
            up_mult += t;

            if (!z_zero && t2 <= t1) {
                // full step: add the row. Rotate d2 onto e_q (Givens from the bottom), carry J2.
#pragma unroll
                for (int k = D - 1; k >= 1; --k) {
                    if (k > g.q) {
                        const double a = d[k - 1], b = d[k];
                        const double hh = sqrt(fma(a, a, b * b));
                        if (hh > 0.0) {
                            const double ih = 1.0 / hh;
                            const double cs = a * ih, sn = b * ih;
                            d[k - 1] = hh;
                            d[k] = 0.0;
#pragma unroll
                            for (int i = 0; i < D; ++i) {
                                const double a1 = g.J[i][k - 1], a2 = g.J[i][k];
                                g.J[i][k - 1] = fma(cs, a1, sn * a2);
                                g.J[i][k] = fma(-sn, a1, cs * a2);
                            }
                        }
                    }
                }
#pragma unroll
                for (int col = 0; col < D; ++col)
                    if (col == g.q) {
#pragma unroll
                        for (int i = 0; i <= col; ++i) g.Rm[i][col] = d[i];
                        g.u[col] = up_mult;
                    }
                ++g.q;
                break;
            }
            // partial (or dual-only) step: drop active row lpos
#pragma unroll
            for (int col = 0; col < D - 1; ++col)
                if (col >= lpos) {
#pragma unroll
                    for (int i = 0; i <= col + 1; ++i) g.Rm[i][col] = g.Rm[i][col + 1];  // Hessenberg part
                    g.u[col] = g.u[col + 1];
                }
            g.u[D - 1] = 0.0;
            --g.q;
#pragma unroll
            for (int k = 0; k < D - 1; ++k)
                if (k >= lpos && k < g.q) {
                    const double a = g.Rm[k][k], b = g.Rm[k + 1][k];
                    const double hh = sqrt(fma(a, a, b * b));
                    if (hh > 0.0) {
                        const double ih = 1.0 / hh;
                        const double cs = a * ih, sn = b * ih;
#pragma unroll
                        for (int col = 0; col < D; ++col)
                            if (col >= k) {
                                // (the sub-diagonal entry b is annihilated; it is never read again)
                                const double a1 = g.Rm[k][col], a2 = (col > k) ? g.Rm[k + 1][col] : b;
                                g.Rm[k][col] = fma(cs, a1, sn * a2);
                                if (col > k) g.Rm[k + 1][col] = fma(-sn, a1, cs * a2);
                            }
#pragma unroll
                        for (int i = 0; i < D; ++i) {
                            const double a1 = g.J[i][k], a2 = g.J[i][k + 1];
                            g.J[i][k] = fma(cs, a1, sn * a2);
                            g.J[i][k + 1] = fma(-sn, a1, cs * a2);
                        }
                    }
                }
        }
    }
}

}  // namespace r1
}  // namespace mpccbf
