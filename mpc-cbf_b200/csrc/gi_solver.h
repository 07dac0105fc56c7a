// Dual active-set (Goldfarb-Idnani, least-distance form) on a D-dimensional projection problem
//      min 1/2 |v - v0|^2   s.t.   n_i . v >= rhs_i
// with the iterate, the multipliers and the ids of the active rows in registers.
// Used by the thread-per-agent fused kernel with D = 4 (x/y axes, coupled by the CBF rows) and
// D = 2 (yaw axis, decoupled). The state is resumable: gi_run can be called again after more rows
// became visible to the row provider (the iterate of a sub-problem is a valid dual-feasible
// starting point for the larger problem), and it is nothing but (v, u, ids) — which is how the stage-A
// solution of an agent is restored before a later IMPC iteration.
//
// SOFT rows (slack mode, ConnectivityIMPCCBF.cpp:76-100 + MPCCBFQPGeneratorBase.cpp:28-38,80-135):
// the safety rows of neighbour j share one slack variable s_j >= 0 with linear cost w_j. With the
// slacks eliminated the QP is
//      min 1/2 |v - v0|^2 + sum_j w_j max(0, max_k g_jk(v)),   g_jk(v) = rhs_jk - n_jk . v,
// whose optimality conditions are those of the hard-row QP with the multipliers of one group
// capped: lambda_jk >= 0, sum_k lambda_jk <= w_j. The group's pieces are its rows plus the zero
// function ("none"); one piece is the group's LEADER (the one attaining the max; "none" while the
// slack is zero) and carries the weight that is not on the other pieces. Every other piece k is
// the generalised row  (n_k - n_lead) . v >= rhs_k - rhs_lead. The dual active-set iteration is
// the usual one with a third step-length rule: the step stops where a group's multiplier sum
// reaches its weight; there the leadership passes to an active row of that group (which leaves
// the active set: its multiplier is now implicit), or to the entering row itself when the group
// has no active row — the row is then "absorbed" by its slack and the major iteration ends.
// The reverse face (a slack returning to zero) is the piece "none" of a group whose leader is a
// row. This removes the separate slack-variable QP: same minimiser v, s_j = max(0, g_lead(v)).
//
// Replaces qpcpp::CPLEXSolver::solve (qpcpp/src/solvers/CPLEX.cpp:35-178) for the reduced QP.
#pragma once
#include <cmath>

#include "../../include/mpccbf.h"

namespace mpccbf {

#ifndef MPCCBF_HD
#define MPCCBF_HD __host__ __device__ __forceinline__
#endif

template <int D>
struct GIState {
    double u[D];      // multipliers of the active rows
    double v[D];      // current iterate
    int id[D];        // row ids of the active set
    int q;            // number of active rows
};

// group bookkeeping of the active rows (SOFT rows only)
template <int D>
struct GISoft {
    int grp[D];      // group of each active row (-1: hard row)
    double cap[D];   // weight of that group
};

template <int D>
MPCCBF_HD void gi_init(GIState<D>& g, const double* v0) {
#pragma unroll
    for (int i = 0; i < D; ++i) {
        g.u[i] = 0.0;
        g.v[i] = v0[i];
        g.id[i] = -1;
    }
    g.q = 0;
}

// removes active row lpos from the lists
template <int D, bool SOFT>
MPCCBF_HD void gi_remove(GIState<D>& g, GISoft<D>& gs, int lpos) {
#pragma unroll
    for (int c = 0; c < D - 1; ++c)
        if (c >= lpos) {
            g.u[c] = g.u[c + 1];
            g.id[c] = g.id[c + 1];
            if constexpr (SOFT) {
                gs.grp[c] = gs.grp[c + 1];
                gs.cap[c] = gs.cap[c + 1];
            }
        }
    g.u[D - 1] = 0.0;
    g.id[D - 1] = -1;
    if constexpr (SOFT) gs.grp[D - 1] = -1;
    --g.q;
}

// Rows must provide:
//   void scan(const double (&v)[D], double& best, int& best_id) const;   best_id < 0: feasible;
//        best = +inf flags a violated row that does not depend on v (infeasible)
//   void normal(int id, double (&np)[D], double& rhs) const;             row  np.v >= rhs
// and, for SOFT = true,
//   int group(int id) const;            group of a row (-1: hard row)
//   double weight(int group) const;     w_j
//   void set_leader(int group, int id) const;   the piece `id` becomes the leader of its group
// Returns a SolveStatus code; n_iter is incremented per active-set step.
//
// The state carried between steps is only (v, u, ids): every step factorises the <= D active normals
// afresh (modified Gram-Schmidt, Q R = [n_1 .. n_q]) instead of updating an orthogonal J and a triangular
// R — with D <= 4 that costs about what the Givens updates did, there is no rounding drift, and the
// 2 D^2 doubles of J and R are live only inside the step, not across the row scans (which is what kept
// the thread-per-agent kernel in local memory).
template <int D, bool SOFT, class Rows>
__host__ __device__ int gi_run(GIState<D>& g, GISoft<D>& gs, const Rows& rows, int max_iter, int& n_iter) {
    const double kInfD = HUGE_VAL;
    for (;;) {
        double best = 0.0;
        int best_id = -1;
        rows.scan(g.v, best, best_id);
        if (best_id < 0) return MPCCBF_STATUS_OPTIMAL;
        if (!(best < kInfD)) return MPCCBF_STATUS_INFEASIBLE;
        double np[D], rhs;
        rows.normal(best_id, np, rhs);
        int pgrp = -1;
        double pcap = kInfD;
        if constexpr (SOFT) {
            pgrp = rows.group(best_id);
            if (pgrp >= 0) pcap = rows.weight(pgrp);
        }

        double up_mult = 0.0;
        for (;;) {
            if (++n_iter > max_iter) return MPCCBF_STATUS_ERROR;
            // ---- Q R = active normals (MGS), d1 = Q' n+, z = n+ - Q d1 ---------------------------------
            double Q[D][D], R[D][D], d1[D], z[D];
#pragma unroll
            for (int i = 0; i < D; ++i) z[i] = np[i];
            double dall = 0.0;
#pragma unroll
            for (int i = 0; i < D; ++i) dall = fma(np[i], np[i], dall);
#pragma unroll
            for (int c = 0; c < D; ++c) {
                d1[c] = 0.0;
                if (c < g.q) {
                    double w[D], unused;
                    rows.normal(g.id[c], w, unused);
#pragma unroll
                    for (int p = 0; p < c; ++p) {
                        double s = 0.0;
#pragma unroll
                        for (int i = 0; i < D; ++i) s = fma(Q[i][p], w[i], s);
                        R[p][c] = s;
#pragma unroll
                        for (int i = 0; i < D; ++i) w[i] = fma(-s, Q[i][p], w[i]);
                    }
                    double nn = 0.0;
#pragma unroll
                    for (int i = 0; i < D; ++i) nn = fma(w[i], w[i], nn);
                    const double nrm = sqrt(nn), inv = 1.0 / nrm;
                    R[c][c] = nrm;
#pragma unroll
                    for (int i = 0; i < D; ++i) Q[i][c] = w[i] * inv;
                    // projection of the entering normal (z is already orthogonal to the earlier columns)
                    double s = 0.0;
#pragma unroll
                    for (int i = 0; i < D; ++i) s = fma(Q[i][c], z[i], s);
                    d1[c] = s;
#pragma unroll
                    for (int i = 0; i < D; ++i) z[i] = fma(-s, Q[i][c], z[i]);
                }
            }
            // one re-orthogonalisation pass of z (twice is enough)
#pragma unroll
            for (int c = 0; c < D; ++c)
                if (c < g.q) {
                    double s = 0.0;
#pragma unroll
                    for (int i = 0; i < D; ++i) s = fma(Q[i][c], z[i], s);
                    d1[c] += s;
#pragma unroll
                    for (int i = 0; i < D; ++i) z[i] = fma(-s, Q[i][c], z[i]);
                }
            double ztn = 0.0;
#pragma unroll
            for (int i = 0; i < D; ++i) ztn = fma(z[i], z[i], ztn);
            // r = R^-1 d1
            double rr[D];
#pragma unroll
            for (int i = D - 1; i >= 0; --i) {
                double s = d1[i];
#pragma unroll
                for (int k = i + 1; k < D; ++k)
                    if (k < g.q) s = fma(-R[i][k], rr[k], s);
                rr[i] = (i < g.q) ? s / R[i][i] : 0.0;
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
            // third rule: a group's multiplier sum reaches its weight
            double t3 = kInfD;
            int g3 = -1;
            if constexpr (SOFT) {
                // entries: the entering row (rate +1, multiplier up_mult) and the active rows (rate -rr)
#pragma unroll
                for (int e = -1; e < D; ++e) {
                    const int ge = (e < 0) ? pgrp : ((e < g.q) ? gs.grp[e] : -1);
                    if (ge < 0) continue;
                    bool first = !(e >= 0 && pgrp == ge);
#pragma unroll
                    for (int f = 0; f < D; ++f)
                        if (f < e && f < g.q && gs.grp[f] == ge) first = false;
                    if (!first) continue;
                    double sum = (pgrp == ge) ? up_mult : 0.0, rate = (pgrp == ge) ? 1.0 : 0.0;
#pragma unroll
                    for (int f = 0; f < D; ++f)
                        if (f < g.q && gs.grp[f] == ge) {
                            sum += g.u[f];
                            rate -= rr[f];
                        }
                    const double w = (e < 0) ? pcap : gs.cap[e];
                    if (rate > 0.0) {
                        const double tt = fmax(w - sum, 0.0) / rate;
                        if (tt < t3) {
                            t3 = tt;
                            g3 = ge;
                        }
                    }
                }
            }
            const double t = fmin(fmin(t1, t2), t3);
            if (!(t < kInfD)) return MPCCBF_STATUS_INFEASIBLE;
            if (!z_zero) {
#pragma unroll
                for (int i = 0; i < D; ++i) g.v[i] = fma(t, z[i], g.v[i]);
            }
#pragma unroll
            for (int j = 0; j < D; ++j)
                if (j < g.q) g.u[j] = fma(-t, rr[j], g.u[j]);
            up_mult += t;

            if (!z_zero && t2 <= t1 && t2 <= t3) {
                // full step: the row joins the active set
#pragma unroll
                for (int col = 0; col < D; ++col)
                    if (col == g.q) {
                        g.u[col] = up_mult;
                        g.id[col] = best_id;
                        if constexpr (SOFT) {
                            gs.grp[col] = pgrp;
                            gs.cap[col] = pcap;
                        }
                    }
                ++g.q;
                break;
            }
            if constexpr (SOFT) if (t3 <= t1) {
                // group g3 is at its weight: the leadership passes to the active row of the group with the
                // largest multiplier, or to the entering row when the group has no active row
                int a = -1;
                double ua = -1.0;
#pragma unroll
                for (int c = 0; c < D; ++c)
                    if (c < g.q && gs.grp[c] == g3 && g.u[c] > ua) {
                        a = c;
                        ua = g.u[c];
                    }
                if (a < 0) {  // no active row in the group: it is the entering row's own group
                    rows.set_leader(g3, best_id);  // absorbed by its slack
                    break;
                }
                int aid = -1;
#pragma unroll
                for (int c = 0; c < D; ++c)
                    if (c == a) aid = g.id[c];
                rows.set_leader(g3, aid);
                gi_remove<D, true>(g, gs, a);   // its multiplier is implicit from now on
                // (the group's other active rows and, if it belongs to g3, the entering row are expressed against
                // the new leader from here on: normal() reads the leader)
                if (pgrp == g3) rows.normal(best_id, np, rhs);
                continue;
            }
            // partial (or dual-only) step: active row lpos leaves
            gi_remove<D, SOFT>(g, gs, lpos);
        }
    }
}

// hard rows only
template <int D, class Rows>
__host__ __device__ int gi_run(GIState<D>& g, const Rows& rows, int max_iter, int& n_iter) {
    GISoft<D> gs;
    return gi_run<D, false, Rows>(g, gs, rows, max_iter, n_iter);
}

}  // namespace mpccbf
