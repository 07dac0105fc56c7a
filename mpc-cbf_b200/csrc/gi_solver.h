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
#include <type_traits>

#include "../../include/mpccbf.h"

namespace mpccbf {

#ifndef MPCCBF_HD
#define MPCCBF_HD __host__ __device__ __forceinline__
#endif

// Reciprocal and reciprocal square root of the active-set step. On the device a / b compiles to a CALLED division
// routine (~35 instructions and spills around the call); rcp.rn.f64 and rsqrt stay inline. The quotients formed with
// them are within 1 ulp of the correctly rounded ones, which the iteration does not depend on.
#ifdef __CUDA_ARCH__
#define GI_RCP(x) __drcp_rn(x)
#define GI_RSQRT(x) rsqrt(x)
#else
#define GI_RCP(x) (1.0 / (x))
#define GI_RSQRT(x) (1.0 / std::sqrt(x))
#endif

// Divergence guard. The iterate of the dual method is the minimiser over a SUBSET of the rows, so its distance to v0
// never exceeds that of the true minimiser; the bound rows confine every feasible v to a ball of a radius comparable
// to |v0|. An iterate beyond vmax2 = 1e14 (1 + |v0|^2) therefore proves infeasibility — and it is where a step along an
// almost dependent normal (t2 ~ 1e15) would otherwise leave the range in which doubles resolve the rows and cycle.
template <int D>
MPCCBF_HD bool gi_diverged(const double (&v)[D], double vmax2) {
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < D; ++i) s = fma(v[i], v[i], s);
    return !(s <= vmax2);
}
template <int D>
MPCCBF_HD double gi_vmax2(const double* v0) {
    double s = 1.0;
#pragma unroll
    for (int i = 0; i < D; ++i) s = fma(v0[i], v0[i], s);
    return 1e14 * s;
}

template <int D>
struct GIState {
    double u[D];      // multipliers of the active rows
    double v[D];      // current iterate
    int id[D];        // row ids of the active set
    int q;            // number of active rows
    // Optional cache of the active rows' normals, used when the row provider declares kCacheNormals: nc[(c D + i) kNcStride]
    // = component i of active row c (storage of the caller: a lane-strided shared-memory block in the fused kernel).
    // It is rebuilt from the ids by gi_refresh whenever the state was set from (v, u, ids).
    double* nc;
};
constexpr int kNcStride = 32;

template <class R, class = void>
struct gi_caches_normals : std::false_type {};
template <class R>
struct gi_caches_normals<R, std::void_t<decltype(R::kCacheNormals)>> : std::true_type {};

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

// normals of the active rows from their ids (providers that cache them)
template <int D, class Rows>
MPCCBF_HD void gi_refresh(GIState<D>& g, const Rows& rows) {
    if constexpr (gi_caches_normals<Rows>::value) {
#pragma unroll 1
        for (int c = 0; c < g.q; ++c) {
            int id = -1;
#pragma unroll
            for (int k = 0; k < D; ++k)
                if (k == c) id = g.id[k];
            double w[D], unused;
            rows.normal(id, w, unused);
#pragma unroll
            for (int i = 0; i < D; ++i) g.nc[(c * D + i) * kNcStride] = w[i];
        }
    }
}
// the cache follows gi_remove: columns above lpos move down (g.q is already the new count)
template <int D>
MPCCBF_HD void gi_nc_shift(GIState<D>& g, int lpos) {
#pragma unroll
    for (int c = 0; c < D - 1; ++c)
        if (c >= lpos && c < g.q) {
#pragma unroll
            for (int i = 0; i < D; ++i) g.nc[(c * D + i) * kNcStride] = g.nc[((c + 1) * D + i) * kNcStride];
        }
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
__host__ __device__ int gi_run(GIState<D>& g, GISoft<D>& gs, const Rows& rows, int max_iter, int& n_iter, double vmax2) {
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
            double Q[D][D], R[D][D], Rinv[D], d1[D], z[D];
#pragma unroll
            for (int i = 0; i < D; ++i) z[i] = np[i];
            double dall = 0.0;
#pragma unroll
            for (int i = 0; i < D; ++i) dall = fma(np[i], np[i], dall);
#pragma unroll
            for (int c = 0; c < D; ++c) {
                d1[c] = 0.0;
                if (c < g.q) {
                    double w[D];
                    if constexpr (gi_caches_normals<Rows>::value) {
#pragma unroll
                        for (int i = 0; i < D; ++i) w[i] = g.nc[(c * D + i) * kNcStride];
                    } else {
                        double unused;
                        rows.normal(g.id[c], w, unused);
                    }
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
                    const double inv = GI_RSQRT(nn);   // nn = 0 (dependent normal): inf, as 1 / sqrt(0) before
                    Rinv[c] = inv;
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
                rr[i] = (i < g.q) ? s * Rinv[i] : 0.0;
            }
            const double zt = 1e4 * 2.220446049250313e-16;
            const bool z_zero = !(ztn > zt * zt * dall);
            // ratio test min_j u_j / rr_j over rr_j > 0 by cross-multiplication (u_j >= 0): one reciprocal at the end
            double t1 = kInfD, un = 0.0, rd = 0.0;
            int lpos = -1;
#pragma unroll
            for (int j = 0; j < D; ++j)
                if (j < g.q && rr[j] > 0.0) {
                    if (lpos < 0 || g.u[j] * rd < un * rr[j]) {
                        un = g.u[j];
                        rd = rr[j];
                        lpos = j;
                    }
                }
            if (lpos >= 0) t1 = un * GI_RCP(rd);
            double sp = -rhs;
#pragma unroll
            for (int i = 0; i < D; ++i) sp = fma(np[i], g.v[i], sp);
            const double t2 = z_zero ? kInfD : (-sp * GI_RCP(ztn));
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
            if (gi_diverged<D>(g.v, vmax2)) return MPCCBF_STATUS_INFEASIBLE;

            if (!z_zero && t2 <= t1 && t2 <= t3) {
                // full step: the row joins the active set
#pragma unroll
                for (int col = 0; col < D; ++col)
                    if (col == g.q) {
                        g.u[col] = up_mult;
                        g.id[col] = best_id;
                        if constexpr (gi_caches_normals<Rows>::value) {
#pragma unroll
                            for (int i = 0; i < D; ++i) g.nc[(col * D + i) * kNcStride] = np[i];
                        }
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
                gi_refresh<D>(g, rows);         // the group's other active rows are expressed against the new leader
                // (the group's other active rows and, if it belongs to g3, the entering row are expressed against
                // the new leader from here on: normal() reads the leader)
                if (pgrp == g3) rows.normal(best_id, np, rhs);
                continue;
            }
            // partial (or dual-only) step: active row lpos leaves
            gi_remove<D, SOFT>(g, gs, lpos);
            if constexpr (gi_caches_normals<Rows>::value) gi_nc_shift<D>(g, lpos);
        }
    }
}

// hard rows only
template <int D, class Rows>
__host__ __device__ int gi_run(GIState<D>& g, const Rows& rows, int max_iter, int& n_iter, double vmax2 = HUGE_VAL) {
    GISoft<D> gs;
    return gi_run<D, false, Rows>(g, gs, rows, max_iter, n_iter, vmax2);
}

// The fused kernel reaches the solver through GICore<D, CORE>. CORE = 0 is the re-factorising iteration above. A second
// core that kept a persistent orthogonal / triangular factorisation updated by Givens rotations (round 1's formulation)
// was measured 1.7x slower (profiles/r2_kernel_ab.md) and has been removed; the parameter stays for experiments.
template <int D, int CORE>
struct GICore;
template <int D>
struct GICore<D, 0> {
    typedef GIState<D> State;
    MPCCBF_HD static void init(State& g, const double* v0) { gi_init<D>(g, v0); }
    template <bool SOFT, class Rows>
    MPCCBF_HD static int run(State& g, GISoft<D>& gs, const Rows& rows, int max_iter, int& n_iter, double vmax2) {
        return gi_run<D, SOFT, Rows>(g, gs, rows, max_iter, n_iter, vmax2);
    }
    template <class Rows>
    MPCCBF_HD static void refactor(State& g, const Rows& rows) { gi_refresh<D>(g, rows); }
};

}  // namespace mpccbf
