// Kernel-side view of the constant tables (see tables.hpp). Passed BY VALUE as a kernel
// parameter; the arrays it points to live in device global memory and are staged into shared
// memory by each CTA.
#pragma once
#include <stdint.h>

namespace mpccbf {

constexpr int kTmplStride = 6;  // {Ct0, Ct1, Ot0, Ot1, inv_norm, pad}

struct DevTables {
    int32_t K, p, c, m1, n;
    int32_t N;          // agents per scenario
    int32_t H;          // cbf_horizon
    int32_t iters;      // impc_iter
    int32_t max_gi;     // active-set iteration cap per QP
    double lim_lo[2][3];  // [0] = acceleration rows, [1] = velocity rows (insertion order of
    double lim_hi[2][3];  //       ConnectivityIMPCCBF.cpp:196-197)
    double dmin2;       // d_min^2
    double gamma;       // class-K gain (5.0)
    double feas_tol;
    int32_t slack;      // slack mode: safety rows of one neighbour share a slack of weight slack_cost * slack_decay^rank
    double slack_cost, slack_decay;
    int32_t conn_mode;  // MPC-lifted connectivity HOCBF / CLF rows (mpccbf_config::conn_mode)
    double dmax;        // d_max: range of the proximity graph
    double Vs[2][2];    // v0 = Vs [pos;vel] + Vt target      (per axis)
    double Vt[2];
    double Kc[2][2];    // c0 = s'Kc s + target (Kt.s)        (per axis)
    double Kt[2];
    // Single-cubic structure of the reducible case: along the horizon the acceleration rows are
    // affine and the velocity rows quadratic in the sample time t_k,
    //   acc(t) = p0 + p1 t,   vel(t) = p2 + p3 t + p4 t^2,   p_i = pq[i][0..1].v + pq[i][2..3].[pos; vel].
    // poly = 1 when the tables verify this to rounding; the row scan then evaluates a row value with
    // one or two FMAs instead of a 4-term dot product (t_k sits in slot 5 of the acc templates).
    int32_t poly;
    double pq[5][4];
    const double* tmpl;  // [3][K][kTmplStride]  d = 0 pos, 1 vel, 2 acc
    const double* yw;    // [m1][4]  {Y1(.,0), Y1(.,1), W1(.,0), W1(.,1)}
    const double* Vr;    // [2][K]   general (per-step) reference
    const double* Kr;    // [2][K]
};

}  // namespace mpccbf
