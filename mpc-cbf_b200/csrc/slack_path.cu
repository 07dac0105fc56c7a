// Slack-mode controller path — sm_100a.
//
// ConnectivityIMPCCBF::optimize with cbf_params.slack_mode = true (ConnectivityIMPCCBF.cpp:76-100): one
// slack variable s_j >= 0 per neighbour (MPCCBFQPGeneratorBase.cpp:28-38) enters every safety row of
// that neighbour with coefficient -1 (ConnectivityMPCCBFQPGenerator.cpp:33-44) and the objective with
// the linear cost slack_cost * slack_decay^rank_j, rank by distance to the ego agent (closest first;
// MPCCBFQPGeneratorBase.cpp:121-135). The Hessian is then only PSD, so the strictly convex active-set
// kernels do not apply; instead each QP is assembled on the device in the reduced coordinates of
// tables.hpp (6 whitened free dof v + N-1 slacks, no equality rows)
//      min 1/2 |v - v0|^2 + sum_j w_j s_j
//      s.t. -(a_jk . acc_k(v)) - s_j <= b_jk,   lo <= Ct[d][k] v + Ot[d][k] s0 <= hi,   s >= 0
// and handed to the one-CTA-per-QP ADMM kernel (dense_qp.cu) through its device-pointer entry. Per
// IMPC iteration: slack_assemble_kernel -> dense_qp_kernel -> slack_finish_kernel (control points,
// predicted ego states for the next iteration, status bookkeeping of ConnectivityIMPCCBF.cpp:203-211).
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <string>

#include "../../include/mpccbf.h"
#include "kernels.h"

namespace mpccbf {

namespace {

constexpr int kTA = 128;
constexpr double kFree = 1e30;  // the dense kernel reads |bound| >= 1e20 as "no bound"

struct SlackLaunch {
    const double* states;
    const double* targets;
    int ref_mode;
    int g0, count;  // entries [g0, g0 + count) of `list` (or QPs g0.. themselves when list is NULL)
    const int32_t* list;
    int agent_begin, agent_count;
    int it, hit, nv, m, J;
    double slack_cost, slack_decay;
    double *P, *q, *A, *l, *u, *c0;
    const double* pred;  // [count][H][6] predicted ego states (it >= 1)
    const int32_t* nok;  // [count] iterates that were OPTIMAL so far
};

// closed form of ConnectivityCBF::initSafetyCBF (cbf/src/detail/ConnectivityCBF.cpp:152-198):
// h = |dp|^2 - dmin^2, alpha(s) = gamma s^3; row a.u + b >= 0 with a = (2dx, 2dy, 0)
__device__ __forceinline__ void safety_hocbf(double px, double py, double vx, double vy, double qx, double qy, double wx,
                                             double wy, double dmin2, double gamma, double& ax, double& ay, double& b) {
    const double dx = px - qx, dy = py - qy;
    const double dvx = vx - wx, dvy = vy - wy;
    const double h = dx * dx + dy * dy - dmin2;
    const double lfh = 2.0 * (dx * dvx + dy * dvy);
    const double lf2h = 2.0 * (dvx * dvx + dvy * dvy);
    const double lfa = 3.0 * gamma * h * h * (2.0 * dx * vx + 2.0 * dy * vy);
    const double psi1 = lfh + gamma * h * h * h;
    ax = 2.0 * dx;
    ay = 2.0 * dy;
    b = lf2h + lfa + gamma * psi1 * psi1 * psi1;
}

// slack_cost * slack_decay^rank, rank of neighbour j among the J neighbours by distance to the ego agent
// (closest first, ties by index; ConnectivityIMPCCBF.cpp:76-100)
__device__ double slack_weight(const double* st, int self, int j, int J, double cost, double decay) {
    const double* me = st + self * 6;
    const int other = j + (j >= self ? 1 : 0);
    // near-ties decide ranks in symmetric formations: the same double arithmetic as the host code of the
    // reference (separately rounded products and sum, no fma), exact ties by index
    const double dx = st[other * 6] - me[0], dy = st[other * 6 + 1] - me[1];
    const double dj = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
    int rank = 0;
    for (int i = 0; i < J; ++i) {
        const int oi = i + (i >= self ? 1 : 0);
        const double ex = st[oi * 6] - me[0], ey = st[oi * 6 + 1] - me[1];
        const double di = sqrt(__dadd_rn(__dmul_rn(ex, ex), __dmul_rn(ey, ey)));
        rank += (di < dj || (di == dj && i < j)) ? 1 : 0;
    }
    return cost * pow(decay, static_cast<double>(rank));
}

// One CTA per QP. Row order: safety rows (neighbour-major, horizon step inside), acceleration bound
// rows [k][axis], velocity bound rows [k][axis], slack >= 0 rows.
__global__ void __launch_bounds__(kTA) slack_assemble_kernel(const DevTables T, const SlackLaunch a) {
    extern __shared__ double st[];  // [N][6] states of the scenario
    const int K = T.K, N = T.N, J = a.J, nv = a.nv, m = a.m, hit = a.hit, t = threadIdx.x;
    const double* tm = T.tmpl;
    for (int b = blockIdx.x; b < a.count; b += gridDim.x) {
        const int g = a.list ? a.list[a.g0 + b] : a.g0 + b;
        const int scen = g / a.agent_count;
        const int self = a.agent_begin + (g - scen * a.agent_count);
        __syncthreads();
        for (int i = t; i < N * 6; i += kTA) st[i] = __ldg(a.states + static_cast<size_t>(scen) * N * 6 + i);
        double* P = a.P + static_cast<size_t>(b) * nv * nv;
        double* A = a.A + static_cast<size_t>(b) * m * nv;
        double* q = a.q + static_cast<size_t>(b) * nv;
        double* l = a.l + static_cast<size_t>(b) * m;
        double* u = a.u + static_cast<size_t>(b) * m;
        for (int e = t; e < nv * nv; e += kTA) P[e] = 0.0;
        for (int e = t; e < m * nv; e += kTA) A[e] = 0.0;
        __syncthreads();
        const bool alive = a.nok[b] == a.it;  // an earlier iterate failed: trivial QP, result ignored
        const double* me = st + self * 6;
        // ---- objective: x'Px + q'x with P = I/2 on v, q = (-v0 | w) ---------------------------------
        if (t < 3) {
            const int ax = t;
            const double ps = me[ax], vs = me[3 + ax];
            double r0, r1, kk;
            if (a.ref_mode == 0) {
                const double tgt = __ldg(a.targets + static_cast<size_t>(g) * 3 + ax);
                r0 = T.Vt[0] * tgt;
                r1 = T.Vt[1] * tgt;
                kk = tgt * (T.Kt[0] * ps + T.Kt[1] * vs);
            } else {
                const double* rf = a.targets + static_cast<size_t>(g) * 3 * K;
                double k0 = 0.0, k1 = 0.0;
                r0 = 0.0;
                r1 = 0.0;
                for (int k = 0; k < K; ++k) {
                    const double rk = __ldg(rf + 3 * k + ax);
                    r0 = fma(T.Vr[k], rk, r0);
                    r1 = fma(T.Vr[K + k], rk, r1);
                    k0 = fma(T.Kr[k], rk, k0);
                    k1 = fma(T.Kr[K + k], rk, k1);
                }
                kk = k0 * ps + k1 * vs;
            }
            q[2 * ax] = -(T.Vs[0][0] * ps + T.Vs[0][1] * vs + r0);
            q[2 * ax + 1] = -(T.Vs[1][0] * ps + T.Vs[1][1] * vs + r1);
            P[(2 * ax) * nv + 2 * ax] = 0.5;
            P[(2 * ax + 1) * nv + 2 * ax + 1] = 0.5;
            // constant part of the reference objective, summed over the axes by the finish kernel
            a.c0[static_cast<size_t>(b) * 3 + ax] =
                ps * (T.Kc[0][0] * ps + T.Kc[0][1] * vs) + vs * (T.Kc[1][0] * ps + T.Kc[1][1] * vs) + kk;
        }
        // ---- slack weights by distance rank (ConnectivityIMPCCBF.cpp:76-100) and s_j >= 0 rows -------
        for (int j = t; j < J; j += kTA) {
            q[6 + j] = slack_weight(st, self, j, J, a.slack_cost, a.slack_decay);
            const int r = J * hit + 6 * K + j;
            A[static_cast<size_t>(r) * nv + 6 + j] = 1.0;
            l[r] = 0.0;
            u[r] = kFree;
        }
        // ---- safety rows -----------------------------------------------------------------------------
        for (int r = t; r < J * hit; r += kTA) {
            const int nb = r / hit, k = r - nb * hit;
            const int other = nb + (nb >= self ? 1 : 0);
            if (!alive) {
                l[r] = -kFree;
                u[r] = kFree;
                continue;
            }
            double px, py, vx, vy;
            if (a.it == 0) {
                px = me[0]; py = me[1]; vx = me[3]; vy = me[4];
            } else {
                const double* ps = a.pred + (static_cast<size_t>(b) * T.H + k) * 6;
                px = ps[0]; py = ps[1]; vx = ps[3]; vy = ps[4];
            }
            const double* o = st + other * 6;
            double ax, ay, bb;
            safety_hocbf(px, py, vx, vy, o[0], o[1], o[3], o[4], T.dmin2, T.gamma, ax, ay, bb);
            const double* tk = tm + (2 * K + k) * kTmplStride;  // acceleration sample k
            double* row = A + static_cast<size_t>(r) * nv;
            row[0] = -ax * tk[0];
            row[1] = -ax * tk[1];
            row[2] = -ay * tk[0];
            row[3] = -ay * tk[1];
            row[6 + nb] = -1.0;
            l[r] = -kFree;
            u[r] = bb + ax * (tk[2] * me[0] + tk[3] * me[3]) + ay * (tk[2] * me[1] + tk[3] * me[4]);
        }
        // ---- bound rows (ConnectivityIMPCCBF.cpp:196-197: acceleration block, then velocity block) ---
        for (int e = t; e < 6 * K; e += kTA) {
            const int dsel = e >= 3 * K ? 1 : 0;
            const int rr = e - dsel * 3 * K;
            const int k = rr / 3, ax = rr - 3 * k;
            const int r = J * hit + e;
            const double* tk = tm + ((2 - dsel) * K + k) * kTmplStride;
            if (!alive) {
                l[r] = -kFree;
                u[r] = kFree;
                continue;
            }
            double* row = A + static_cast<size_t>(r) * nv;
            row[2 * ax] = tk[0];
            row[2 * ax + 1] = tk[1];
            const double off = tk[2] * me[ax] + tk[3] * me[3 + ax];
            l[r] = T.lim_lo[dsel][ax] - off;
            u[r] = T.lim_hi[dsel][ax] - off;
        }
    }
}

struct SlackFinish {
    const double* states;
    const int32_t* list;
    int g0, count, agent_begin, agent_count, it, nv;
    int m, hit, J;
    const double *x, *c0, *A, *u, *q;  // the assembled QP of this iteration (rows: see slack_assemble_kernel)
    const int32_t *dstat, *dit;
    double *pred, *vlast, *objlast;
    int32_t* nok;
    // StepArgs outputs (indexed by the global QP index g)
    double *ctrl_pts, *ctrl_all, *objective;
    int32_t *status, *n_curves, *gi_iters;
};

__device__ __forceinline__ double pick3(int ax, double a, double b, double c) { return ax == 0 ? a : ax == 1 ? b : c; }

// One warp per QP: status bookkeeping (first non-OPTIMAL iterate stops the agent, later ones are
// SKIPPED), control points x_axis = Y1 [pos; vel] + W1 v, predicted ego states at t_k, k < H.
__global__ void __launch_bounds__(kTA) slack_finish_kernel(const DevTables T, const SlackFinish a) {
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * kTA) >> 5;
    for (int b = (blockIdx.x * kTA + threadIdx.x) >> 5; b < a.count; b += warps) {
        const int g = a.list ? a.list[a.g0 + b] : a.g0 + b;
        const int scen = g / a.agent_count;
        const int self = a.agent_begin + (g - scen * a.agent_count);
        const double* me = a.states + (static_cast<size_t>(scen) * T.N + self) * 6;
        const double s0x = __ldg(me), s0y = __ldg(me + 1), s0w = __ldg(me + 2);
        const double s1x = __ldg(me + 3), s1y = __ldg(me + 4), s1w = __ldg(me + 5);
        const int nok_in = a.nok[b];
        const bool alive = nok_in == a.it;
        const int stat = alive ? a.dstat[b] : MPCCBF_STATUS_SKIPPED;
        const bool ok = alive && stat == MPCCBF_STATUS_OPTIMAL;
        double v[6];
        if (ok) {
#pragma unroll
            for (int i = 0; i < 6; ++i) v[i] = a.x[static_cast<size_t>(b) * a.nv + i];
        } else {
#pragma unroll
            for (int i = 0; i < 6; ++i) v[i] = a.vlast[static_cast<size_t>(b) * 6 + i];
        }
        __syncwarp();
        const int nok = nok_in + (ok ? 1 : 0);
        // objective of the reference's QP at the returned v with the slacks at their optimal values
        // s_j = max(0, max_k violation of neighbour j's rows) (exact for w_j > 0; the ADMM iterate's own
        // slacks carry its tolerance times weights of up to slack_cost)
        double objv = 0.0;
        if (ok) {
            const double* A = a.A + static_cast<size_t>(b) * a.m * a.nv;
            const double* u = a.u + static_cast<size_t>(b) * a.m;
            const double* q = a.q + static_cast<size_t>(b) * a.nv;
            double pen = 0.0;
            for (int j = lane; j < a.J; j += 32) {
                double sj = 0.0;
                for (int k = 0; k < a.hit; ++k) {
                    const int r = j * a.hit + k;
                    const double* row = A + static_cast<size_t>(r) * a.nv;
                    sj = fmax(sj, row[0] * v[0] + row[1] * v[1] + row[2] * v[2] + row[3] * v[3] - u[r]);
                }
                pen = fma(q[6 + j], sj, pen);
            }
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) pen += __shfl_xor_sync(0xffffffffu, pen, o);
            double dd = 0.0, nn = 0.0;
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const double e = v[i] + q[i];  // q = -v0
                dd = fma(e, e, dd);
                nn = fma(q[i], q[i], nn);
            }
            objv = 0.5 * dd - 0.5 * nn + pen +
                   (a.c0[static_cast<size_t>(b) * 3] + a.c0[static_cast<size_t>(b) * 3 + 1] + a.c0[static_cast<size_t>(b) * 3 + 2]);
        }
        if (lane == 0) {
            a.status[static_cast<size_t>(g) * T.iters + a.it] = stat;
            if (a.gi_iters) a.gi_iters[static_cast<size_t>(g) * T.iters + a.it] = alive ? a.dit[b] : 0;
            if (ok) {
                a.nok[b] = nok;
#pragma unroll
                for (int i = 0; i < 6; ++i) a.vlast[static_cast<size_t>(b) * 6 + i] = v[i];
                a.objlast[b] = objv;
            }
        }
        const bool last = a.it + 1 == T.iters;
        if ((ok && a.ctrl_all) || last) {
            for (int i = lane; i < T.n; i += 32) {
                const int piece = i / (3 * T.c), rem = i - piece * 3 * T.c;
                const int ax = rem / T.c, cp = rem - ax * T.c;
                const double* y = T.yw + (piece * T.c + cp) * 4;
                const double val = y[0] * pick3(ax, s0x, s0y, s0w) + y[1] * pick3(ax, s1x, s1y, s1w) +
                                   y[2] * pick3(ax, v[0], v[2], v[4]) + y[3] * pick3(ax, v[1], v[3], v[5]);
                if (ok && a.ctrl_all) a.ctrl_all[(static_cast<size_t>(g) * T.iters + a.it) * T.n + i] = val;
                if (last) a.ctrl_pts[static_cast<size_t>(g) * T.n + i] = nok > 0 ? val : 0.0;
            }
        }
        if (ok && !last) {  // ConnectivityIMPCCBF.cpp:160-168
            for (int e = lane; e < T.H * 6; e += 32) {
                const int k = e / 6, comp = e - 6 * k;
                const int d = comp / 3, ax = comp - 3 * d;
                const double* tk = T.tmpl + (d * T.K + k) * kTmplStride;
                a.pred[static_cast<size_t>(b) * T.H * 6 + e] =
                    tk[0] * pick3(ax, v[0], v[2], v[4]) + tk[1] * pick3(ax, v[1], v[3], v[5]) +
                    tk[2] * pick3(ax, s0x, s0y, s0w) + tk[3] * pick3(ax, s1x, s1y, s1w);
            }
        }
        if (last && lane == 0) {
            if (a.n_curves) a.n_curves[g] = nok;
            if (a.objective) a.objective[g] = nok > 0 ? a.objlast[b] : 0.0;  // written by this lane above
        }
    }
}

// ---- first tier: is the hard-row optimum already the slack-mode optimum? --------------------------------
// The slack QP min 1/2|v - v0|^2 + sum_j w_j s_j has the same minimiser as the hard-row QP (with s = 0) iff
// the hard optimum's multipliers on neighbour j's safety rows sum to at most w_j (KKT of the slack QP with
// mu_j = w_j - sum_k lambda_jk >= 0 on s_j >= 0). One thread per agent re-derives, per IMPC iterate of the
// fused hard-row kernels (their whitened iterates come through StepArgs::v_all), the active x/y rows, solves the <= 4-column least-squares
// system v0 - v = sum_i lambda_i n_i, and accepts the agent when stationarity holds to 1e-7, lambda >= 0 and
// every cap holds with margin. Everything else (infeasible hard QP, a cap exceeded, degenerate or more
// than 4 active rows) goes to the ADMM tier through `list`.
struct SlackVerify {
    const double* states;
    const double* targets;
    int ref_mode;
    int num_qp, agent_begin, agent_count;
    double slack_cost, slack_decay;
    const double* v_all;     // [B][iters][6] iterates of the hard-row solve in whitened coordinates
    const int32_t* status;   // [B][iters]
    int32_t* list;
    int32_t* count;
};
constexpr int kVT = 128;  // threads per CTA, one agent each

__global__ void __launch_bounds__(kVT) slack_verify_kernel(const DevTables T, const SlackVerify a) {
    extern __shared__ double tm[];  // [3][K][kTmplStride] row templates
    const int N = T.N, K = T.K, J = N - 1;
    for (int i = threadIdx.x; i < 3 * K * kTmplStride; i += kVT) tm[i] = T.tmpl[i];
    __syncthreads();
    const double tol = 1e-8;
    for (int g = blockIdx.x * kVT + threadIdx.x; g < a.num_qp; g += gridDim.x * kVT) {
        const int scen = g / a.agent_count;
        const int self = a.agent_begin + (g - scen * a.agent_count);
        const double* st = a.states + static_cast<size_t>(scen) * N * 6;  // neighbours' states: L1/L2 hits
        double me[6];
#pragma unroll
        for (int c = 0; c < 6; ++c) me[c] = __ldg(st + self * 6 + c);
        bool pass = true;
        double vprev[6] = {0, 0, 0, 0, 0, 0};
        for (int it = 0; it < T.iters && pass; ++it) {
            if (a.status[static_cast<size_t>(g) * T.iters + it] != MPCCBF_STATUS_OPTIMAL) {
                pass = false;
                break;
            }
            // whitened free dof of this iterate, and v0 - v on the x / y axes
            double v[6], r4[4];
#pragma unroll
            for (int c = 0; c < 6; ++c) v[c] = a.v_all[(static_cast<size_t>(g) * T.iters + it) * 6 + c];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int ax = c >> 1, r = c & 1;
                double rr = 0.0;
                if (a.ref_mode == 0) {
                    rr = T.Vt[r] * __ldg(a.targets + static_cast<size_t>(g) * 3 + ax);
                } else {
                    const double* rf = a.targets + static_cast<size_t>(g) * 3 * K;
                    for (int k = 0; k < K; ++k) rr = fma(T.Vr[r * K + k], __ldg(rf + 3 * k + ax), rr);
                }
                r4[c] = T.Vs[r][0] * me[ax] + T.Vs[r][1] * me[3 + ax] + rr - v[c];
            }
            // ---- active x / y rows at v ------------------------------------------------------------------
            double nrm[4][4];
            int nbi[4] = {-1, -1, -1, -1};
            int cnt = 0;
            for (int e = 0; e < 4 * K; ++e) {  // bound rows: acceleration block, then velocity block
                const int dsel = e >= 2 * K ? 1 : 0;
                const int rr = e - dsel * 2 * K;
                const int k = rr >> 1, ax = rr & 1;
                const double* tk = tm + ((2 - dsel) * K + k) * kTmplStride;
                if (tk[0] == 0.0 && tk[1] == 0.0) continue;
                const double val = tk[0] * (ax ? v[2] : v[0]) + tk[1] * (ax ? v[3] : v[1]) + tk[2] * me[ax] + tk[3] * me[3 + ax];
                const double lo = T.lim_lo[dsel][ax], hi = T.lim_hi[dsel][ax];
                double sgn = 0.0;
                if (hi < 1e20 && fabs(val - hi) <= tol * (1.0 + fabs(hi))) sgn = 1.0;
                else if (lo > -1e20 && fabs(val - lo) <= tol * (1.0 + fabs(lo))) sgn = -1.0;
                if (sgn != 0.0) {
                    if (cnt < 4) {
                        nrm[cnt][0] = ax ? 0.0 : sgn * tk[0];
                        nrm[cnt][1] = ax ? 0.0 : sgn * tk[1];
                        nrm[cnt][2] = ax ? sgn * tk[0] : 0.0;
                        nrm[cnt][3] = ax ? sgn * tk[1] : 0.0;
                        nbi[cnt] = -1;
                    }
                    ++cnt;
                }
            }
            const int hit = it == 0 ? 1 : T.H;
            for (int k = 0; k < hit; ++k) {  // safety rows of horizon step k
                double px = me[0], py = me[1], vx = me[3], vy = me[4];
                if (it > 0) {  // predicted ego state at t_k from the previous iterate
                    const double* t0 = tm + (0 * K + k) * kTmplStride;
                    const double* t1 = tm + (1 * K + k) * kTmplStride;
                    px = t0[0] * vprev[0] + t0[1] * vprev[1] + t0[2] * me[0] + t0[3] * me[3];
                    py = t0[0] * vprev[2] + t0[1] * vprev[3] + t0[2] * me[1] + t0[3] * me[4];
                    vx = t1[0] * vprev[0] + t1[1] * vprev[1] + t1[2] * me[0] + t1[3] * me[3];
                    vy = t1[0] * vprev[2] + t1[1] * vprev[3] + t1[2] * me[1] + t1[3] * me[4];
                }
                const double* tk = tm + (2 * K + k) * kTmplStride;
                const double ux = tk[0] * v[0] + tk[1] * v[1] + tk[2] * me[0] + tk[3] * me[3];  // acceleration at t_k
                const double uy = tk[0] * v[2] + tk[1] * v[3] + tk[2] * me[1] + tk[3] * me[4];
                for (int nb = 0; nb < J; ++nb) {
                    const double* o = st + (nb + (nb >= self ? 1 : 0)) * 6;
                    double ax, ay, bb;
                    safety_hocbf(px, py, vx, vy, __ldg(o), __ldg(o + 1), __ldg(o + 3), __ldg(o + 4), T.dmin2, T.gamma, ax, ay, bb);
                    const double slack = ax * ux + ay * uy + bb;  // row: a.u + b >= 0
                    if (fabs(slack) <= tol * (1.0 + fabs(bb))) {
                        if (cnt < 4) {
                            nrm[cnt][0] = -ax * tk[0];
                            nrm[cnt][1] = -ax * tk[1];
                            nrm[cnt][2] = -ay * tk[0];
                            nrm[cnt][3] = -ay * tk[1];
                            nbi[cnt] = nb;
                        }
                        ++cnt;
                    }
                }
            }
            if (cnt > 4) {
                pass = false;
                break;
            }
            // ---- multipliers: least squares on the active normals ----------------------------------------
            double G[4][4], rhs[4], lam[4] = {0, 0, 0, 0};
            for (int i = 0; i < cnt; ++i) {
                rhs[i] = nrm[i][0] * r4[0] + nrm[i][1] * r4[1] + nrm[i][2] * r4[2] + nrm[i][3] * r4[3];
                for (int j = 0; j <= i; ++j)
                    G[i][j] = nrm[i][0] * nrm[j][0] + nrm[i][1] * nrm[j][1] + nrm[i][2] * nrm[j][2] + nrm[i][3] * nrm[j][3];
            }
            for (int j = 0; j < cnt && pass; ++j) {  // Cholesky G = L L'
                double d = G[j][j];
                for (int k = 0; k < j; ++k) d -= G[j][k] * G[j][k];
                if (!(d > 1e-10 * G[j][j])) {
                    pass = false;  // dependent active normals: multipliers not unique
                    break;
                }
                G[j][j] = sqrt(d);
                for (int i = j + 1; i < cnt; ++i) {
                    double s2 = G[i][j];
                    for (int k = 0; k < j; ++k) s2 -= G[i][k] * G[j][k];
                    G[i][j] = s2 / G[j][j];
                }
            }
            if (!pass) break;
            for (int i = 0; i < cnt; ++i) {
                double s2 = rhs[i];
                for (int k = 0; k < i; ++k) s2 -= G[i][k] * lam[k];
                lam[i] = s2 / G[i][i];
            }
            for (int i = cnt - 1; i >= 0; --i) {
                double s2 = lam[i];
                for (int k = i + 1; k < cnt; ++k) s2 -= G[k][i] * lam[k];
                lam[i] = s2 / G[i][i];
            }
            double rmax = 0.0, res = 0.0, lmax = 0.0;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                double e = r4[c];
                for (int i = 0; i < cnt; ++i) e -= lam[i] * nrm[i][c];
                res = fmax(res, fabs(e));
                rmax = fmax(rmax, fabs(r4[c]));
            }
            for (int i = 0; i < cnt; ++i) lmax = fmax(lmax, fabs(lam[i]));
            if (!(res <= 1e-7 * (1.0 + rmax))) pass = false;
            for (int i = 0; i < cnt && pass; ++i) {
                if (lam[i] < -1e-7 * (1.0 + lmax)) pass = false;
                if (nbi[i] < 0) continue;
                bool first = true;
                double sum = 0.0;
                for (int j = 0; j < cnt; ++j)
                    if (nbi[j] == nbi[i]) {
                        if (j < i) first = false;
                        sum += fmax(lam[j], 0.0);
                    }
                if (first && !(sum <= (1.0 - 1e-6) * slack_weight(st, self, nbi[i], J, a.slack_cost, a.slack_decay)))
                    pass = false;
            }
#pragma unroll
            for (int c = 0; c < 6; ++c) vprev[c] = v[c];
        }
        if (!pass) a.list[atomicAdd(a.count, 1)] = g;
    }
}

__global__ void slack_reset_kernel(int32_t* nok, double* vlast, int count) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) nok[i] = 0;
    if (i < count * 6) vlast[i] = 0.0;
}

template <typename P>
bool grow(P** ptr, size_t elems) {
    if (*ptr) cudaFree(*ptr);
    *ptr = nullptr;
    return cudaMalloc(reinterpret_cast<void**>(ptr), sizeof(P) * std::max<size_t>(elems, 1)) == cudaSuccess;
}

}  // namespace

static void free_admm(SlackWorkspace& w) {  // the ADMM tier's buffers only (they are re-sized on demand)
    void* ptrs[] = {w.P, w.q, w.A, w.l, w.u, w.wsP, w.wsA, w.x, w.obj, w.pred, w.vlast, w.objlast, w.c0, w.stat, w.it, w.nok};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    w.P = w.q = w.A = w.l = w.u = w.wsP = w.wsA = w.x = w.obj = w.pred = w.vlast = w.objlast = w.c0 = nullptr;
    w.stat = w.it = w.nok = nullptr;
    w.cap_qps = 0;
}

int slack_prepare_hard(const DevTables& T, int num_qp, SlackWorkspace& w) {
    if (w.hard_cap >= static_cast<size_t>(num_qp)) return MPCCBF_OK;
    bool ok = grow(&w.hard_all, static_cast<size_t>(num_qp) * T.iters * 6) && grow(&w.list, static_cast<size_t>(num_qp));
    if (!w.count) ok = ok && grow(&w.count, 1);
    if (!w.h_count) ok = ok && cudaMallocHost(reinterpret_cast<void**>(&w.h_count), sizeof(int32_t)) == cudaSuccess;
    if (!ok) {
        set_error("slack mode: device workspace allocation failed");
        return MPCCBF_ERR_ALLOC;
    }
    w.hard_cap = num_qp;
    return MPCCBF_OK;
}

int slack_verify(const DevTables& T, const mpccbf_config& cfg, const StepArgs& a, SlackWorkspace& w, int sm_count,
                 cudaStream_t stream, int* n_fallback) {
    SlackVerify V;
    V.states = a.states; V.targets = a.targets; V.ref_mode = a.ref_mode;
    V.num_qp = a.num_qp; V.agent_begin = a.agent_begin; V.agent_count = a.agent_count;
    V.slack_cost = cfg.slack_cost; V.slack_decay = cfg.slack_decay;
    V.v_all = a.v_all; V.status = a.status; V.list = w.list; V.count = w.count;
    const size_t smem = sizeof(double) * 3 * T.K * kTmplStride;
    cudaError_t e = cudaMemsetAsync(w.count, 0, sizeof(int32_t), stream);
    if (e == cudaSuccess) {
        const int grid = std::min((a.num_qp + kVT - 1) / kVT, 8 * std::max(1, sm_count));
        slack_verify_kernel<<<grid, kVT, smem, stream>>>(T, V);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(w.h_count, w.count, sizeof(int32_t), cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) {
        set_error(std::string("slack verification: ") + cudaGetErrorString(e));
        return MPCCBF_ERR_CUDA;
    }
    *n_fallback = *w.h_count;
    return MPCCBF_OK;
}

void slack_free(SlackWorkspace& w) {
    void* ptrs[] = {w.P, w.q, w.A, w.l, w.u, w.wsP, w.wsA, w.x, w.obj, w.pred, w.vlast, w.objlast, w.c0, w.stat, w.it, w.nok,
                    w.hard_all, w.list, w.count};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    if (w.h_count) cudaFreeHost(w.h_count);
    w = SlackWorkspace();
}

// One control step of the slack-mode controller for the QPs of `a`, in chunks that fit the workspace.
int slack_step(const DevTables& T, const mpccbf_config& cfg, const StepArgs& a, const DenseOpts& o, SlackWorkspace& w,
               const int32_t* list, int list_count, int64_t* launches) {
    const int total_qp = list ? list_count : a.num_qp;
    const int N = T.N, J = N - 1, H = std::max(1, static_cast<int>(T.H)), K = T.K, iters = T.iters;
    const int nv = 6 + J;
    const int m_max = J * H + 6 * K + J;
    if (nv > 128) {
        set_error("slack mode: more than 122 neighbours exceed the 128 variables of the one-CTA-per-QP kernel");
        return MPCCBF_ERR_UNSUPPORTED;
    }
    const bool need_wsA = dense_qp_needs_wsA(nv, m_max);
    const size_t per_qp = sizeof(double) * (2 * static_cast<size_t>(nv) * nv + static_cast<size_t>(m_max) * nv +
                                            (need_wsA ? static_cast<size_t>(m_max) * (nv | 1) : 0) + 2 * nv + 2 * m_max +
                                            H * 6 + 12) + 12;
    size_t cap_mb = 1536;
    if (const char* env = std::getenv("MPCCBF_SLACK_WS_MB")) cap_mb = std::max(1, std::atoi(env));
    const size_t want = std::min<size_t>(total_qp, std::max<size_t>(1, (cap_mb << 20) / per_qp));
    if (w.cap_qps < want || w.nv != nv || w.m_max != m_max) {
        free_admm(w);
        bool ok = true;
        ok &= grow(&w.P, want * nv * nv) && grow(&w.wsP, want * nv * nv) && grow(&w.A, want * m_max * nv);
        ok &= grow(&w.wsA, need_wsA ? want * m_max * (nv | 1) : 1);
        ok &= grow(&w.q, want * nv) && grow(&w.l, want * m_max) && grow(&w.u, want * m_max) && grow(&w.x, want * nv);
        ok &= grow(&w.obj, want) && grow(&w.pred, want * H * 6) && grow(&w.vlast, want * 6) && grow(&w.objlast, want);
        ok &= grow(&w.c0, want * 3) && grow(&w.stat, want) && grow(&w.it, want) && grow(&w.nok, want);
        if (!ok) {
            free_admm(w);
            set_error("slack mode: device workspace allocation failed (lower MPCCBF_SLACK_WS_MB)");
            return MPCCBF_ERR_ALLOC;
        }
        w.cap_qps = want;
        w.nv = nv;
        w.m_max = m_max;
    }
    cudaStream_t st = o.stream;
    for (int g0 = 0; g0 < total_qp; g0 += static_cast<int>(w.cap_qps)) {
        const int count = std::min<int>(static_cast<int>(w.cap_qps), total_qp - g0);
        slack_reset_kernel<<<(count * 6 + 255) / 256, 256, 0, st>>>(w.nok, w.vlast, count);
        for (int it = 0; it < iters; ++it) {
            const int hit = it == 0 ? 1 : H;
            const int m = J * hit + 6 * K + J;
            SlackLaunch L;
            L.states = a.states; L.targets = a.targets; L.ref_mode = a.ref_mode;
            L.g0 = g0; L.count = count; L.list = list; L.agent_begin = a.agent_begin; L.agent_count = a.agent_count;
            L.it = it; L.hit = hit; L.nv = nv; L.m = m; L.J = J;
            L.slack_cost = cfg.slack_cost; L.slack_decay = cfg.slack_decay;
            L.P = w.P; L.q = w.q; L.A = w.A; L.l = w.l; L.u = w.u; L.c0 = w.c0;
            L.pred = w.pred; L.nok = w.nok;
            const int grid_a = std::min(count, 16 * std::max(1, o.sm_count));
            slack_assemble_kernel<<<grid_a, kTA, sizeof(double) * N * 6, st>>>(T, L);
            if (cudaGetLastError() != cudaSuccess) {
                set_error("slack_assemble_kernel launch failed");
                return MPCCBF_ERR_CUDA;
            }
            const int rc = dense_qp_solve_device(o, count, nv, m, w.P, w.q, w.A, w.l, w.u, w.wsP, w.wsA, w.x, w.obj, w.stat, w.it);
            if (rc != MPCCBF_OK) return rc;
            SlackFinish F;
            F.states = a.states; F.list = list; F.g0 = g0; F.count = count; F.agent_begin = a.agent_begin; F.agent_count = a.agent_count;
            F.it = it; F.nv = nv; F.m = m; F.hit = hit; F.J = J; F.x = w.x; F.c0 = w.c0; F.A = w.A; F.u = w.u; F.q = w.q; F.dstat = w.stat; F.dit = w.it;
            F.pred = w.pred; F.vlast = w.vlast; F.objlast = w.objlast; F.nok = w.nok;
            F.ctrl_pts = a.ctrl_pts; F.ctrl_all = a.ctrl_all; F.objective = a.objective;
            F.status = a.status; F.n_curves = a.n_curves; F.gi_iters = a.gi_iters;
            const int grid_f = std::min((count * 32 + kTA - 1) / kTA, 32 * std::max(1, o.sm_count));
            slack_finish_kernel<<<grid_f, kTA, 0, st>>>(T, F);
            if (cudaGetLastError() != cudaSuccess) {
                set_error("slack_finish_kernel launch failed");
                return MPCCBF_ERR_CUDA;
            }
            if (launches) *launches += 3;
        }
    }
    return MPCCBF_OK;
}

}  // namespace mpccbf
