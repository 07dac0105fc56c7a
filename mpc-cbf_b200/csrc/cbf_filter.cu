// CBF-only safety / connectivity filter — sm_100a. One WARP per scenario, one lane per agent.
//
// Replaces cbf::ConnectivityControl::optimize (cbf/src/controller/ConnectivityControl.cpp:22-99,
// slack_mode = false) for every agent of every scenario:
//   1. algebraic connectivity: Laplacian of the exp-weighted proximity graph and its second smallest
//      eigenpair (ConnectivityCBF::getLambda2, cbf/src/detail/ConnectivityCBF.cpp:388-424) by a cyclic
//      Jacobi iteration on the N x N matrix held in shared memory — the rotation angle is computed
//      redundantly by every lane, lane k rotates row/column k;
//   2. per agent the rows of its 3-variable QP  min |u - u_des|^2 :
//        safety HOCBF rows (closed form of initSafetyCBF :152-198), velocity CBF rows (:251-284),
//        the connectivity HOCBF row when lambda2 > 0.1 (initConnCBF :472-519 in closed form),
//        else one distance-CLF row per neighbour (initCLFCBF :200-243);
//   3. the QP itself: yaw only meets its two velocity rows (a clamp); x/y is a 2-dof projection
//      solved by the same dual active-set iteration as the IMPC kernel (gi_solver.h), rows
//      recomputed from shared memory on every scan.
#include <cuda_runtime.h>

#include <cmath>

#include "../../include/mpccbf.h"
#include "gi_solver.h"
#include "kernels.h"

namespace mpccbf {

namespace {

struct FilterRows {
    const double* st;   // scenario states [N][6] (shared memory)
    int N, self;
    double dmin2, gamma, tol;
    double vlo[2], vhi[2];  // rows  u_d >= vlo[d],  u_d <= vhi[d]   (d = x, y)
    int use_conn;
    double gx, gy, bconn;   // connectivity row  g.u + bconn >= 0

    // row id -> np, rhs with the meaning np.u >= rhs
    __device__ __forceinline__ void row(int id, double (&np)[2], double& rhs) const {
        const double* e = st + self * 6;
        if (id < N - 1) {  // safety HOCBF, neighbour id (skipping self)
            const double* o = st + (id + (id >= self ? 1 : 0)) * 6;
            const double dx = e[0] - o[0], dy = e[1] - o[1];
            const double dvx = e[3] - o[3], dvy = e[4] - o[4];
            const double h = dx * dx + dy * dy - dmin2;
            const double lfh = 2.0 * (dx * dvx + dy * dvy);
            const double lf2h = 2.0 * (dvx * dvx + dvy * dvy);
            const double lfa = 3.0 * gamma * h * h * (2.0 * dx * e[3] + 2.0 * dy * e[4]);
            const double psi1 = lfh + gamma * h * h * h;
            np[0] = 2.0 * dx;
            np[1] = 2.0 * dy;
            rhs = -(lf2h + lfa + gamma * psi1 * psi1 * psi1);
        } else if (id < N + 3) {  // velocity CBF rows: x min, y min, x max, y max
            const int r = id - (N - 1), d = r & 1, up = r >> 1;
            np[0] = (d == 0) ? (up ? -1.0 : 1.0) : 0.0;
            np[1] = (d == 1) ? (up ? -1.0 : 1.0) : 0.0;
            rhs = up ? -(d ? vhi[1] : vhi[0]) : (d ? vlo[1] : vlo[0]);
        } else if (use_conn) {
            np[0] = gx;
            np[1] = gy;
            rhs = -bconn;
        } else {  // distance CLF, neighbour id - (N + 3):  Ac.u <= -Bc
            const int j = id - (N + 3);
            const double* o = st + (j + (j >= self ? 1 : 0)) * 6;
            const double dx = e[0] - o[0], dy = e[1] - o[1];
            const double d2 = dx * dx + dy * dy, dist = sqrt(d2), err = dist - 2.0;
            const double gxx = 2.0 * err * dx / dist, gyy = 2.0 * err * dy / dist;
            const double vx = e[3], vy = e[4];
            const double i1 = 1.0 / dist, i2 = 1.0 / d2, i3 = i1 * i2;
            const double hxx = 2.0 * (dx * dx * i2 + err * (i1 - dx * dx * i3));
            const double hxy = 2.0 * (dx * dy * i2 + err * (-dx * dy * i3));
            const double hyy = 2.0 * (dy * dy * i2 + err * (i1 - dy * dy * i3));
            const double lfv = gxx * vx + gyy * vy;
            const double lf2v = vx * (hxx * vx + hxy * vy) + vy * (hxy * vx + hyy * vy);
            np[0] = -gxx;
            np[1] = -gyy;
            rhs = lf2v + 5.0 * lfv + 2.0 * err * err;
        }
    }
    __device__ __forceinline__ int num_rows() const { return (N - 1) + 4 + (use_conn ? 1 : (N - 1)); }
    __device__ __forceinline__ void scan(const double (&v)[2], double& best, int& best_id) const {
        const int m = num_rows();
        for (int id = 0; id < m; ++id) {
            double np[2], rhs;
            row(id, np, rhs);
            const double viol = rhs - fma(np[0], v[0], np[1] * v[1]);
            if (viol > tol * fmax(1.0, fabs(rhs)) && rhs > -1e20) {   // only an upper bound >= 1e20 (rhs <= -1e20 here) is "no bound"
                const double n2 = fma(np[0], np[0], np[1] * np[1]);
                const double score = (n2 > 0.0) ? viol * rsqrt(n2) : HUGE_VAL;
                if (score > best) {
                    best = score;
                    best_id = id;
                }
            }
        }
    }
    __device__ __forceinline__ void normal(int id, double (&np)[2], double& rhs) const { row(id, np, rhs); }
};

// Slack mode (ConnectivityControl.cpp:30-45, CBFQPGeneratorBase.cpp:9-30,59-70): N slacks s_g >= 0 with linear cost
// w_g = slack_cost * slack_decay^g; safety row g and CLF row g share s_g (g < N - 1), the connectivity row has s_{N-1};
// the velocity rows stay hard. With the slacks eliminated this is the capped-multiplier problem of gi_solver.h:
// group g = {safety row, CLF row, "none"} (or {connectivity row, "none"}), one piece of it the leader.
// Row ids: 0..3 the velocity rows (hard), 4 + 3 g + p with p = 0 safety / connectivity, 1 CLF, 2 "none".
struct FilterRowsSoft {
    FilterRows r;
    unsigned char* lead;  // [N] leaders of this agent's groups: 0 none, 1 + p
    double slack_cost, slack_decay;

    // piece p < 2 of group g as a row np.u >= rhs; false when the group has no such piece
    __device__ __forceinline__ bool piece(int g, int p, double (&np)[2], double& rhs) const {
        const int N = r.N;
        if (g == N - 1) {
            if (p != 0 || !r.use_conn) return false;
            r.row(N + 3, np, rhs);
            return true;
        }
        if (p == 0) {
            r.row(g, np, rhs);
            return true;
        }
        if (r.use_conn) return false;
        r.row(N + 3 + g, np, rhs);
        return true;
    }
    __device__ __forceinline__ void check_hard(int id, const double (&v)[2], double& best, int& best_id) const {
        double np[2], rhs;
        r.row(r.N - 1 + id, np, rhs);
        const double viol = rhs - fma(np[0], v[0], np[1] * v[1]);
        if (viol > r.tol * fmax(1.0, fabs(rhs)) && rhs > -1e20) {
            const double score = viol;  // unit normals
            if (score > best) {
                best = score;
                best_id = id;
            }
        }
    }
    __device__ __forceinline__ void scan(const double (&v)[2], double& best, int& best_id) const {
        for (int id = 0; id < 4; ++id) check_hard(id, v, best, best_id);
        for (int g = 0; g < r.N; ++g) {
            const int ld = lead[g];
            double nl[2] = {0.0, 0.0}, rl = 0.0, gl = 0.0;
            if (ld) {
                piece(g, ld - 1, nl, rl);
                gl = rl - fma(nl[0], v[0], nl[1] * v[1]);
            }
            for (int p = 0; p < 2; ++p) {
                double np[2], rhs;
                if (!piece(g, p, np, rhs)) continue;
                if (!(rhs > -1e20)) continue;   // upper bound >= 1e20 in the reference's orientation: no row
                const double gp = rhs - fma(np[0], v[0], np[1] * v[1]);
                double viol, thr = r.tol * fmax(1.0, fabs(rhs)), nn;
                int id = 4 + 3 * g + p;
                if (ld == 0 || ld == p + 1) {
                    viol = ld ? -gp : gp;   // the leader itself: piece "none" is violated once the row holds again
                    if (ld) id = 4 + 3 * g + 2;
                    nn = fma(np[0], np[0], np[1] * np[1]);
                } else {
                    viol = gp - gl;
                    thr = r.tol * fmax(fmax(1.0, fabs(rhs)), fabs(rl));
                    const double d0 = np[0] - nl[0], d1 = np[1] - nl[1];
                    nn = fma(d0, d0, d1 * d1);
                }
                if (viol > thr) {
                    const double score = (nn > 0.0) ? viol * rsqrt(nn) : 1e150;  // independent of u: taken up by the slack
                    if (score > best) {
                        best = score;
                        best_id = id;
                    }
                }
            }
        }
    }
    __device__ __forceinline__ void normal(int id, double (&np)[2], double& rhs) const {
        if (id < 4) {
            r.row(r.N - 1 + id, np, rhs);
            return;
        }
        const int g = (id - 4) / 3, p = (id - 4) - 3 * g;
        np[0] = np[1] = 0.0;
        rhs = 0.0;
        if (p < 2) piece(g, p, np, rhs);
        const int ld = lead[g];
        if (ld) {
            double nl[2], rl;
            piece(g, ld - 1, nl, rl);
            np[0] -= nl[0];
            np[1] -= nl[1];
            rhs -= rl;
        }
    }
    __device__ __forceinline__ int group(int id) const { return id < 4 ? -1 : (id - 4) / 3; }
    // the reference's cost is |u - u_d|^2 + sum w s (no 1/2): against 1/2 |u - u_d|^2 the weights are halved
    __device__ __forceinline__ double weight(int g) const { return 0.5 * slack_cost * pow(slack_decay, static_cast<double>(g)); }
    __device__ __forceinline__ void set_leader(int g, int id) const {
        const int p = (id - 4) - 3 * g;
        lead[g] = static_cast<unsigned char>(p < 2 ? p + 1 : 0);
    }
};

struct FilterArgs {
    const double* states;     // [S][N][6]
    const double* desired_u;  // [S][N][3]   (may be NULL when only rows are wanted)
    double* u;                // [S][N][3]   (may be NULL)
    int32_t* status;          // [S][N]      (may be NULL)
    double* lambda2;          // [S]         (may be NULL)
    double* conn_rows;        // [S][N][4] (gx, gy, 0, b), may be NULL
    double* fiedler;          // [S][N] normalised eigenvector of lambda2, may be NULL
    int S, N, NP;             // NP: padded row length (odd)
    double dmin2, dmax, gamma, tol;
    double vmin[3], vmax[3];
    int max_gi;
    int slack;                // slack_mode
    double slack_cost, slack_decay;
};

// Everything after the eigen-solve for ONE agent: its connectivity HOCBF row (closed form), the safety / velocity / CLF
// rows and the 3-variable QP (2-dof active-set projection for x / y, a clamp for yaw). st: the scenario's states [N][6],
// fv: the normalised eigenvector of lambda2, lead: N bytes of scratch for this agent (slack mode).
__device__ __forceinline__ void filter_agent(const FilterArgs& a, const double* st, const double* fv, double lam2, double R2,
                                             double sigma, int s, int self, unsigned char* lead) {
    const int N = a.N;
    const size_t g = static_cast<size_t>(s) * N + self;
    const double* e = st + self * 6;
    // connectivity HOCBF row (closed form; eigenvector held constant, no range gating)
    double gx = 0.0, gy = 0.0, hxx = 0.0, hxy = 0.0, hyy = 0.0;
    for (int j = 0; j < N; ++j) {
        if (j == self) continue;
        const double dx = e[0] - st[j * 6], dy = e[1] - st[j * 6 + 1];
        const double diff = R2 - (dx * dx + dy * dy);
        const double E = exp(diff * diff / sigma);
        const double dvj = fv[self] - fv[j], wj = dvj * dvj;
        const double c = -4.0 * E * diff / sigma;
        const double cp2 = 2.0 * (4.0 * E / sigma * (2.0 * diff * diff / sigma + 1.0));
        gx = fma(c * wj, dx, gx);
        gy = fma(c * wj, dy, gy);
        hxx += wj * (c + cp2 * dx * dx);
        hxy += wj * (cp2 * dx * dy);
        hyy += wj * (c + cp2 * dy * dy);
    }
    const double vx = e[3], vy = e[4];
    const double lfh = gx * vx + gy * vy;
    const double lf2h = vx * (hxx * vx + hxy * vy) + vy * (hxy * vx + hyy * vy);
    const double bconn = lf2h + a.gamma * lfh + a.gamma * (lfh + a.gamma * (lam2 - 0.1));
    if (a.conn_rows) {
        double* r = a.conn_rows + g * 4;
        r[0] = gx; r[1] = gy; r[2] = 0.0; r[3] = bconn;
    }
    if (a.u && a.desired_u) {
        const double* du = a.desired_u + g * 3;
        FilterRows rows;
        rows.st = st; rows.N = N; rows.self = self;
        rows.dmin2 = a.dmin2; rows.gamma = a.gamma; rows.tol = a.tol;
        rows.vlo[0] = -(e[3] - a.vmin[0]); rows.vhi[0] = a.vmax[0] - e[3];
        rows.vlo[1] = -(e[4] - a.vmin[1]); rows.vhi[1] = a.vmax[1] - e[4];
        rows.use_conn = lam2 > 0.1;
        rows.gx = gx; rows.gy = gy; rows.bconn = bconn;
        GIState<2> gi;
        const double v0[2] = {__ldg(du), __ldg(du + 1)};
        gi_init<2>(gi, v0);
        int iters = 0;
        int stat;
        if (a.slack) {
            // leaders of this agent's groups: N bytes provided by the caller
            FilterRowsSoft soft;
            soft.r = rows;
            soft.lead = lead;
            soft.slack_cost = a.slack_cost;
            soft.slack_decay = a.slack_decay;
            for (int gq = 0; gq < N; ++gq) soft.lead[gq] = 0;
            GISoft<2> gs;
            gs.grp[0] = gs.grp[1] = -1;
            gs.cap[0] = gs.cap[1] = HUGE_VAL;
            stat = gi_run<2, true, FilterRowsSoft>(gi, gs, soft, a.max_gi, iters, gi_vmax2<2>(v0));
        } else {
            stat = gi_run<2>(gi, rows, a.max_gi, iters);
        }
        // yaw: -(w - wmin) <= u <= wmax - w
        const double ylo = -(e[5] - a.vmin[2]), yhi = a.vmax[2] - e[5];
        if (ylo > yhi + a.tol * fmax(1.0, fabs(yhi))) stat = (stat == MPCCBF_STATUS_OPTIMAL) ? MPCCBF_STATUS_INFEASIBLE : stat;
        const double uy = fmin(fmax(__ldg(du + 2), ylo), yhi);
        const bool ok = stat == MPCCBF_STATUS_OPTIMAL;
        a.u[g * 3 + 0] = ok ? gi.v[0] : 0.0;
        a.u[g * 3 + 1] = ok ? gi.v[1] : 0.0;
        a.u[g * 3 + 2] = ok ? uy : 0.0;
        if (a.status) a.status[g] = stat;
    }
}

__global__ void __launch_bounds__(128) cbf_filter_kernel(const FilterArgs a) {
    extern __shared__ __align__(16) double fsm[];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int N = a.N, NP = a.NP;
    const int per_warp = N * 6 + 2 * N * NP + N;
    double* st = fsm + static_cast<size_t>(wid) * per_warp;
    double* Lm = st + N * 6;
    double* Vm = Lm + N * NP;
    double* fv = Vm + N * NP;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    const double R2 = a.dmax * a.dmax, sigma = R2 * R2 / log(2.0);

    for (int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < a.S; s += warps) {
        const double* src = a.states + static_cast<size_t>(s) * N * 6;
        for (int i = lane; i < N * 6; i += 32) st[i] = __ldg(src + i);
        __syncwarp();
        // ---- Laplacian (lane = row), V = I ---------------------------------------------------
        if (lane < N) {
            double deg = 0.0;
            for (int j = 0; j < N; ++j) {
                double w = 0.0;
                if (j != lane) {
                    const double dx = st[lane * 6] - st[j * 6], dy = st[lane * 6 + 1] - st[j * 6 + 1];
                    const double d2 = dx * dx + dy * dy;
                    if (d2 <= R2) w = exp((R2 - d2) * (R2 - d2) / sigma) - 1.0;
                }
                Lm[lane * NP + j] = -w;
                Vm[lane * NP + j] = (j == lane) ? 1.0 : 0.0;
                deg += w;
            }
            Lm[lane * NP + lane] = deg;
        }
        __syncwarp();
        // ---- cyclic Jacobi ---------------------------------------------------------------------
        for (int sweep = 0; sweep < 60 && N > 1; ++sweep) {
            double off = 0.0, dg = 0.0;
            if (lane < N)
                for (int j = 0; j < N; ++j) {
                    const double x = Lm[lane * NP + j];
                    if (j == lane) dg = x * x; else off = fma(x, x, off);
                }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                off += __shfl_xor_sync(full, off, o);
                dg += __shfl_xor_sync(full, dg, o);
            }
            if (!(off > 4.930380657631324e-32 * dg)) break;  // eps^2
            for (int p = 0; p < N - 1; ++p)
                for (int q = p + 1; q < N; ++q) {
                    const double apq = Lm[p * NP + q];
                    if (apq == 0.0) continue;  // warp-uniform
                    const double app = Lm[p * NP + p], aqq = Lm[q * NP + q];
                    const double theta = (aqq - app) / (2.0 * apq);
                    const double t = copysign(1.0, theta) / (fabs(theta) + sqrt(fma(theta, theta, 1.0)));
                    const double c = rsqrt(fma(t, t, 1.0)), sn = t * c;
                    __syncwarp();
                    if (lane < N) {
                        const int k = lane;
                        if (k != p && k != q) {
                            const double akp = Lm[k * NP + p], akq = Lm[k * NP + q];
                            const double nkp = c * akp - sn * akq, nkq = sn * akp + c * akq;
                            Lm[k * NP + p] = nkp;
                            Lm[p * NP + k] = nkp;
                            Lm[k * NP + q] = nkq;
                            Lm[q * NP + k] = nkq;
                        }
                        const double vkp = Vm[k * NP + p], vkq = Vm[k * NP + q];
                        Vm[k * NP + p] = c * vkp - sn * vkq;
                        Vm[k * NP + q] = sn * vkp + c * vkq;
                    }
                    __syncwarp();
                    if (lane == 0) {
                        Lm[p * NP + p] = app - t * apq;
                        Lm[q * NP + q] = aqq + t * apq;
                        Lm[p * NP + q] = 0.0;
                        Lm[q * NP + p] = 0.0;
                    }
                    __syncwarp();
                }
        }
        // ---- second smallest eigenvalue and its (normalised) eigenvector ---------------------------
        double lam2 = 0.0;
        if (N > 1) {
            const double dv = (lane < N) ? Lm[lane * NP + lane] : HUGE_VAL;
            double m1 = dv;
            int i1 = lane;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double om = __shfl_xor_sync(full, m1, o);
                const int oi = __shfl_xor_sync(full, i1, o);
                if (om < m1 || (om == m1 && oi < i1)) { m1 = om; i1 = oi; }
            }
            double m2 = (lane == i1) ? HUGE_VAL : dv;
            int i2 = lane;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double om = __shfl_xor_sync(full, m2, o);
                const int oi = __shfl_xor_sync(full, i2, o);
                if (om < m2 || (om == m2 && oi < i2)) { m2 = om; i2 = oi; }
            }
            lam2 = m2;
            double x = (lane < N) ? Vm[lane * NP + i2] : 0.0;
            double nn = x * x;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) nn += __shfl_xor_sync(full, nn, o);
            if (lane < N) fv[lane] = x * rsqrt(nn);
        } else if (lane < N) {
            fv[lane] = 0.0;
        }
        __syncwarp();
        if (lane == 0 && a.lambda2) a.lambda2[s] = lam2;
        if (a.fiedler && lane < N) a.fiedler[static_cast<size_t>(s) * N + lane] = fv[lane];
        if (!a.conn_rows && !a.u) {  // only the eigenpair was wanted
            __syncwarp();
            continue;
        }

        // ---- per agent: rows and the 3-variable QP --------------------------------------------------
        // (slack mode: the leaders of an agent's groups take N bytes of the eigenvector matrix's storage, free by now)
        if (lane < N) filter_agent(a, st, fv, lam2, R2, sigma, s, lane, reinterpret_cast<unsigned char*>(Vm) + lane * 32);
        __syncwarp();
    }
}


// ---- swarms of 33 .. 256 agents: one CTA of 256 threads per scenario ----------------------------------------------------
// The Laplacian and the (transposed) eigenvector matrix no longer fit in shared memory (2 x 8 N^2 bytes): they live in a
// per-CTA global workspace, which stays in L2. The eigen-solve is the same Jacobi iteration in its PARALLEL ordering: a
// sweep is N - 1 rounds of N / 2 rotations on disjoint index pairs (round-robin tournament), each round = the angles from
// the 2 x 2 blocks, then J' L and J' V' as row operations (coalesced) and L J as a column operation. Same convergence test.
constexpr int kWideT = 256;

// sum over the CTA of two values per thread; every thread returns the totals
__device__ __forceinline__ void wide_sum2(double& x, double& y, double* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        x += __shfl_xor_sync(0xffffffffu, x, o);
        y += __shfl_xor_sync(0xffffffffu, y, o);
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) {
        red[2 * w] = x;
        red[2 * w + 1] = y;
    }
    __syncthreads();
    x = 0.0;
    y = 0.0;
    for (int i = 0; i < kWideT / 32; ++i) {
        x += red[2 * i];
        y += red[2 * i + 1];
    }
}
// arg-min over the CTA (ties: smaller index); every thread returns the winner
__device__ __forceinline__ void wide_argmin(double& v, int& idx, double* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ov < v || (ov == v && oi < idx)) { v = ov; idx = oi; }
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) {
        red[2 * w] = v;
        red[2 * w + 1] = static_cast<double>(idx);
    }
    __syncthreads();
    v = red[0];
    idx = static_cast<int>(red[1]);
    for (int i = 1; i < kWideT / 32; ++i) {
        const double ov = red[2 * i];
        const int oi = static_cast<int>(red[2 * i + 1]);
        if (ov < v || (ov == v && oi < idx)) { v = ov; idx = oi; }
    }
}

__global__ void __launch_bounds__(kWideT) cbf_filter_wide_kernel(const FilterArgs a, double* ws_mats, unsigned char* ws_lead) {
    extern __shared__ __align__(16) double wsm[];
    const int N = a.N, t = threadIdx.x;
    const int Ne = (N + 1) & ~1, half = Ne / 2;
    double* st = wsm;            // [N][6]
    double* fv = st + N * 6;     // [N]
    double* rc = fv + N;         // [half] cosines of the round
    double* rs = rc + half;      // [half] sines
    double* red = rs + half;     // reductions
    int* rp = reinterpret_cast<int*>(red + 2 * (kWideT / 32));  // [half] pair (p, q) of the round, p < 0: no rotation
    int* rq = rp + half;
    double* Lm = ws_mats + static_cast<size_t>(blockIdx.x) * 2 * N * N;
    double* Vt = Lm + static_cast<size_t>(N) * N;   // V transposed: row i = eigenvector of the i-th diagonal entry
    unsigned char* lead = ws_lead ? ws_lead + static_cast<size_t>(blockIdx.x) * N * N : nullptr;
    const double R2 = a.dmax * a.dmax, sigma = R2 * R2 / log(2.0);

    for (int s = blockIdx.x; s < a.S; s += gridDim.x) {
        __syncthreads();
        const double* src = a.states + static_cast<size_t>(s) * N * 6;
        for (int i = t; i < N * 6; i += kWideT) st[i] = __ldg(src + i);
        __syncthreads();
        for (int r = t; r < N; r += kWideT) {   // Laplacian row r, V = I
            double deg = 0.0;
            for (int j = 0; j < N; ++j) {
                double w = 0.0;
                if (j != r) {
                    const double dx = st[r * 6] - st[j * 6], dy = st[r * 6 + 1] - st[j * 6 + 1];
                    const double d2 = dx * dx + dy * dy;
                    if (d2 <= R2) w = exp((R2 - d2) * (R2 - d2) / sigma) - 1.0;
                }
                Lm[static_cast<size_t>(r) * N + j] = -w;
                Vt[static_cast<size_t>(r) * N + j] = (j == r) ? 1.0 : 0.0;
                deg += w;
            }
            Lm[static_cast<size_t>(r) * N + r] = deg;
        }
        __syncthreads();
        for (int sweep = 0; sweep < 60; ++sweep) {
            double off = 0.0, dg = 0.0;
            for (int r = t; r < N; r += kWideT)
                for (int j = 0; j < N; ++j) {
                    const double x = Lm[static_cast<size_t>(r) * N + j];
                    if (j == r) dg = fma(x, x, dg); else off = fma(x, x, off);
                }
            wide_sum2(off, dg, red);
            if (!(off > 4.930380657631324e-32 * dg)) break;  // eps^2, CTA-uniform
            for (int round = 0; round < Ne - 1; ++round) {
                if (t < half) {   // pair t of this round (circle method), its rotation from the 2 x 2 block
                    int p0, q0;
                    if (t == 0) { p0 = Ne - 1; q0 = round; }
                    else { p0 = (round + t) % (Ne - 1); q0 = (round - t + Ne - 1) % (Ne - 1); }
                    int p = min(p0, q0), q = max(p0, q0);
                    double c = 1.0, sn = 0.0;
                    if (q >= N) {
                        p = -1;   // the bye of an odd N
                    } else {
                        const double apq = Lm[static_cast<size_t>(p) * N + q];
                        if (apq == 0.0) {
                            p = -1;
                        } else {
                            const double app = Lm[static_cast<size_t>(p) * N + p], aqq = Lm[static_cast<size_t>(q) * N + q];
                            const double theta = (aqq - app) / (2.0 * apq);
                            const double tt = copysign(1.0, theta) / (fabs(theta) + sqrt(fma(theta, theta, 1.0)));
                            c = rsqrt(fma(tt, tt, 1.0));
                            sn = tt * c;
                        }
                    }
                    rc[t] = c; rs[t] = sn; rp[t] = p; rq[t] = q;
                }
                __syncthreads();
                for (int e = t; e < half * N; e += kWideT) {   // rows: L <- J' L, V' <- J' V'
                    const int i = e / N, k = e - i * N;
                    const int p = rp[i];
                    if (p < 0) continue;
                    const int q = rq[i];
                    const double c = rc[i], sn = rs[i];
                    const size_t ip = static_cast<size_t>(p) * N + k, iq = static_cast<size_t>(q) * N + k;
                    const double lp = Lm[ip], lq = Lm[iq];
                    Lm[ip] = c * lp - sn * lq;
                    Lm[iq] = sn * lp + c * lq;
                    const double vp = Vt[ip], vq = Vt[iq];
                    Vt[ip] = c * vp - sn * vq;
                    Vt[iq] = sn * vp + c * vq;
                }
                __syncthreads();
                for (int e = t; e < half * N; e += kWideT) {   // columns: L <- L J
                    const int k = e / half, i = e - k * half;
                    const int p = rp[i];
                    if (p < 0) continue;
                    const int q = rq[i];
                    const double c = rc[i], sn = rs[i];
                    const size_t ip = static_cast<size_t>(k) * N + p, iq = static_cast<size_t>(k) * N + q;
                    const double lp = Lm[ip], lq = Lm[iq];
                    Lm[ip] = c * lp - sn * lq;
                    Lm[iq] = sn * lp + c * lq;
                }
                __syncthreads();
                if (t < half && rp[t] >= 0) {   // the annihilated entries are exactly zero
                    Lm[static_cast<size_t>(rp[t]) * N + rq[t]] = 0.0;
                    Lm[static_cast<size_t>(rq[t]) * N + rp[t]] = 0.0;
                }
                __syncthreads();
            }
        }
        // ---- second smallest eigenvalue and its (normalised) eigenvector --------------------------------------------
        const double dv = (t < N) ? Lm[static_cast<size_t>(t) * N + t] : HUGE_VAL;
        double m1 = dv;
        int i1 = t;
        wide_argmin(m1, i1, red);
        double m2 = (t == i1) ? HUGE_VAL : dv;
        int i2 = t;
        wide_argmin(m2, i2, red);
        const double lam2 = m2;
        double x = (t < N) ? Vt[static_cast<size_t>(i2) * N + t] : 0.0, nn = x * x, unused = 0.0;
        wide_sum2(nn, unused, red);
        if (t < N) fv[t] = x * rsqrt(nn);
        __syncthreads();
        if (t == 0 && a.lambda2) a.lambda2[s] = lam2;
        if (a.fiedler && t < N) a.fiedler[static_cast<size_t>(s) * N + t] = fv[t];
        if (!a.conn_rows && !a.u) continue;
        if (t < N) filter_agent(a, st, fv, lam2, R2, sigma, s, t, lead ? lead + static_cast<size_t>(t) * N : nullptr);
    }
}

}  // namespace

// Returns a cudaError_t; cudaErrorInvalidValue when N > 256. Swarms of more than 32 agents need `wide_ws`: at least
// cbf_filter_wide_ws_bytes(cfg, blocks) bytes for `blocks` resident CTAs (the grid is min(S, blocks)).
size_t cbf_filter_wide_ws_bytes(const mpccbf_config& cfg, int blocks) {
    const size_t N = cfg.num_agents;
    return static_cast<size_t>(blocks) * (2 * N * N * sizeof(double) + N * N);
}

int launch_cbf_filter(const mpccbf_config& cfg, int S, const double* states, const double* desired_u, double* u,
                      int32_t* status, double* lambda2, double* conn_rows, int sm_count, cudaStream_t stream, double* fiedler,
                      void* wide_ws, int wide_blocks) {
    if (S <= 0) return cudaSuccess;
    const int N = cfg.num_agents;
    if (N > 256 || N < 1) return cudaErrorInvalidValue;
    FilterArgs a;
    a.states = states; a.desired_u = desired_u; a.u = u; a.status = status; a.lambda2 = lambda2; a.conn_rows = conn_rows; a.fiedler = fiedler;
    a.S = S; a.N = N; a.NP = N | 1;
    a.dmin2 = cfg.d_min * cfg.d_min; a.dmax = cfg.d_max;
    a.gamma = cfg.cbf_gamma > 0 ? cfg.cbf_gamma : 5.0;
    a.tol = cfg.feas_tol > 0 ? cfg.feas_tol : 1e-9;
    for (int d = 0; d < 3; ++d) { a.vmin[d] = cfg.v_min[d]; a.vmax[d] = cfg.v_max[d]; }
    a.max_gi = 50 * (2 * N + 4) + 100;
    a.slack = cfg.slack_mode ? 1 : 0;
    a.slack_cost = cfg.slack_cost;
    a.slack_decay = cfg.slack_decay;
    if (N > 32) {
        if (!wide_ws || wide_blocks < 1) return cudaErrorInvalidValue;
        const int half = ((N + 1) & ~1) / 2;
        const size_t smem = sizeof(double) * (N * 6 + N + 2 * half + 2 * (kWideT / 32)) + sizeof(int) * 2 * half;
        double* mats = static_cast<double*>(wide_ws);
        unsigned char* lead = reinterpret_cast<unsigned char*>(mats + static_cast<size_t>(wide_blocks) * 2 * N * N);
        const int grid = S < wide_blocks ? S : wide_blocks;
        cbf_filter_wide_kernel<<<grid, kWideT, smem, stream>>>(a, mats, lead);
        return cudaGetLastError();
    }
    const int threads = 128;
    const size_t per_warp = sizeof(double) * (N * 6 + 2 * N * a.NP + N);
    const size_t smem = per_warp * (threads / 32);
    cudaError_t e = cudaFuncSetAttribute(cbf_filter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    int per_sm = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, cbf_filter_kernel, threads, smem);
    if (e != cudaSuccess) return e;
    const int need = (S + threads / 32 - 1) / (threads / 32);
    int grid = sm_count * (per_sm < 1 ? 1 : per_sm);  // persistent: a multiple of the SM count
    if (need < grid) grid = need;
    cbf_filter_kernel<<<grid, threads, smem, stream>>>(a);
    return cudaGetLastError();
}

}  // namespace mpccbf
