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
    int g0, count;  // QP range [g0, g0 + count) of the StepArgs batch handled by this chunk
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

// One CTA per QP. Row order: safety rows (neighbour-major, horizon step inside), acceleration bound
// rows [k][axis], velocity bound rows [k][axis], slack >= 0 rows.
__global__ void __launch_bounds__(kTA) slack_assemble_kernel(const DevTables T, const SlackLaunch a) {
    extern __shared__ double st[];  // [N][6] states of the scenario
    const int K = T.K, N = T.N, J = a.J, nv = a.nv, m = a.m, hit = a.hit, t = threadIdx.x;
    const double* tm = T.tmpl;
    for (int b = blockIdx.x; b < a.count; b += gridDim.x) {
        const int g = a.g0 + b;
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
            const int other = j + (j >= self ? 1 : 0);
            const double dx = st[other * 6] - me[0], dy = st[other * 6 + 1] - me[1];
            const double dj = sqrt(dx * dx + dy * dy);
            int rank = 0;
            for (int i = 0; i < J; ++i) {
                const int oi = i + (i >= self ? 1 : 0);
                const double ex = st[oi * 6] - me[0], ey = st[oi * 6 + 1] - me[1];
                const double di = sqrt(ex * ex + ey * ey);
                rank += (di < dj || (di == dj && i < j)) ? 1 : 0;
            }
            q[6 + j] = a.slack_cost * pow(a.slack_decay, static_cast<double>(rank));
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
        const int g = a.g0 + b;
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

void slack_free(SlackWorkspace& w) {
    void* ptrs[] = {w.P, w.q, w.A, w.l, w.u, w.wsP, w.wsA, w.x, w.obj, w.pred, w.vlast, w.objlast, w.c0, w.stat, w.it, w.nok};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    w = SlackWorkspace();
}

// One control step of the slack-mode controller for the QPs of `a`, in chunks that fit the workspace.
int slack_step(const DevTables& T, const mpccbf_config& cfg, const StepArgs& a, const DenseOpts& o, SlackWorkspace& w,
               int64_t* launches) {
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
    const size_t want = std::min<size_t>(a.num_qp, std::max<size_t>(1, (cap_mb << 20) / per_qp));
    if (w.cap_qps < want || w.nv != nv || w.m_max != m_max) {
        slack_free(w);
        bool ok = true;
        ok &= grow(&w.P, want * nv * nv) && grow(&w.wsP, want * nv * nv) && grow(&w.A, want * m_max * nv);
        ok &= grow(&w.wsA, need_wsA ? want * m_max * (nv | 1) : 1);
        ok &= grow(&w.q, want * nv) && grow(&w.l, want * m_max) && grow(&w.u, want * m_max) && grow(&w.x, want * nv);
        ok &= grow(&w.obj, want) && grow(&w.pred, want * H * 6) && grow(&w.vlast, want * 6) && grow(&w.objlast, want);
        ok &= grow(&w.c0, want * 3) && grow(&w.stat, want) && grow(&w.it, want) && grow(&w.nok, want);
        if (!ok) {
            slack_free(w);
            set_error("slack mode: device workspace allocation failed (lower MPCCBF_SLACK_WS_MB)");
            return MPCCBF_ERR_ALLOC;
        }
        w.cap_qps = want;
        w.nv = nv;
        w.m_max = m_max;
    }
    cudaStream_t st = o.stream;
    for (int g0 = 0; g0 < a.num_qp; g0 += static_cast<int>(w.cap_qps)) {
        const int count = std::min<int>(static_cast<int>(w.cap_qps), a.num_qp - g0);
        slack_reset_kernel<<<(count * 6 + 255) / 256, 256, 0, st>>>(w.nok, w.vlast, count);
        for (int it = 0; it < iters; ++it) {
            const int hit = it == 0 ? 1 : H;
            const int m = J * hit + 6 * K + J;
            SlackLaunch L;
            L.states = a.states; L.targets = a.targets; L.ref_mode = a.ref_mode;
            L.g0 = g0; L.count = count; L.agent_begin = a.agent_begin; L.agent_count = a.agent_count;
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
            F.states = a.states; F.g0 = g0; F.count = count; F.agent_begin = a.agent_begin; F.agent_count = a.agent_count;
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
