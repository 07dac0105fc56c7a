// Fused IMPC kernel for configurations that do not collapse to one cubic per axis — sm_100a.
//
// bezier_continuity_upto_degree and num_control_points are configuration values of the reference
// (common/include/common/parsing.hpp:20-28, experiments/config/base_config.json:72). Whatever they are, the
// initial-state and continuity rows (PiecewiseBezierMPCQPGenerator.cpp:167-226) are constant and leave
// r = m1 - me1 free variables per axis; tables.cpp derives the same whitened null-space coordinates as for the
// cubic / C^3 case (r = 2, served by impc_thread.cu): objective 1/2 |v - v0|^2 + const, bound rows
// lo <= Ct[d][k] . v_axis + Ot[d][k] . [pos; vel] <= hi, safety rows coupling x and y through the acceleration at t_k.
// This kernel runs ConnectivityIMPCCBF::optimize (mpc_cbf/src/controller/ConnectivityIMPCCBF.cpp:47-215) for any
// 1 <= r <= 8: one thread per (scenario, agent), the yaw axis as an r-dof projection, x/y as a 2r-dof projection,
// the same dual active-set iteration as gi_solver.h (state = iterate, multipliers, active row ids; every step
// re-factorises the active normals by modified Gram-Schmidt) with run-time dimensions. The row families have no
// polynomial structure here, so a scan evaluates every row (r FMAs each); the working arrays of a step live in
// local memory. It is the correctness path for non-default configurations, not the throughput path.
// Slack mode and conn_mode are not served here.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <vector>

#include "../../include/mpccbf.h"
#include "kernels.h"

namespace mpccbf {

#ifndef MPCCBF_HD
#define MPCCBF_HD __host__ __device__ __forceinline__
#endif

namespace {

constexpr int GD = 2 * kGenMaxR;  // largest projection dimension (x and y together)

struct GenRows {
    const GenTables* T;
    int D;           // projection dimension: r (yaw) or 2 r (x / y)
    int nax;         // axes in this projection: 1 (yaw) or 2 (x, y)
    int ax0;         // first axis
    double s[2][2];  // [axis][pos, vel]
    // safety rows (x / y only)
    const double* st;  // scenario states [N][6]
    int self, ncbf, hit;
    double pred[8][4];  // ego (px, py, vx, vy) per CBF horizon step

    MPCCBF_HD int nfix() const { return nax * 2 * T->K * 2; }  // (dsel, k, axis, side)
    // row id -> normal (D) and rhs with the meaning np . v >= rhs; returns false for "no bound"
    __host__ __device__ bool row(int id, double* np, double& rhs, double& bound_abs) const {
        const int r = T->r, K = T->K;
        for (int i = 0; i < D; ++i) np[i] = 0.0;
        if (id < nfix()) {
            const int up = id & 1;
            int q = id >> 1;
            const int ax = q % nax;
            q /= nax;
            const int dsel = q / K, k = q - dsel * K;  // dsel 0: acceleration rows, 1: velocity rows
            const double* t = T->gt + (static_cast<size_t>(2 - dsel) * K + k) * (r + 2);
            const double bound = up ? T->lim_hi[dsel][ax0 + ax] : T->lim_lo[dsel][ax0 + ax];
            bound_abs = fabs(bound);
            if (!(bound_abs < 1e20)) return false;
            const double sg = up ? -1.0 : 1.0;
            for (int j = 0; j < r; ++j) np[ax * r + j] = sg * t[j];
            rhs = sg * (bound - (t[r] * s[ax][0] + t[r + 1] * s[ax][1]));
            return true;
        }
        const int j = id - nfix();
        const int nb = j / hit, k = j - nb * hit;
        const double* o = st + (nb + (nb >= self ? 1 : 0)) * 6;
        // closed form of ConnectivityCBF::initSafetyCBF (cbf/src/detail/ConnectivityCBF.cpp:152-198)
        const double dx = pred[k][0] - o[0], dy = pred[k][1] - o[1];
        const double dvx = pred[k][2] - o[3], dvy = pred[k][3] - o[4];
        const double h = dx * dx + dy * dy - T->dmin2;
        const double lfh = 2.0 * (dx * dvx + dy * dvy);
        const double lf2h = 2.0 * (dvx * dvx + dvy * dvy);
        const double lfa = 3.0 * T->gamma * h * h * (2.0 * dx * pred[k][2] + 2.0 * dy * pred[k][3]);
        const double psi1 = lfh + T->gamma * h * h * h;
        const double ax = 2.0 * dx, ay = 2.0 * dy, b = lf2h + lfa + T->gamma * psi1 * psi1 * psi1;
        bound_abs = fabs(b);
        if (!(b < 1e20)) return false;  // one-sided: an upper bound >= 1e20 is "no bound"
        const double* t = T->gt + (static_cast<size_t>(2) * K + k) * (r + 2);  // acceleration at t_k
        for (int i = 0; i < r; ++i) {
            np[i] = ax * t[i];
            np[r + i] = ay * t[i];
        }
        const double offx = t[r] * s[0][0] + t[r + 1] * s[0][1];
        const double offy = t[r] * s[1][0] + t[r + 1] * s[1][1];
        rhs = -(b + ax * offx + ay * offy);
        return true;
    }
};

struct GenState {
    double v[GD], u[GD];
    int id[GD];
    int q;
};

// Dual active-set iteration (Goldfarb-Idnani, least-distance form) with run-time dimension D <= GD.
__host__ __device__ int gen_solve(GenState& g, const GenRows& rows, int nrows, int max_iter, int& n_iter) {
    const int D = rows.D;
    // divergence guard (see gi_solver.h): an iterate 1e7 times farther out than v0 proves infeasibility
    double vmax2 = 1.0;
    for (int i = 0; i < D; ++i) vmax2 = fma(g.v[i], g.v[i], vmax2);
    vmax2 *= 1e14;
    const double tol = rows.T->feas_tol;
    double np[GD], z[GD], d1[GD], rr[GD], w[GD];
    double Q[GD][GD], R[GD][GD];
    for (;;) {
        // ---- scan: most violated row in whitened distance ---------------------------------------------------
        double best = 0.0;
        int best_id = -1;
        for (int id = 0; id < nrows; ++id) {
            double rhs, babs;
            if (!rows.row(id, w, rhs, babs)) continue;
            double val = -rhs, nn = 0.0;
            for (int i = 0; i < D; ++i) {
                val = fma(w[i], g.v[i], val);
                nn = fma(w[i], w[i], nn);
            }
            if (-val > tol * fmax(1.0, babs)) {
                const double score = (nn > 0.0) ? -val / sqrt(nn) : HUGE_VAL;
                if (score > best) {
                    best = score;
                    best_id = id;
                }
            }
        }
        if (best_id < 0) return MPCCBF_STATUS_OPTIMAL;
        if (!(best < HUGE_VAL)) return MPCCBF_STATUS_INFEASIBLE;
        double rhs, babs;
        rows.row(best_id, np, rhs, babs);
        double up_mult = 0.0;
        for (;;) {
            if (++n_iter > max_iter) return MPCCBF_STATUS_ERROR;
            double dall = 0.0;
            for (int i = 0; i < D; ++i) {
                z[i] = np[i];
                dall = fma(np[i], np[i], dall);
            }
            for (int c = 0; c < g.q; ++c) {
                double r2, b2;
                rows.row(g.id[c], w, r2, b2);
                for (int p = 0; p < c; ++p) {
                    double s = 0.0;
                    for (int i = 0; i < D; ++i) s = fma(Q[p][i], w[i], s);
                    R[p][c] = s;
                    for (int i = 0; i < D; ++i) w[i] = fma(-s, Q[p][i], w[i]);
                }
                double nn = 0.0;
                for (int i = 0; i < D; ++i) nn = fma(w[i], w[i], nn);
                const double nrm = sqrt(nn), inv = 1.0 / nrm;
                R[c][c] = nrm;
                double s = 0.0;
                for (int i = 0; i < D; ++i) {
                    Q[c][i] = w[i] * inv;
                    s = fma(Q[c][i], z[i], s);
                }
                d1[c] = s;
                for (int i = 0; i < D; ++i) z[i] = fma(-s, Q[c][i], z[i]);
            }
            for (int c = 0; c < g.q; ++c) {  // second orthogonalisation pass
                double s = 0.0;
                for (int i = 0; i < D; ++i) s = fma(Q[c][i], z[i], s);
                d1[c] += s;
                for (int i = 0; i < D; ++i) z[i] = fma(-s, Q[c][i], z[i]);
            }
            double ztn = 0.0;
            for (int i = 0; i < D; ++i) ztn = fma(z[i], z[i], ztn);
            for (int i = g.q - 1; i >= 0; --i) {
                double s = d1[i];
                for (int k = i + 1; k < g.q; ++k) s = fma(-R[i][k], rr[k], s);
                rr[i] = s / R[i][i];
            }
            const double zt = 1e4 * 2.220446049250313e-16;
            const bool z_zero = !(ztn > zt * zt * dall);
            double t1 = HUGE_VAL;
            int lpos = -1;
            for (int j = 0; j < g.q; ++j)
                if (rr[j] > 0.0) {
                    const double t = g.u[j] / rr[j];
                    if (t < t1) {
                        t1 = t;
                        lpos = j;
                    }
                }
            double sp = -rhs;
            for (int i = 0; i < D; ++i) sp = fma(np[i], g.v[i], sp);
            const double t2 = z_zero ? HUGE_VAL : (-sp / ztn);
            const double t = fmin(t1, t2);
            if (!(t < HUGE_VAL)) return MPCCBF_STATUS_INFEASIBLE;
            if (!z_zero)
                for (int i = 0; i < D; ++i) g.v[i] = fma(t, z[i], g.v[i]);
            for (int j = 0; j < g.q; ++j) g.u[j] = fma(-t, rr[j], g.u[j]);
            up_mult += t;
            {
                double vv = 0.0;
                for (int i = 0; i < D; ++i) vv = fma(g.v[i], g.v[i], vv);
                if (!(vv <= vmax2)) return MPCCBF_STATUS_INFEASIBLE;
            }
            if (!z_zero && t2 <= t1) {
                g.u[g.q] = up_mult;
                g.id[g.q] = best_id;
                ++g.q;
                break;
            }
            for (int c = lpos; c + 1 < g.q; ++c) {
                g.u[c] = g.u[c + 1];
                g.id[c] = g.id[c + 1];
            }
            --g.q;
        }
    }
}

// Everything optimize() does for agent-QP g.
__host__ __device__ void general_agent(const GenTables& T, const StepArgs& a, int g, int self, const double* st) {
    const int K = T.K, r = T.r, N = T.N, H = T.H, n = T.n, c = T.c, m1 = T.m1;
    double s[3][2];
    for (int ax = 0; ax < 3; ++ax) {
        s[ax][0] = st[self * 6 + ax];
        s[ax][1] = st[self * 6 + 3 + ax];
    }
    // unconstrained minimiser (whitened) and the constant part of the objective
    double v0[3][kGenMaxR], c0 = 0.0;
    for (int ax = 0; ax < 3; ++ax) {
        const double ps = s[ax][0], vs = s[ax][1];
        double kk = 0.0;
        if (a.ref_mode == 0) {
            const double tgt = a.targets[static_cast<size_t>(g) * 3 + ax];
            for (int j = 0; j < r; ++j) v0[ax][j] = T.vs[2 * j] * ps + T.vs[2 * j + 1] * vs + T.vt[j] * tgt;
            kk = tgt * (T.Kt[0] * ps + T.Kt[1] * vs);
        } else {
            const double* rf = a.targets + static_cast<size_t>(g) * 3 * K;
            for (int j = 0; j < r; ++j) {
                double acc = 0.0;
                for (int k = 0; k < K; ++k) acc = fma(T.vr[j * K + k], rf[3 * k + ax], acc);
                v0[ax][j] = T.vs[2 * j] * ps + T.vs[2 * j + 1] * vs + acc;
            }
            double k0 = 0.0, k1 = 0.0;
            for (int k = 0; k < K; ++k) {
                k0 = fma(T.kr[k], rf[3 * k + ax], k0);
                k1 = fma(T.kr[K + k], rf[3 * k + ax], k1);
            }
            kk = k0 * ps + k1 * vs;
        }
        c0 += ps * (T.Kc[0][0] * ps + T.Kc[0][1] * vs) + vs * (T.Kc[1][0] * ps + T.Kc[1][1] * vs) + kk;
    }
    // ---- yaw: r-dof projection, shared by every IMPC iteration ----------------------------------------------
    GenRows ry;
    ry.T = &T; ry.D = r; ry.nax = 1; ry.ax0 = 2;
    ry.s[0][0] = s[2][0]; ry.s[0][1] = s[2][1];
    ry.s[1][0] = ry.s[1][1] = 0.0;
    ry.st = st; ry.self = self; ry.ncbf = 0; ry.hit = 1;
    GenState gy;
    gy.q = 0;
    for (int j = 0; j < r; ++j) gy.v[j] = v0[2][j];
    int yaw_iter = 0;
    int base_status = gen_solve(gy, ry, ry.nfix(), T.max_gi, yaw_iter);
    // ---- x / y stage A: bound rows only ------------------------------------------------------------------------
    GenRows rx;
    rx.T = &T; rx.D = 2 * r; rx.nax = 2; rx.ax0 = 0;
    for (int ax = 0; ax < 2; ++ax) {
        rx.s[ax][0] = s[ax][0];
        rx.s[ax][1] = s[ax][1];
    }
    rx.st = st; rx.self = self; rx.ncbf = 0; rx.hit = 1;
    GenState gx, snap;
    gx.q = 0;
    for (int j = 0; j < r; ++j) {
        gx.v[j] = v0[0][j];
        gx.v[r + j] = v0[1][j];
    }
    int stage_iter = 0;
    if (base_status == MPCCBF_STATUS_OPTIMAL) base_status = gen_solve(gx, rx, rx.nfix(), T.max_gi, stage_iter);
    snap = gx;
    const size_t sbase = static_cast<size_t>(g) * T.iters;
    int n_ok = 0;
    double vlast[3][kGenMaxR];
    for (int it = 0; it < T.iters; ++it) {
        int stat = base_status, n2 = 0;
        if (base_status == MPCCBF_STATUS_OPTIMAL) {
            const int hit = (it == 0) ? 1 : H;
            for (int k = 0; k < hit; ++k) {  // ego states the rows are built at (ConnectivityIMPCCBF.cpp:135-141,160-175)
                if (it == 0) {
                    rx.pred[0][0] = s[0][0]; rx.pred[0][1] = s[1][0]; rx.pred[0][2] = s[0][1]; rx.pred[0][3] = s[1][1];
                } else {
                    for (int d = 0; d < 2; ++d) {
                        const double* t = T.gt + (static_cast<size_t>(d) * K + k) * (r + 2);
                        for (int ax = 0; ax < 2; ++ax) {
                            double val = t[r] * s[ax][0] + t[r + 1] * s[ax][1];
                            for (int j = 0; j < r; ++j) val = fma(t[j], gx.v[ax * r + j], val);
                            rx.pred[k][d * 2 + ax] = val;
                        }
                    }
                }
            }
            gx = snap;
            rx.hit = hit;
            rx.ncbf = (N - 1) * hit;
            stat = gen_solve(gx, rx, rx.nfix() + rx.ncbf, T.max_gi, n2);
        }
        a.status[sbase + it] = stat;
        if (a.gi_iters) a.gi_iters[sbase + it] = n2 + (it == 0 ? yaw_iter + stage_iter : 0);
        if (stat != MPCCBF_STATUS_OPTIMAL) {
            for (int q = it + 1; q < T.iters; ++q) {
                a.status[sbase + q] = MPCCBF_STATUS_SKIPPED;
                if (a.gi_iters) a.gi_iters[sbase + q] = 0;
            }
            break;
        }
        ++n_ok;
        for (int j = 0; j < r; ++j) {
            vlast[0][j] = gx.v[j];
            vlast[1][j] = gx.v[r + j];
            vlast[2][j] = gy.v[j];
        }
        if (a.ctrl_all) {
            double* out = a.ctrl_all + (sbase + it) * n;
            for (int i = 0; i < m1; ++i) {
                const int piece = i / c, cp = i - piece * c;
                const double* y = T.yw + static_cast<size_t>(i) * (2 + r);
                for (int ax = 0; ax < 3; ++ax) {
                    double val = y[0] * s[ax][0] + y[1] * s[ax][1];
                    for (int j = 0; j < r; ++j) val = fma(y[2 + j], vlast[ax][j], val);
                    out[piece * 3 * c + ax * c + cp] = val;
                }
            }
        }
    }
    if (a.ctrl_pts) {
        double* out = a.ctrl_pts + static_cast<size_t>(g) * n;
        for (int i = 0; i < m1; ++i) {
            const int piece = i / c, cp = i - piece * c;
            const double* y = T.yw + static_cast<size_t>(i) * (2 + r);
            for (int ax = 0; ax < 3; ++ax) {
                double val = 0.0;
                if (n_ok > 0) {
                    val = y[0] * s[ax][0] + y[1] * s[ax][1];
                    for (int j = 0; j < r; ++j) val = fma(y[2 + j], vlast[ax][j], val);
                }
                out[piece * 3 * c + ax * c + cp] = val;
            }
        }
    }
    if (a.n_curves) a.n_curves[g] = n_ok;
    if (a.objective) {
        double obj = 0.0;
        if (n_ok > 0) {
            double dd = 0.0, nn = 0.0;
            for (int ax = 0; ax < 3; ++ax)
                for (int j = 0; j < r; ++j) {
                    const double e = vlast[ax][j] - v0[ax][j];
                    dd = fma(e, e, dd);
                    nn = fma(v0[ax][j], v0[ax][j], nn);
                }
            obj = 0.5 * dd - 0.5 * nn + c0;
        }
        a.objective[g] = obj;
    }
}

__global__ void __launch_bounds__(64) impc_general_kernel(const __grid_constant__ GenTables T, const __grid_constant__ StepArgs a) {
    for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < a.num_qp; g += gridDim.x * blockDim.x) {
        const int scen = g / a.agent_count;
        const int self = a.agent_begin + (g - scen * a.agent_count);
        general_agent(T, a, g, self, a.states + static_cast<size_t>(scen) * T.N * 6);
    }
}

}  // namespace

#ifdef MPCCBF_HOST_EMUL
void emul_impc_general_host(const GenTables& T, const StepArgs& a) {
    for (int g = 0; g < a.num_qp; ++g) {
        const int scen = g / a.agent_count;
        const int self = a.agent_begin + (g - scen * a.agent_count);
        general_agent(T, a, g, self, a.states + static_cast<size_t>(scen) * T.N * 6);
    }
}
#endif

int launch_impc_general(const GenTables& T, const StepArgs& a, int sm_count, cudaStream_t stream) {
    if (a.num_qp <= 0) return cudaSuccess;
    if (T.r < 1 || T.r > kGenMaxR || T.H > 8) return cudaErrorInvalidValue;
    const int threads = 64;
    const int need = (a.num_qp + threads - 1) / threads;
    const int grid = std::min(need, std::max(1, sm_count) * 16);
    impc_general_kernel<<<grid, threads, 0, stream>>>(T, a);
    return cudaGetLastError();
}

}  // namespace mpccbf
