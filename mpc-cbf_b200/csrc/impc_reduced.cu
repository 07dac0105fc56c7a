// Fused IMPC kernel, reduced (null-space) formulation — sm_100a.
//
// One lane-group of W lanes per (scenario, agent): it stages the scenario's agent states in
// shared memory, builds the pairwise safety-HOCBF rows, solves the 6-dof strictly convex QP by a
// dual active-set method, rolls the iterate out to the predicted ego states, rebuilds the CBF rows
// and solves again — i.e. the whole body of ConnectivityIMPCCBF::optimize
// (mpc_cbf/src/controller/ConnectivityIMPCCBF.cpp:47-215) without leaving the SM.
//
// Formulation (tables.hpp): with C^3 cubic pieces each axis has exactly two free dof after the
// equality rows; in whitened coordinates v (Hessian = I) the QP is the projection of the
// unconstrained minimiser v0 onto the polytope of bound rows and CBF rows:
//      min 1/2 |v - v0|^2   s.t.   lo <= Ct[d][k].v_ax + Ot[d][k].s_ax <= hi     (acc, vel rows)
//                                  ax*acc_x(k) + ay*acc_y(k) + b >= 0            (CBF rows)
// The active-set iteration keeps an orthogonal J (6x6) and triangular R of the active normals in
// registers, replicated across the group's lanes; lanes split the row scan and meet in a
// shuffle arg-max.
#include <cuda_runtime.h>
#include <cmath>

#include "../../include/mpccbf.h"
#include <algorithm>
#include <string>
#include <vector>

#include "gi_solver.h"
#include "kernels.h"

namespace mpccbf {

constexpr int R = 6;
#define kInf (HUGE_VAL)

template <int W>
struct Grp {
    __host__ __device__ static __forceinline__ unsigned mask(int lane) {
        if (W == 32) return 0xffffffffu;
        return ((1u << W) - 1u) << (lane & ~(W - 1));
    }
    __host__ __device__ static __forceinline__ void sync(unsigned m) {
#ifdef __CUDA_ARCH__
        if (W > 1) __syncwarp(m);
#endif
    }
};

// closed form of ConnectivityCBF::initSafetyCBF (cbf/src/detail/ConnectivityCBF.cpp:152-198):
// h = |dp|^2 - dmin^2, alpha(s) = gamma s^3, row  a.u + b >= 0 with a = (2dx, 2dy, 0).
__host__ __device__ __forceinline__ void safety_row(double px, double py, double vx, double vy, double qx,
                                           double qy, double wx, double wy, double dmin2,
                                           double gamma, double& ax, double& ay, double& b) {
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

struct QPData {
    const double* tm;    // shared: templates [3][K][kTmplStride]
    const double* cbf;   // shared: rows [ncbf][4] = {ax, ay, b, inv_norm}
    int K, ncbf, hit;
    double s[3][2];      // own (pos, vel) per axis
    double lo[2][3], hi[2][3];
    double tol;
};

// value of bound row r (0 <= r < 6K) at v, and its template pointer
__host__ __device__ __forceinline__ double fixed_row_value(const QPData& qp, int r, const double (&v)[R],
                                                  const double*& t, int& dsel, int& ax) {
    const int K3 = 3 * qp.K;
    dsel = (r >= K3) ? 1 : 0;          // 0: acceleration block, 1: velocity block
    const int rr = r - dsel * K3;
    const int k = rr / 3;
    ax = rr - 3 * k;
    t = qp.tm + ((2 - dsel) * qp.K + k) * kTmplStride;
    // static selection of the axis pair (keeps v in registers)
    const double va = (ax == 0) ? v[0] : (ax == 1) ? v[2] : v[4];
    const double vb = (ax == 0) ? v[1] : (ax == 1) ? v[3] : v[5];
    const double sa = (ax == 0) ? qp.s[0][0] : (ax == 1) ? qp.s[1][0] : qp.s[2][0];
    const double sb = (ax == 0) ? qp.s[0][1] : (ax == 1) ? qp.s[1][1] : qp.s[2][1];
    return t[0] * va + t[1] * vb + (t[2] * sa + t[3] * sb);
}

__host__ __device__ __forceinline__ double sel3(int ax, double a, double b, double c) {
    return (ax == 0) ? a : (ax == 1) ? b : c;
}

// Dual active-set (Goldfarb-Idnani scheme, least-distance form). Returns a SolveStatus code.
template <int W>
__host__ __device__ int solve_projection(const QPData& qp, const double (&v0)[R], double (&v)[R],
                                int max_iter, int& n_iter, unsigned gmask, int gl) {
    double J[R][R], Rm[R][R], u[R];
    int act[R];
#pragma unroll
    for (int i = 0; i < R; ++i) {
#pragma unroll
        for (int j = 0; j < R; ++j) {
            J[i][j] = (i == j) ? 1.0 : 0.0;
            Rm[i][j] = 0.0;
        }
        u[i] = 0.0;
        act[i] = -1;
        v[i] = v0[i];
    }
    int q = 0;
    n_iter = 0;
    const int nfix = 6 * qp.K;
    const int nid_fix = 2 * nfix;

    for (;;) {
        // ---- scan: most violated row (distance to its hyperplane in v-space) -------------------
        double best = 0.0;
        int best_id = -1;
        for (int r = gl; r < nfix; r += W) {
            const double* t;
            int dsel, ax;
            const double val = fixed_row_value(qp, r, v, t, dsel, ax);
            const double lo = sel3(ax, qp.lo[dsel][0], qp.lo[dsel][1], qp.lo[dsel][2]);
            const double hi = sel3(ax, qp.hi[dsel][0], qp.hi[dsel][1], qp.hi[dsel][2]);
            const double vlo = lo - val, vhi = val - hi;
            const bool up = vhi > vlo;
            const double viol = up ? vhi : vlo;
            const double bound = up ? hi : lo;
            if (viol > qp.tol * fmax(1.0, fabs(bound)) && fabs(bound) < 1e20) {
                const double score = viol * t[4];
                if (score > best) {
                    best = score;
                    best_id = 2 * r + (up ? 1 : 0);
                }
            }
        }
        if (qp.ncbf > 0) {
            for (int j = gl; j < qp.ncbf; j += W) {
                const int k = (qp.hit == 1) ? 0 : (j % qp.hit);
                const double* t = qp.tm + (2 * qp.K + k) * kTmplStride;
                const double accx = t[0] * v[0] + t[1] * v[1] + (t[2] * qp.s[0][0] + t[3] * qp.s[0][1]);
                const double accy = t[0] * v[2] + t[1] * v[3] + (t[2] * qp.s[1][0] + t[3] * qp.s[1][1]);
                const double* c = qp.cbf + 4 * j;
                const double viol = -(c[0] * accx + c[1] * accy + c[2]);
                if (viol > qp.tol * fmax(1.0, fabs(c[2])) && c[2] < 1e20) {
                    const double score = viol * c[3];
                    if (score > best) {
                        best = score;
                        best_id = nid_fix + j;
                    }
                }
            }
        }
#ifdef __CUDA_ARCH__
        if (W > 1) {
#pragma unroll
            for (int off = W / 2; off > 0; off >>= 1) {
                const double ob = __shfl_xor_sync(gmask, best, off, W);
                const int oi = __shfl_xor_sync(gmask, best_id, off, W);
                const bool take = (ob > best) || (ob == best && oi >= 0 && (best_id < 0 || oi < best_id));
                if (take) {
                    best = ob;
                    best_id = oi;
                }
            }
        }
#endif
        if (best_id < 0) return MPCCBF_STATUS_OPTIMAL;
        if (!(best < kInf)) return MPCCBF_STATUS_INFEASIBLE;  // a w-independent row is violated

        // ---- normal n+ and right-hand side of the chosen row:  n+.v >= rhs ---------------------
        double np[R] = {0, 0, 0, 0, 0, 0};
        double rhs;
        if (best_id < nid_fix) {
            const int r = best_id >> 1;
            const bool up = best_id & 1;
            const double* t;
            int dsel, ax;
            const double dummy = fixed_row_value(qp, r, v, t, dsel, ax);
            (void)dummy;
            const double sg = up ? -1.0 : 1.0;
            const double bound = up ? sel3(ax, qp.hi[dsel][0], qp.hi[dsel][1], qp.hi[dsel][2])
                                    : sel3(ax, qp.lo[dsel][0], qp.lo[dsel][1], qp.lo[dsel][2]);
            const double sa = sel3(ax, qp.s[0][0], qp.s[1][0], qp.s[2][0]);
            const double sb = sel3(ax, qp.s[0][1], qp.s[1][1], qp.s[2][1]);
            const double off = t[2] * sa + t[3] * sb;
            const double n0 = sg * t[0], n1 = sg * t[1];
            np[0] = (ax == 0) ? n0 : 0.0;
            np[1] = (ax == 0) ? n1 : 0.0;
            np[2] = (ax == 1) ? n0 : 0.0;
            np[3] = (ax == 1) ? n1 : 0.0;
            np[4] = (ax == 2) ? n0 : 0.0;
            np[5] = (ax == 2) ? n1 : 0.0;
            rhs = sg * (bound - off);
        } else {
            const int j = best_id - nid_fix;
            const int k = (qp.hit == 1) ? 0 : (j % qp.hit);
            const double* t = qp.tm + (2 * qp.K + k) * kTmplStride;
            const double* c = qp.cbf + 4 * j;
            np[0] = c[0] * t[0];
            np[1] = c[0] * t[1];
            np[2] = c[1] * t[0];
            np[3] = c[1] * t[1];
            const double offx = t[2] * qp.s[0][0] + t[3] * qp.s[0][1];
            const double offy = t[2] * qp.s[1][0] + t[3] * qp.s[1][1];
            rhs = -(c[2] + c[0] * offx + c[1] * offy);
        }

        double up_mult = 0.0;
        for (;;) {
            if (++n_iter > max_iter) return MPCCBF_STATUS_ERROR;
            // d = J' n+
            double d[R];
#pragma unroll
            for (int j = 0; j < R; ++j) {
                double sacc = 0.0;
#pragma unroll
                for (int i = 0; i < R; ++i) sacc = fma(J[i][j], np[i], sacc);
                d[j] = sacc;
            }
            // z = J2 d2 ; |d2|^2 ; |d|^2
            double z[R] = {0, 0, 0, 0, 0, 0};
            double ztn = 0.0, dall = 0.0;
#pragma unroll
            for (int j = 0; j < R; ++j) {
                dall = fma(d[j], d[j], dall);
                if (j >= q) {
                    ztn = fma(d[j], d[j], ztn);
#pragma unroll
                    for (int i = 0; i < R; ++i) z[i] = fma(J[i][j], d[j], z[i]);
                }
            }
            // rr = R^-1 d1
            double rr[R];
#pragma unroll
            for (int i = R - 1; i >= 0; --i) {
                double sacc = d[i];
#pragma unroll
                for (int k = i + 1; k < R; ++k)
                    if (k < q) sacc = fma(-Rm[i][k], rr[k], sacc);
                rr[i] = (i < q) ? sacc * GI_RCP(Rm[i][i]) : 0.0;
            }
            const double zt = 1e4 * 2.220446049250313e-16;
            const bool z_zero = !(ztn > zt * zt * dall);
            // ratio test by cross-multiplication (u_j >= 0, rr_j > 0); no called division routine in the step (gi_solver.h)
            double t1 = kInf, un = 0.0, rd = 0.0;
            int lpos = -1;
#pragma unroll
            for (int j = 0; j < R; ++j)
                if (j < q && rr[j] > 0.0) {
                    if (lpos < 0 || u[j] * rd < un * rr[j]) {
                        un = u[j];
                        rd = rr[j];
                        lpos = j;
                    }
                }
            if (lpos >= 0) t1 = un * GI_RCP(rd);
            double sp = -rhs;
#pragma unroll
            for (int i = 0; i < R; ++i) sp = fma(np[i], v[i], sp);
            const double t2 = z_zero ? kInf : (-sp * GI_RCP(ztn));
            const double t = fmin(t1, t2);
            if (!(t < kInf)) return MPCCBF_STATUS_INFEASIBLE;
            if (!z_zero) {
#pragma unroll
                for (int i = 0; i < R; ++i) v[i] = fma(t, z[i], v[i]);
            }
#pragma unroll
            for (int j = 0; j < R; ++j)
                if (j < q) u[j] = fma(-t, rr[j], u[j]);
            up_mult += t;

            if (!z_zero && t2 <= t1) {
                // full step: add the row. Rotate d2 onto e_q (Givens from the bottom), carry J2.
#pragma unroll
                for (int k = R - 1; k >= 1; --k) {
                    if (k > q) {
                        const double a = d[k - 1], b = d[k];
                        const double h2 = fma(a, a, b * b);
                        if (h2 > 0.0) {
                            const double ih = GI_RSQRT(h2), hh = h2 * ih;
                            const double cs = a * ih, sn = b * ih;
                            d[k - 1] = hh;
                            d[k] = 0.0;
#pragma unroll
                            for (int i = 0; i < R; ++i) {
                                const double t1j = J[i][k - 1], t2j = J[i][k];
                                J[i][k - 1] = fma(cs, t1j, sn * t2j);
                                J[i][k] = fma(-sn, t1j, cs * t2j);
                            }
                        }
                    }
                }
#pragma unroll
                for (int col = 0; col < R; ++col)
                    if (col == q) {
#pragma unroll
                        for (int i = 0; i < R; ++i) Rm[i][col] = (i <= col) ? d[i] : 0.0;
                        u[col] = up_mult;
                        act[col] = best_id;
                    }
                ++q;
                break;
            }
            // partial (or dual-only) step: drop active row lpos
#pragma unroll
            for (int col = 0; col < R - 1; ++col)
                if (col >= lpos) {
#pragma unroll
                    for (int i = 0; i < R; ++i) Rm[i][col] = Rm[i][col + 1];
                    u[col] = u[col + 1];
                    act[col] = act[col + 1];
                }
#pragma unroll
            for (int i = 0; i < R; ++i) Rm[i][R - 1] = 0.0;
            u[R - 1] = 0.0;
            act[R - 1] = -1;
            --q;
#pragma unroll
            for (int k = 0; k < R - 1; ++k)
                if (k >= lpos && k < q) {
                    const double a = Rm[k][k], b = Rm[k + 1][k];
                    const double h2 = fma(a, a, b * b);
                    if (h2 > 0.0) {
                        const double ih = GI_RSQRT(h2);
                        const double cs = a * ih, sn = b * ih;
#pragma unroll
                        for (int col = 0; col < R; ++col)
                            if (col >= k) {
                                const double t1r = Rm[k][col], t2r = Rm[k + 1][col];
                                Rm[k][col] = fma(cs, t1r, sn * t2r);
                                Rm[k + 1][col] = fma(-sn, t1r, cs * t2r);
                            }
#pragma unroll
                        for (int i = 0; i < R; ++i) {
                            const double t1j = J[i][k], t2j = J[i][k + 1];
                            J[i][k] = fma(cs, t1j, sn * t2j);
                            J[i][k + 1] = fma(-sn, t1j, cs * t2j);
                        }
                    }
                }
        }
    }
}

#ifdef __CUDA_ARCH__
#define MPCCBF_LDG(p) __ldg(p)
#else
#define MPCCBF_LDG(p) (*(p))
#endif

// Everything optimize() does for agent-QP `g`, executed by one lane group (lane gl of W).
// tm / yw / vr / kr: CTA-shared tables; st / rows / pred: the group's scratch.
template <int W>
__host__ __device__ void impc_agent(const DevTables& T, const StepArgs& a, int g, int gl, unsigned gmask,
                                    const double* tm, const double* yw, const double* vr, const double* kr,
                                    double* st, double* rows, double* pred) {
    const int K = T.K, N = T.N, H = T.H;
    QPData qp;
    qp.tm = tm;
    qp.cbf = rows;
    qp.K = K;
    qp.tol = T.feas_tol;
#pragma unroll
    for (int d = 0; d < 2; ++d)
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            qp.lo[d][ax] = T.lim_lo[d][ax];
            qp.hi[d][ax] = T.lim_hi[d][ax];
        }
    const int scen = g / a.agent_count;
    const int self = a.agent_begin + (g - scen * a.agent_count);
    // ---- stage the scenario's states (coalesced 16-byte loads) ----------------------------
    {
        const double2* src = reinterpret_cast<const double2*>(a.states + static_cast<size_t>(scen) * N * 6);
        double2* dst = reinterpret_cast<double2*>(st);
        for (int i = gl; i < N * 3; i += W) dst[i] = MPCCBF_LDG(src + i);
    }
    Grp<W>::sync(gmask);
#pragma unroll
    for (int ax = 0; ax < 3; ++ax) {
        qp.s[ax][0] = st[self * 6 + ax];
        qp.s[ax][1] = st[self * 6 + 3 + ax];
    }
    // ---- unconstrained minimiser and constant objective part ------------------------------
    double v0[R], c0 = 0.0;
    if (a.ref_mode == 0) {
        const double* tg = a.targets + static_cast<size_t>(g) * 3;
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            const double tgt = MPCCBF_LDG(tg + ax);
            const double ps = qp.s[ax][0], vs = qp.s[ax][1];
            v0[2 * ax] = T.Vs[0][0] * ps + T.Vs[0][1] * vs + T.Vt[0] * tgt;
            v0[2 * ax + 1] = T.Vs[1][0] * ps + T.Vs[1][1] * vs + T.Vt[1] * tgt;
            c0 += ps * (T.Kc[0][0] * ps + T.Kc[0][1] * vs) + vs * (T.Kc[1][0] * ps + T.Kc[1][1] * vs) +
                  tgt * (T.Kt[0] * ps + T.Kt[1] * vs);
        }
    } else {
        const double* rf = a.targets + static_cast<size_t>(g) * 3 * K;
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            const double ps = qp.s[ax][0], vs = qp.s[ax][1];
            double r0 = 0.0, r1 = 0.0, k0 = 0.0, k1 = 0.0;
            for (int k = 0; k < K; ++k) {
                const double rk = MPCCBF_LDG(rf + 3 * k + ax);
                r0 = fma(vr[k], rk, r0);
                r1 = fma(vr[K + k], rk, r1);
                k0 = fma(kr[k], rk, k0);
                k1 = fma(kr[K + k], rk, k1);
            }
            v0[2 * ax] = T.Vs[0][0] * ps + T.Vs[0][1] * vs + r0;
            v0[2 * ax + 1] = T.Vs[1][0] * ps + T.Vs[1][1] * vs + r1;
            c0 += ps * (T.Kc[0][0] * ps + T.Kc[0][1] * vs) + vs * (T.Kc[1][0] * ps + T.Kc[1][1] * vs) +
                  (k0 * ps + k1 * vs);
        }
    }

    double v[R], vlast[R] = {0, 0, 0, 0, 0, 0};
    int n_ok = 0;
    for (int it = 0; it < T.iters; ++it) {
        const int hit = (it == 0) ? 1 : H;
        const int ncbf = (N - 1) * hit;
        // ---- CBF rows: lane j <-> (neighbour, horizon step) -------------------------------
        for (int j = gl; j < ncbf; j += W) {
            const int nb = j / hit, k = j - nb * hit;
            const int other = nb + (nb >= self ? 1 : 0);
            double px, py, vx, vy;
            if (it == 0) {
                px = qp.s[0][0]; py = qp.s[1][0]; vx = qp.s[0][1]; vy = qp.s[1][1];
            } else {
                px = pred[k * 6 + 0]; py = pred[k * 6 + 1]; vx = pred[k * 6 + 3]; vy = pred[k * 6 + 4];
            }
            const double* o = st + other * 6;
            double ax, ay, b;
            safety_row(px, py, vx, vy, o[0], o[1], o[3], o[4], T.dmin2, T.gamma, ax, ay, b);
            const double* t = tm + (2 * K + k) * kTmplStride;
            const double an2 = (ax * ax + ay * ay) * (t[0] * t[0] + t[1] * t[1]);
            double* c = rows + 4 * j;
            c[0] = ax;
            c[1] = ay;
            c[2] = b;
            c[3] = (an2 > 0.0) ? GI_RSQRT(an2) : kInf;   // only ranks the violated rows (whitened distance)
        }
        Grp<W>::sync(gmask);
        qp.ncbf = ncbf;
        qp.hit = hit;
        int n_iter = 0;
        const int stat = solve_projection<W>(qp, v0, v, T.max_gi, n_iter, gmask, gl);
        if (gl == 0) {
            a.status[static_cast<size_t>(g) * T.iters + it] = stat;
            if (a.gi_iters) a.gi_iters[static_cast<size_t>(g) * T.iters + it] = n_iter;
        }
        if (stat != MPCCBF_STATUS_OPTIMAL) {
            if (gl == 0)
                for (int r = it + 1; r < T.iters; ++r) {
                    a.status[static_cast<size_t>(g) * T.iters + r] = MPCCBF_STATUS_SKIPPED;
                    if (a.gi_iters) a.gi_iters[static_cast<size_t>(g) * T.iters + r] = 0;
                }
            break;
        }
        ++n_ok;
#pragma unroll
        for (int i = 0; i < R; ++i) vlast[i] = v[i];
        if (a.ctrl_all) {
            double* out = a.ctrl_all + (static_cast<size_t>(g) * T.iters + it) * T.n;
            for (int i = gl; i < T.n; i += W) {
                const int piece = i / (3 * T.c), rem = i - piece * 3 * T.c;
                const int ax = rem / T.c, cp = rem - ax * T.c;
                const double* y = yw + (piece * T.c + cp) * 4;
                out[i] = fma(y[0], sel3(ax, qp.s[0][0], qp.s[1][0], qp.s[2][0]),
                             fma(y[1], sel3(ax, qp.s[0][1], qp.s[1][1], qp.s[2][1]),
                                 fma(y[2], sel3(ax, v[0], v[2], v[4]), y[3] * sel3(ax, v[1], v[3], v[5]))));
            }
        }
        if (it + 1 < T.iters) {
            // ---- rollout: predicted ego states at t_k, k < H (ConnectivityIMPCCBF.cpp:160-168)
            Grp<W>::sync(gmask);
            for (int e = gl; e < H * 6; e += W) {
                const int k = e / 6, comp = e - 6 * k;
                const int d = comp / 3, ax = comp - 3 * d;
                const double* t = tm + (d * K + k) * kTmplStride;
                pred[e] = t[0] * sel3(ax, v[0], v[2], v[4]) + t[1] * sel3(ax, v[1], v[3], v[5]) +
                          t[2] * sel3(ax, qp.s[0][0], qp.s[1][0], qp.s[2][0]) +
                          t[3] * sel3(ax, qp.s[0][1], qp.s[1][1], qp.s[2][1]);
            }
            Grp<W>::sync(gmask);
        }
    }
    // ---- outputs --------------------------------------------------------------------------
    {
        if (a.v_out && gl == 0) {
#pragma unroll
            for (int i = 0; i < R; ++i) a.v_out[static_cast<size_t>(g) * R + i] = (n_ok > 0) ? vlast[i] : 0.0;
        }
        double* out = a.ctrl_pts + static_cast<size_t>(g) * T.n;
        for (int i = gl; a.ctrl_pts && i < T.n; i += W) {
            double val = 0.0;
            if (n_ok > 0) {
                const int piece = i / (3 * T.c), rem = i - piece * 3 * T.c;
                const int ax = rem / T.c, cp = rem - ax * T.c;
                const double* y = yw + (piece * T.c + cp) * 4;
                val = fma(y[0], sel3(ax, qp.s[0][0], qp.s[1][0], qp.s[2][0]),
                          fma(y[1], sel3(ax, qp.s[0][1], qp.s[1][1], qp.s[2][1]),
                              fma(y[2], sel3(ax, vlast[0], vlast[2], vlast[4]), y[3] * sel3(ax, vlast[1], vlast[3], vlast[5]))));
            }
            out[i] = val;
        }
        if (gl == 0) {
            if (a.n_curves) a.n_curves[g] = n_ok;
            if (a.objective) {
                double obj = 0.0;
                if (n_ok > 0) {
                    double dd = 0.0, nn = 0.0;
#pragma unroll
                    for (int i = 0; i < R; ++i) {
                        const double e = vlast[i] - v0[i];
                        dd = fma(e, e, dd);
                        nn = fma(v0[i], v0[i], nn);
                    }
                    obj = 0.5 * dd - 0.5 * nn + c0;
                }
                a.objective[g] = obj;
            }
        }
    }
}

// shared memory per CTA:  tmpl [3*K*6] | yw [m1*4] | vr [2K] | kr [2K] | per group: states [N*6], rows [(N-1)*H*4], pred [H*6]
template <int W>
__global__ void __launch_bounds__(128) impc_reduced_kernel(const DevTables T, const StepArgs a) {
    extern __shared__ double smem[];
    const int K = T.K, N = T.N, H = T.H, m1 = T.m1;
    double* tm = smem;
    double* yw = tm + 3 * K * kTmplStride;
    double* vr = yw + m1 * 4;
    double* kr = vr + 2 * K;
    double* gbase = kr + 2 * K;
    const int per_group = N * 6 + (N - 1) * H * 4 + H * 6;
    const int groups_per_cta = blockDim.x / W;
    const int gid = threadIdx.x / W;
    const int gl = threadIdx.x % W;
    const int lane = threadIdx.x & 31;
    const unsigned gmask = Grp<W>::mask(lane);
    double* st = gbase + gid * per_group;
    double* rows = st + N * 6;
    double* pred = rows + (N - 1) * H * 4;

    for (int i = threadIdx.x; i < 3 * K * kTmplStride; i += blockDim.x) tm[i] = T.tmpl[i];
    for (int i = threadIdx.x; i < m1 * 4; i += blockDim.x) yw[i] = T.yw[i];
    for (int i = threadIdx.x; i < 2 * K; i += blockDim.x) {
        vr[i] = T.Vr[i];
        kr[i] = T.Kr[i];
    }
    __syncthreads();

    const int total_groups = gridDim.x * groups_per_cta;
    for (int g = blockIdx.x * groups_per_cta + gid; g < a.num_qp; g += total_groups) {
        impc_agent<W>(T, a, g, gl, gmask, tm, yw, vr, kr, st, rows, pred);
        Grp<W>::sync(gmask);  // group smem is reused by the next agent
    }
}

#ifdef MPCCBF_HOST_EMUL
// Test-only (tests/emul): runs the SAME per-agent function on the host, one lane per QP, so the
// CPU-only CI can check the active-set logic against the oracle. Never compiled into libmpccbf.so.
void emul_impc_reduced_host(const DevTables& T, const StepArgs& a) {
    std::vector<double> st(static_cast<size_t>(T.N) * 6), rows(static_cast<size_t>(std::max(1, T.N - 1)) * T.H * 4),
        pred(static_cast<size_t>(T.H) * 6);
    for (int g = 0; g < a.num_qp; ++g)
        impc_agent<1>(T, a, g, 0, 1u, T.tmpl, T.yw, T.Vr, T.Kr, st.data(), rows.data(), pred.data());
}
#endif

size_t impc_reduced_smem_bytes(const DevTables& T, int groups_per_cta) {
    const size_t shared_tables = static_cast<size_t>(3) * T.K * kTmplStride + static_cast<size_t>(T.m1) * 4 + 4 * T.K;
    const size_t per_group = static_cast<size_t>(T.N) * 6 + static_cast<size_t>(T.N - 1) * T.H * 4 + static_cast<size_t>(T.H) * 6;
    return sizeof(double) * (shared_tables + per_group * groups_per_cta);
}

// Launches the fused kernel. `width` = lanes per QP (32, 8 or 1). Returns a cudaError_t.
int launch_impc_reduced(const DevTables& T, const StepArgs& a, int width, int sm_count, cudaStream_t stream) {
    if (a.num_qp <= 0) return cudaSuccess;
    int threads = 128;
    auto pick = [&](auto kernel, int W) -> int {
        int gpc = threads / W;
        size_t smem = impc_reduced_smem_bytes(T, gpc);
        while (smem > 200 * 1024 && threads > W && threads > 32) {
            threads /= 2;
            gpc = threads / W;
            smem = impc_reduced_smem_bytes(T, gpc);
        }
        if (smem > 227 * 1024) return cudaErrorInvalidConfiguration;
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return e;
        int per_sm = 1;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) per_sm = 1;
        const int need = (a.num_qp + gpc - 1) / gpc;
        int grid = sm_count * per_sm;  // persistent: a multiple of the SM count
        if (need < grid) grid = need;
        kernel<<<grid, threads, smem, stream>>>(T, a);
        return cudaGetLastError();
    };
    switch (width) {
        case 32: return pick(impc_reduced_kernel<32>, 32);
        case 8: return pick(impc_reduced_kernel<8>, 8);
        case 1: return pick(impc_reduced_kernel<1>, 1);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace mpccbf
