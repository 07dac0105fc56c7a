// EXPERIMENT ONLY (tools/build_variant.sh -DMPCCBF_WITH_R1): round 1 fused kernel (commit dff6a0c) in namespace mpccbf::r1,
// linked next to the current one for same-harness A/B runs (profiles/r2_kernel_ab.md). Not part of the product library.
// Fused IMPC kernel, thread-per-agent variant (throughput path) — sm_100a.
//
// One THREAD runs the whole body of ConnectivityIMPCCBF::optimize
// (mpc_cbf/src/controller/ConnectivityIMPCCBF.cpp:47-215) for one (scenario, agent); a warp owns
// 32 consecutive agents and its own slice of shared memory (no CTA-wide barriers after the
// constant tables are staged). Structure exploited (tables.hpp): in whitened null-space
// coordinates the QP is a projection, the yaw axis is decoupled from x/y (only the safety rows
// couple x and y), and the cost and the bound rows are the same in every IMPC iteration. So
//   1. yaw: 2-dof projection onto its acceleration/velocity rows (once, shared by all iterations);
//   2. x/y stage A: 4-dof projection onto the bound rows only (shared by all iterations);
//   3. per IMPC iteration: build the safety-HOCBF rows at the (predicted) ego states, keep the
//      ones that can be violated inside the acceleration box (exact pruning), and resume the
//      active-set iteration from stage A with those rows visible.
// Rows are recomputed from the neighbour states in shared memory every scan (nothing per-row is
// stored), so shared memory per agent is ~250 B whatever the swarm size.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "mpccbf.h"
#include "gi_solver_r1.h"
#include "kernels.h"

namespace mpccbf {
namespace r1 {

#ifndef MPCCBF_THREAD_MIN_CTAS
#define MPCCBF_THREAD_MIN_CTAS 4
#endif
constexpr int kCandMax = 32;   // candidate safety rows kept per agent (more: scan all rows)
constexpr int kSolStride = 13; // per-agent result record: s[6] (pos xyz, vel xyz), v[6], n_ok

#ifdef __CUDA_ARCH__
#define TLDG(p) __ldg(p)
#define TRSQRT(x) rsqrt(x)
#else
#define TLDG(p) (*(p))
#define TRSQRT(x) (1.0 / std::sqrt(x))
#endif

MPCCBF_HD void hocbf_row(double px, double py, double vx, double vy, double qx, double qy, double wx,
                         double wy, double dmin2, double gamma, double& ax, double& ay, double& b) {
    // closed form of ConnectivityCBF::initSafetyCBF (cbf/src/detail/ConnectivityCBF.cpp:152-198)
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

// one two-sided bound row: value `val`, normalisation `inv`; updates the arg-max
MPCCBF_HD void check_two_sided(double val, double lo, double hi, double inv, double tol, int id_base, double& best,
                               int& best_id) {
    const double vlo = lo - val, vhi = val - hi;
    const bool up = vhi > vlo;
    const double viol = up ? vhi : vlo;
    const double bound = up ? hi : lo;
    if (viol > tol * fmax(1.0, fabs(bound)) && fabs(bound) < 1e20) {
        const double score = viol * inv;
        if (score > best) {
            best = score;
            best_id = id_base + (up ? 1 : 0);
        }
    }
}

// Polynomial row scan of ONE axis (DevTables::poly): acc(t_k) = p0 + p1 t_k, vel(t_k) = p2 + p3 t_k +
// p4 t_k^2 with p_i linear in this axis' (v, s). The score of a two-sided row is its excess over the
// feasibility threshold in whitened distance, max(val - (hi + thr), (lo - thr) - val) * inv_norm;
// positive means violated. hiE / loE are the thresholded bounds (+-inf when there is no bound).
// A violated row that does not depend on v (inv_norm = inf) scores +inf: infeasible.
struct AxisPoly {
    double p[5];
};
MPCCBF_HD AxisPoly axis_poly(const double (*pq)[4], double va, double vb, double s0, double s1) {
    AxisPoly a;
#pragma unroll
    for (int i = 0; i < 5; ++i) a.p[i] = fma(pq[i][0], va, fma(pq[i][1], vb, fma(pq[i][2], s0, pq[i][3] * s1)));
    return a;
}
// Key scan. With finite two-sided bounds the score of a row is (|val - mid| - half) * inv_norm
// (mid / half of the thresholded bounds); it only serves to PICK the pivot, so it is compared through
// the high word of its FP64 pattern (monotone for positive doubles, negative for satisfied rows) with
// the horizon index packed into the low 7 bits: one integer max per row keeps both the maximum and the
// arg-max, and the acceleration / velocity families are independent dependency chains. Whether any row
// is violated is still decided exactly (sign of the score). scn[k] = {t_k, inv_norm_acc, inv_norm_vel,
// -}; rows that do not depend on v carry inv_norm = 1e300 (violated: key >= kKeyInf, infeasible).
constexpr int kKeyMask = ~127;
constexpr int kKeyInf = 0x69000000;  // high word of ~1e200
#ifdef __CUDA_ARCH__
#define HIWORD(x) __double2hiint(x)
#else
static inline int hiword_host(double x) {
    long long b;
    std::memcpy(&b, &x, 8);
    return static_cast<int>(b >> 32);
}
#define HIWORD(x) hiword_host(x)
#endif
MPCCBF_HD void scan_axis_keys(const double* scn, int K, const AxisPoly& a, double midA, double halfA, double midV,
                              double halfV, int& keyA, int& keyV) {
    const double a0 = a.p[0] - midA, v0 = a.p[2] - midV;
#pragma unroll 4
    for (int k = 0; k < K; ++k) {
        const double tk = scn[4 * k], ia = scn[4 * k + 1], iv = scn[4 * k + 2];
        const double sa = (fabs(fma(a.p[1], tk, a0)) - halfA) * ia;
        const double sv = (fabs(fma(fma(a.p[4], tk, a.p[3]), tk, v0)) - halfV) * iv;
        keyA = max(keyA, (HIWORD(sa) & kKeyMask) | k);
        keyV = max(keyV, (HIWORD(sv) & kKeyMask) | k);
    }
}
// both x/y axes in one pass over k: four independent chains, one load of scn[k]
MPCCBF_HD void scan_axes_keys2(const double* scn, int K, const AxisPoly& ax, const AxisPoly& ay, const double* lim, int (&key)[4]) {
    const double xa0 = ax.p[0] - lim[0], xv0 = ax.p[2] - lim[2];
    const double ya0 = ay.p[0] - lim[4], yv0 = ay.p[2] - lim[6];
    const double hxa = lim[1], hxv = lim[3], hya = lim[5], hyv = lim[7];
#pragma unroll 2
    for (int k = 0; k < K; ++k) {
        const double tk = scn[4 * k], ia = scn[4 * k + 1], iv = scn[4 * k + 2];
        const double s0 = (fabs(fma(ax.p[1], tk, xa0)) - hxa) * ia;
        const double s1 = (fabs(fma(fma(ax.p[4], tk, ax.p[3]), tk, xv0)) - hxv) * iv;
        const double s2 = (fabs(fma(ay.p[1], tk, ya0)) - hya) * ia;
        const double s3 = (fabs(fma(fma(ay.p[4], tk, ay.p[3]), tk, yv0)) - hyv) * iv;
        key[0] = max(key[0], (HIWORD(s0) & kKeyMask) | k);
        key[1] = max(key[1], (HIWORD(s1) & kKeyMask) | k);
        key[2] = max(key[2], (HIWORD(s2) & kKeyMask) | k);
        key[3] = max(key[3], (HIWORD(s3) & kKeyMask) | k);
    }
}
MPCCBF_HD double bound_hi_eff(double hi, double tol) { return fabs(hi) < 1e20 ? hi + tol * fmax(1.0, fabs(hi)) : HUGE_VAL; }
MPCCBF_HD double bound_lo_eff(double lo, double tol) { return fabs(lo) < 1e20 ? lo - tol * fmax(1.0, fabs(lo)) : -HUGE_VAL; }

// ---- yaw: acceleration and velocity rows of one axis, 2 dof ------------------------------------
struct RowsAxis {
    const double* tm;  // templates [3][K][kTmplStride]
    int K;
    double s0, s1;       // own pos, vel on this axis
    double lo[2], hi[2]; // [0] acceleration rows, [1] velocity rows
    double tol;
    const double (*pq)[4];  // polynomial row structure (DevTables::poly)
    const double* scn;      // [K][4] key-scan table
    const double* lim;      // {midA, halfA, midV, halfV} of this axis

    MPCCBF_HD void scan(const double (&v)[2], double& best, int& best_id) const {
        const AxisPoly a = axis_poly(pq, v[0], v[1], s0, s1);
        int keyA = 0, keyV = 0;
        scan_axis_keys(scn, K, a, lim[0], lim[1], lim[2], lim[3], keyA, keyV);
        const int key = max(keyA, keyV);
        if (key < 128) return;  // no row violated
        const int dsel = (keyV > keyA) ? 1 : 0, k = key & 127;
        best = (key >= kKeyInf) ? HUGE_VAL : 1.0;
        best_id = (dsel * K + k) << 1;
        // which side of the winning row is violated
        const double* t = tm + ((2 - dsel) * K + k) * kTmplStride;
        const double val = fma(t[0], v[0], fma(t[1], v[1], fma(t[2], s0, t[3] * s1)));
        if (val - hi[dsel] > lo[dsel] - val) best_id |= 1;
    }
    MPCCBF_HD void normal(int id, double (&np)[2], double& rhs) const {
        const bool up = id & 1;
        const int r = id >> 1;
        const int dsel = (r >= K) ? 1 : 0;
        const int k = r - dsel * K;
        const double* t = tm + ((2 - dsel) * K + k) * kTmplStride;
        const double sg = up ? -1.0 : 1.0;
        const double bound = up ? (dsel ? hi[1] : hi[0]) : (dsel ? lo[1] : lo[0]);
        np[0] = sg * t[0];
        np[1] = sg * t[1];
        rhs = sg * (bound - fma(t[2], s0, t[3] * s1));
    }
};

// ---- x/y: bound rows of both axes + the candidate safety rows, 4 dof -----------------------------
struct RowsXY {
    const double* tm;
    int K;
    double s[2][2];       // [axis][pos, vel]
    double lo[2][2], hi[2][2];  // [dsel][axis]
    double tol;
    // safety rows
    const double* st;     // scenario states (shared memory) [N][6]
    const double* pred;   // ego states per horizon step: pred[(k*4 + c) * pstride], c = px,py,vx,vy
    const unsigned short* cand;  // candidate row ids: cand[i * cstride]
    int pstride, cstride;
    int ncand;            // < 0: scan all `ncbf` rows
    int ncbf, hit, self;
    double dmin2, gamma;
    int nid_fix;          // 8K
    const double (*pq)[4];  // polynomial row structure (DevTables::poly)
    const double* scn;      // [K][4] key-scan table
    const double* lim;      // [axis x, y]{midA, halfA, midV, halfV}

    MPCCBF_HD void cbf_row(int j, double& ax, double& ay, double& b, int& k) const {
        const int nb = j / hit;
        k = j - nb * hit;
        const int other = nb + (nb >= self ? 1 : 0);
        const double* o = st + other * 6;
        const double* e = pred + (k * 4) * pstride;
        hocbf_row(e[0], e[pstride], e[2 * pstride], e[3 * pstride], o[0], o[1], o[3], o[4], dmin2, gamma, ax, ay, b);
    }
    MPCCBF_HD void check_cbf(int j, const double (&v)[4], int& key, int& jbest) const {
        double ax, ay, b;
        int k;
        cbf_row(j, ax, ay, b, k);
        const double* t = tm + (2 * K + k) * kTmplStride;
        const double accx = fma(t[0], v[0], fma(t[1], v[1], fma(t[2], s[0][0], t[3] * s[0][1])));
        const double accy = fma(t[0], v[2], fma(t[1], v[3], fma(t[2], s[1][0], t[3] * s[1][1])));
        const double viol = -(fma(ax, accx, fma(ay, accy, b)));
        if (viol > tol * fmax(1.0, fabs(b)) && fabs(b) < 1e20) {
            const double an2 = fma(ax, ax, ay * ay);
            const double score = (an2 > 0.0) ? viol * t[4] * TRSQRT(an2) : HUGE_VAL;
            const int kk = HIWORD(score) & kKeyMask;
            if (kk > key) {
                key = kk;
                jbest = j;
            }
        }
    }
    MPCCBF_HD void scan(const double (&v)[4], double& best, int& best_id) const {
        int key[4] = {0, 0, 0, 0};  // [ax * 2 + dsel]
        {
            const AxisPoly px = axis_poly(pq, v[0], v[1], s[0][0], s[0][1]);
            const AxisPoly py = axis_poly(pq, v[2], v[3], s[1][0], s[1][1]);
            scan_axes_keys2(scn, K, px, py, lim, key);
        }
        int kb = 0, jb = -1;
        if (ncand < 0) {
            for (int j = 0; j < ncbf; ++j) check_cbf(j, v, kb, jb);
        } else {
            for (int i = 0; i < ncand; ++i) check_cbf(cand[i * cstride], v, kb, jb);
        }
        const int kx = max(key[0], key[1]), ky = max(key[2], key[3]);
        const int kfix = max(kx, ky);
        if (kfix < 128 && jb < 0) return;  // feasible
        if (jb >= 0 && kb >= (kfix & kKeyMask)) {
            best = (kb >= kKeyInf) ? HUGE_VAL : 1.0;
            best_id = nid_fix + jb;
            return;
        }
        const int ax = (ky > kx) ? 1 : 0;
        const int dsel = ax ? ((key[3] > key[2]) ? 1 : 0) : ((key[1] > key[0]) ? 1 : 0);
        const int k = kfix & 127;
        best = (kfix >= kKeyInf) ? HUGE_VAL : 1.0;
        best_id = ((dsel * K + k) * 2 + ax) << 1;
        // which side of the winning row is violated
        const double* t = tm + ((2 - dsel) * K + k) * kTmplStride;
        const double va = ax ? v[2] : v[0], vb = ax ? v[3] : v[1];
        const double sa = ax ? s[1][0] : s[0][0], sb = ax ? s[1][1] : s[0][1];
        const double val = fma(t[0], va, fma(t[1], vb, fma(t[2], sa, t[3] * sb)));
        const double l = dsel ? (ax ? lo[1][1] : lo[1][0]) : (ax ? lo[0][1] : lo[0][0]);
        const double h = dsel ? (ax ? hi[1][1] : hi[1][0]) : (ax ? hi[0][1] : hi[0][0]);
        if (val - h > l - val) best_id |= 1;
    }
    MPCCBF_HD void normal(int id, double (&np)[4], double& rhs) const {
        if (id < nid_fix) {
            const bool up = id & 1;
            const int r = id >> 1;
            const int ax = r & 1;
            const int dk = r >> 1;
            const int dsel = (dk >= K) ? 1 : 0;
            const int k = dk - dsel * K;
            const double* t = tm + ((2 - dsel) * K + k) * kTmplStride;
            const double sg = up ? -1.0 : 1.0;
            const double bhi = dsel ? (ax ? hi[1][1] : hi[1][0]) : (ax ? hi[0][1] : hi[0][0]);
            const double blo = dsel ? (ax ? lo[1][1] : lo[1][0]) : (ax ? lo[0][1] : lo[0][0]);
            const double bound = up ? bhi : blo;
            const double sa = ax ? s[1][0] : s[0][0], sb = ax ? s[1][1] : s[0][1];
            const double n0 = sg * t[0], n1 = sg * t[1];
            np[0] = ax ? 0.0 : n0;
            np[1] = ax ? 0.0 : n1;
            np[2] = ax ? n0 : 0.0;
            np[3] = ax ? n1 : 0.0;
            rhs = sg * (bound - fma(t[2], sa, t[3] * sb));
        } else {
            double ax, ay, b;
            int k;
            cbf_row(id - nid_fix, ax, ay, b, k);
            const double* t = tm + (2 * K + k) * kTmplStride;
            np[0] = ax * t[0];
            np[1] = ax * t[1];
            np[2] = ay * t[0];
            np[3] = ay * t[1];
            const double offx = fma(t[2], s[0][0], t[3] * s[0][1]);
            const double offy = fma(t[2], s[1][0], t[3] * s[1][1]);
            rhs = -(b + ax * offx + ay * offy);
        }
    }
};

struct AgentCtx {
    const double *tm, *yw, *vr, *kr;  // constant tables (shared memory on the device)
    const double (*pq)[4];            // polynomial row structure (DevTables::poly)
    const double* scn;                // [K][4] key-scan table {t_k, inv_norm_acc, inv_norm_vel, -}
    const double* lim;                // [3 axes]{midA, halfA, midV, halfV} of the thresholded bounds
    const double* st;                 // this agent's scenario states [N][6]
    double* pred;                     // [(k*4 + c) * pstride]
    unsigned short* cand;             // [i * cstride]
    int pstride, cstride;
    double* sol;                      // [kSolStride] result record of this agent
};

// keeps the safety rows that can be violated somewhere inside the acceleration box (a row that
// holds on the whole box holds at every point the bound rows admit, so dropping it is exact)
MPCCBF_HD void build_candidates(RowsXY& rx, const DevTables& T, unsigned short* cand, int cstride) {
    int nc = 0;
    const bool wide = rx.ncbf > 65535;
    for (int j = 0; j < rx.ncbf && !wide; ++j) {
        double ax, ay, b;
        int k;
        rx.cbf_row(j, ax, ay, b, k);
        if (!(fabs(b) < 1e20)) continue;  // CPLEX: |bound| >= 1e20 is "no bound"
        const double mn = fmin(ax * T.lim_lo[0][0], ax * T.lim_hi[0][0]) + fmin(ay * T.lim_lo[0][1], ay * T.lim_hi[0][1]);
        if (!(mn + b > 1e-7 * (1.0 + fabs(ax) + fabs(ay)))) {
            if (nc < kCandMax) cand[nc * cstride] = static_cast<unsigned short>(j);
            ++nc;
        }
    }
    rx.ncand = (wide || nc > kCandMax) ? -1 : nc;
}

MPCCBF_HD double ctrl_point(const double* y, int ax, const double* sol) {
    // x[piece, ax, cp] = Y1 (pos, vel) + W1 (va, vb)
    return y[0] * sol[ax] + y[1] * sol[3 + ax] + y[2] * sol[6 + 2 * ax] + y[3] * sol[7 + 2 * ax];
}

// Everything optimize() does for agent-QP g (one thread).
__host__ __device__ void agent_thread(const DevTables& T, const StepArgs& a, int g, int self, const AgentCtx& c) {
    const int K = T.K, N = T.N, H = T.H;
    double s[3][2];
#pragma unroll
    for (int ax = 0; ax < 3; ++ax) {
        s[ax][0] = c.st[self * 6 + ax];
        s[ax][1] = c.st[self * 6 + 3 + ax];
    }
    // ---- unconstrained minimiser (whitened) and constant objective part ----------------------
    double v0[6], c0 = 0.0;
    if (a.ref_mode == 0) {
        const double* tg = a.targets + static_cast<size_t>(g) * 3;
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            const double tgt = TLDG(tg + ax);
            const double ps = s[ax][0], vs = s[ax][1];
            v0[2 * ax] = T.Vs[0][0] * ps + T.Vs[0][1] * vs + T.Vt[0] * tgt;
            v0[2 * ax + 1] = T.Vs[1][0] * ps + T.Vs[1][1] * vs + T.Vt[1] * tgt;
            c0 += ps * (T.Kc[0][0] * ps + T.Kc[0][1] * vs) + vs * (T.Kc[1][0] * ps + T.Kc[1][1] * vs) +
                  tgt * (T.Kt[0] * ps + T.Kt[1] * vs);
        }
    } else {
        const double* rf = a.targets + static_cast<size_t>(g) * 3 * K;
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            const double ps = s[ax][0], vs = s[ax][1];
            double r0 = 0.0, r1 = 0.0, k0 = 0.0, k1 = 0.0;
            for (int k = 0; k < K; ++k) {
                const double rk = TLDG(rf + 3 * k + ax);
                r0 = fma(c.vr[k], rk, r0);
                r1 = fma(c.vr[K + k], rk, r1);
                k0 = fma(c.kr[k], rk, k0);
                k1 = fma(c.kr[K + k], rk, k1);
            }
            v0[2 * ax] = T.Vs[0][0] * ps + T.Vs[0][1] * vs + r0;
            v0[2 * ax + 1] = T.Vs[1][0] * ps + T.Vs[1][1] * vs + r1;
            c0 += ps * (T.Kc[0][0] * ps + T.Kc[0][1] * vs) + vs * (T.Kc[1][0] * ps + T.Kc[1][1] * vs) + (k0 * ps + k1 * vs);
        }
    }
    int total_iter = 0;
    // ---- yaw: 2-dof projection, shared by every IMPC iteration --------------------------------
    GIState<2> gy;
    gi_init<2>(gy, v0 + 4);
    int base_status;
    {
        RowsAxis ry;
        ry.tm = c.tm;
        ry.K = K;
        ry.s0 = s[2][0];
        ry.s1 = s[2][1];
        ry.lo[0] = T.lim_lo[0][2]; ry.hi[0] = T.lim_hi[0][2];
        ry.lo[1] = T.lim_lo[1][2]; ry.hi[1] = T.lim_hi[1][2];
        ry.tol = T.feas_tol;
        ry.pq = c.pq;
        ry.scn = c.scn;
        ry.lim = c.lim + 8;
        base_status = gi_run<2>(gy, ry, T.max_gi, total_iter);
    }
    // ---- x/y stage A: bound rows only ---------------------------------------------------------
    RowsXY rx;
    rx.tm = c.tm;
    rx.K = K;
#pragma unroll
    for (int ax = 0; ax < 2; ++ax) {
        rx.s[ax][0] = s[ax][0];
        rx.s[ax][1] = s[ax][1];
#pragma unroll
        for (int d = 0; d < 2; ++d) {
            rx.lo[d][ax] = T.lim_lo[d][ax];
            rx.hi[d][ax] = T.lim_hi[d][ax];
        }
    }
    rx.tol = T.feas_tol;
    rx.st = c.st;
    rx.pred = c.pred;
    rx.cand = c.cand;
    rx.pstride = c.pstride;
    rx.cstride = c.cstride;
    rx.ncand = 0;
    rx.ncbf = 0;
    rx.hit = 1;
    rx.self = self;
    rx.dmin2 = T.dmin2;
    rx.gamma = T.gamma;
    rx.nid_fix = 8 * K;
    rx.pq = c.pq;
    rx.scn = c.scn;
    rx.lim = c.lim;
    GIState<4> gx;
    gi_init<4>(gx, v0);
    GIState<4> snap;  // stage-A factorisation; its address escapes so that it lives in local memory
#ifdef __CUDA_ARCH__
    asm volatile("" ::"l"(&snap) : "memory");
#endif

    // Phase machine with ONE gi_run<4> call site (the unrolled 4-dof active-set code is the bulk of
    // the kernel's instruction footprint; inlining it once keeps the kernel inside the instruction
    // cache):  kStageA -> kIter(0) -> kIter(1) -> ...   Before an iteration whose predecessor put
    // safety rows into the active set, the stage-A factorisation is restored from `snap`.
    enum { kStageA = 0, kIter = 2 };
    int n_ok = 0;
    bool dirty = false;  // gx has safety rows in its active set (not reusable as stage A)
    double vlast[6] = {0, 0, 0, 0, 0, 0};
    const size_t sbase = static_cast<size_t>(g) * T.iters;
    int mode = kStageA, it = 0, it_iter = 0;
    while (it < T.iters) {
        int stat = base_status;
        bool run = false;
        if (base_status == MPCCBF_STATUS_OPTIMAL) {
            if (mode == kIter) {
                // ego states the rows are built at: current state (iter 0), else the previous iterate
                // rolled out to t_k, k < H (ConnectivityIMPCCBF.cpp:135-141,160-175)
                const int hit = (it == 0) ? 1 : H;
                if (it == 0) {
                    c.pred[0] = s[0][0];
                    c.pred[c.pstride] = s[1][0];
                    c.pred[2 * c.pstride] = s[0][1];
                    c.pred[3 * c.pstride] = s[1][1];
                } else {
                    for (int k = 0; k < H; ++k) {
#pragma unroll
                        for (int d = 0; d < 2; ++d) {
                            const double* t = c.tm + (d * K + k) * kTmplStride;
#pragma unroll
                            for (int ax = 0; ax < 2; ++ax)
                                c.pred[(k * 4 + d * 2 + ax) * c.pstride] =
                                    fma(t[0], gx.v[2 * ax], fma(t[1], gx.v[2 * ax + 1], fma(t[2], s[ax][0], t[3] * s[ax][1])));
                        }
                    }
                }
                if (dirty) gx = snap;
                dirty = false;
                rx.hit = hit;
                rx.ncbf = (N - 1) * hit;
                build_candidates(rx, T, c.cand, c.cstride);
                run = rx.ncand != 0;
            } else {
                run = true;
            }
        }
        int n2 = 0;
        if (run) stat = gi_run<4>(gx, rx, T.max_gi, n2);
        it_iter += n2;
        if (mode == kStageA) {
            base_status = stat;
            it_iter += total_iter;  // yaw iterations are booked on IMPC iteration 0
            snap = gx;
            mode = kIter;
            continue;
        }
        dirty = n2 > 0;
        a.status[sbase + it] = stat;
        if (a.gi_iters) a.gi_iters[sbase + it] = it_iter;
        it_iter = 0;
        if (stat != MPCCBF_STATUS_OPTIMAL) {
            for (int r = it + 1; r < T.iters; ++r) {
                a.status[sbase + r] = MPCCBF_STATUS_SKIPPED;
                if (a.gi_iters) a.gi_iters[sbase + r] = 0;
            }
            break;
        }
        ++n_ok;
#pragma unroll
        for (int i = 0; i < 4; ++i) vlast[i] = gx.v[i];
        vlast[4] = gy.v[0];
        vlast[5] = gy.v[1];
        if (a.ctrl_all) {  // every iterate (mpccbf_optimize_one): small batches, direct stores
#pragma unroll
            for (int ax = 0; ax < 3; ++ax) {
                c.sol[ax] = s[ax][0];
                c.sol[3 + ax] = s[ax][1];
            }
#pragma unroll
            for (int i = 0; i < 6; ++i) c.sol[6 + i] = vlast[i];
            double* out = a.ctrl_all + (sbase + it) * T.n;
            for (int i = 0; i < T.n; ++i) {
                const int piece = i / (3 * T.c), rem = i - piece * 3 * T.c;
                const int ax = rem / T.c, cp = rem - ax * T.c;
                out[i] = ctrl_point(c.yw + (piece * T.c + cp) * 4, ax, c.sol);
            }
        }
        ++it;
    }
    // ---- result record (expanded to control points cooperatively by the warp) ------------------
#pragma unroll
    for (int ax = 0; ax < 3; ++ax) {
        c.sol[ax] = s[ax][0];
        c.sol[3 + ax] = s[ax][1];
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) c.sol[6 + i] = vlast[i];
    c.sol[12] = static_cast<double>(n_ok);
    if (a.n_curves) a.n_curves[g] = n_ok;
    if (a.objective) {
        double obj = 0.0;
        if (n_ok > 0) {
            double dd = 0.0, nn = 0.0;
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const double e = vlast[i] - v0[i];
                dd = fma(e, e, dd);
                nn = fma(v0[i], v0[i], nn);
            }
            obj = 0.5 * dd - 0.5 * nn + c0;
        }
        a.objective[g] = obj;
    }
}

// lim[axis] = {midA, halfA, midV, halfV} of the thresholded bounds; scn[k] = {t_k, inv_norm of the
// acceleration row, inv_norm of the velocity row, 0} with 1e300 standing for "row independent of v"
__host__ __device__ inline void fill_scan_tables(const DevTables& T, const double* tm, double* lim, double* scn, int tid, int nthr) {
    for (int i = tid; i < 6; i += nthr) {
        const int axis = i >> 1, dsel = i & 1;
        const double hi = bound_hi_eff(T.lim_hi[dsel][axis], T.feas_tol), lo = bound_lo_eff(T.lim_lo[dsel][axis], T.feas_tol);
        lim[axis * 4 + dsel * 2 + 0] = 0.5 * (hi + lo);
        lim[axis * 4 + dsel * 2 + 1] = 0.5 * (hi - lo);
    }
    for (int k = tid; k < T.K; k += nthr) {
        scn[4 * k + 0] = tm[(2 * T.K + k) * kTmplStride + 5];
        scn[4 * k + 1] = fmin(tm[(2 * T.K + k) * kTmplStride + 4], 1e300);
        scn[4 * k + 2] = fmin(tm[(1 * T.K + k) * kTmplStride + 4], 1e300);
        scn[4 * k + 3] = 0.0;
    }
}

struct ThreadSmemLayout {
    int tm, yw, vr, kr, vq, omap_bytes;  // doubles (omap in bytes), CTA-shared
    int st, pred, sol, cand_bytes;   // per warp: doubles (cand in bytes)
    int nscen_max;
    size_t shared_bytes, warp_bytes;
};

static ThreadSmemLayout thread_layout(const DevTables& T, int agent_count) {
    ThreadSmemLayout L;
    L.tm = 3 * T.K * kTmplStride;
    L.yw = T.m1 * 4;
    L.vr = 2 * T.K;
    L.kr = 2 * T.K;
    L.vq = 20 + 12 + 4 * T.K;  // pq, lim[3][4], scn[K][4]
    L.omap_bytes = ((T.n * 2 + 15) / 16) * 16;
    L.nscen_max = (32 - 1) / std::max(agent_count, 1) + 2;
    L.st = L.nscen_max * T.N * 6;
    L.pred = std::max(T.H, 1) * 4 * 32;
    L.sol = 32 * kSolStride;
    L.cand_bytes = kCandMax * 32 * 2;
    L.shared_bytes = sizeof(double) * (L.tm + L.yw + L.vr + L.kr + L.vq) + L.omap_bytes;
    L.warp_bytes = sizeof(double) * (L.st + L.pred + L.sol) + L.cand_bytes;
    return L;
}

__global__ void __launch_bounds__(128, MPCCBF_THREAD_MIN_CTAS) impc_thread_kernel(const DevTables T, const StepArgs a, const ThreadSmemLayout L) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* tm = reinterpret_cast<double*>(smem_raw);
    double* yw = tm + L.tm;
    double* vr = yw + L.yw;
    double* kr = vr + L.vr;
    double* svq = kr + L.kr;
    unsigned short* omap = reinterpret_cast<unsigned short*>(svq + L.vq);
    unsigned char* wbase = smem_raw + L.shared_bytes + static_cast<size_t>(threadIdx.x >> 5) * L.warp_bytes;
    double* wst = reinterpret_cast<double*>(wbase);
    double* wpred = wst + L.st;
    double* wsol = wpred + L.pred;
    unsigned short* wcand = reinterpret_cast<unsigned short*>(wsol + L.sol);

    for (int i = threadIdx.x; i < L.tm; i += blockDim.x) tm[i] = T.tmpl[i];
    for (int i = threadIdx.x; i < L.yw; i += blockDim.x) yw[i] = T.yw[i];
    for (int i = threadIdx.x; i < L.vr; i += blockDim.x) {
        vr[i] = T.Vr[i];
        kr[i] = T.Kr[i];
    }
    if (threadIdx.x < 20) svq[threadIdx.x] = T.pq[threadIdx.x >> 2][threadIdx.x & 3];
    fill_scan_tables(T, tm, svq + 20, svq + 32, threadIdx.x, blockDim.x);
    for (int i = threadIdx.x; i < T.n; i += blockDim.x) {
        const int piece = i / (3 * T.c), rem = i - piece * 3 * T.c;
        const int ax = rem / T.c, cp = rem - ax * T.c;
        omap[i] = static_cast<unsigned short>(((piece * T.c + cp) << 2) | ax);
    }
    __syncthreads();

    const int lane = threadIdx.x & 31;
    const int warps_per_cta = blockDim.x >> 5;
    const int gwarp = blockIdx.x * warps_per_cta + (threadIdx.x >> 5);
    const int total_warps = gridDim.x * warps_per_cta;
    const int ntiles = (a.num_qp + 31) >> 5;
    const int N = T.N, n = T.n;

    for (int tile = gwarp; tile < ntiles; tile += total_warps) {
        const int g0 = tile << 5;
        const int cnt = min(32, a.num_qp - g0);
        const int s_lo = g0 / a.agent_count;
        const int s_hi = (g0 + cnt - 1) / a.agent_count;
        {   // stage the tile's scenarios (coalesced 16-byte loads)
            const double2* src = reinterpret_cast<const double2*>(a.states + static_cast<size_t>(s_lo) * N * 6);
            double2* dst = reinterpret_cast<double2*>(wst);
            const int words = (s_hi - s_lo + 1) * N * 3;
            for (int i = lane; i < words; i += 32) dst[i] = __ldg(src + i);
        }
        __syncwarp();
        if (lane < cnt) {
            const int g = g0 + lane;
            const int scen = g / a.agent_count;
            const int self = a.agent_begin + (g - scen * a.agent_count);
            AgentCtx c;
            c.tm = tm; c.yw = yw; c.vr = vr; c.kr = kr;
            c.pq = reinterpret_cast<const double (*)[4]>(svq);
            c.lim = svq + 20;
            c.scn = svq + 32;
            c.st = wst + (scen - s_lo) * N * 6;
            c.pred = wpred + lane;
            c.pstride = 32;
            c.cand = wcand + lane;
            c.cstride = 32;
            c.sol = wsol + lane * kSolStride;
            agent_thread(T, a, g, self, c);
        }
        __syncwarp();
        {   // control points of the last successful iterate: coalesced stores, whole warp
            double* out = a.ctrl_pts + static_cast<size_t>(g0) * n;
            const int total = cnt * n;
            for (int idx = lane; idx < total; idx += 32) {
                const int ag = idx / n, i = idx - ag * n;
                const double* rec = wsol + ag * kSolStride;
                const unsigned m = omap[i];
                out[idx] = (rec[12] > 0.0) ? ctrl_point(yw + (m >> 2) * 4, m & 3, rec) : 0.0;
            }
        }
        __syncwarp();
    }
}



// Launches the thread-per-agent kernel. Returns a cudaError_t.
int launch_impc_thread_r1(const DevTables& T, const StepArgs& a, int sm_count, cudaStream_t stream) {
    if (a.num_qp <= 0) return cudaSuccess;
    const ThreadSmemLayout L = thread_layout(T, a.agent_count);
    int threads = 128;
    size_t smem = L.shared_bytes + L.warp_bytes * (threads / 32);
    while (smem > 100 * 1024 && threads > 32) {
        threads /= 2;
        smem = L.shared_bytes + L.warp_bytes * (threads / 32);
    }
    if (smem > 227 * 1024) return cudaErrorInvalidConfiguration;
    cudaError_t e = cudaFuncSetAttribute(impc_thread_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    int per_sm = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, impc_thread_kernel, threads, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) per_sm = 1;
    const int ntiles = (a.num_qp + 31) / 32;
    const int need = (ntiles + threads / 32 - 1) / (threads / 32);
    int grid = sm_count * per_sm;  // persistent: a multiple of the SM count
    if (need < grid) grid = need;
    impc_thread_kernel<<<grid, threads, smem, stream>>>(T, a, L);
    return cudaGetLastError();
}

}  // namespace r1
}  // namespace mpccbf
