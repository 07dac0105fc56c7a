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
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/mpccbf.h"
#include "gi_solver.h"
#include "kernels.h"

namespace mpccbf {

#ifndef MPCCBF_THREAD_CORE
#define MPCCBF_THREAD_CORE 0
#endif
#ifndef MPCCBF_THREAD_CFG
#define MPCCBF_THREAD_CFG 4
#endif
#ifndef MPCCBF_CAND_MAX
#define MPCCBF_CAND_MAX 16
#endif
constexpr int kCandMaxSmall = MPCCBF_CAND_MAX;   // candidate safety rows kept per agent (more: scan all rows) for swarms of <= 32
constexpr int kCandMaxLarge = 32;                // agents, and for larger ones (their scenario states dominate the block anyway)
constexpr int kSolStride = 7;  // per-agent result record: v[6] (whitened free dof of the last successful iterate), n_ok
constexpr int kNcRows = 16;    // normals of the <= 4 active rows of the x / y problem (gi_solver.h: GIState::nc)
// Per-agent scratch is LANE-STRIDED: element i of an array belongs to lane l at [i * kLS + l] (doubles), and the arrays
// of a warp follow each other at compile-time offsets, so every access is one per-thread base plus an immediate (the
// tests' host build uses the same layout with lane 0).
constexpr int kLS = 32;
static_assert(kLS == kNcStride, "the normal cache of gi_solver.h is lane-strided like the rest");
constexpr int kRow = kLS * 8;                          // bytes of one lane-strided row
constexpr int kOffSol = 0;
constexpr int kOffNc = kOffSol + kSolStride * kRow;
constexpr int kOffPred = kOffNc + kNcRows * kRow;      // H * 4 rows; then the candidate list (cand_max rows of kLS unsigned shorts),
                                                       // (conn_mode) H * 3 rows, Fiedler vectors; scenario states
// CTA-shared tables at compile-time offsets from the start of shared memory (doubles): pq[5][4], lim[3][8], scn[K][4];
// the horizon-sized tables follow (ThreadSmemLayout)
constexpr int kOffLim = 20, kOffScn = 44;

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

// Distance CLF row as the MPC-lifted reference inserts it (ConnectivityMPCCBFQPOperations.cpp:72-86, :152-173 with
// initCLFCBF cbf/src/detail/ConnectivityCBF.cpp:200-243): V = (|p - p_n| - 2)^2, row  -(a.u) <= -Bc  with a = grad V,
// Bc = Lf^2 V + 5 Lf V + 2 V (only the EGO velocity enters). Returned in the common form a.u + b >= 0, b = -Bc.
MPCCBF_HD void clf_row_lifted(double px, double py, double vx, double vy, double qx, double qy, double& ax, double& ay, double& b) {
    const double dx = px - qx, dy = py - qy;
    const double d2 = dx * dx + dy * dy, dist = sqrt(d2), err = dist - 2.0;
    const double i1 = 1.0 / dist, i2 = 1.0 / d2, i3 = i1 * i2;
    ax = 2.0 * err * dx * i1;
    ay = 2.0 * err * dy * i1;
    const double hxx = 2.0 * (dx * dx * i2 + err * (i1 - dx * dx * i3));
    const double hxy = 2.0 * (dx * dy * i2 + err * (-dx * dy * i3));
    const double hyy = 2.0 * (dy * dy * i2 + err * (i1 - dy * dy * i3));
    const double lfv = ax * vx + ay * vy;
    const double lf2v = vx * (hxx * vx + hxy * vy) + vy * (hxy * vx + hyy * vy);
    b = -(lf2v + 5.0 * lfv + 2.0 * err * err);
}

// Connectivity HOCBF row of agent `self` with its own state replaced by (px, py, vx, vy)
// (ConnectivityMPCCBFQPOperations.cpp:45-70, :111-150; initConnCBF / compute_full_grad_h
// cbf/src/detail/ConnectivityCBF.cpp:439-519 in closed form): h = lambda2 - 0.1, gradient and Hessian w.r.t. the
// agent's own position with the eigenvector fv held constant, no range gating, alpha linear with gain gamma.
MPCCBF_HD void conn_row_lifted(const double* st, int N, int self, const double* fv, double lam2, double dmax, double gamma,
                               double px, double py, double vx, double vy, double& gx, double& gy, double& b) {
    const double R2 = dmax * dmax, sigma = R2 * R2 / 0.6931471805599453;
    gx = 0.0;
    gy = 0.0;
    double hxx = 0.0, hxy = 0.0, hyy = 0.0;
    for (int j = 0; j < N; ++j) {
        if (j == self) continue;
        const double dx = px - st[j * 6], dy = py - st[j * 6 + 1];
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
    const double lfh = gx * vx + gy * vy;
    const double lf2h = vx * (hxx * vx + hxy * vy) + vy * (hxy * vx + hyy * vy);
    b = lf2h + gamma * lfh + gamma * (lfh + gamma * (lam2 - 0.1));
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
// The scan does not walk the horizon. Along the horizon the acceleration row value is affine in t_k and
// the velocity row value quadratic, and the thresholded bounds are the same at every k, so
// max_k |val_k - mid| is attained at an end sample or (velocity) at a sample next to the parabola's
// vertex: "is any row violated" is decided exactly by 2 + 5 samples per axis instead of 2 K. Only a velocity
// family that has a violated row is then searched along the horizon for the pivot (largest violation in whitened
// distance).
// scn[k] = {t_k, inv_norm_acc, inv_norm_vel, -}; kinv = (K - 1) / t_{K-1} maps a time to a sample index.
MPCCBF_HD int key_acc(const double* scn, int k, const AxisPoly& a, double a0, double halfA) {
    const double tk = scn[4 * k], ia = scn[4 * k + 1];
    const double sa = (fabs(fma(a.p[1], tk, a0)) - halfA) * ia;
    return (HIWORD(sa) & kKeyMask) | k;
}
MPCCBF_HD int key_vel(const double* scn, int k, const AxisPoly& a, double v0, double halfV) {
    const double tk = scn[4 * k], iv = scn[4 * k + 2];
    const double sv = (fabs(fma(fma(a.p[4], tk, a.p[3]), tk, v0)) - halfV) * iv;
    return (HIWORD(sv) & kKeyMask) | k;
}
// sample index next to the vertex of vel(t) = p2 + p3 t + p4 t^2, clamped to [1, K - 2] (float precision
// is ample: the three samples kv - 1, kv, kv + 1 are all evaluated)
MPCCBF_HD int vertex_sample(const AxisPoly& a, float kinv, int K) {
    const float tv = -0.5f * static_cast<float>(a.p[3]) / static_cast<float>(a.p[4]);  // +-inf / NaN when p4 = 0
    const float kf = fminf(fmaxf(tv * kinv, 1.0f), static_cast<float>(K - 2));         // NaN -> 1
    return static_cast<int>(kf + 0.5f);
}
MPCCBF_HD void scan_axis_keys(const double* scn, int K, float kinv, const AxisPoly& a, double midA, double halfA, double midV,
                              double halfV, int& keyA, int& keyV) {
    const double a0 = a.p[0] - midA, v0 = a.p[2] - midV;
    int kA = max(key_acc(scn, 0, a, a0, halfA), key_acc(scn, K - 1, a, a0, halfA));
    const int kv = vertex_sample(a, kinv, K);
    int kV = max(key_vel(scn, 0, a, v0, halfV), key_vel(scn, K - 1, a, v0, halfV));
    kV = max(kV, max(key_vel(scn, kv - 1, a, v0, halfV), max(key_vel(scn, kv, a, v0, halfV), key_vel(scn, kv + 1, a, v0, halfV))));
    // A velocity family with a violated row: its pivot candidate is the row with the largest violation in WHITENED
    // distance over the whole horizon — the row norms grow along the horizon, and chasing the largest raw violation
    // (which sits at the vertex) costs 74 % more active-set steps (profiles/r2_warm_start_and_pivot_rule.md). The score is
    // smooth in k, so a stride-s walk refined around its best sample finds the same pivots as the full walk
    // (measured: identical step counts). The acceleration family is affine: its end samples suffice (measured likewise).
    // Families without a violated row — the common case, and every final scan — stop at the 2 + 5 samples above.
    if (kV >= 128) {
        const int s = K >= 30 ? 5 : 3;
#pragma unroll 2
        for (int k = s; k < K - 1; k += s) kV = max(kV, key_vel(scn, k, a, v0, halfV));
        const int kb = kV & 127;
        for (int dk = 1; dk < s; ++dk) {
            if (kb - dk >= 0) kV = max(kV, key_vel(scn, kb - dk, a, v0, halfV));
            if (kb + dk < K) kV = max(kV, key_vel(scn, kb + dk, a, v0, halfV));
        }
    }
    keyA = max(keyA, kA);
    keyV = max(keyV, kV);
}
// both x/y axes: key[ax * 2 + dsel]
MPCCBF_HD void scan_axes_keys2(const double* scn, int K, float kinv, const AxisPoly& ax, const AxisPoly& ay, const double* lim, int (&key)[4]) {
    scan_axis_keys(scn, K, kinv, ax, lim[0], lim[1], lim[2], lim[3], key[0], key[1]);
    scan_axis_keys(scn, K, kinv, ay, lim[8], lim[9], lim[10], lim[11], key[2], key[3]);
}
MPCCBF_HD double bound_hi_eff(double hi, double tol) { return fabs(hi) < 1e20 ? hi + tol * fmax(1.0, fabs(hi)) : HUGE_VAL; }
MPCCBF_HD double bound_lo_eff(double lo, double tol) { return fabs(lo) < 1e20 ? lo - tol * fmax(1.0, fabs(lo)) : -HUGE_VAL; }

// ---- yaw: acceleration and velocity rows of one axis, 2 dof ------------------------------------
struct RowsAxis {
    const double* tm;  // templates [3][K][kTmplStride]
    int K;
    double s0, s1;       // own pos, vel on this axis
    double tol;
    const double (*pq)[4];  // polynomial row structure (DevTables::poly)
    const double* scn;      // [K][4] key-scan table
    float kinv;             // (K - 1) / t_{K-1}
    const double* lim;      // this axis: {midA, halfA, midV, halfV, loA, hiA, loV, hiV} (mid / half of the thresholded bounds)

    MPCCBF_HD void scan(const double (&v)[2], double& best, int& best_id) const {
        const AxisPoly a = axis_poly(pq, v[0], v[1], s0, s1);
        int keyA = 0, keyV = 0;
        scan_axis_keys(scn, K, kinv, a, lim[0], lim[1], lim[2], lim[3], keyA, keyV);
        const int key = max(keyA, keyV);
        if (key < 128) return;  // no row violated
        const int dsel = (keyV > keyA) ? 1 : 0, k = key & 127;
        best = (key >= kKeyInf) ? HUGE_VAL : 1.0;
        best_id = (dsel * K + k) << 1;
        // which side of the winning row is violated
        const double* t = tm + ((2 - dsel) * K + k) * kTmplStride;
        const double val = fma(t[0], v[0], fma(t[1], v[1], fma(t[2], s0, t[3] * s1)));
        if (val - lim[5 + 2 * dsel] > lim[4 + 2 * dsel] - val) best_id |= 1;
    }
    MPCCBF_HD void normal(int id, double (&np)[2], double& rhs) const {
        const bool up = id & 1;
        const int r = id >> 1;
        const int dsel = (r >= K) ? 1 : 0;
        const int k = r - dsel * K;
        const double* t = tm + ((2 - dsel) * K + k) * kTmplStride;
        const double sg = up ? -1.0 : 1.0;
        const double bound = lim[4 + 2 * dsel + (up ? 1 : 0)];
        np[0] = sg * t[0];
        np[1] = sg * t[1];
        rhs = sg * (bound - fma(t[2], s0, t[3] * s1));
    }
};

// slack_cost * slack_decay^rank, rank of neighbour j among the J neighbours by distance to the ego agent
// (closest first, ties by index; ConnectivityIMPCCBF.cpp:76-100). Near-ties decide ranks in symmetric
// formations: the same double arithmetic as the host code of the reference (separately rounded products
// and sum, no fma), exact ties by index.
#ifdef __CUDA_ARCH__
#define TSQSUM(x, y) __dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y))
#else
static inline double tsqsum_host(double x, double y) {
    volatile double a = x * x, b = y * y;
    return a + b;
}
#define TSQSUM(x, y) tsqsum_host(x, y)
#endif
MPCCBF_HD double slack_weight(const double* st, int self, int j, int J, double cost, double decay) {
    const double* me = st + self * 6;
    const int other = j + (j >= self ? 1 : 0);
    const double dx = st[other * 6] - me[0], dy = st[other * 6 + 1] - me[1];
    const double dj = sqrt(TSQSUM(dx, dy));
    int rank = 0;
    for (int i = 0; i < J; ++i) {
        const int oi = i + (i >= self ? 1 : 0);
        const double ex = st[oi * 6] - me[0], ey = st[oi * 6 + 1] - me[1];
        const double di = sqrt(TSQSUM(ex, ey));
        rank += (di < dj || (di == dj && i < j)) ? 1 : 0;
    }
    return cost * pow(decay, static_cast<double>(rank));
}

// ---- x/y: bound rows of both axes + the candidate safety rows, 4 dof -----------------------------
// SOFT (slack mode): the safety rows of neighbour nb form a group with one slack (gi_solver.h). Row ids
// are then nid_fix + nb * (hit + 1) + kk with kk < hit the row of horizon step kk and kk == hit the
// piece "none"; lead[nb] = 0 while the group's slack is zero, else 1 + the horizon step of its leader.
template <bool SOFT, bool CONN>
struct RowsXY {
    const double* tm;
    int K;
    double s[2][2];       // [axis][pos, vel]
    double tol;
    // safety rows
    const double* st;     // scenario states (shared memory) [N][6]
    const double* pred;   // ego states per horizon step: pred[(k*4 + c) * kLS], c = px,py,vx,vy
    const unsigned short* cand;  // candidate row ids: cand[i * kLS]
    int ncand;            // < 0: scan all `ncbf` rows
    int cand_max;         // capacity of the candidate list
    int ncbf, hit, self;  // ncbf pairwise rows: [0, nsafe) safety HOCBF, [nsafe, ncbf) distance CLF (conn_mode, lambda2 <= 0.1)
    int nsafe;            // (N - 1) * hit
    int nconn;            // connectivity HOCBF rows (conn_mode, lambda2 > 0.1): row ids ncbf + k, values in conn[]
    const double* conn;   // conn[(k * 3 + i) * kLS] = (gx, gy, b) of horizon step k
    double dmin2, gamma;
    int nid_fix;          // 8K
    const double (*pq)[4];  // polynomial row structure (DevTables::poly)
    const double* scn;      // [K][4] key-scan table
    float kinv;             // (K - 1) / t_{K-1}
    const double* lim;      // [axis x, y][8] = {midA, halfA, midV, halfV, loA, hiA, loV, hiV}
    // slack mode
    unsigned char* lead;  // [N-1] leaders of this agent's groups
    int nnb;              // N - 1
    double slack_cost, slack_decay;

    static constexpr bool kCacheNormals = true;   // gi_solver.h keeps the active rows' normals in GIState::nc

    MPCCBF_HD void cbf_row(int j, double& ax, double& ay, double& b, int& k) const {
        if constexpr (CONN) {
            if (j >= ncbf) {  // connectivity HOCBF row of horizon step j - ncbf (computed once per IMPC iteration)
                k = j - ncbf;
                ax = conn[(k * 3) * kLS];
                ay = conn[(k * 3 + 1) * kLS];
                b = conn[(k * 3 + 2) * kLS];
                return;
            }
        }
        bool clf = false;
        int jj = j;
        if constexpr (CONN) {
            clf = j >= nsafe;
            if (clf) jj = j - nsafe;
        }
        // hit is 1 (IMPC iteration 0) or cbf_horizon (2 in every shipped config): no general integer division
        const int nb = (hit == 1) ? jj : ((hit == 2) ? (jj >> 1) : jj / hit);
        k = jj - nb * hit;
        const int other = nb + (nb >= self ? 1 : 0);
        const double* o = st + other * 6;
        const double* e = pred + (k * 4) * kLS;
        if (clf)
            clf_row_lifted(e[0], e[kLS], e[2 * kLS], e[3 * kLS], o[0], o[1], ax, ay, b);
        else
            hocbf_row(e[0], e[kLS], e[2 * kLS], e[3 * kLS], o[0], o[1], o[3], o[4], dmin2, gamma, ax, ay, b);
    }
    // g(v) = rhs - n.v of safety row j: positive when violated
    MPCCBF_HD double cbf_g(const double (&v)[4], double ax, double ay, double b, int k) const {
        const double* t = tm + (2 * K + k) * kTmplStride;
        const double accx = fma(t[0], v[0], fma(t[1], v[1], fma(t[2], s[0][0], t[3] * s[0][1])));
        const double accy = fma(t[0], v[2], fma(t[1], v[3], fma(t[2], s[1][0], t[3] * s[1][1])));
        return -(fma(ax, accx, fma(ay, accy, b)));
    }
    MPCCBF_HD void check_cbf(int j, const double (&v)[4], int& key, int& jbest) const {
        double ax, ay, b;
        int k;
        cbf_row(j, ax, ay, b, k);
        const double* t = tm + (2 * K + k) * kTmplStride;
        const double g = cbf_g(v, ax, ay, b, k);
        if (!(b < 1e20)) return;
        if constexpr (!SOFT) {
            if (g > tol * fmax(1.0, fabs(b))) {
                const double an2 = fma(ax, ax, ay * ay);
                const double score = (an2 > 0.0) ? g * t[4] * TRSQRT(an2) : HUGE_VAL;
                const int kk = HIWORD(score) & kKeyMask;
                if (kk > key) {
                    key = kk;
                    jbest = j;
                }
            }
            return;
        }
        // a violated soft row that does not depend on v is taken up by its slack: large finite score
        const int nb = (hit == 1) ? j : ((hit == 2) ? (j >> 1) : j / hit);
        const int ld = lead[nb];
        double viol, thr = tol * fmax(1.0, fabs(b)), score;
        int id = nb * (hit + 1) + k;
        if (ld == 0 || ld == k + 1) {
            viol = ld ? -g : g;  // the leader itself: piece "none" is violated once the row holds again
            if (ld) id = nb * (hit + 1) + hit;
            const double an2 = fma(ax, ax, ay * ay);
            score = (an2 > 0.0) ? viol * t[4] * TRSQRT(an2) : 1e150;
        } else {
            double axl, ayl, bl;
            int kl;
            cbf_row(nb * hit + ld - 1, axl, ayl, bl, kl);
            const double* tl = tm + (2 * K + kl) * kTmplStride;
            viol = g - cbf_g(v, axl, ayl, bl, kl);
            thr = tol * fmax(fmax(1.0, fabs(b)), fabs(bl));
            const double d0 = ax * t[0] - axl * tl[0], d1 = ax * t[1] - axl * tl[1];
            const double d2 = ay * t[0] - ayl * tl[0], d3 = ay * t[1] - ayl * tl[1];
            const double nn = fma(d0, d0, fma(d1, d1, fma(d2, d2, d3 * d3)));
            score = (nn > 0.0) ? viol * TRSQRT(nn) : 1e150;
        }
        if (viol > thr) {
            const int kk = HIWORD(score) & kKeyMask;
            if (kk > key) {
                key = kk;
                jbest = id;
            }
        }
    }
    MPCCBF_HD void scan(const double (&v)[4], double& best, int& best_id) const {
        int key[4] = {0, 0, 0, 0};  // [ax * 2 + dsel]
        {
            const AxisPoly px = axis_poly(pq, v[0], v[1], s[0][0], s[0][1]);
            const AxisPoly py = axis_poly(pq, v[2], v[3], s[1][0], s[1][1]);
            scan_axes_keys2(scn, K, kinv, px, py, lim, key);
        }
        int kb = 0, jb = -1;
        if (ncand < 0) {
            for (int j = 0; j < ncbf; ++j) check_cbf(j, v, kb, jb);
        } else {
            for (int i = 0; i < ncand; ++i) check_cbf(cand[i * kLS], v, kb, jb);
        }
        if constexpr (CONN)
            for (int k = 0; k < nconn; ++k) check_cbf(ncbf + k, v, kb, jb);
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
        const double* lm = lim + 8 * ax + 4 + 2 * dsel;
        if (val - lm[1] > lm[0] - val) best_id |= 1;
    }
    // sg * (normal, rhs) of safety row j accumulated into (np, rhs)
    MPCCBF_HD void add_cbf_normal(int j, double sg, double (&np)[4], double& rhs) const {
        double ax, ay, b;
        int k;
        cbf_row(j, ax, ay, b, k);
        const double* t = tm + (2 * K + k) * kTmplStride;
        np[0] = fma(sg * ax, t[0], np[0]);
        np[1] = fma(sg * ax, t[1], np[1]);
        np[2] = fma(sg * ay, t[0], np[2]);
        np[3] = fma(sg * ay, t[1], np[3]);
        const double offx = fma(t[2], s[0][0], t[3] * s[0][1]);
        const double offy = fma(t[2], s[1][0], t[3] * s[1][1]);
        rhs -= sg * (b + ax * offx + ay * offy);
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
            const double bound = lim[8 * ax + 4 + 2 * dsel + (up ? 1 : 0)];
            const double sa = ax ? s[1][0] : s[0][0], sb = ax ? s[1][1] : s[0][1];
            const double n0 = sg * t[0], n1 = sg * t[1];
            np[0] = ax ? 0.0 : n0;
            np[1] = ax ? 0.0 : n1;
            np[2] = ax ? n0 : 0.0;
            np[3] = ax ? n1 : 0.0;
            rhs = sg * (bound - fma(t[2], sa, t[3] * sb));
        } else {
            np[0] = np[1] = np[2] = np[3] = 0.0;
            rhs = 0.0;
            if constexpr (!SOFT) {
                add_cbf_normal(id - nid_fix, 1.0, np, rhs);
            } else {
                const int jj = id - nid_fix;
                const int nb = jj / (hit + 1), kk = jj - nb * (hit + 1);
                const int ld = lead[nb];
                if (kk < hit) add_cbf_normal(nb * hit + kk, 1.0, np, rhs);
                if (ld) add_cbf_normal(nb * hit + ld - 1, -1.0, np, rhs);
            }
        }
    }
    // ---- group interface of gi_run<.., SOFT = true> ------------------------------------------------------
    MPCCBF_HD int group(int id) const { return id < nid_fix ? -1 : (id - nid_fix) / (hit + 1); }
    MPCCBF_HD double weight(int grp) const { return slack_weight(st, self, grp, nnb, slack_cost, slack_decay); }
    MPCCBF_HD void set_leader(int grp, int id) const {
        const int kk = (id - nid_fix) - grp * (hit + 1);
        lead[grp] = static_cast<unsigned char>(kk < hit ? kk + 1 : 0);
    }
    // sum_j w_j s_j at v, s_j = max(0, g of the group's leader) (groups without a leader have s_j = 0)
    MPCCBF_HD double penalty(const double (&v)[4]) const {
        double pen = 0.0;
        for (int nb = 0; nb < nnb; ++nb) {
            const int ld = lead[nb];
            if (!ld) continue;
            double ax, ay, b;
            int k;
            cbf_row(nb * hit + ld - 1, ax, ay, b, k);
            pen = fma(weight(nb), fmax(cbf_g(v, ax, ay, b, k), 0.0), pen);
        }
        return pen;
    }
};

struct AgentCtx {
    const double *tm, *yw, *vr, *kr;  // constant tables (shared memory on the device)
    const double (*pq)[4];            // polynomial row structure (DevTables::poly)
    const double* scn;                // [K][4] key-scan table {t_k, inv_norm_acc, inv_norm_vel, -}
    const double* lim;                // [3 axes][8]: mid / half of the thresholded bounds, raw lo / hi
    const double* st;                 // this agent's scenario states [N][6]
    double* pred;                     // [(k*4 + c) * kLS]
    unsigned short* cand;             // [i * kLS]
    int cand_max;
    double* sol;                      // [kSolStride * kLS] result record of this agent
    double* nc;                       // [kNcRows * kLS] normals of the active rows (GIState::nc)
    unsigned char* lead;              // [N-1] group leaders of this agent (slack mode), else NULL
    // conn_mode: algebraic connectivity and Fiedler vector of this agent's scenario, scratch for its connectivity rows
    double lam2;
    const double* fied;
    double* conn;                     // [(k * 3 + i) * kLS]
};

// keeps the safety rows that can be violated somewhere inside the acceleration box (a row that
// holds on the whole box holds at every point the bound rows admit, so dropping it is exact)
template <bool SOFT, bool CONN>
MPCCBF_HD void build_candidates(RowsXY<SOFT, CONN>& rx, const DevTables& T, unsigned short* cand, unsigned long long& k0mask,
                                 bool& k0valid) {
    // The predicted ego state at t_0 = 0 is the current state (the curve is pinned there by the initial-state rows), so
    // the k = 0 safety row of neighbour nb is the same row in every IMPC iteration: its candidate test is done once
    // (iteration 0) and remembered in a bit mask while the neighbours fit one.
    int nc = 0;
    const bool wide = rx.ncbf > 65535;
    const bool fill = rx.hit == 1 && rx.nnb <= 64;
    const bool reuse = k0valid && !fill;
    if (fill) k0mask = 0ull;
    for (int j = 0; j < rx.ncbf && !wide; ++j) {
        bool keep;
        const int jnb = (rx.hit == 2) ? (j >> 1) : j / rx.hit;
        if (reuse && j < rx.nsafe && j == jnb * rx.hit) {
            keep = (k0mask >> jnb) & 1ull;
        } else {
            double ax, ay, b;
            int k;
            rx.cbf_row(j, ax, ay, b, k);
            if (!(b < 1e20)) continue;  // CPLEX: an upper bound >= 1e20 is "no bound" (b <= -1e20 still binds)
            const double mn = fmin(ax * T.lim_lo[0][0], ax * T.lim_hi[0][0]) + fmin(ay * T.lim_lo[0][1], ay * T.lim_hi[0][1]);
            keep = !(mn + b > 1e-7 * (1.0 + fabs(ax) + fabs(ay)));
            if (fill && keep && j < rx.nsafe) k0mask |= 1ull << j;
        }
        if (keep) {
            if (nc < rx.cand_max) cand[nc * kLS] = static_cast<unsigned short>(j);
            ++nc;
        }
    }
    if (fill) k0valid = true;
    rx.ncand = (wide || nc > rx.cand_max) ? -1 : nc;
}

MPCCBF_HD double ctrl_point(const double* y, double ps, double vs, double va, double vb) {
    // x[piece, ax, cp] = Y1 (pos, vel) + W1 (va, vb)
    // (explicit fma chain: mpccbf_expand_ctrl_pts evaluates the same expression on the host, bit for bit)
    return fma(y[0], ps, fma(y[1], vs, fma(y[2], va, y[3] * vb)));
}

// Unconstrained minimiser v0 (whitened coordinates) and, when WANT_C0, the constant part c0 of the objective.
// Evaluated twice per agent (before the solves, and again for the objective value at the end) instead of
// keeping v0 live across the active-set iterations.
struct AgentCtx;
template <bool WANT_C0>
MPCCBF_HD void unconstrained_minimiser(const DevTables& T, const StepArgs& a, int g, const double (&s)[3][2], const AgentCtx& c,
                                       double (&v0)[6], double& c0);

// Everything optimize() does for agent-QP g (one thread). CORE selects the active-set core (gi_solver.h).
template <bool SOFT, bool CONN, int CORE>
__host__ __device__ void agent_thread(const DevTables& T, const StepArgs& a, int g, int self, const AgentCtx& c) {
    const int K = T.K, N = T.N, H = T.H;
    double s[3][2];
#pragma unroll
    for (int ax = 0; ax < 3; ++ax) {
        s[ax][0] = c.st[self * 6 + ax];
        s[ax][1] = c.st[self * 6 + 3 + ax];
    }
    double v0[6];
    {
        double c0_unused;
        unconstrained_minimiser<false>(T, a, g, s, c, v0, c0_unused);
    }
    const float kinv = static_cast<float>(c.scn[3]);  // (K - 1) / t_{K-1}
    int total_iter = 0;
    // ---- bound rows only: the axes decouple, so yaw (final) and the stage-A solution of x / y are three 2-dof
    // projections through ONE rolled loop (a 2-dof step and scan cost about a third of the coupled 4-dof ones; the coupled
    // state is assembled from the two axis solutions: same minimiser, the problem is separable) -------------------------
    typename GICore<2, CORE>::State gy;
    int base_status = MPCCBF_STATUS_OPTIMAL;
    double axv[3][2], axu[2][2];
    int axid[2][2], axq[2] = {0, 0};
    constexpr bool kAxisStageA = (CORE == 0);
    {
        RowsAxis ry;
        ry.tm = c.tm;
        ry.K = K;
        ry.tol = T.feas_tol;
        ry.pq = c.pq;
        ry.scn = c.scn;
        ry.kinv = kinv;
        GISoft<2> gsy;
#pragma unroll 1
        for (int axi = (kAxisStageA ? 0 : 2); axi < 3; ++axi) {
            // yaw first would let an infeasible yaw skip x / y; the order only matters for which status is reported
            const int ax = (axi == 0) ? 2 : axi - 1;
            if (base_status != MPCCBF_STATUS_OPTIMAL) break;
            const double va = (ax == 0) ? v0[0] : ((ax == 1) ? v0[2] : v0[4]);
            const double vb = (ax == 0) ? v0[1] : ((ax == 1) ? v0[3] : v0[5]);
            const double vv[2] = {va, vb};
            ry.s0 = (ax == 0) ? s[0][0] : ((ax == 1) ? s[1][0] : s[2][0]);
            ry.s1 = (ax == 0) ? s[0][1] : ((ax == 1) ? s[1][1] : s[2][1]);
            ry.lim = c.lim + 8 * ax;
            GICore<2, CORE>::init(gy, vv);
            base_status = GICore<2, CORE>::template run<false>(gy, gsy, ry, T.max_gi, total_iter, gi_vmax2<2>(vv));
            if constexpr (kAxisStageA) {
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    if (ax == 0) { axv[0][i] = gy.v[i]; axu[0][i] = gy.u[i]; axid[0][i] = gy.id[i]; }
                    if (ax == 1) { axv[1][i] = gy.v[i]; axu[1][i] = gy.u[i]; axid[1][i] = gy.id[i]; }
                    if (ax == 2) axv[2][i] = gy.v[i];
                }
                if (ax == 0) axq[0] = gy.q;
                if (ax == 1) axq[1] = gy.q;
            } else {
                axv[2][0] = gy.v[0];
                axv[2][1] = gy.v[1];
            }
        }
    }
    // ---- x/y stage A: bound rows only ---------------------------------------------------------
    RowsXY<SOFT, CONN> rx;
    rx.tm = c.tm;
    rx.K = K;
#pragma unroll
    for (int ax = 0; ax < 2; ++ax) {
        rx.s[ax][0] = s[ax][0];
        rx.s[ax][1] = s[ax][1];
    }
    rx.tol = T.feas_tol;
    rx.st = c.st;
    rx.pred = c.pred;
    rx.cand = c.cand;
    rx.cand_max = c.cand_max;
    rx.ncand = 0;
    rx.ncbf = 0;
    rx.nsafe = 0;
    rx.nconn = 0;
    rx.conn = c.conn;
    rx.hit = 1;
    rx.self = self;
    rx.dmin2 = T.dmin2;
    rx.gamma = T.gamma;
    rx.nid_fix = 8 * K;
    rx.pq = c.pq;
    rx.scn = c.scn;
    rx.kinv = kinv;
    rx.lim = c.lim;
    rx.lead = c.lead;
    rx.nnb = N - 1;
    rx.slack_cost = T.slack_cost;
    rx.slack_decay = T.slack_decay;
    typename GICore<4, CORE>::State gx;
    GISoft<4> gs;
    const double vmax2 = gi_vmax2<4>(v0);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        gs.grp[i] = -1;
        gs.cap[i] = HUGE_VAL;
    }
    if constexpr (CORE == 0) gx.nc = c.nc;
    GICore<4, CORE>::init(gx, v0);
    if constexpr (kAxisStageA) {
        // coupled stage-A state from the two axis solutions: iterate, then the active rows of x followed by those of y
        // (axis row id (dsel K + k) << 1 | side  ->  ((dsel K + k) 2 + axis) << 1 | side)
        if (base_status == MPCCBF_STATUS_OPTIMAL) {
            gx.v[0] = axv[0][0]; gx.v[1] = axv[0][1]; gx.v[2] = axv[1][0]; gx.v[3] = axv[1][1];
            int q = 0;
#pragma unroll
            for (int ax = 0; ax < 2; ++ax)
#pragma unroll
                for (int i = 0; i < 2; ++i)
                    if (i < axq[ax]) {
                        const int id = ((((axid[ax][i] >> 1) << 1) + ax) << 1) | (axid[ax][i] & 1);
#pragma unroll
                        for (int col = 0; col < 4; ++col)
                            if (col == q) {
                                gx.u[col] = axu[ax][i];
                                gx.id[col] = id;
                            }
                        ++q;
                    }
            gx.q = q;
            GICore<4, CORE>::refactor(gx, rx);   // normals of the assembled active rows
        }
    }
    // stage-A solution: iterate, multipliers and ids of its active bound rows (the factorisation is
    // rebuilt from the ids when a later IMPC iteration has to start from it again)

    // Phase machine with ONE gi_run<4> call site (the unrolled 4-dof active-set code is the bulk of
    // the kernel's instruction footprint; inlining it once keeps the kernel inside the instruction
    // cache):  kStageA -> kIter(0) -> kIter(1) -> ...   Before an iteration whose predecessor put
    // safety rows into the active set, the stage-A factorisation is restored from `snap`.
    enum { kStageA = 0, kIter = 2 };
    int n_ok = 0;
    unsigned long long k0mask = 0ull;  // candidate flags of the k = 0 safety rows (same rows in every IMPC iteration)
    bool k0valid = false;
    bool dirty = false;  // gx has moved away from the stage-A solution
    // The rows of IMPC iteration i that exist in iteration i + 1 as well: the bound rows and the k = 0 safety rows (the
    // predicted state at t_0 is the current state in every iteration). When the active set of a finished iteration holds
    // only such rows, its (v, multipliers, active set) is a solution pair of a SUB-problem of the next iteration — a
    // valid starting point of the dual method — so the next iteration continues from it instead of from stage A and
    // only adds what its new rows (k >= 1, rebuilt at the new predicted states) require. Same minimiser (it is
    // unique); measured on the bench workload: 4.9 -> 4.0 active-set steps in iteration 1 (its k = 1 rows displace most of
    // what iteration 0 found).
    bool reuse = false;
    double pen_last = 0.0;  // slack mode: sum_j w_j s_j of the last successful iterate
    const size_t sbase = static_cast<size_t>(g) * T.iters;
    int mode = kStageA, it = 0, it_iter = 0;
    // stage-A solution every iteration can restart from: 8 doubles + 5 ints that are written once and read at most once
    // per IMPC iteration, so they may live wherever the compiler puts them (registers or its local-memory spill space)
    double snv[4] = {0, 0, 0, 0}, snu[4] = {0, 0, 0, 0};
    int snid[4] = {-1, -1, -1, -1}, snq = 0;
    if constexpr (kAxisStageA) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            snv[i] = gx.v[i];
            snu[i] = gx.u[i];
            snid[i] = gx.id[i];
        }
        snq = gx.q;
        it_iter = total_iter;
        mode = kIter;
    }
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
                    c.pred[kLS] = s[1][0];
                    c.pred[2 * kLS] = s[0][1];
                    c.pred[3 * kLS] = s[1][1];
                } else {
                    for (int k = 0; k < H; ++k) {
#pragma unroll
                        for (int d = 0; d < 2; ++d) {
                            const double* t = c.tm + (d * K + k) * kTmplStride;
#pragma unroll
                            for (int ax = 0; ax < 2; ++ax)
                                c.pred[(k * 4 + d * 2 + ax) * kLS] =
                                    fma(t[0], gx.v[2 * ax], fma(t[1], gx.v[2 * ax + 1], fma(t[2], s[ax][0], t[3] * s[ax][1])));
                        }
                    }
                }
                if (dirty && reuse) {
                    // row ids of the k = 0 safety rows in the numbering of this iteration (hit rows per neighbour)
                    if (rx.hit != hit) {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            if (gx.id[i] >= rx.nid_fix) {
                                const int j = gx.id[i] - rx.nid_fix;
                                const int nb = (rx.hit == 1) ? j : ((rx.hit == 2) ? (j >> 1) : j / rx.hit);
                                gx.id[i] = rx.nid_fix + nb * hit;
                            }
                    }
                } else if (dirty) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        gx.v[i] = snv[i];
                        gx.u[i] = snu[i];
                        gx.id[i] = snid[i];
                        if constexpr (SOFT) gs.grp[i] = -1;
                    }
                    gx.q = snq;
                    GICore<4, CORE>::refactor(gx, rx);
                }
                if (!reuse) dirty = false;
                rx.hit = hit;
                rx.nsafe = (N - 1) * hit;
                rx.ncbf = rx.nsafe;
                rx.nconn = 0;
                if constexpr (CONN) {
                    if (T.conn_mode) {  // ConnectivityIMPCCBF.cpp:143-157 / :177-192 (commented out in the reference)
                        if (c.lam2 > 0.1) {
                            for (int k = 0; k < hit; ++k) {
                                double gx2, gy2, bb;
                                conn_row_lifted(c.st, N, self, c.fied, c.lam2, T.dmax, T.gamma, c.pred[(k * 4) * kLS],
                                                c.pred[(k * 4 + 1) * kLS], c.pred[(k * 4 + 2) * kLS],
                                                c.pred[(k * 4 + 3) * kLS], gx2, gy2, bb);
                                c.conn[(k * 3) * kLS] = gx2;
                                c.conn[(k * 3 + 1) * kLS] = gy2;
                                c.conn[(k * 3 + 2) * kLS] = bb;
                            }
                            rx.nconn = hit;
                        } else {
                            rx.ncbf = 2 * rx.nsafe;  // one distance-CLF row per neighbour and step after the safety rows
                        }
                    }
                }
                if constexpr (SOFT)
                    for (int nb = 0; nb < N - 1; ++nb) c.lead[nb] = 0;
                build_candidates<SOFT, CONN>(rx, T, c.cand, k0mask, k0valid);
                run = rx.ncand != 0 || rx.nconn > 0;
            } else {
                run = true;
            }
        }
        int n2 = 0;
        if (run) stat = GICore<4, CORE>::template run<SOFT>(gx, gs, rx, T.max_gi, n2, vmax2);
        it_iter += n2;
        if (mode == kStageA) {
            base_status = stat;
            it_iter += total_iter;  // yaw iterations are booked on IMPC iteration 0
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                snv[i] = gx.v[i];
                snu[i] = gx.u[i];
                snid[i] = gx.id[i];
            }
            snq = gx.q;
            mode = kIter;
            continue;
        }
        dirty = dirty || n2 > 0;
        reuse = false;
        if constexpr (!SOFT && !CONN && CORE == 0) {
            if (dirty && stat == MPCCBF_STATUS_OPTIMAL) {
                reuse = true;
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if (i < gx.q && gx.id[i] >= rx.nid_fix) {
                        const int j = gx.id[i] - rx.nid_fix;
                        const int k = (rx.hit == 1) ? 0 : ((rx.hit == 2) ? (j & 1) : j % rx.hit);
                        if (k != 0) reuse = false;
                    }
            }
        }
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
        for (int i = 0; i < 4; ++i) c.sol[i * kLS] = gx.v[i];
        c.sol[4 * kLS] = axv[2][0];
        c.sol[5 * kLS] = axv[2][1];
        if constexpr (SOFT) pen_last = (run && a.objective) ? rx.penalty(gx.v) : 0.0;
        if (a.ctrl_all) {  // every iterate (mpccbf_optimize_one): small batches, direct stores
            double* out = a.ctrl_all + (sbase + it) * T.n;
            for (int i = 0; i < T.n; ++i) {
                const int piece = i / (3 * T.c), rem = i - piece * 3 * T.c;
                const int ax = rem / T.c, cp = rem - ax * T.c;
                out[i] = ctrl_point(c.yw + (piece * T.c + cp) * 4, c.st[self * 6 + ax], c.st[self * 6 + 3 + ax], c.sol[2 * ax * kLS], c.sol[(2 * ax + 1) * kLS]);
            }
        }
        ++it;
    }
    // ---- result record (expanded to control points cooperatively by the warp) ------------------
    if (n_ok == 0) {
#pragma unroll
        for (int i = 0; i < 6; ++i) c.sol[i * kLS] = 0.0;
    }
    c.sol[6 * kLS] = static_cast<double>(n_ok);
    if (a.n_curves) a.n_curves[g] = n_ok;
    if (a.objective) {
        double obj = 0.0;
        if (n_ok > 0) {
            double c0;
            unconstrained_minimiser<true>(T, a, g, s, c, v0, c0);
            double dd = 0.0, nn = 0.0;
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const double e = c.sol[i * kLS] - v0[i];
                dd = fma(e, e, dd);
                nn = fma(v0[i], v0[i], nn);
            }
            obj = 0.5 * dd - 0.5 * nn + c0 + pen_last;
        }
        a.objective[g] = obj;
    }
}

template <bool WANT_C0>
MPCCBF_HD void unconstrained_minimiser(const DevTables& T, const StepArgs& a, int g, const double (&s)[3][2], const AgentCtx& c,
                                       double (&v0)[6], double& c0) {
    const int K = T.K;
    c0 = 0.0;
    if (a.ref_mode == 0) {
        const double* tg = a.targets + static_cast<size_t>(g) * 3;
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            const double tgt = TLDG(tg + ax);
            const double ps = s[ax][0], vs = s[ax][1];
            v0[2 * ax] = T.Vs[0][0] * ps + T.Vs[0][1] * vs + T.Vt[0] * tgt;
            v0[2 * ax + 1] = T.Vs[1][0] * ps + T.Vs[1][1] * vs + T.Vt[1] * tgt;
            if (WANT_C0)
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
                if (WANT_C0) {
                    k0 = fma(c.kr[k], rk, k0);
                    k1 = fma(c.kr[K + k], rk, k1);
                }
            }
            v0[2 * ax] = T.Vs[0][0] * ps + T.Vs[0][1] * vs + r0;
            v0[2 * ax + 1] = T.Vs[1][0] * ps + T.Vs[1][1] * vs + r1;
            if (WANT_C0) c0 += ps * (T.Kc[0][0] * ps + T.Kc[0][1] * vs) + vs * (T.Kc[1][0] * ps + T.Kc[1][1] * vs) + (k0 * ps + k1 * vs);
        }
    }
}

// lim[axis] = {midA, halfA, midV, halfV (of the thresholded bounds), loA, hiA, loV, hiV}; scn[k] = {t_k, inv_norm of the
// acceleration row, inv_norm of the velocity row, 0} with 1e300 standing for "row independent of v"; scn[3] = (K - 1) / t_{K-1}
__host__ __device__ inline void fill_scan_tables(const DevTables& T, const double* tm, double* lim, double* scn, int tid, int nthr) {
    for (int i = tid; i < 6; i += nthr) {
        const int axis = i >> 1, dsel = i & 1;
        const double hi = bound_hi_eff(T.lim_hi[dsel][axis], T.feas_tol), lo = bound_lo_eff(T.lim_lo[dsel][axis], T.feas_tol);
        lim[axis * 8 + dsel * 2 + 0] = 0.5 * (hi + lo);
        lim[axis * 8 + dsel * 2 + 1] = 0.5 * (hi - lo);
        lim[axis * 8 + 4 + dsel * 2 + 0] = T.lim_lo[dsel][axis];
        lim[axis * 8 + 4 + dsel * 2 + 1] = T.lim_hi[dsel][axis];
    }
    for (int k = tid; k < T.K; k += nthr) {
        scn[4 * k + 0] = tm[(2 * T.K + k) * kTmplStride + 5];
        scn[4 * k + 1] = fmin(tm[(2 * T.K + k) * kTmplStride + 4], 1e300);
        scn[4 * k + 2] = fmin(tm[(1 * T.K + k) * kTmplStride + 4], 1e300);
        scn[4 * k + 3] = (k == 0) ? (T.K - 1) / tm[(2 * T.K + T.K - 1) * kTmplStride + 5] : 0.0;  // [3]: sample index per unit time
    }
}

struct ThreadSmemLayout {
    // CTA-shared tables (doubles from the start of shared memory): pq, lim, scn at the compile-time offsets 0 / kOffLim /
    // kOffScn, then tm [3][K][kTmplStride], vr [2][K], kr [2][K], yw [m1][4], omap (unsigned shorts)
    int off_tm, off_vr, off_kr, off_yw, off_omap;
    int n_tm, n_yw, n_vr;
    // per warp (bytes from the warp's block): the fixed-offset arrays (kOffSol .. kOffPred), then
    int off_cand, cand_max;          // candidate list and its capacity
    int off_conn, off_fied, off_st;  // conn_mode rows and Fiedler vectors (empty otherwise), scenario states
    int n_fied;
    int apw;                         // agents per warp (a tile): 32, or fewer so that a small batch still fills the SMs
    int nscen_max;
    int shared_bytes, warp_bytes;
};

static ThreadSmemLayout thread_layout(const DevTables& T, int agent_count, int apw) {
    ThreadSmemLayout L;
    L.apw = apw;
    L.n_tm = 3 * T.K * kTmplStride;
    L.n_vr = 2 * T.K;
    L.n_yw = T.m1 * 4;
    L.off_tm = kOffScn + 4 * T.K;
    L.off_vr = L.off_tm + L.n_tm;
    L.off_kr = L.off_vr + L.n_vr;
    L.off_yw = L.off_kr + L.n_vr;
    L.off_omap = L.off_yw + L.n_yw;
    L.shared_bytes = (8 * L.off_omap + T.n * 2 + 15) / 16 * 16;
    // scenarios a tile of 32 consecutive agents can touch
    const int ac = std::max(agent_count, 1);
    L.nscen_max = (apw % ac == 0) ? apw / ac : (ac % apw == 0 ? 1 : (apw - 1) / ac + 2);
    const int H = std::max(T.H, 1);
    L.cand_max = (T.N <= 32) ? kCandMaxSmall : kCandMaxLarge;
    L.off_cand = kOffPred + H * 4 * kRow;
    L.off_conn = L.off_cand + L.cand_max * kLS * 2;
    L.off_fied = L.off_conn + (T.conn_mode ? H * 3 * kRow : 0);
    L.n_fied = T.conn_mode ? L.nscen_max * T.N : 0;
    L.off_st = (L.off_fied + 8 * L.n_fied + 15) / 16 * 16;
    L.warp_bytes = (L.off_st + 8 * L.nscen_max * T.N * 6 + 15) / 16 * 16;
    return L;
}

// Launch shapes (threads per CTA, CTAs per SM the register budget is set for). All keep >= 14 warps per SM, which is what
// one wave of the 4096 x 16 workload needs on 148 SMs:  0: 128 x 4 (128 registers / thread)   2: 224 x 2 (144)
// 3: 64 x 7 (144)   4: 224 x 2 at 128 registers — the measured best and the default: two CTAs of 7 warps need 159 KB of
// shared memory per SM instead of 193 KB, so the driver picks the 164 KB carve-out and the L1, which holds the kernel's
// local memory, is 64 KB instead of 32 KB (0.185 against 0.205 ms; profiles/r2_kernel_ab.md section 6).
// Shape 1 (448 x 1) was dropped: its launch exceeds the register file.
template <int CFG> struct ThreadCfg;
template <> struct ThreadCfg<0> { static constexpr int kThreads = 128, kCtas = 4; };
template <> struct ThreadCfg<2> { static constexpr int kThreads = 224, kCtas = 2; };
template <> struct ThreadCfg<3> { static constexpr int kThreads = 64, kCtas = 7; };

// ptxas derives 128 registers from __launch_bounds__(448, 1) and (224, 2) alike (it budgets for 512 threads), although
// 14 warps x 32 x 144 registers fit the 64 K register file: shapes 1 and 2 state their budget with __maxnreg__ instead.
template <bool SOFT, bool CONN, int CORE>
__device__ __forceinline__ void thread_kernel_body(const DevTables& T, const StepArgs& a, const ThreadSmemLayout& L);

template <bool SOFT, bool CONN, int CORE, int CFG>
__global__ void __launch_bounds__(128, 4)
impc_thread_kernel(const __grid_constant__ DevTables T, const __grid_constant__ StepArgs a, const __grid_constant__ ThreadSmemLayout L) {
    thread_kernel_body<SOFT, CONN, CORE>(T, a, L);
}
template <bool SOFT, bool CONN, int CORE, int CFG>
__global__ void __launch_bounds__(224, 2)
impc_thread_kernel_b224(const __grid_constant__ DevTables T, const __grid_constant__ StepArgs a, const __grid_constant__ ThreadSmemLayout L) {
    thread_kernel_body<SOFT, CONN, CORE>(T, a, L);
}
template <bool SOFT, bool CONN, int CORE, int CFG>
__global__ void __maxnreg__(144)
impc_thread_kernel_r144(const __grid_constant__ DevTables T, const __grid_constant__ StepArgs a, const __grid_constant__ ThreadSmemLayout L) {
    thread_kernel_body<SOFT, CONN, CORE>(T, a, L);
}

template <bool SOFT, bool CONN, int CORE>
__device__ __forceinline__ void thread_kernel_body(const DevTables& T, const StepArgs& a, const ThreadSmemLayout& L) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* svq = reinterpret_cast<double*>(smem_raw);   // pq | lim | scn at compile-time offsets
    double* tm = svq + L.off_tm;
    double* vr = svq + L.off_vr;
    double* kr = svq + L.off_kr;
    double* yw = svq + L.off_yw;
    unsigned short* omap = reinterpret_cast<unsigned short*>(svq + L.off_omap);
    unsigned char* wbase = smem_raw + L.shared_bytes + (threadIdx.x >> 5) * L.warp_bytes;
    double* wsol = reinterpret_cast<double*>(wbase + kOffSol);
    double* wfied = reinterpret_cast<double*>(wbase + L.off_fied);
    double* wst = reinterpret_cast<double*>(wbase + L.off_st);

    for (int i = threadIdx.x; i < L.n_tm; i += blockDim.x) tm[i] = T.tmpl[i];
    for (int i = threadIdx.x; i < L.n_yw; i += blockDim.x) yw[i] = T.yw[i];
    for (int i = threadIdx.x; i < L.n_vr; i += blockDim.x) {
        vr[i] = T.Vr[i];
        kr[i] = T.Kr[i];
    }
    if (threadIdx.x < 20) svq[threadIdx.x] = T.pq[threadIdx.x >> 2][threadIdx.x & 3];
    __syncthreads();   // fill_scan_tables reads tm
    fill_scan_tables(T, tm, svq + kOffLim, svq + kOffScn, threadIdx.x, blockDim.x);
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
    const int apw = L.apw;
    const int ntiles = (a.num_qp + apw - 1) / apw;
    const int N = T.N, n = T.n;

    for (int tile = gwarp; tile < ntiles; tile += total_warps) {
        const int g0 = tile * apw;
        const int cnt = min(apw, a.num_qp - g0);
        const int s_lo = g0 / a.agent_count;
        const int s_hi = (g0 + cnt - 1) / a.agent_count;
        {   // stage the tile's scenarios (coalesced 16-byte loads)
            const double2* src = reinterpret_cast<const double2*>(a.states + static_cast<size_t>(s_lo) * N * 6);
            double2* dst = reinterpret_cast<double2*>(wst);
            const int words = (s_hi - s_lo + 1) * N * 3;
            for (int i = lane; i < words; i += 32) dst[i] = __ldg(src + i);
            if (L.n_fied)
                for (int i = lane; i < (s_hi - s_lo + 1) * N; i += 32) wfied[i] = __ldg(a.fiedler + static_cast<size_t>(s_lo) * N + i);
        }
        __syncwarp();
        if (lane < cnt) {
            const int g = g0 + lane;
            const int scen = g / a.agent_count;
            const int self = a.agent_begin + (g - scen * a.agent_count);
            AgentCtx c;
            c.tm = tm; c.yw = yw; c.vr = vr; c.kr = kr;
            c.pq = reinterpret_cast<const double (*)[4]>(svq);
            c.lim = svq + kOffLim;
            c.scn = svq + kOffScn;
            c.st = wst + (scen - s_lo) * N * 6;
            unsigned char* tb = wbase + lane * 8;   // every lane-strided array of this agent: tb + immediate
            c.sol = reinterpret_cast<double*>(tb + kOffSol);
            c.nc = reinterpret_cast<double*>(tb + kOffNc);
            c.cand = reinterpret_cast<unsigned short*>(wbase + L.off_cand) + lane;
            c.cand_max = L.cand_max;
            c.pred = reinterpret_cast<double*>(tb + kOffPred);
            c.conn = reinterpret_cast<double*>(tb + L.off_conn);
            c.lam2 = L.n_fied ? __ldg(a.lambda2 + scen) : 0.0;
            c.fied = wfied + (scen - s_lo) * N;
            c.lead = SOFT ? a.lead_ws + static_cast<size_t>(g) * (N - 1) : nullptr;
            agent_thread<SOFT, CONN, CORE>(T, a, g, self, c);
        }
        __syncwarp();
        if (a.v_out) {  // compact result: the 6 whitened free dof per agent, coalesced
            double* vo = a.v_out + static_cast<size_t>(g0) * 6;
            for (int idx = lane; idx < cnt * 6; idx += 32) {
                const int ag = __float2int_rz((idx + 0.5f) * (1.0f / 6.0f)), i = idx - ag * 6;
                vo[idx] = wsol[i * kLS + ag];   // zeros when the agent has no curve
            }
        }
        if (a.ctrl_pts) {   // control points of the last successful iterate: coalesced stores, whole warp
            // one agent after the other (its state, record and scenario are warp-uniform); lane l owns the control
            // points l, l + 32, ... of every agent, so their table entries are looked up once per tile
            double* out = a.ctrl_pts + static_cast<size_t>(g0) * n;
            int sc = s_lo, a0 = g0 - s_lo * a.agent_count;   // scenario and agent-in-shard of the tile's first agent
            for (int i0 = 0; i0 < n; i0 += 32) {
                const int i = i0 + lane;
                const unsigned m = omap[min(i, n - 1)];
                const int ax = m & 3;
                const double* y = yw + (m >> 2) * 4;
                const double y0 = y[0], y1 = y[1], y2 = y[2], y3 = y[3];
                int scc = sc, aa = a0;
                for (int ag = 0; ag < cnt; ++ag) {
                    const double* sta = wst + ((scc - s_lo) * N + a.agent_begin + aa) * 6;
                    const double val = fma(y0, sta[ax], fma(y1, sta[3 + ax], fma(y2, wsol[(2 * ax) * kLS + ag], y3 * wsol[(2 * ax + 1) * kLS + ag])));
                    if (i < n) out[static_cast<size_t>(ag) * n + i] = (wsol[6 * kLS + ag] > 0.0) ? val : 0.0;
                    if (++aa == a.agent_count) {
                        aa = 0;
                        ++scc;
                    }
                }
            }
        }
        __syncwarp();
    }
}

#ifdef MPCCBF_HOST_EMUL
// Test-only (tests/emul): the SAME per-agent function on the host, one agent at a time.
void emul_impc_thread_host(const DevTables& T, const StepArgs& a) {
    // the same lane-strided scratch layout as a warp of the kernel, used by lane 0
    const int H = std::max(T.H, 1);
    std::vector<double> pred(static_cast<size_t>(H) * 4 * kLS), sol(static_cast<size_t>(kSolStride) * kLS);
    std::vector<unsigned short> cand(static_cast<size_t>(kCandMaxLarge) * kLS);
    std::vector<double> ncache(static_cast<size_t>(kNcRows) * kLS), connrows(static_cast<size_t>(H) * 3 * kLS);
    std::vector<double> lim(24), scn(static_cast<size_t>(4) * T.K);
    std::vector<unsigned char> lead(static_cast<size_t>(std::max(T.N, 1)));
    fill_scan_tables(T, T.tmpl, lim.data(), scn.data(), 0, 1);
    for (int g = 0; g < a.num_qp; ++g) {
        const int scen = g / a.agent_count;
        const int self = a.agent_begin + (g - scen * a.agent_count);
        AgentCtx c;
        c.tm = T.tmpl; c.yw = T.yw; c.vr = T.Vr; c.kr = T.Kr;
        c.pq = T.pq;
        c.lim = lim.data();
        c.scn = scn.data();
        c.st = a.states + static_cast<size_t>(scen) * T.N * 6;
        c.pred = pred.data();
        c.cand = cand.data();
        c.cand_max = (T.N <= 32) ? kCandMaxSmall : kCandMaxLarge;
        c.sol = sol.data();
        c.nc = ncache.data();
        c.lam2 = (T.conn_mode && a.lambda2) ? a.lambda2[scen] : 0.0;
        c.fied = (T.conn_mode && a.fiedler) ? a.fiedler + static_cast<size_t>(scen) * T.N : nullptr;
        c.conn = connrows.data();
        c.lead = lead.data();
        if (T.slack) {
            agent_thread<true, false, 0>(T, a, g, self, c);
        } else if (T.conn_mode) {
            agent_thread<false, true, 0>(T, a, g, self, c);
        } else {
            agent_thread<false, false, 0>(T, a, g, self, c);
        }
        if (a.v_out)
            for (int i = 0; i < 6; ++i) a.v_out[static_cast<size_t>(g) * 6 + i] = sol[i * kLS];
        double* out = a.ctrl_pts + static_cast<size_t>(g) * T.n;
        for (int i = 0; a.ctrl_pts && i < T.n; ++i) {
            const int piece = i / (3 * T.c), rem = i - piece * 3 * T.c;
            const int ax = rem / T.c, cp = rem - ax * T.c;
            out[i] = (sol[6 * kLS] > 0.0) ? ctrl_point(T.yw + (piece * T.c + cp) * 4, c.st[self * 6 + ax], c.st[self * 6 + 3 + ax], sol[2 * ax * kLS], sol[(2 * ax + 1) * kLS]) : 0.0;
        }
    }
}
#endif

// Launches the thread-per-agent kernel. Returns a cudaError_t.
typedef void (*ThreadKernel)(const DevTables, const StepArgs, const ThreadSmemLayout);

template <int CORE, int CFG>
static ThreadKernel pick_kernel(const DevTables& T) {
    if (CFG == 0) {
        if (T.slack) return impc_thread_kernel<true, false, CORE, 0>;
        if (T.conn_mode) return impc_thread_kernel<false, true, CORE, 0>;
        return impc_thread_kernel<false, false, CORE, 0>;
    }
    if (T.slack) return impc_thread_kernel_r144<true, false, CORE, 0>;
    if (T.conn_mode) return impc_thread_kernel_r144<false, true, CORE, 0>;
    return impc_thread_kernel_r144<false, false, CORE, 0>;
}

#ifdef MPCCBF_WITH_R1
namespace r1 { int launch_impc_thread_r1(const DevTables& T, const StepArgs& a, int sm_count, cudaStream_t stream); }
#endif

int launch_impc_thread(const DevTables& T, const StepArgs& a, int sm_count, cudaStream_t stream) {
    if (a.num_qp <= 0) return cudaSuccess;
#ifdef MPCCBF_WITH_R1
    {   // experiment builds only: the round-1 kernel on the same inputs (MPCCBF_THREAD_CORE=9)
        static const bool use_r1 = [] { const char* e = std::getenv("MPCCBF_THREAD_CORE"); return e && std::atoi(e) == 9; }();
        if (use_r1 && !T.slack && !T.conn_mode && a.ctrl_pts) return r1::launch_impc_thread_r1(T, a, sm_count, stream);
    }
#endif
    if (T.slack && !a.lead_ws) return cudaErrorInvalidValue;
    if (T.conn_mode && (!a.lambda2 || !a.fiedler)) return cudaErrorInvalidValue;
    // Launch shape: measured default (profiles/r2_kernel_ab.md);
    // MPCCBF_THREAD_CFG = 0 / 2 / 3 / 4 override them for experiments.
    static const int cfgi = [] { const char* e = std::getenv("MPCCBF_THREAD_CFG"); const int v = e ? std::atoi(e) : MPCCBF_THREAD_CFG; return (v >= 2 && v <= 4) ? v : 0; }();
    ThreadKernel kern;
    int threads;
    if (cfgi == 2) {
        kern = pick_kernel<0, 2>(T);
        threads = ThreadCfg<2>::kThreads;
    } else if (cfgi == 3) {
        kern = pick_kernel<0, 2>(T);
        threads = ThreadCfg<3>::kThreads;
    } else if (cfgi == 4) {   // 224 x 2 at 128 registers: 159 KB of shared memory per SM, i.e. the 164 KB carve-out (L1 64 KB)
        kern = T.slack ? impc_thread_kernel_b224<true, false, 0, 0> : (T.conn_mode ? impc_thread_kernel_b224<false, true, 0, 0> : impc_thread_kernel_b224<false, false, 0, 0>);
        threads = 224;
    } else {
        kern = pick_kernel<0, 0>(T);
        threads = ThreadCfg<0>::kThreads;
    }
    // Agents per warp. The kernel is latency-bound (its duration is that of its slowest warp, ~5 us per active-set step of
    // one thread), and a warp executes the union of what its agents need. Measured (tools/apw_sweep.sh): fewer agents per
    // warp only pay when every agent can have a warp of its own in one wave (2048 agents: 0.127 vs 0.145 ms); in
    // between, 32 agents per warp win (more warps per SM cost more than shorter warps gain). One agent's arithmetic
    // does not depend on it. MPCCBF_THREAD_APW overrides (experiments).
    static const int apw_env = [] { const char* e = std::getenv("MPCCBF_THREAD_APW"); return e ? std::atoi(e) : 0; }();
    int apw = (a.num_qp <= 16 * sm_count) ? 1 : 32;
    if (apw_env >= 1 && apw_env <= 32 && (apw_env & (apw_env - 1)) == 0) apw = apw_env;
    const ThreadSmemLayout L = thread_layout(T, a.agent_count, apw);
    const int ntiles = (a.num_qp + apw - 1) / apw;
    // small batches: no more warps per CTA than there are tiles per SM (keeps the grid spread over the SMs)
    const int want = std::max(1, (ntiles + sm_count - 1) / std::max(1, sm_count));
    while (threads > 32 && threads / 32 > want && (threads / 32) % 2 == 0 && threads > 128) threads /= 2;
    size_t smem = L.shared_bytes + L.warp_bytes * (threads / 32);
    while (smem > 227 * 1024 && threads > 32) {
        threads = std::max(32, (threads / 64) * 32);
        smem = L.shared_bytes + L.warp_bytes * (threads / 32);
    }
    if (smem > 227 * 1024) return cudaErrorInvalidConfiguration;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    int per_sm = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, smem);
    {   // experiments: shared-memory carve-out in percent of 228 KB (MPCCBF_THREAD_CARVEOUT); default: the driver's choice,
        // the smallest that fits — the L1 it leaves holds the kernel's local memory (profiles/r2_kernel_ab.md section 6)
        static const int co = [] { const char* v = std::getenv("MPCCBF_THREAD_CARVEOUT"); return v ? std::atoi(v) : -1; }();
        if (co >= 0) cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, co);
    }
    if (e != cudaSuccess) return e;
    if (per_sm < 1) per_sm = 1;
    const int need = (ntiles + threads / 32 - 1) / (threads / 32);
    int grid = sm_count * per_sm;  // persistent: a multiple of the SM count
    if (need < grid) grid = need;
    kern<<<grid, threads, smem, stream>>>(T, a, L);
    return cudaGetLastError();
}

}  // namespace mpccbf
