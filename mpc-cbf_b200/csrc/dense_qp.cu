// K2 (full-space form): batched dense convex QP, one CTA per QP — sm_100a.
//
// Replaces qpcpp::CPLEXSolver::solve (qpcpp/src/solvers/CPLEX.cpp:35-178) at the S-qp seam
// (qpcpp/include/qpcpp/solvers/Solver.h:27-37) for a generic
//      min x'Px + q'x   s.t.  l <= Ax <= u        (reference objective convention,
//                                                   qpcpp/src/Problem.cpp:89-117)
// Method: OSQP-style ADMM (Stellato et al. 2020) on the Ruiz-equilibrated problem with a
// per-problem Cholesky factor of  P~ + sigma I + A' diag(rho) A  cached in shared memory
// (P~ = P + P'), refactorised when rho adapts; residual norms by warp-shuffle reductions;
// primal/dual infeasibility certificates; and an active-set polish (equality-constrained solves
// on the identified working set by the proximal method of multipliers, with add/drop repair)
// that takes the ADMM iterate to the exact vertex so that controls match an exact solver to
// ~1e-8. tests/admm_model.py is the step-for-step numpy specification of this kernel.
//
// Layout: the factor (n x n, odd row stride) and every n- and m-vector live in shared memory; the
// scaled constraint matrix A (m x n, odd row stride) too when it fits in the 227 KB, else it stays
// in a per-QP global workspace (L2-resident). The scaled P~ lives in the global workspace.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <string>
#include <vector>

#include "../../include/mpccbf.h"
#include "kernels.h"

namespace mpccbf {

namespace {

constexpr int kT = 256;          // threads per CTA
constexpr int kWarps = kT / 32;
constexpr double kInfBound = 1e20;  // CPLEX: |bound| >= 1e20 is "no bound"
constexpr int kMaxN = 128;

struct DenseParams {
    int n, m, batch;
    int lda, ldl;        // row strides (doubles) of A and of the factor
    int a_in_smem;
    const double *P, *q, *A, *l, *u;   // [batch] inputs, reference convention
    double* wsP;         // [batch][n*n]   scaled P~ (symmetric)
    double* wsA;         // [batch][m*lda] scaled A when it does not fit in shared memory
    double* x;           // [batch][n]
    double* obj;         // [batch] or NULL
    int32_t* status;     // [batch]
    int32_t* iters;      // [batch] or NULL
    int max_iter, check_every, ruiz_iters;
    double eps, rho0, sigma, alpha, eps_inf;
};

// ---- block-wide reductions: warp shuffles, then one shared-memory hop across the 8 warps ------
template <int K, bool SUM>
__device__ __forceinline__ void block_reduce(double (&v)[K], double* red) {
#pragma unroll
    for (int k = 0; k < K; ++k) {
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            const double other = __shfl_xor_sync(0xffffffffu, v[k], o);
            v[k] = SUM ? v[k] + other : fmax(v[k], other);
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < K; ++k) red[warp * K + k] = v[k];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < K; ++k) {
        double r = red[k];
        for (int w = 1; w < kWarps; ++w) r = SUM ? r + red[w * K + k] : fmax(r, red[w * K + k]);
        v[k] = r;
    }
    __syncthreads();
}

struct Ctx {
    int n, m, lda, ldl;
    double* A;      // scaled A (shared or global), row stride lda
    double* P;      // scaled P~ (global), row stride n
    double* L;      // factor (shared), row stride ldl
    double* invd;   // 1 / L_jj
    double* part;   // [kT] scratch for column-wise partial sums
    double* red;    // [kWarps * 8] reduction scratch
};

// out[j] = sum_i A[i][j] * w[i]  (j < n). Thread (g, j): rows g, g + ng, ...
__device__ void mat_tvec(const Ctx& c, const double* w, double* out) {
    const int n = c.n, m = c.m;
    const int ng = kT / n;
    const int g = threadIdx.x / n, j = threadIdx.x - g * n;
    if (g < ng) {
        double s = 0.0;
        for (int i = g; i < m; i += ng) s = fma(c.A[static_cast<size_t>(i) * c.lda + j], w[i], s);
        c.part[threadIdx.x] = s;
    }
    __syncthreads();
    if (threadIdx.x < n) {
        double s = 0.0;
        for (int gg = 0; gg < ng; ++gg) s += c.part[gg * n + threadIdx.x];
        out[threadIdx.x] = s;
    }
    __syncthreads();
}

// out[i] = sum_j A[i][j] * x[j]  (i < m)
__device__ void mat_vec(const Ctx& c, const double* x, double* out, bool a_in_smem) {
    const int n = c.n, m = c.m;
    if (a_in_smem) {  // thread per row; the odd row stride keeps the 32 rows on distinct banks
        for (int i = threadIdx.x; i < m; i += kT) {
            const double* a = c.A + static_cast<size_t>(i) * c.lda;
            double s = 0.0;
            for (int j = 0; j < n; ++j) s = fma(a[j], x[j], s);
            out[i] = s;
        }
    } else {  // warp per row, coalesced
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        for (int i = warp; i < m; i += kWarps) {
            const double* a = c.A + static_cast<size_t>(i) * c.lda;
            double s = 0.0;
            for (int j = lane; j < n; j += 32) s = fma(a[j], x[j], s);
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0) out[i] = s;
        }
    }
    __syncthreads();
}

// out[i] = sum_j P[j][i] x[j]  (P symmetric: column access is coalesced)
__device__ void sym_vec(const Ctx& c, const double* x, double* out) {
    const int n = c.n;
    const int ng = kT / n;
    const int g = threadIdx.x / n, i = threadIdx.x - g * n;
    if (g < ng) {
        double s = 0.0;
        for (int j = g; j < n; j += ng) s = fma(c.P[static_cast<size_t>(j) * n + i], x[j], s);
        c.part[threadIdx.x] = s;
    }
    __syncthreads();
    if (threadIdx.x < n) {
        double s = 0.0;
        for (int gg = 0; gg < ng; ++gg) s += c.part[gg * n + threadIdx.x];
        out[threadIdx.x] = s;
    }
    __syncthreads();
}

// L (lower) := chol( P + shift I + A' diag(wv) A ). Returns false (uniformly) when a pivot is not
// positive.
__device__ bool factorize(const Ctx& c, const double* wv, double shift) {
    const int n = c.n, m = c.m;
    for (int e = threadIdx.x; e < n * n; e += kT) {
        const int i = e / n, j = e - i * n;
        if (j > i) continue;
        double s = c.P[static_cast<size_t>(i) * n + j] + (i == j ? shift : 0.0);
        const double* ai = c.A + i;
        const double* aj = c.A + j;
        for (int k = 0; k < m; ++k) {
            const size_t o = static_cast<size_t>(k) * c.lda;
            s = fma(wv[k] * ai[o], aj[o], s);
        }
        c.L[i * c.ldl + j] = s;
    }
    bool ok = true;
    for (int j = 0; j < n; ++j) {
        __syncthreads();
        const double d = c.L[j * c.ldl + j];
        if (!(d > 0.0)) {
            ok = false;
            break;
        }
        const double inv = 1.0 / sqrt(d);
        __syncthreads();  // everyone has read the pivot
        const int t = threadIdx.x;
        if (t == j) {
            c.L[j * c.ldl + j] = d * inv;
            c.invd[j] = inv;
        } else if (t > j && t < n) {
            c.L[t * c.ldl + j] *= inv;
        }
        __syncthreads();
        const int rem = n - j - 1;
        for (int e = t; e < rem * rem; e += kT) {
            const int a = e / rem, b = e - a * rem;
            if (b <= a) {
                const int i = j + 1 + a, k = j + 1 + b;
                c.L[i * c.ldl + k] = fma(-c.L[i * c.ldl + j], c.L[k * c.ldl + j], c.L[i * c.ldl + k]);
            }
        }
    }
    __syncthreads();
    return ok;
}

// Explicit inverse for the ADMM loop. The triangular solves of chol_solve are a serial chain of 2n
// steps run by one warp while the other seven wait (42 % of the kernel's stall samples in
// profiles/r1g_subkernels_ncu_summary.md); the loop applies the SAME matrix hundreds of times between
// refactorisations, so it pays to form Minv = (L L')^-1 once (W = L^-1 in place, then Minv = W'W by
// all threads) and turn every solve into a dense n x n mat-vec. Minv overwrites c.L:
// callers that need the factor again (polish) refactorise first, which they already do.
__device__ void invert_factor(const Ctx& c) {
    const int n = c.n, t = threadIdx.x;
    // W = L^-1 in place, last column first: W[j+1:, j] = -W[j+1:, j+1:] L[j+1:, j] / L_jj  (dtrti2)
    for (int j = n - 1; j >= 0; --j) {
        double s = 0.0;
        if (t > j && t < n) {
            const double* wi = c.L + t * c.ldl;
            for (int k = j + 1; k <= t; ++k) s = fma(wi[k], c.L[k * c.ldl + j], s);
        }
        __syncthreads();  // column j of L has been read by everyone
        if (t > j && t < n) c.L[t * c.ldl + j] = -s * c.invd[j];
        if (t == j) c.L[j * c.ldl + j] = c.invd[j];
        __syncthreads();
    }
    // Minv = W'W: strictly-lower entries go to the (free) upper triangle, the diagonal to invd, then
    // the upper triangle is mirrored down once nobody reads W any more
    for (int e = t; e < n * n; e += kT) {
        const int a = e / n, b = e - a * n;
        if (b > a) continue;
        double s = 0.0;
        for (int k = a; k < n; ++k) s = fma(c.L[k * c.ldl + a], c.L[k * c.ldl + b], s);
        if (a == b) c.part[a] = s; else c.L[b * c.ldl + a] = s;
    }
    __syncthreads();
    for (int e = t; e < n * n; e += kT) {
        const int a = e / n, b = e - a * n;
        if (b < a) c.L[a * c.ldl + b] = c.L[b * c.ldl + a];
        else if (a == b) c.L[a * c.ldl + a] = c.part[a];
    }
    __syncthreads();
}

// b := Minv b with Minv (symmetric) in c.L. Thread (g, i): columns g, g + ng, ...
__device__ void minv_apply(const Ctx& c, double* b) {
    const int n = c.n;
    const int ng = kT / n;
    const int g = threadIdx.x / n, i = threadIdx.x - g * n;
    if (g < ng) {
        double s = 0.0;
        for (int j = g; j < n; j += ng) s = fma(c.L[j * c.ldl + i], b[j], s);
        c.part[threadIdx.x] = s;
    }
    __syncthreads();
    if (threadIdx.x < n) {
        double s = 0.0;
        for (int gg = 0; gg < ng; ++gg) s += c.part[gg * n + threadIdx.x];
        b[threadIdx.x] = s;
    }
    __syncthreads();
}

// b := (L L')^-1 b, warp 0 only (the other warps wait at the barrier). Rows are dealt to lanes
// round-robin (row i -> lane i & 31, slot i >> 5); the pivot value travels by shuffle.
__device__ void chol_solve(const Ctx& c, double* b) {
    const int n = c.n;
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        double r[4];
#pragma unroll
        for (int s = 0; s < 4; ++s) r[s] = (lane + 32 * s < n) ? b[lane + 32 * s] : 0.0;
        // forward: L y = b
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            for (int jl = 0; jl < 32; ++jl) {
                const int j = 32 * s + jl;
                if (j >= n) break;
                const double xj = __shfl_sync(0xffffffffu, r[s], jl) * c.invd[j];
                if (lane == jl) r[s] = xj;
#pragma unroll
                for (int ss = 0; ss < 4; ++ss) {
                    if (ss < s) continue;
                    const int i = lane + 32 * ss;
                    if (i > j && i < n) r[ss] = fma(-c.L[i * c.ldl + j], xj, r[ss]);
                }
            }
        }
        // backward: L' x = y
#pragma unroll
        for (int s = 3; s >= 0; --s) {
            for (int jl = 31; jl >= 0; --jl) {
                const int j = 32 * s + jl;
                if (j >= n) continue;
                const double xj = __shfl_sync(0xffffffffu, r[s], jl) * c.invd[j];
                if (lane == jl) r[s] = xj;
#pragma unroll
                for (int ss = 0; ss < 4; ++ss) {
                    if (ss > s) continue;
                    const int i = lane + 32 * ss;
                    if (i < j) r[ss] = fma(-c.L[j * c.ldl + i], xj, r[ss]);
                }
            }
        }
#pragma unroll
        for (int s = 0; s < 4; ++s)
            if (lane + 32 * s < n) b[lane + 32 * s] = r[s];
    }
    __syncthreads();
}

// ---- Ruiz equilibration helpers ---------------------------------------------------------------
// out[j] = max( max_i |P[i][j]| (if with_P), max_i |A[i][j]| (if with_A) )
__device__ void col_maxabs(const Ctx& c, bool with_P, bool with_A, double* out) {
    const int n = c.n, m = c.m;
    const int ng = kT / n;
    const int g = threadIdx.x / n, j = threadIdx.x - g * n;
    if (g < ng) {
        double s = 0.0;
        if (with_P)
            for (int i = g; i < n; i += ng) s = fmax(s, fabs(c.P[static_cast<size_t>(i) * n + j]));
        if (with_A)
            for (int i = g; i < m; i += ng) s = fmax(s, fabs(c.A[static_cast<size_t>(i) * c.lda + j]));
        c.part[threadIdx.x] = s;
    }
    __syncthreads();
    if (threadIdx.x < n) {
        double s = 0.0;
        for (int gg = 0; gg < ng; ++gg) s = fmax(s, c.part[gg * n + threadIdx.x]);
        out[threadIdx.x] = s;
    }
    __syncthreads();
}

__device__ __forceinline__ double ruiz_clamp(double v) { return v < 1e-4 ? 1.0 : fmin(v, 1e4); }

// ---- active-set polish --------------------------------------------------------------------------
// Working set from the ADMM iterate (OSQP's rule: slack smaller than the multiplier), then
// equality-constrained solves by the proximal method of multipliers on
//      M = P + delta I + rho_p A_W' A_W
// and repair (add violated rows, drop rows with a wrong-signed multiplier) until the KKT
// conditions hold on the full problem. Returns true with the polished x in `xk`.
struct PolishBuf {
    double *xk, *rhs, *tn, *q;        // n
    double *lam, *wv, *w, *r, *l, *u; // m
    const double *y, *z;              // m (ADMM iterate)
    signed char* act;                 // m: 0 inactive, -1 lower, +1 upper, 2 equality
};

__device__ bool polish(const Ctx& c, const PolishBuf& b, bool a_in_smem, double delta, double rho_p, int max_rounds,
                       int max_inner, int* rounds_out) {
    const int n = c.n, m = c.m;
    for (int i = threadIdx.x; i < m; i += kT) {
        if (b.act[i] == 2) {
            b.lam[i] = b.y[i];
            continue;
        }
        const double yi = b.y[i], zi = b.z[i];
        signed char a = 0;
        if (b.l[i] > -HUGE_VAL && zi - b.l[i] < -yi) a = -1;
        else if (b.u[i] < HUGE_VAL && b.u[i] - zi < yi) a = 1;
        b.act[i] = a;
        b.lam[i] = a < 0 ? fmin(yi, 0.0) : (a > 0 ? fmax(yi, 0.0) : 0.0);
    }
    __syncthreads();
    for (int rnd = 1; rnd <= max_rounds; ++rnd) {
        *rounds_out = rnd;
        for (int i = threadIdx.x; i < m; i += kT) b.wv[i] = b.act[i] != 0 ? rho_p : 0.0;
        __syncthreads();
        if (!factorize(c, b.wv, delta)) return false;
        double bmax = 0.0, qmax = 0.0;
        for (int i = threadIdx.x; i < m; i += kT)
            if (b.act[i] != 0) bmax = fmax(bmax, fabs(b.act[i] == 1 ? b.u[i] : b.l[i]));
        for (int j = threadIdx.x; j < n; j += kT) qmax = fmax(qmax, fabs(b.q[j]));
        {
            double v[2] = {bmax, qmax};
            block_reduce<2, false>(v, c.red);
            bmax = v[0];
            qmax = v[1];
        }
        // Newton steps on the augmented Lagrangian in residual-correction form: the right-hand side
        // is the (small) gradient, so the rounding error of the ill-conditioned solve scales with the
        // step and the iterate converges to full precision.
        mat_vec(c, b.xk, b.r, a_in_smem);
        for (int in = 0; in < max_inner; ++in) {
            for (int i = threadIdx.x; i < m; i += kT) {
                const signed char a = b.act[i];
                const double bi = a == 0 ? 0.0 : (a == 1 ? b.u[i] : b.l[i]);
                b.w[i] = a == 0 ? 0.0 : fma(rho_p, b.r[i] - bi, b.lam[i]);
            }
            __syncthreads();
            mat_tvec(c, b.w, b.tn);
            sym_vec(c, b.xk, b.rhs);
            if (threadIdx.x < n) b.rhs[threadIdx.x] += b.q[threadIdx.x] + b.tn[threadIdx.x];
            __syncthreads();
            chol_solve(c, b.rhs);
            if (threadIdx.x < n) b.xk[threadIdx.x] -= b.rhs[threadIdx.x];
            __syncthreads();
            mat_vec(c, b.xk, b.r, a_in_smem);
            double rmax = 0.0;
            for (int i = threadIdx.x; i < m; i += kT) {
                const signed char a = b.act[i];
                if (a != 0) {
                    const double ri = b.r[i] - (a == 1 ? b.u[i] : b.l[i]);
                    b.lam[i] = fma(rho_p, ri, b.lam[i]);
                    rmax = fmax(rmax, fabs(ri));
                }
            }
            {
                double v[1] = {rmax};
                block_reduce<1, false>(v, c.red);
                rmax = v[0];
            }
            if (rmax < 1e-13 * fmax(1.0, bmax)) {  // stationarity with the updated multipliers
                mat_tvec(c, b.lam, b.tn);
                sym_vec(c, b.xk, b.rhs);
                double v[1] = {threadIdx.x < n ? fabs(b.rhs[threadIdx.x] + b.q[threadIdx.x] + b.tn[threadIdx.x]) : 0.0};
                block_reduce<1, false>(v, c.red);
                if (v[0] < 1e-12 * fmax(1.0, qmax)) break;
            }
        }
        // KKT check on the full problem (b.r = A xk from the last inner iteration)
        double lmax = 0.0;
        for (int i = threadIdx.x; i < m; i += kT)
            if (b.act[i] != 0) lmax = fmax(lmax, fabs(b.lam[i]));
        {
            double v[1] = {lmax};
            block_reduce<1, false>(v, c.red);
            lmax = v[0];
        }
        const double ltol = 1e-10 * fmax(1.0, lmax);
        double changed = 0.0;
        for (int i = threadIdx.x; i < m; i += kT) {
            signed char a = b.act[i];
            if (a == 2) continue;
            const double axi = b.r[i];
            const double tolp = 1e-10 * fmax(1.0, fabs(axi));
            if (a == 0) {
                if (b.l[i] - axi > tolp) a = -1;
                else if (axi - b.u[i] > tolp) a = 1;
            } else if (a == -1) {
                if (b.lam[i] > ltol) a = 0;
            } else {
                if (b.lam[i] < -ltol) a = 0;
            }
            if (a != b.act[i]) {
                changed = 1.0;
                b.act[i] = a;
                if (a == 0) b.lam[i] = 0.0;
            }
        }
        {
            double v[1] = {changed};
            block_reduce<1, false>(v, c.red);
            changed = v[0];
        }
        if (changed == 0.0) return true;
    }
    return false;
}

// ---- the kernel -----------------------------------------------------------------------------------
struct SmemPlan {
    int off_L, off_invd, off_nvec, off_mvec, off_part, off_red, off_A, off_act;  // in doubles (act: bytes from base)
    size_t bytes;
};
constexpr int kNVecs = 8;   // x, xt, rhs, q, D, dx, tn, xk
constexpr int kMVecs = 11;  // z, y, zt, l, u, E, rv, irv, w, dy, lam

__host__ __device__ inline SmemPlan plan_smem(int n, int m, int lda, int ldl, bool a_in_smem) {
    SmemPlan s;
    int o = 0;
    s.off_L = o; o += n * ldl;
    s.off_invd = o; o += n;
    s.off_nvec = o; o += kNVecs * n;
    s.off_mvec = o; o += kMVecs * (m > 0 ? m : 1);
    s.off_part = o; o += kT;
    s.off_red = o; o += kWarps * 8;
    s.off_A = o; if (a_in_smem) o += m * lda;
    s.off_act = o * 8;
    s.bytes = static_cast<size_t>(o) * 8 + ((m + 15) / 16) * 16 + 16;
    return s;
}

__global__ void __launch_bounds__(kT) dense_qp_kernel(const DenseParams p) {
    extern __shared__ __align__(16) unsigned char smraw[];
    double* sm = reinterpret_cast<double*>(smraw);
    const int n = p.n, m = p.m, t = threadIdx.x;
    const SmemPlan pl = plan_smem(n, m, p.lda, p.ldl, p.a_in_smem != 0);
    const int ms = m > 0 ? m : 1;
    double* nv = sm + pl.off_nvec;
    double *x = nv, *xt = nv + n, *rhs = nv + 2 * n, *q = nv + 3 * n, *D = nv + 4 * n, *dx = nv + 5 * n,
           *tn = nv + 6 * n, *xk = nv + 7 * n;
    double* mv = sm + pl.off_mvec;
    double *z = mv, *y = mv + ms, *zt = mv + 2 * ms, *l = mv + 3 * ms, *u = mv + 4 * ms, *E = mv + 5 * ms,
           *rv = mv + 6 * ms, *irv = mv + 7 * ms, *w = mv + 8 * ms, *dy = mv + 9 * ms, *lam = mv + 10 * ms;
    signed char* act = reinterpret_cast<signed char*>(smraw + pl.off_act);
    const bool a_sm = p.a_in_smem != 0;

    for (int qp = blockIdx.x; qp < p.batch; qp += gridDim.x) {
        __syncthreads();
        Ctx c;
        c.n = n; c.m = m; c.lda = p.lda; c.ldl = p.ldl;
        c.A = a_sm ? sm + pl.off_A : p.wsA + static_cast<size_t>(qp) * m * p.lda;
        c.P = p.wsP + static_cast<size_t>(qp) * n * n;
        c.L = sm + pl.off_L;
        c.invd = sm + pl.off_invd;
        c.part = sm + pl.off_part;
        c.red = sm + pl.off_red;
        const double* Pin = p.P + static_cast<size_t>(qp) * n * n;
        const double* Ain = p.A + static_cast<size_t>(qp) * m * n;
        // ---- load: P~ = P + P', A, q, bounds --------------------------------------------------
        for (int e = t; e < n * n; e += kT) {
            const int i = e / n, j = e - i * n;
            c.P[e] = Pin[e] + Pin[j * n + i];
        }
        for (int e = t; e < m * n; e += kT) {
            const int i = e / n, j = e - i * n;
            c.A[static_cast<size_t>(i) * p.lda + j] = Ain[e];
        }
        for (int j = t; j < n; j += kT) {
            q[j] = p.q[static_cast<size_t>(qp) * n + j];
            D[j] = 1.0;
            x[j] = 0.0;
            dx[j] = 0.0;
        }
        for (int i = t; i < m; i += kT) {
            const double li = p.l[static_cast<size_t>(qp) * m + i], ui = p.u[static_cast<size_t>(qp) * m + i];
            l[i] = (li <= -kInfBound) ? -HUGE_VAL : li;
            u[i] = (ui >= kInfBound) ? HUGE_VAL : ui;
            E[i] = 1.0;
            z[i] = 0.0;
            y[i] = 0.0;
            dy[i] = 0.0;
        }
        __syncthreads();
        // ---- Ruiz equilibration (OSQP section 5.1) ----------------------------------------------
        double cs = 1.0;
        for (int it = 0; it < p.ruiz_iters; ++it) {
            col_maxabs(c, true, m > 0, tn);
            if (t < n) tn[t] = 1.0 / sqrt(ruiz_clamp(tn[t]));
            for (int i = t; i < m; i += kT) {  // row norms of A (before this sweep's column scaling)
                const double* a = c.A + static_cast<size_t>(i) * p.lda;
                double s = 0.0;
                for (int j = 0; j < n; ++j) s = fmax(s, fabs(a[j]));
                w[i] = 1.0 / sqrt(ruiz_clamp(s));
            }
            __syncthreads();
            for (int e = t; e < n * n; e += kT) {
                const int i = e / n, j = e - i * n;
                c.P[e] *= tn[i] * tn[j];
            }
            for (int e = t; e < m * n; e += kT) {
                const int i = e / n, j = e - i * n;
                c.A[static_cast<size_t>(i) * p.lda + j] *= w[i] * tn[j];
            }
            if (t < n) {
                q[t] *= tn[t];
                D[t] *= tn[t];
            }
            for (int i = t; i < m; i += kT) E[i] *= w[i];
            __syncthreads();
            col_maxabs(c, true, false, xt);
            double v[2] = {t < n ? xt[t] : 0.0, 0.0};
            double qm[1] = {t < n ? fabs(q[t]) : 0.0};
            block_reduce<2, true>(v, c.red);
            block_reduce<1, false>(qm, c.red);
            const double g = 1.0 / ruiz_clamp(fmax(v[0] / n, qm[0]));
            for (int e = t; e < n * n; e += kT) c.P[e] *= g;
            if (t < n) q[t] *= g;
            cs *= g;
            __syncthreads();
        }
        // ---- scaled bounds, row classes, rho vector ----------------------------------------------
        for (int i = t; i < m; i += kT) {
            if (l[i] > -HUGE_VAL) l[i] *= E[i];
            if (u[i] < HUGE_VAL) u[i] *= E[i];
            const bool fin = l[i] > -HUGE_VAL && u[i] < HUGE_VAL;
            act[i] = (fin && (u[i] - l[i]) <= 1e-12 * fmax(1.0, fabs(u[i]))) ? 2 : 0;
        }
        __syncthreads();
        double rho = p.rho0;
        auto set_rho = [&](double r) {
            for (int i = t; i < m; i += kT) {
                const bool fr = !(l[i] > -HUGE_VAL) && !(u[i] < HUGE_VAL);
                const double ri = act[i] == 2 ? 1e3 * r : (fr ? 1e-6 : r);
                rv[i] = ri;
                irv[i] = 1.0 / ri;
            }
            __syncthreads();
        };
        set_rho(rho);
        int status = MPCCBF_STATUS_UNKNOWN;
        bool have_x = false;
        int iter = 0;
        bool fact_ok = factorize(c, rv, p.sigma);
        if (fact_ok) invert_factor(c);
        if (!fact_ok) status = MPCCBF_STATUS_ERROR;
        PolishBuf pb;
        pb.xk = xk; pb.rhs = rhs; pb.tn = tn; pb.q = q;
        pb.lam = lam; pb.wv = w; pb.w = zt; pb.r = dy; pb.l = l; pb.u = u; pb.y = y; pb.z = z; pb.act = act;
        // NOTE: polish borrows w (weights), zt (rhs weights) and dy (A xk); the ADMM loop rebuilds
        // w and zt every iteration and dy is only read at check iterations after being rewritten.
        int since_polish = 0;
        int polish_gap = 2 * p.check_every;
        while (fact_ok && iter < p.max_iter) {
            ++iter;
            const bool check = (iter % p.check_every) == 0 || iter == p.max_iter;
            // x~ = M^-1 (sigma x - q + A'(rho z - y))
            for (int i = t; i < m; i += kT) w[i] = fma(rv[i], z[i], -y[i]);
            __syncthreads();
            if (m > 0) mat_tvec(c, w, rhs);
            if (t < n) rhs[t] = (m > 0 ? rhs[t] : 0.0) + fma(p.sigma, x[t], -q[t]);
            __syncthreads();
            minv_apply(c, rhs);  // rhs now holds x~
            if (m > 0) mat_vec(c, rhs, zt, a_sm);
            if (t < n) {
                const double xn = fma(p.alpha, rhs[t], (1.0 - p.alpha) * x[t]);
                dx[t] = xn - x[t];
                x[t] = xn;
            }
            for (int i = t; i < m; i += kT) {
                const double zr = fma(p.alpha, zt[i], (1.0 - p.alpha) * z[i]);
                const double zn = fmin(fmax(fma(y[i], irv[i], zr), l[i]), u[i]);
                const double yn = fma(rv[i], zr - zn, y[i]);
                dy[i] = yn - y[i];
                z[i] = zn;
                y[i] = yn;
            }
            __syncthreads();
            if (!check) continue;
            // ---- residuals (unscaled, OSQP section 5.1) -------------------------------------------
            if (m > 0) mat_vec(c, x, zt, a_sm);       // zt := A x
            sym_vec(c, x, rhs);                        // rhs := P x
            if (m > 0) mat_tvec(c, y, xt);             // xt := A'y
            double v[6] = {0, 0, 0, 0, 0, 0};          // rp, |Ax|, |z|, rd, max(|Px|,|A'y|), |q|
            for (int i = t; i < m; i += kT) {
                const double ie = 1.0 / E[i];
                v[0] = fmax(v[0], fabs(zt[i] - z[i]) * ie);
                v[1] = fmax(v[1], fabs(zt[i]) * ie);
                v[2] = fmax(v[2], fabs(z[i]) * ie);
            }
            if (t < n) {
                const double id = 1.0 / D[t];
                const double aty = m > 0 ? xt[t] : 0.0;
                v[3] = fabs(rhs[t] + q[t] + aty) * id;
                v[4] = fmax(fabs(rhs[t]), fabs(aty)) * id;
                v[5] = fabs(q[t]) * id;
            }
            block_reduce<6, false>(v, c.red);
            const double ics = 1.0 / cs;
            const double rp = v[0], np_ = fmax(v[1], v[2]);
            const double rd = v[3] * ics, nd = fmax(v[4], v[5]) * ics;
            const bool converged = rp <= p.eps + p.eps * np_ && rd <= p.eps + p.eps * nd;
            since_polish += p.check_every;
            const bool loose = rp <= 1e-2 * (1.0 + np_) && rd <= 1e-2 * (1.0 + nd);
            if (converged || (loose && since_polish >= polish_gap) || iter == p.max_iter) {
                since_polish = 0;
                polish_gap = min(2 * polish_gap, 64 * p.check_every);  // a failed attempt costs ~100 iterations: back off
                if (t < n) xk[t] = x[t];
                __syncthreads();
                int rounds = 0;
                const bool ok = m > 0 ? polish(c, pb, a_sm, 1e-7, 1e4, 12, 12, &rounds) : false;
                if (ok) {
                    if (t < n) x[t] = xk[t];
                    __syncthreads();
                    status = MPCCBF_STATUS_OPTIMAL;
                    have_x = true;
                    break;
                }
                if (converged) {  // unpolished ADMM solution at tolerance eps
                    status = MPCCBF_STATUS_OPTIMAL;
                    have_x = true;
                    break;
                }
                if (iter == p.max_iter) break;
                fact_ok = factorize(c, rv, p.sigma);  // polish overwrote the cached factor
                if (fact_ok) invert_factor(c);
                if (!fact_ok) {
                    status = MPCCBF_STATUS_ERROR;
                    break;
                }
                continue;
            }
            // ---- primal infeasibility certificate (delta y) ----------------------------------------
            if (m > 0) {
                mat_tvec(c, dy, xt);  // xt := A' dy
                double s[2] = {0.0, 0.0};   // |E dy|_inf, |A'dy / D|_inf
                double sup[2] = {0.0, 0.0}; // support function, "bad sign on an infinite bound" flag
                for (int i = t; i < m; i += kT) s[0] = fmax(s[0], fabs(E[i] * dy[i]));
                if (t < n) s[1] = fabs(xt[t]) / D[t];
                block_reduce<2, false>(s, c.red);
                const double thr = p.eps_inf * s[0];
                for (int i = t; i < m; i += kT) {
                    const double d = dy[i];
                    if (d > 0.0) {
                        if (u[i] < HUGE_VAL) sup[0] = fma(u[i], d, sup[0]);
                        else if (E[i] * d > thr) sup[1] += 1.0;
                    } else if (d < 0.0) {
                        if (l[i] > -HUGE_VAL) sup[0] = fma(l[i], d, sup[0]);
                        else if (-E[i] * d > thr) sup[1] += 1.0;
                    }
                }
                block_reduce<2, true>(sup, c.red);
                if (s[0] > 1e-30 && s[1] <= thr && sup[0] < -thr && sup[1] == 0.0) {
                    status = MPCCBF_STATUS_INFEASIBLE;
                    break;
                }
            }
            // ---- dual infeasibility certificate (delta x) ------------------------------------------
            {
                sym_vec(c, dx, rhs);                    // rhs := P dx
                if (m > 0) mat_vec(c, dx, zt, a_sm);    // zt := A dx
                double s[2] = {0.0, 0.0};               // |D dx|_inf, |P dx / D|_inf
                double g[2] = {0.0, 0.0};               // q'dx, recession-cone violations
                if (t < n) {
                    s[0] = fabs(D[t] * dx[t]);
                    s[1] = fabs(rhs[t]) / D[t];
                    g[0] = q[t] * dx[t];
                }
                block_reduce<2, false>(s, c.red);
                const double thr = p.eps_inf * s[0];
                for (int i = t; i < m; i += kT) {
                    const double a = zt[i] / E[i];
                    if ((u[i] < HUGE_VAL && a > thr) || (l[i] > -HUGE_VAL && a < -thr)) g[1] += 1.0;
                }
                block_reduce<2, true>(g, c.red);
                if (s[0] > 1e-30 && s[1] <= thr * cs && g[0] < -thr * cs && g[1] == 0.0) {
                    status = MPCCBF_STATUS_UNBOUNDED;
                    break;
                }
            }
            // ---- adaptive rho ------------------------------------------------------------------------
            if ((iter % (5 * p.check_every)) == 0) {
                const double a = rp / fmax(np_, 1e-30), b = rd / fmax(nd, 1e-30);
                double nr = rho * sqrt(a / fmax(b, 1e-30));
                nr = fmin(fmax(nr, 1e-6), 1e6);
                if (nr > 5.0 * rho || nr < 0.2 * rho) {
                    rho = nr;
                    set_rho(rho);
                    fact_ok = factorize(c, rv, p.sigma);
                    if (fact_ok) invert_factor(c);
                    if (!fact_ok) status = MPCCBF_STATUS_ERROR;
                }
            }
        }
        // ---- outputs (unscaled) ---------------------------------------------------------------------
        double obj = 0.0;
        if (have_x) {
            sym_vec(c, x, rhs);
            double v[1] = {t < n ? x[t] * fma(0.5, rhs[t], q[t]) : 0.0};
            block_reduce<1, true>(v, c.red);
            obj = v[0] / cs;
        }
        if (t < n) p.x[static_cast<size_t>(qp) * n + t] = have_x ? D[t] * x[t] : 0.0;
        if (t == 0) {
            p.status[qp] = status;
            if (p.obj) p.obj[qp] = obj;
            if (p.iters) p.iters[qp] = iter;
        }
    }
}

}  // namespace

#define DQ_TRY(expr)                                                                \
    do {                                                                            \
        cudaError_t _e = (expr);                                                    \
        if (_e != cudaSuccess) {                                                    \
            set_error(std::string(#expr) + ": " + cudaGetErrorString(_e));          \
            rc = MPCCBF_ERR_CUDA;                                                   \
            goto done;                                                              \
        }                                                                           \
    } while (0)

// Launch geometry and solver settings shared by the host- and device-pointer entries.
static bool plan_launch(const DenseOpts& o, int batch, int n, int mt, DenseParams& p, SmemPlan& pl) {
    p.n = n; p.m = mt; p.batch = batch;
    p.lda = n | 1;
    p.ldl = n | 1;
    pl = plan_smem(n, mt, p.lda, p.ldl, true);
    p.a_in_smem = pl.bytes <= 200 * 1024 ? 1 : 0;
    if (!p.a_in_smem) pl = plan_smem(n, mt, p.lda, p.ldl, false);
    if (pl.bytes > 227 * 1024) return false;
    p.max_iter = o.max_iter > 0 ? o.max_iter : 20000;
    p.check_every = o.check_every > 0 ? o.check_every : 10;
    p.ruiz_iters = 15;
    p.eps = o.eps > 0 ? o.eps : 1e-7;
    p.rho0 = o.rho > 0 ? o.rho : 0.1;
    p.sigma = o.sigma > 0 ? o.sigma : 1e-6;
    p.alpha = 1.6;
    p.eps_inf = 1e-5;
    return true;
}

static int launch_kernel(const DenseOpts& o, const DenseParams& p, const SmemPlan& pl) {
    cudaError_t e = cudaFuncSetAttribute(dense_qp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(pl.bytes));
    int per_sm = 1;
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, dense_qp_kernel, kT, pl.bytes);
    if (e == cudaSuccess) {
        // one CTA per QP: iteration counts vary by 100x between QPs, so the hardware block scheduler
        // balances the SMs instead of a static grid-stride assignment
        (void)per_sm;
        dense_qp_kernel<<<p.batch, kT, pl.bytes, o.stream>>>(p);
        e = cudaGetLastError();
    }
    if (e != cudaSuccess) {
        set_error(std::string("dense_qp_kernel launch: ") + cudaGetErrorString(e));
        return MPCCBF_ERR_CUDA;
    }
    return MPCCBF_OK;
}

// Device-pointer entry: QPs already assembled in HBM (slack_path.cu). Bytes of per-QP scratch the
// caller provides: wsP n*n doubles, wsA m*(n|1) doubles when dense_qp_needs_wsA says so.
bool dense_qp_needs_wsA(int n, int m) {
    const SmemPlan pl = plan_smem(n, m, n | 1, n | 1, true);
    return pl.bytes > 200 * 1024;
}

int dense_qp_solve_device(const DenseOpts& o, int32_t batch, int32_t n, int32_t m, const double* P, const double* q,
                          const double* A, const double* l, const double* u, double* wsP, double* wsA, double* x,
                          double* objective, int32_t* status, int32_t* iters) {
    if (batch == 0) return MPCCBF_OK;
    DenseParams p;
    SmemPlan pl;
    if (n > kMaxN || !plan_launch(o, batch, n, m, p, pl)) {
        set_error("dense QP kernel: more than 128 variables or too many rows for one CTA's shared memory");
        return MPCCBF_ERR_UNSUPPORTED;
    }
    p.P = P; p.q = q; p.A = A; p.l = l; p.u = u;
    p.wsP = wsP; p.wsA = wsA; p.x = x; p.obj = objective; p.status = status; p.iters = iters;
    return launch_kernel(o, p, pl);
}

// Host-pointer entry (mpccbf_solve_qp_dense). Variable bounds become identity rows.
int dense_qp_solve_host(const DenseOpts& o, int32_t batch, int32_t n, int32_t m, const double* P,
                        const double* q, const double* A, const double* l, const double* u,
                        const double* lb, const double* ub, double* x, double* objective,
                        int32_t* status, int32_t* iters) {
    if (batch == 0) return MPCCBF_OK;
    if (n > kMaxN) {
        set_error("mpccbf_solve_qp_dense: n > 128 decision variables is not supported by the one-CTA-per-QP kernel");
        return MPCCBF_ERR_UNSUPPORTED;
    }
    // variables with a finite bound in any QP of the batch get a row e_j
    std::vector<int> bvars;
    if (lb && ub)
        for (int j = 0; j < n; ++j) {
            bool fin = false;
            for (int b = 0; b < batch && !fin; ++b)
                fin = std::fabs(lb[static_cast<size_t>(b) * n + j]) < kInfBound || std::fabs(ub[static_cast<size_t>(b) * n + j]) < kInfBound;
            if (fin) bvars.push_back(j);
        }
    const int mt = m + static_cast<int>(bvars.size());
    std::vector<double> Ax, lx, ux;
    const double *Ah = A, *lh = l, *uh = u;
    if (!bvars.empty()) {
        Ax.assign(static_cast<size_t>(batch) * mt * n, 0.0);
        lx.resize(static_cast<size_t>(batch) * mt);
        ux.resize(static_cast<size_t>(batch) * mt);
        for (int b = 0; b < batch; ++b) {
            double* Ab = Ax.data() + static_cast<size_t>(b) * mt * n;
            if (m > 0) std::copy(A + static_cast<size_t>(b) * m * n, A + static_cast<size_t>(b + 1) * m * n, Ab);
            for (int i = 0; i < m; ++i) {
                lx[static_cast<size_t>(b) * mt + i] = l[static_cast<size_t>(b) * m + i];
                ux[static_cast<size_t>(b) * mt + i] = u[static_cast<size_t>(b) * m + i];
            }
            for (size_t r = 0; r < bvars.size(); ++r) {
                const int j = bvars[r];
                Ab[static_cast<size_t>(m + r) * n + j] = 1.0;
                lx[static_cast<size_t>(b) * mt + m + r] = lb[static_cast<size_t>(b) * n + j];
                ux[static_cast<size_t>(b) * mt + m + r] = ub[static_cast<size_t>(b) * n + j];
            }
        }
        Ah = Ax.data();
        lh = lx.data();
        uh = ux.data();
    }
    DenseParams p;
    SmemPlan pl;
    if (!plan_launch(o, batch, n, mt, p, pl)) {
        set_error("mpccbf_solve_qp_dense: too many rows for the shared-memory vectors of one CTA");
        return MPCCBF_ERR_UNSUPPORTED;
    }
    int rc = MPCCBF_OK;
    double *dP = nullptr, *dq = nullptr, *dA = nullptr, *dl = nullptr, *du = nullptr, *dwP = nullptr, *dwA = nullptr,
           *dx = nullptr, *dobj = nullptr;
    int32_t *dst = nullptr, *dit = nullptr;
    const size_t B = batch;
    cudaStream_t st = o.stream;
    {
        DQ_TRY(cudaMalloc(&dP, sizeof(double) * B * n * n));
        DQ_TRY(cudaMalloc(&dq, sizeof(double) * B * n));
        DQ_TRY(cudaMalloc(&dA, sizeof(double) * std::max<size_t>(B * mt * n, 1)));
        DQ_TRY(cudaMalloc(&dl, sizeof(double) * std::max<size_t>(B * mt, 1)));
        DQ_TRY(cudaMalloc(&du, sizeof(double) * std::max<size_t>(B * mt, 1)));
        DQ_TRY(cudaMalloc(&dwP, sizeof(double) * B * n * n));
        DQ_TRY(cudaMalloc(&dwA, sizeof(double) * std::max<size_t>(p.a_in_smem ? 1 : B * mt * p.lda, 1)));
        DQ_TRY(cudaMalloc(&dx, sizeof(double) * B * n));
        DQ_TRY(cudaMalloc(&dobj, sizeof(double) * B));
        DQ_TRY(cudaMalloc(&dst, sizeof(int32_t) * B));
        DQ_TRY(cudaMalloc(&dit, sizeof(int32_t) * B));
        DQ_TRY(cudaMemcpyAsync(dP, P, sizeof(double) * B * n * n, cudaMemcpyHostToDevice, st));
        DQ_TRY(cudaMemcpyAsync(dq, q, sizeof(double) * B * n, cudaMemcpyHostToDevice, st));
        if (mt > 0) {
            DQ_TRY(cudaMemcpyAsync(dA, Ah, sizeof(double) * B * mt * n, cudaMemcpyHostToDevice, st));
            DQ_TRY(cudaMemcpyAsync(dl, lh, sizeof(double) * B * mt, cudaMemcpyHostToDevice, st));
            DQ_TRY(cudaMemcpyAsync(du, uh, sizeof(double) * B * mt, cudaMemcpyHostToDevice, st));
        }
        p.P = dP; p.q = dq; p.A = dA; p.l = dl; p.u = du;
        p.wsP = dwP; p.wsA = dwA; p.x = dx; p.obj = dobj; p.status = dst; p.iters = dit;
        rc = launch_kernel(o, p, pl);
        if (rc != MPCCBF_OK) goto done;
        DQ_TRY(cudaMemcpyAsync(x, dx, sizeof(double) * B * n, cudaMemcpyDeviceToHost, st));
        DQ_TRY(cudaMemcpyAsync(status, dst, sizeof(int32_t) * B, cudaMemcpyDeviceToHost, st));
        if (objective) DQ_TRY(cudaMemcpyAsync(objective, dobj, sizeof(double) * B, cudaMemcpyDeviceToHost, st));
        if (iters) DQ_TRY(cudaMemcpyAsync(iters, dit, sizeof(int32_t) * B, cudaMemcpyDeviceToHost, st));
        DQ_TRY(cudaStreamSynchronize(st));
    }
done:
    for (void* ptr : {static_cast<void*>(dP), static_cast<void*>(dq), static_cast<void*>(dA), static_cast<void*>(dl),
                      static_cast<void*>(du), static_cast<void*>(dwP), static_cast<void*>(dwA), static_cast<void*>(dx),
                      static_cast<void*>(dobj), static_cast<void*>(dst), static_cast<void*>(dit)})
        if (ptr) cudaFree(ptr);
    return rc;
}

}  // namespace mpccbf

// ---- allocation-free entry (mpccbf_solve_qp_dense_noalloc) ----------------------------------------------------------
// For callers that may not call cudaMalloc (CUDA graph capture, serving loops, harnesses that forbid it): the
// kernel's per-QP scratch (scaled P~, and scaled A when it does not fit in shared memory) lives in a file-scope device
// array of fixed size instead of an allocation. Every GPU that loads the library holds the array until exit; every call
// shares it, so calls that overlap on the GPU race. The kernel writes each QP's scratch before it reads it; the used
// part is cleared first all the same, so a result cannot depend on what the previous call left behind.
namespace {

constexpr size_t kNoAllocDoubles = MPCCBF_NOALLOC_SCRATCH_BYTES / sizeof(double);
static __device__ __align__(256) double g_noalloc_scratch[kNoAllocDoubles];

constexpr int kNoAllocMaxDevices = 64;
double* g_noalloc_ptr[kNoAllocMaxDevices];  // device address of g_noalloc_scratch per device (the name is a host stand-in)

// Once per device, with that device current: the device address of the array.
int noalloc_scratch(double** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kNoAllocMaxDevices) {
        mpccbf::set_error(std::string("mpccbf_solve_qp_dense_noalloc: cudaGetDevice: ") +
                          (e != cudaSuccess ? cudaGetErrorString(e) : "device ordinal out of range"));
        return MPCCBF_ERR_CUDA;
    }
    double* p = __atomic_load_n(&g_noalloc_ptr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        void* sym = nullptr;
        e = cudaGetSymbolAddress(&sym, g_noalloc_scratch);
        if (e != cudaSuccess || !sym) {
            mpccbf::set_error(std::string("mpccbf_solve_qp_dense_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
            return MPCCBF_ERR_CUDA;
        }
        p = static_cast<double*>(sym);
        __atomic_store_n(&g_noalloc_ptr[dev], p, __ATOMIC_RELEASE);  // every thread computes the same value
    }
    *out = p;
    return MPCCBF_OK;
}

}  // namespace

extern "C" int mpccbf_solve_qp_dense_noalloc(const mpccbf_config* cfg, int32_t batch, int32_t n, int32_t m, const double* P,
                                             const double* q, const double* A, const double* l, const double* u, double* x,
                                             double* objective, int32_t* status, int32_t* iters, void* cuda_stream) {
    using namespace mpccbf;
    if (batch < 0 || n < 1 || m < 0 || !P || !q || (m > 0 && (!A || !l || !u)) || !x || !objective || !status || !iters) {
        set_error("mpccbf_solve_qp_dense_noalloc: invalid argument");
        return MPCCBF_ERR_INVALID;
    }
    if (batch == 0) return MPCCBF_OK;
    if (n > kMaxN) {
        set_error("mpccbf_solve_qp_dense_noalloc: n > 128 decision variables is not supported by the one-CTA-per-QP kernel");
        return MPCCBF_ERR_UNSUPPORTED;
    }
    // scratch: [batch][n*n] scaled P~, then (256-byte aligned) [batch][m*(n|1)] scaled A if it does not fit in shared memory
    const size_t B = static_cast<size_t>(batch);
    const size_t nP = (B * n * n + 31) / 32 * 32;
    const size_t nA = dense_qp_needs_wsA(n, m) ? B * m * (n | 1) : 0;
    if (nP + nA > kNoAllocDoubles) {
        set_error("mpccbf_solve_qp_dense_noalloc: batch * n * n (+ batch * m * (n|1) when A does not fit in shared memory) = " +
                  std::to_string(nP + nA) + " doubles of scratch exceed the fixed " + std::to_string(kNoAllocDoubles) +
                  " (MPCCBF_NOALLOC_SCRATCH_BYTES); split the batch or use mpccbf_solve_qp_dense");
        return MPCCBF_ERR_UNSUPPORTED;
    }
    double* scratch = nullptr;
    int rc = noalloc_scratch(&scratch);
    if (rc != MPCCBF_OK) return rc;
    DenseOpts o;
    o.max_iter = cfg ? cfg->admm_max_iter : 0;
    o.check_every = cfg ? cfg->admm_check_every : 0;
    o.eps = cfg ? cfg->admm_eps : 0.0;
    o.rho = cfg ? cfg->admm_rho : 0.0;
    o.sigma = cfg ? cfg->admm_sigma : 0.0;
    o.sm_count = 0;
    o.stream = static_cast<cudaStream_t>(cuda_stream);
    cudaError_t e = cudaMemsetAsync(scratch, 0, sizeof(double) * (nP + nA), o.stream);
    if (e != cudaSuccess) {
        set_error(std::string("mpccbf_solve_qp_dense_noalloc: cudaMemsetAsync: ") + cudaGetErrorString(e));
        return MPCCBF_ERR_CUDA;
    }
    return dense_qp_solve_device(o, batch, n, m, P, q, A, l, u, scratch, nA ? scratch + nP : nullptr, x, objective, status, iters);
}
