// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// Part of the CPU oracle (see oracle/README.md). Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may build, link or call this.
//
// Minimal dense row-major linear algebra, templated on the scalar so the same
// restatement runs in `long double` (oracle accuracy) and `double` (timed CPU baseline,
// the analogue of the reference's Eigen<double> path, lib/math/include/math/Types.h:8-30).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <stdexcept>
#include <vector>

namespace oracle {

template <typename R>
struct Mat {
    int rows = 0, cols = 0;
    std::vector<R> a;
    Mat() = default;
    Mat(int r, int c) : rows(r), cols(c), a(static_cast<size_t>(r) * c, R(0)) {}
    R& operator()(int i, int j) { return a[static_cast<size_t>(i) * cols + j]; }
    const R& operator()(int i, int j) const { return a[static_cast<size_t>(i) * cols + j]; }
    R* row(int i) { return a.data() + static_cast<size_t>(i) * cols; }
    const R* row(int i) const { return a.data() + static_cast<size_t>(i) * cols; }
    static Mat identity(int n) {
        Mat m(n, n);
        for (int i = 0; i < n; ++i) m(i, i) = R(1);
        return m;
    }
};

template <typename R>
using Vec = std::vector<R>;

template <typename R>
Mat<R> matmul(const Mat<R>& A, const Mat<R>& B) {
    if (A.cols != B.rows) throw std::runtime_error("matmul: shape mismatch");
    Mat<R> C(A.rows, B.cols);
    for (int i = 0; i < A.rows; ++i)
        for (int k = 0; k < A.cols; ++k) {
            const R aik = A(i, k);
            if (aik == R(0)) continue;
            const R* b = B.row(k);
            R* c = C.row(i);
            for (int j = 0; j < B.cols; ++j) c[j] += aik * b[j];
        }
    return C;
}

template <typename R>
Mat<R> transpose(const Mat<R>& A) {
    Mat<R> T(A.cols, A.rows);
    for (int i = 0; i < A.rows; ++i)
        for (int j = 0; j < A.cols; ++j) T(j, i) = A(i, j);
    return T;
}

template <typename R>
Vec<R> matvec(const Mat<R>& A, const Vec<R>& x) {
    if (A.cols != static_cast<int>(x.size())) throw std::runtime_error("matvec: shape mismatch");
    Vec<R> y(A.rows, R(0));
    for (int i = 0; i < A.rows; ++i) {
        R s = 0;
        const R* a = A.row(i);
        for (int j = 0; j < A.cols; ++j) s += a[j] * x[j];
        y[i] = s;
    }
    return y;
}

template <typename R>
Vec<R> matTvec(const Mat<R>& A, const Vec<R>& x) {
    if (A.rows != static_cast<int>(x.size())) throw std::runtime_error("matTvec: shape mismatch");
    Vec<R> y(A.cols, R(0));
    for (int i = 0; i < A.rows; ++i) {
        const R xi = x[i];
        if (xi == R(0)) continue;
        const R* a = A.row(i);
        for (int j = 0; j < A.cols; ++j) y[j] += a[j] * xi;
    }
    return y;
}

template <typename R>
R dot(const R* a, const R* b, int n) {
    R s = 0;
    for (int i = 0; i < n; ++i) s += a[i] * b[i];
    return s;
}

template <typename R>
R norm_inf(const Vec<R>& v) {
    R m = 0;
    for (const R& x : v) m = std::max(m, std::fabs(x));
    return m;
}

// In-place lower Cholesky A = L L^T of a symmetric matrix; returns false when a pivot
// is <= tol (not positive definite at working precision).
template <typename R>
bool cholesky_lower(Mat<R>& A, R tol) {
    const int n = A.rows;
    for (int j = 0; j < n; ++j) {
        R d = A(j, j);
        for (int k = 0; k < j; ++k) d -= A(j, k) * A(j, k);
        if (!(d > tol)) return false;
        d = std::sqrt(d);
        A(j, j) = d;
        for (int i = j + 1; i < n; ++i) {
            R s = A(i, j);
            for (int k = 0; k < j; ++k) s -= A(i, k) * A(j, k);
            A(i, j) = s / d;
        }
        for (int i = 0; i < j; ++i) A(i, j) = R(0);
    }
    return true;
}

// Solve L y = b (forward) and L^T x = y (backward) for lower-triangular L.
template <typename R>
void solve_lower(const Mat<R>& L, Vec<R>& b) {
    const int n = L.rows;
    for (int i = 0; i < n; ++i) {
        R s = b[i];
        for (int k = 0; k < i; ++k) s -= L(i, k) * b[k];
        b[i] = s / L(i, i);
    }
}
template <typename R>
void solve_lower_T(const Mat<R>& L, Vec<R>& b) {
    const int n = L.rows;
    for (int i = n - 1; i >= 0; --i) {
        R s = b[i];
        for (int k = i + 1; k < n; ++k) s -= L(k, i) * b[k];
        b[i] = s / L(i, i);
    }
}

// Householder QR with column pivoting of an (m x n) matrix M (overwritten by R);
// accumulates Q (m x m) explicitly. Returns the numerical rank.
// Used for the null-space of the equality block (columns of M = equality rows).
template <typename R>
int qr_colpiv(Mat<R>& M, Mat<R>& Q, std::vector<int>& perm, R rel_tol) {
    const int m = M.rows, n = M.cols;
    Q = Mat<R>::identity(m);
    perm.resize(n);
    for (int j = 0; j < n; ++j) perm[j] = j;
    Vec<R> cn(n);
    R max0 = 0;
    for (int j = 0; j < n; ++j) {
        R s = 0;
        for (int i = 0; i < m; ++i) s += M(i, j) * M(i, j);
        cn[j] = s;
        max0 = std::max(max0, s);
    }
    int rank = 0;
    const int kmax = std::min(m, n);
    Vec<R> v(m);
    for (int k = 0; k < kmax; ++k) {
        // pivot: recompute the remaining column norms exactly (sizes are tiny)
        int piv = k;
        R best = -1;
        for (int j = k; j < n; ++j) {
            R s = 0;
            for (int i = k; i < m; ++i) s += M(i, j) * M(i, j);
            cn[j] = s;
            if (s > best) { best = s; piv = j; }
        }
        if (!(best > rel_tol * rel_tol * std::max(max0, R(1e-300)))) break;
        if (piv != k) {
            for (int i = 0; i < m; ++i) std::swap(M(i, k), M(i, piv));
            std::swap(perm[k], perm[piv]);
        }
        R alpha = std::sqrt(best);
        if (M(k, k) > 0) alpha = -alpha;
        for (int i = 0; i < m; ++i) v[i] = 0;
        for (int i = k; i < m; ++i) v[i] = M(i, k);
        v[k] -= alpha;
        R vv = 0;
        for (int i = k; i < m; ++i) vv += v[i] * v[i];
        if (vv > 0) {
            for (int j = k; j < n; ++j) {
                R s = 0;
                for (int i = k; i < m; ++i) s += v[i] * M(i, j);
                s = 2 * s / vv;
                for (int i = k; i < m; ++i) M(i, j) -= s * v[i];
            }
            for (int r = 0; r < m; ++r) {  // Q <- Q * Hk
                R s = 0;
                for (int i = k; i < m; ++i) s += Q(r, i) * v[i];
                s = 2 * s / vv;
                for (int i = k; i < m; ++i) Q(r, i) -= s * v[i];
            }
        }
        ++rank;
    }
    return rank;
}

}  // namespace oracle
