// TEST INFRASTRUCTURE — NOT PRODUCT CODE (see oracle/README.md).
//
// CPU restatement of the reference's spline / model / combinatorics layer, following the
// reference's own loops (same summation order) so the constant tables agree to the last
// few ulps. Each function cites the reference file:line it restates
// (paths relative to /root/reference/workspace/lib).
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>

#include "linalg.hpp"

namespace oracle {

// math/src/Combinatorics.cpp:11-20 — fac(n), throws for n > 20.
inline uint64_t fac(uint64_t n) {
    if (n > 20) throw std::runtime_error("warning: fac overflow");
    uint64_t r = 1;
    for (uint64_t i = 2; i <= n; ++i) r *= i;
    return r;
}
// math/src/Combinatorics.cpp:22-32 — comb(n,k) = top/bottom with k := min(k, n-k).
inline uint64_t comb(uint64_t n, uint64_t k) {
    if (k > n) return 0;
    k = std::min(k, n - k);
    uint64_t top = 1, bottom = 1;
    for (uint64_t i = 0; i < k; ++i) {
        bottom *= (i + 1);
        top *= (n - i);
    }
    return top / bottom;
}
// math/src/Combinatorics.cpp:34-41 — perm(n,k) = n!/(n-k)!.
inline uint64_t perm(uint64_t n, uint64_t k) {
    if (k > n) return 0;
    uint64_t r = 1;
    for (uint64_t i = n - k + 1; i <= n; ++i) r *= i;
    return r;
}
// math/src/Combinatorics.cpp:43-52 — pow with 0^0 = 1.
template <typename R>
R ipow(R base, uint64_t e) {
    if (base == R(0) && e == 0) return R(1);
    return std::pow(base, static_cast<R>(e));
}

// splines/src/detail/BezierOperations.cpp:11-50 — row of d-th derivative Bernstein basis
// functions of a degree-`deg` Bezier on [0, T], evaluated at `t` (monomial expansion).
template <typename R>
Vec<R> bernstein_basis(uint64_t deg, R T, R t, uint64_t d) {
    if (t < 0 || t > T)
        throw std::runtime_error("bernsteinBasis:: given parameter is outside of [0, max_parameter].");
    Vec<R> out(deg + 1, R(0));
    if (T == 0) {
        if (d == 0) out[0] = 1;
        return out;
    }
    const R oneOverA = R(1) / T;
    for (uint64_t i = 0; i <= deg; ++i) {
        R base = 0, mult = 1;
        for (uint64_t j = 0; j + d <= deg; ++j, mult *= t) {
            if (j + d >= i) {
                const uint64_t c = comb(deg - i, j + d - i);
                const uint64_t p = perm(j + d, d);
                const R pw = ipow<R>(oneOverA, j + d);
                base += R(c) * pw * R(p) * mult * R(((j + d - i) % 2 == 0) ? 1 : -1);
            }
        }
        base *= R(comb(deg, i));
        out[i] = base;
    }
    return out;
}

// splines/src/detail/BezierOperations.cpp:54-121 — (deg+1)x(deg+1) matrix B_d with
// basis_d(t)[i] = sum_j B_d(i,j) t^j.
template <typename R>
Mat<R> bernstein_coefficient_matrix(uint64_t deg, R T, uint64_t d) {
    if (T < 0) throw std::runtime_error("bernsteinCoefficientMatrix: max_parameter is negative.");
    const int n1 = static_cast<int>(deg) + 1;
    Mat<R> B(n1, n1);
    if (T == 0) {
        if (d == 0) B(0, 0) = 1;
        return B;
    }
    uint64_t dcombi = 1;
    const R oneOverA = R(1) / T;
    for (uint64_t i = 0; i < deg + 1; ++i) {
        uint64_t dmini = 1;
        R min1 = 1;
        R pw = ipow<R>(oneOverA, i);
        for (uint64_t j = i; j < deg + 1; ++j, min1 *= -1, pw *= oneOverA) {
            B(i, j) = R(dcombi) * R(dmini) * min1 * pw;
            dmini *= (deg - j);
            dmini /= (j + 1 - i);
        }
        dcombi *= (deg - i);
        dcombi /= (i + 1);
    }
    Mat<R> D(n1, n1);
    uint64_t jpermk = fac(d);
    for (uint64_t j = d; j < deg + 1; ++j) {
        D(j, j - d) = R(jpermk);
        jpermk *= (j + 1);
        jpermk /= (j + 1 - d);
    }
    return matmul(B, D);
}

// splines/src/optimization/BezierQPOperations.cpp:73-111 — per-axis c x c block
// lambda * B_d * S * B_d^T with S_ij = T^(i+j+1)/(i+j+1); zero when d > degree.
template <typename R>
Mat<R> integrated_squared_derivative_block(int num_ctrl, R T, uint64_t d, R lambda) {
    Mat<R> cost(num_ctrl, num_ctrl);
    if (num_ctrl == 0 || d > static_cast<uint64_t>(num_ctrl - 1)) return cost;
    const Mat<R> B = bernstein_coefficient_matrix<R>(num_ctrl - 1, T, d);
    Mat<R> S(num_ctrl, num_ctrl);
    for (int i = 0; i < num_ctrl; ++i)
        for (int j = 0; j < num_ctrl; ++j) S(i, j) = ipow<R>(T, i + j + 1) / R(i + j + 1);
    Mat<R> lB = B;  // Eigen evaluates lambda*B*S*B^T left to right
    for (R& v : lB.a) v *= lambda;
    return matmul(matmul(lB, S), transpose(B));
}

// Eigen 3.4 LinSpaced(size, low, high) for floating scalars: low + i*step with the last
// element forced to `high` (used at mpc/src/optimization/PiecewiseBezierMPCQPOperations.cpp:33-34
// and mpc_cbf/src/controller/ConnectivityIMPCCBF.cpp:43).
template <typename R>
Vec<R> linspaced(int size, R low, R high) {
    Vec<R> v(size);
    if (size == 1) {
        v[0] = high;
        return v;
    }
    const R step = (high - low) / R(size - 1);
    for (int i = 0; i < size; ++i) v[i] = (i == size - 1) ? high : low + R(i) * step;
    return v;
}

// model/src/DoubleIntegratorXYYaw.cpp:9-21 — A = [I hI; 0 I], B = [h^2/2 I; h I], DIM = 3.
template <typename R>
struct DoubleIntegratorXYYaw {
    Mat<R> A{6, 6}, B{6, 3};
    explicit DoubleIntegratorXYYaw(R ts) {
        for (int i = 0; i < 6; ++i) A(i, i) = 1;
        for (int i = 0; i < 3; ++i) {
            A(i, i + 3) = ts;
            B(i, i) = R(0.5) * std::pow(ts, R(2.0));
            B(i + 3, i) = ts;
        }
    }
    // model/src/DoubleIntegrator.cpp:9-27 — A0.pos_ (3K x 6): block k = top rows of A^(k+1).
    void get_A0(int K, Mat<R>& pos, Mat<R>& vel) const {
        pos = Mat<R>(3 * K, 6);
        vel = Mat<R>(3 * K, 6);
        Mat<R> prev = Mat<R>::identity(6);
        for (int k = 0; k < K; ++k) {
            Mat<R> nr = matmul(A, prev);
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 6; ++j) {
                    pos(3 * k + i, j) = nr(i, j);
                    vel(3 * k + i, j) = nr(3 + i, j);
                }
            prev = nr;
        }
    }
    // model/src/DoubleIntegrator.cpp:30-51 — Lambda (3K x 3K): new = A*prev + [0..B at k..0].
    void get_lambda(int K, Mat<R>& pos, Mat<R>& vel) const {
        pos = Mat<R>(3 * K, 3 * K);
        vel = Mat<R>(3 * K, 3 * K);
        Mat<R> prev(6, 3 * K);
        for (int k = 0; k < K; ++k) {
            Mat<R> nr = matmul(A, prev);
            for (int i = 0; i < 6; ++i)
                for (int j = 0; j < 3; ++j) nr(i, 3 * k + j) += B(i, j);
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3 * K; ++j) {
                    pos(3 * k + i, j) = nr(i, j);
                    vel(3 * k + i, j) = nr(3 + i, j);
                }
            prev = nr;
        }
    }
    // model/src/DoubleIntegrator.cpp:54-63 — x+ = A x + B u.
    void apply_input(const R x[6], const R u[3], R out[6]) const {
        for (int i = 0; i < 6; ++i) {
            R s = 0;
            for (int j = 0; j < 6; ++j) s += A(i, j) * x[j];
            for (int j = 0; j < 3; ++j) s += B(i, j) * u[j];
            out[i] = s;
        }
    }
};

}  // namespace oracle
