// TEST INFRASTRUCTURE — NOT PRODUCT CODE (see oracle/README.md).
//
// lambda2 / Fiedler vector, the connectivity HOCBF row and the distance CLF row in closed form
// (cbf::ConnectivityCBF, cbf/src/detail/ConnectivityCBF.cpp:200-243,388-424,439-519; GiNaC expressions
// there). Used by the CBF-only filter (cbf_filter.hpp) and by the MPC-lifted rows of mpc_cbf.hpp.
// Pinned by the reference's own goldens (cbf/tests/TestInitConnectivity.cpp:103-153).
#pragma once
#include <algorithm>
#include <cmath>
#include <limits>
#include <numeric>
#include <vector>

#include "linalg.hpp"

namespace oracle {

// Cyclic Jacobi eigen-decomposition of a symmetric matrix (stands in for
// Eigen::SelfAdjointEigenSolver, ConnectivityCBF.cpp:417). Eigenvalues ascending in `w`, matching
// unit eigenvectors in the columns of V.
template <typename R>
void jacobi_eigh(Mat<R> A, Vec<R>& w, Mat<R>& V) {
    const int n = A.rows;
    V = Mat<R>::identity(n);
    for (int sweep = 0; sweep < 100; ++sweep) {
        R off = 0, diag = 0;
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j) (i == j ? diag : off) += A(i, j) * A(i, j);
        if (!(off > std::numeric_limits<R>::epsilon() * std::numeric_limits<R>::epsilon() * diag)) break;
        for (int p = 0; p < n - 1; ++p)
            for (int q = p + 1; q < n; ++q) {
                const R apq = A(p, q);
                if (apq == R(0)) continue;
                const R theta = (A(q, q) - A(p, p)) / (2 * apq);
                const R t = (theta >= 0 ? R(1) : R(-1)) / (std::fabs(theta) + std::sqrt(theta * theta + 1));
                const R c = 1 / std::sqrt(t * t + 1), s = t * c;
                for (int k = 0; k < n; ++k) {
                    if (k != p && k != q) {
                        const R akp = A(k, p), akq = A(k, q);
                        A(k, p) = A(p, k) = c * akp - s * akq;
                        A(k, q) = A(q, k) = s * akp + c * akq;
                    }
                    const R vkp = V(k, p), vkq = V(k, q);
                    V(k, p) = c * vkp - s * vkq;
                    V(k, q) = s * vkp + c * vkq;
                }
                A(p, p) -= t * apq;
                A(q, q) += t * apq;
                A(p, q) = A(q, p) = 0;
            }
    }
    std::vector<int> idx(n);
    std::iota(idx.begin(), idx.end(), 0);
    std::sort(idx.begin(), idx.end(), [&](int a, int b) { return A(a, a) < A(b, b); });
    w.resize(n);
    Mat<R> Vs(n, n);
    for (int j = 0; j < n; ++j) {
        w[j] = A(idx[j], idx[j]);
        for (int i = 0; i < n; ++i) Vs(i, j) = V(i, idx[j]);
    }
    V = Vs;
}

// ConnectivityCBF::getLambda2 (:388-424): weights exp((R^2 - d^2)^2 / sigma) - 1 inside range R = dmax,
// sigma = R^4 / ln 2 (:379-381); second smallest eigenpair of the Laplacian, eigenvector normalised.
template <typename R>
R lambda2(const R* states, int N, R dmax, Vec<R>& fiedler) {
    const R R2 = dmax * dmax, sigma = R2 * R2 / std::log(R(2));
    Mat<R> L(N, N);
    for (int i = 0; i < N; ++i)
        for (int j = 0; j < N; ++j) {
            if (i == j) continue;
            const R dx = states[6 * i] - states[6 * j], dy = states[6 * i + 1] - states[6 * j + 1];
            const R d2 = dx * dx + dy * dy;
            if (d2 <= R2) {
                const R wgt = std::exp((R2 - d2) * (R2 - d2) / sigma) - 1;
                L(i, j) = -wgt;
                L(i, i) += wgt;
            }
        }
    fiedler.assign(N, R(0));
    if (N < 2) return R(0);
    Vec<R> w;
    Mat<R> V;
    jacobi_eigh(L, w, V);
    R nn = 0;
    for (int i = 0; i < N; ++i) nn += V(i, 1) * V(i, 1);
    nn = std::sqrt(nn);
    for (int i = 0; i < N; ++i) fiedler[i] = V(i, 1) / nn;
    return w[1];
}

// Connectivity HOCBF row of agent `self` (initConnCBF :472-519 with compute_full_grad_h :439-469):
// h = lambda2 - eps, gradient and Hessian w.r.t. the agent's own position with the eigenvector held
// constant and NO range gating inside the symbolic form; alpha linear with gain gamma (:497).
// Returns Ac (3) and Bc with the meaning  Ac.u + Bc >= 0  (row inserted as -Ac.u <= Bc).
template <typename R>
void conn_cbf_row(const R* states, int N, int self, const Vec<R>& v, R lam2, R dmax, R gamma, R eps, R Ac[3], R& Bc) {
    const R R2 = dmax * dmax, sigma = R2 * R2 / std::log(R(2));
    R g[2] = {0, 0}, H[2][2] = {{0, 0}, {0, 0}};
    for (int j = 0; j < N; ++j) {
        if (j == self) continue;
        const R d[2] = {states[6 * self] - states[6 * j], states[6 * self + 1] - states[6 * j + 1]};
        const R r2 = d[0] * d[0] + d[1] * d[1], diff = R2 - r2;
        const R E = std::exp(diff * diff / sigma);
        const R wj = (v[self] - v[j]) * (v[self] - v[j]);
        const R c = -4 * E * diff / sigma;                        // dA_ij/dp = c * d   (:455-456)
        const R cp = 4 * E / sigma * (2 * diff * diff / sigma + 1);  // dc / d(r2)
        for (int a = 0; a < 2; ++a) {
            g[a] += c * wj * d[a];
            for (int b = 0; b < 2; ++b) H[a][b] += wj * ((a == b ? c : R(0)) + 2 * cp * d[a] * d[b]);
        }
    }
    const R vx = states[6 * self + 3], vy = states[6 * self + 4];
    const R lfh = g[0] * vx + g[1] * vy;                                        // :484
    const R lf2h = vx * (H[0][0] * vx + H[0][1] * vy) + vy * (H[1][0] * vx + H[1][1] * vy);  // :492-493
    const R h = lam2 - eps;
    Ac[0] = g[0];
    Ac[1] = g[1];
    Ac[2] = 0;
    Bc = lf2h + gamma * lfh + gamma * (lfh + gamma * h);                       // :520
}

// Distance CLF row (initCLFCBF :200-243): V = (|p - p_n| - 2)^2, row  Ac.u <= -Bc  with
// Ac = grad V, Bc = Lf^2 V + 5 Lf V + 2 V  (only the EGO velocity enters Lf V).
template <typename R>
void clf_row(const R ego[6], const R nbr[6], R Ac[3], R& Bc) {
    const R dd = 2, beta1 = 5, beta2 = 2;
    const R d[2] = {ego[0] - nbr[0], ego[1] - nbr[1]};
    const R dist = std::sqrt(d[0] * d[0] + d[1] * d[1]), e = dist - dd;
    const R gx = 2 * e * d[0] / dist, gy = 2 * e * d[1] / dist;
    const R vx = ego[3], vy = ego[4];
    R H[2][2];
    for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2; ++b)
            H[a][b] = 2 * (d[a] * d[b] / (dist * dist) + e * ((a == b ? R(1) : R(0)) / dist - d[a] * d[b] / (dist * dist * dist)));
    const R lfv = gx * vx + gy * vy;
    const R lf2v = vx * (H[0][0] * vx + H[0][1] * vy) + vy * (H[1][0] * vx + H[1][1] * vy);
    Ac[0] = gx;
    Ac[1] = gy;
    Ac[2] = 0;
    Bc = lf2v + beta1 * lfv + beta2 * e * e;
}

}  // namespace oracle
