// Host-side constant tables (see tables.hpp). Pure host C++; compiled into libmpccbf.so.
#include "tables.hpp"

#include <algorithm>
#include <cstring>

namespace mpccbf {

XMat xmul(const XMat& A, const XMat& B) {
    XMat C(A.rows, B.cols);
    for (int i = 0; i < A.rows; ++i)
        for (int k = 0; k < A.cols; ++k) {
            const xreal aik = A(i, k);
            if (aik == 0.0L) continue;
            for (int j = 0; j < B.cols; ++j) C(i, j) += aik * B(k, j);
        }
    return C;
}

XMat xtrans(const XMat& A) {
    XMat T(A.cols, A.rows);
    for (int i = 0; i < A.rows; ++i)
        for (int j = 0; j < A.cols; ++j) T(j, i) = A(i, j);
    return T;
}

static xreal binom(int n, int k) {
    if (k < 0 || k > n) return 0.0L;
    xreal r = 1.0L;
    for (int i = 1; i <= k; ++i) r = r * static_cast<xreal>(n - k + i) / static_cast<xreal>(i);
    return std::round(r);
}

// Bernstein polynomial B_{i,m}(s) on [0,1]; zero outside 0 <= i <= m.
static xreal bern(int i, int m, xreal s) {
    if (i < 0 || i > m || m < 0) return 0.0L;
    return binom(m, i) * std::pow(s, static_cast<xreal>(i)) * std::pow(1.0L - s, static_cast<xreal>(m - i));
}

// d/dt^d of B_{i,n}(t/T) = n!/(n-d)! / T^d * sum_j (-1)^(d-j) C(d,j) B_{i-j,n-d}(s).
// Same function the reference evaluates by monomial expansion in
// splines/src/detail/BezierOperations.cpp:11-50 (bernsteinBasis).
void bernstein_deriv(int nctrl, xreal T, xreal t, int d, xreal* out) {
    const int n = nctrl - 1;
    for (int i = 0; i <= n; ++i) out[i] = 0.0L;
    if (d > n) return;
    const xreal s = t / T;
    xreal fall = 1.0L;  // n!/(n-d)!
    for (int j = 0; j < d; ++j) fall *= static_cast<xreal>(n - j);
    const xreal scale = fall / std::pow(T, static_cast<xreal>(d));
    for (int i = 0; i <= n; ++i) {
        xreal acc = 0.0L;
        for (int j = 0; j <= d; ++j) {
            const xreal sign = ((d - j) % 2 == 0) ? 1.0L : -1.0L;
            acc += sign * binom(d, j) * bern(i - j, n - d, s);
        }
        out[i] = scale * acc;
    }
}

namespace {

// Householder QR of M (rows x cols, rows >= cols): M = Q R. Returns Q (rows x rows) explicitly
// and leaves R in the top of M. No pivoting: callers check the diagonal of R for rank.
void householder_qr(XMat& M, XMat& Q) {
    const int m = M.rows, n = M.cols;
    Q = XMat(m, m);
    for (int i = 0; i < m; ++i) Q(i, i) = 1.0L;
    std::vector<xreal> v(m);
    for (int k = 0; k < std::min(m - 1, n); ++k) {
        xreal norm2 = 0.0L;
        for (int i = k; i < m; ++i) norm2 += M(i, k) * M(i, k);
        if (norm2 == 0.0L) continue;
        xreal alpha = std::sqrt(norm2);
        if (M(k, k) > 0.0L) alpha = -alpha;
        for (int i = 0; i < m; ++i) v[i] = 0.0L;
        for (int i = k; i < m; ++i) v[i] = M(i, k);
        v[k] -= alpha;
        xreal vv = 0.0L;
        for (int i = k; i < m; ++i) vv += v[i] * v[i];
        if (vv == 0.0L) continue;
        for (int j = k; j < n; ++j) {
            xreal s = 0.0L;
            for (int i = k; i < m; ++i) s += v[i] * M(i, j);
            s = 2.0L * s / vv;
            for (int i = k; i < m; ++i) M(i, j) -= s * v[i];
        }
        for (int r = 0; r < m; ++r) {
            xreal s = 0.0L;
            for (int i = k; i < m; ++i) s += Q(r, i) * v[i];
            s = 2.0L * s / vv;
            for (int i = k; i < m; ++i) Q(r, i) -= s * v[i];
        }
    }
}

}  // namespace

bool HostTables::build(const mpccbf_config& cf) {
    cfg = cf;
    K = cf.k_hor;
    p = cf.num_pieces;
    c = cf.num_ctrl_pts;
    deg = cf.continuity_degree;
    m1 = p * c;
    n = 3 * m1;
    const xreal T = static_cast<xreal>(cf.piece_max_param);
    const xreal h = static_cast<xreal>(cf.h);

    // ---- sample times: Eigen LinSpaced(K, 0, (K-1)h) evaluated in double like the reference
    // (PiecewiseBezierMPCQPOperations.cpp:33-34), lower_bound piece lookup (:190-223).
    h_samples.assign(K, 0.0);
    {
        const double high = static_cast<double>(K - 1) * cf.h;
        const double step = (K > 1) ? (high - 0.0) / static_cast<double>(K - 1) : 0.0;
        for (int k = 0; k < K; ++k) h_samples[k] = (k == K - 1) ? high : 0.0 + static_cast<double>(k) * step;
    }
    std::vector<double> cum(p);
    for (int i = 0; i < p; ++i) cum[i] = (i == 0) ? cf.piece_max_param : cum[i - 1] + cf.piece_max_param;
    piece_of_k.assign(K, 0);
    local_of_k.assign(K, 0.0);
    for (int k = 0; k < K; ++k) {
        const double t = h_samples[k];
        if (t < 0 || t > cum.back()) {
            error = "MPC sampling range exceeds Bezier curve parameter range";
            return false;
        }
        const int pi = static_cast<int>(std::lower_bound(cum.begin(), cum.end(), t) - cum.begin());
        const double local = (pi == 0) ? t : t - cum[pi - 1];
        piece_of_k[k] = pi;
        local_of_k[k] = std::min(std::max(local, 0.0), cf.piece_max_param);
    }

    // ---- sampling bases for pos / vel / acc
    std::vector<xreal> row(c);
    for (int d = 0; d < 3; ++d) {
        B[d] = XMat(K, m1);
        for (int k = 0; k < K; ++k) {
            bernstein_deriv(c, T, static_cast<xreal>(local_of_k[k]), d, row.data());
            for (int j = 0; j < c; ++j) B[d](k, piece_of_k[k] * c + j) = row[j];
        }
    }

    // ---- equality rows of one axis, reference insertion order
    // (ConnectivityIMPCCBF.cpp:123-132): pos(0), vel(0), then per knot d = 0..deg.
    me1 = 2 + (deg + 1) * (p - 1);
    E1 = XMat(me1, m1);
    {
        int r = 0;
        for (int d = 0; d <= 1; ++d) {
            bernstein_deriv(c, T, 0.0L, d, row.data());
            for (int j = 0; j < c; ++j) E1(r, j) = row[j];
            ++r;
        }
        std::vector<xreal> first(c), second(c);
        for (int piece = 0; piece + 1 < p; ++piece)
            for (int d = 0; d <= deg; ++d) {
                bernstein_deriv(c, T, T, d, first.data());
                bernstein_deriv(c, T, 0.0L, d, second.data());
                for (int j = 0; j < c; ++j) {
                    E1(r, piece * c + j) = first[j];
                    E1(r, (piece + 1) * c + j) = -second[j];
                }
                ++r;
            }
    }

    // ---- double-integrator propagators for one axis (DoubleIntegrator.cpp:9-51 with
    // A = [1 h; 0 1], B = [h^2/2; h]): pos_{k+1} = pos + (k+1)h vel + sum_{j<=k} h^2 (k-j+1/2) u_j
    tau.assign(K, 0.0L);
    XMat Lam(K, K);
    for (int k = 0; k < K; ++k) {
        tau[k] = static_cast<xreal>(k + 1) * h;
        for (int j = 0; j <= k; ++j) Lam(k, j) = h * h * (static_cast<xreal>(k - j) + 0.5L);
    }
    Phi1 = xmul(Lam, B[2]);
    Qd.assign(K, 0.0L);
    for (int k = std::max(0, K - cf.spd_f); k < K; ++k) Qd[k] = static_cast<xreal>(cf.w_pos_err);

    // ---- quadratic cost of one axis: Phi'Q Phi + sum_{d=1..deg} w_u * int_0^T b_d b_d' dt per piece
    P1 = XMat(m1, m1);
    for (int k = 0; k < K; ++k) {
        if (Qd[k] == 0.0L) continue;
        for (int i = 0; i < m1; ++i)
            for (int j = 0; j < m1; ++j) P1(i, j) += Phi1(k, i) * Qd[k] * Phi1(k, j);
    }
    {
        // Gauss-Legendre, 8 nodes: exact for the degree <= 2(c-1) integrands up to c = 8
        static const xreal gx[4] = {0.183434642495649804939476142360L, 0.525532409916328985817739049189L,
                                    0.796666477413626739591553936476L, 0.960289856497536231683560868569L};
        static const xreal gw[4] = {0.362683783378361982965150449277L, 0.313706645877887287337962201987L,
                                    0.222381034453374470544355994426L, 0.101228536290376259152531354310L};
        for (int d = 1; d <= deg && d <= c - 1; ++d) {
            XMat G(c, c);
            for (int q = 0; q < 8; ++q) {
                const xreal xi = (q < 4) ? -gx[q] : gx[q - 4];
                const xreal wq = (q < 4) ? gw[q] : gw[q - 4];
                const xreal t = 0.5L * T * (xi + 1.0L);
                bernstein_deriv(c, T, t, d, row.data());
                for (int i = 0; i < c; ++i)
                    for (int j = 0; j < c; ++j) G(i, j) += 0.5L * T * wq * row[i] * row[j];
            }
            for (int piece = 0; piece < p; ++piece)
                for (int i = 0; i < c; ++i)
                    for (int j = 0; j < c; ++j)
                        P1(piece * c + i, piece * c + j) += static_cast<xreal>(cf.w_u_eff) * G(i, j);
        }
    }

    // ---- null-space parametrisation of the equality block: E1' = Q R
    reducible = false;
    r1 = m1 - me1;
    if (me1 <= m1 && r1 >= 0) {
        XMat Et = xtrans(E1);
        XMat Q;
        householder_qr(Et, Q);
        xreal dmax = 0.0L, dmin = 1e300L;
        for (int i = 0; i < me1; ++i) {
            dmax = std::max(dmax, std::fabs(Et(i, i)));
            dmin = std::min(dmin, std::fabs(Et(i, i)));
        }
        if (me1 == 0 || dmin > 1e-10L * dmax) {
            reducible = true;
            Z1 = XMat(m1, r1);
            for (int i = 0; i < m1; ++i)
                for (int j = 0; j < r1; ++j) Z1(i, j) = Q(i, me1 + j);
            // x_p = Q1 R^-T e with e = [pos; vel; 0...]: column c of Y1 solves R' y = e_c
            Y1 = XMat(m1, 2);
            for (int col = 0; col < 2; ++col) {
                std::vector<xreal> y(me1, 0.0L);
                for (int i = 0; i < me1; ++i) {
                    xreal s = (i == col) ? 1.0L : 0.0L;
                    for (int k = 0; k < i; ++k) s -= Et(k, i) * y[k];
                    y[i] = s / Et(i, i);
                }
                for (int i = 0; i < m1; ++i) {
                    xreal s = 0.0L;
                    for (int k = 0; k < me1; ++k) s += Q(i, k) * y[k];
                    Y1(i, col) = s;
                }
            }
        }
    }
    if (!reducible) {
        r1 = 0;
        return true;  // dense path only
    }

    // ---- reduced Hessian, its Cholesky factor and the whitened basis W1 = Z1 L^-T
    XMat H1 = xmul(xtrans(Z1), xmul(P1, Z1));
    for (int i = 0; i < r1; ++i)
        for (int j = 0; j < r1; ++j) H1(i, j) *= 2.0L;
    for (int i = 0; i < r1; ++i)
        for (int j = i + 1; j < r1; ++j) H1(i, j) = H1(j, i) = 0.5L * (H1(i, j) + H1(j, i));
    L1 = XMat(r1, r1);
    for (int j = 0; j < r1; ++j) {
        xreal dj = H1(j, j);
        for (int k = 0; k < j; ++k) dj -= L1(j, k) * L1(j, k);
        if (!(dj > 0.0L)) {
            reducible = false;  // reduced Hessian not PD: dense path only
            r1 = 0;
            return true;
        }
        L1(j, j) = std::sqrt(dj);
        for (int i = j + 1; i < r1; ++i) {
            xreal s = H1(i, j);
            for (int k = 0; k < j; ++k) s -= L1(i, k) * L1(j, k);
            L1(i, j) = s / L1(j, j);
        }
    }
    // W1 row i solves L w' = Z1 row i'  (W1 = Z1 L^-T  <=>  W1 L' = Z1)
    W1 = XMat(m1, r1);
    for (int i = 0; i < m1; ++i)
        for (int j = 0; j < r1; ++j) {
            xreal s = Z1(i, j);
            for (int k = 0; k < j; ++k) s -= W1(i, k) * L1(j, k);
            W1(i, j) = s / L1(j, j);
        }

    // ---- unconstrained minimiser v0 = -W1'(2 P1 Y1 s + q),  q = 2 Phi'Q (pos 1 + vel tau - ref)
    XMat PhiQ(m1, K);  // Phi' Q
    for (int i = 0; i < m1; ++i)
        for (int k = 0; k < K; ++k) PhiQ(i, k) = Phi1(k, i) * Qd[k];
    XMat onesTau(K, 2);
    for (int k = 0; k < K; ++k) {
        onesTau(k, 0) = 1.0L;
        onesTau(k, 1) = tau[k];
    }
    XMat W1t = xtrans(W1);
    XMat qs = xmul(PhiQ, onesTau);  // (m1 x 2): q = 2 qs s - 2 PhiQ ref
    XMat PY = xmul(P1, Y1);
    XMat lin(m1, 2);
    for (int i = 0; i < m1; ++i)
        for (int j = 0; j < 2; ++j) lin(i, j) = 2.0L * PY(i, j) + 2.0L * qs(i, j);
    Vs = xmul(W1t, lin);
    for (auto& x : Vs.a) x = -x;
    Vr = xmul(W1t, PhiQ);
    for (auto& x : Vr.a) x *= 2.0L;

    // ---- row templates in whitened coordinates: value = Ct.v + Ot.s
    for (int d = 0; d < 3; ++d) {
        Ct[d] = xmul(B[d], W1);
        Ot[d] = xmul(B[d], Y1);
    }

    // ---- constant part of the objective: c0 = s'(Y1'P1Y1) s + q'Y1 s
    XMat Y1t = xtrans(Y1);
    Kc = xmul(Y1t, PY);
    XMat qsY = xmul(xtrans(qs), Y1);  // (2 x 2): (2 qs s)' Y1 s = s' (2 qs' Y1) s
    for (int i = 0; i < 2; ++i)
        for (int j = 0; j < 2; ++j) Kc(i, j) += 2.0L * qsY(i, j);
    Kr = xmul(Y1t, PhiQ);  // (2 x K): -2 (PhiQ ref)' Y1 s = (Kr ref).s
    for (auto& x : Kr.a) x *= -2.0L;
    return true;
}

// ---------------------------------------------------------------------------------------------
std::vector<double> HostTables::full_basis(int d) const {
    std::vector<double> out(static_cast<size_t>(3) * K * n, 0.0);
    for (int k = 0; k < K; ++k)
        for (int ax = 0; ax < 3; ++ax)
            for (int j = 0; j < c; ++j)
                out[static_cast<size_t>(3 * k + ax) * n + var_index(piece_of_k[k], ax, j)] =
                    static_cast<double>(B[d](k, piece_of_k[k] * c + j));
    return out;
}
std::vector<double> HostTables::full_U() const { return full_basis(2); }

std::vector<double> HostTables::full_P() const {
    std::vector<double> out(static_cast<size_t>(n) * n, 0.0);
    for (int ax = 0; ax < 3; ++ax)
        for (int pi = 0; pi < p; ++pi)
            for (int i = 0; i < c; ++i)
                for (int pj = 0; pj < p; ++pj)
                    for (int j = 0; j < c; ++j)
                        out[static_cast<size_t>(var_index(pi, ax, i)) * n + var_index(pj, ax, j)] =
                            static_cast<double>(P1(pi * c + i, pj * c + j));
    return out;
}

std::vector<double> HostTables::full_E() const {
    std::vector<double> out(static_cast<size_t>(3) * me1 * n, 0.0);
    auto put = [&](int row, int ax, int r1row) {
        for (int pi = 0; pi < p; ++pi)
            for (int j = 0; j < c; ++j)
                out[static_cast<size_t>(row) * n + var_index(pi, ax, j)] = static_cast<double>(E1(r1row, pi * c + j));
    };
    int row = 0;
    for (int d = 0; d <= 1; ++d)
        for (int ax = 0; ax < 3; ++ax) put(row++, ax, d);
    for (int piece = 0; piece + 1 < p; ++piece)
        for (int d = 0; d <= deg; ++d)
            for (int ax = 0; ax < 3; ++ax) put(row++, ax, 2 + piece * (deg + 1) + d);
    return out;
}

std::vector<double> HostTables::full_Afix() const {
    std::vector<double> acc = full_basis(2), vel = full_basis(1);
    acc.insert(acc.end(), vel.begin(), vel.end());
    return acc;
}

std::vector<double> HostTables::full_Z() const {
    const int r = 3 * r1;
    std::vector<double> out(static_cast<size_t>(n) * r, 0.0);
    for (int ax = 0; ax < 3; ++ax)
        for (int pi = 0; pi < p; ++pi)
            for (int j = 0; j < c; ++j)
                for (int k = 0; k < r1; ++k)
                    out[static_cast<size_t>(var_index(pi, ax, j)) * r + ax * r1 + k] =
                        static_cast<double>(Z1(pi * c + j, k));
    return out;
}

std::vector<double> HostTables::full_Y() const {
    std::vector<double> out(static_cast<size_t>(n) * 6, 0.0);
    for (int ax = 0; ax < 3; ++ax)
        for (int pi = 0; pi < p; ++pi)
            for (int j = 0; j < c; ++j) {
                out[static_cast<size_t>(var_index(pi, ax, j)) * 6 + ax] = static_cast<double>(Y1(pi * c + j, 0));
                out[static_cast<size_t>(var_index(pi, ax, j)) * 6 + 3 + ax] = static_cast<double>(Y1(pi * c + j, 1));
            }
    return out;
}

std::vector<double> HostTables::full_Hred() const {
    const int r = 3 * r1;
    std::vector<double> out(static_cast<size_t>(r) * r, 0.0);
    for (int ax = 0; ax < 3; ++ax)
        for (int i = 0; i < r1; ++i)
            for (int j = 0; j < r1; ++j) {
                xreal s = 0.0L;
                for (int k = 0; k < r1; ++k) s += L1(i, k) * L1(j, k);
                out[static_cast<size_t>(ax * r1 + i) * r + ax * r1 + j] = static_cast<double>(s);
            }
    return out;
}

void HostTables::linear_cost(const double* x0, const double* ref, double* q) const {
    for (int i = 0; i < n; ++i) q[i] = 0.0;
    for (int ax = 0; ax < 3; ++ax) {
        std::vector<xreal> coef(K);
        for (int k = 0; k < K; ++k)
            coef[k] = 2.0L * Qd[k] * (static_cast<xreal>(x0[ax]) + tau[k] * static_cast<xreal>(x0[3 + ax]) -
                                      static_cast<xreal>(ref[3 * k + ax]));
        for (int pi = 0; pi < p; ++pi)
            for (int j = 0; j < c; ++j) {
                xreal s = 0.0L;
                for (int k = 0; k < K; ++k) s += Phi1(k, pi * c + j) * coef[k];
                q[var_index(pi, ax, j)] = static_cast<double>(s);
            }
    }
}

}  // namespace mpccbf
