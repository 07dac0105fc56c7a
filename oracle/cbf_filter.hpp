// TEST INFRASTRUCTURE — NOT PRODUCT CODE (see oracle/README.md).
//
// CPU restatement of the CBF-only safety / connectivity filter:
//   cbf::ConnectivityControl::optimize        cbf/src/controller/ConnectivityControl.cpp:22-99
//   cbf::ConnectivityQPGenerator              cbf/src/optimization/ConnectivityQPGenerator.cpp:13-150
//   cbf::ConnectivityCBF::getLambda2          cbf/src/detail/ConnectivityCBF.cpp:388-424
//   ... initConnCBF / compute_full_grad_h     :439-519   (GiNaC expressions, here in closed form)
//   ... initCLFCBF                            :200-243
//   ... initVelCBF                            :251-284
// Paths relative to /root/reference/workspace/lib. Pinned by the reference's own goldens
// (cbf/tests/TestInitConnectivity.cpp:103-153) in tests/test_oracle_goldens.py.
#pragma once
#include <cmath>
#include <limits>

#include "mpc_cbf.hpp"

namespace oracle {

template <typename R>
struct FilterResult {
    int status = UNKNOWN;
    R u[3] = {0, 0, 0};
    R lambda2 = 0;
    int used_conn = 0;  // 1: connectivity row, 0: CLF rows
    QPResult<R> qp;
};

// ConnectivityControl::optimize (ConnectivityControl.cpp:22-99), slack_mode = false:
//   min |u - desired_u|^2  s.t.  safety rows (:50-55), velocity CBF rows (:58-59), and the
//   connectivity row when lambda2 > 0.1 else one CLF row per neighbour (:69-84).
template <typename R>
FilterResult<R> cbf_filter(const mpccbf_config& cfg, const R* states, int N, int self, const R desired_u[3], R feas_tol) {
    FilterResult<R> out;
    const R gamma = R(cfg.cbf_gamma > 0 ? cfg.cbf_gamma : 5.0);
    std::vector<std::array<R, 4>> rows;  // a0 a1 a2 | ub   meaning a.u <= ub
    const R* ego = states + 6 * self;
    for (int i = 0; i < N - 1; ++i) {
        const R* nb = states + 6 * (i + (i >= self ? 1 : 0));
        R a[3], b;
        safety_cbf_row<R>(ego, nb, R(cfg.d_min), gamma, a, b);
        rows.push_back({-a[0], -a[1], -a[2], b});
    }
    for (int d = 0; d < 3; ++d) {  // addMinVelConstraints: Ac = e_d, Bc = v_d - vmin_d
        std::array<R, 4> r = {0, 0, 0, ego[3 + d] - R(cfg.v_min[d])};
        r[d] = -1;
        rows.push_back(r);
    }
    for (int d = 0; d < 3; ++d) {  // addMaxVelConstraints: Ac = -e_d, Bc = vmax_d - v_d
        std::array<R, 4> r = {0, 0, 0, R(cfg.v_max[d]) - ego[3 + d]};
        r[d] = 1;
        rows.push_back(r);
    }
    Vec<R> fv;
    out.lambda2 = lambda2<R>(states, N, R(cfg.d_max), fv);
    if (out.lambda2 > R(0.1)) {
        R a[3], b;
        conn_cbf_row<R>(states, N, self, fv, out.lambda2, R(cfg.d_max), gamma, R(0.1), a, b);
        rows.push_back({-a[0], -a[1], -a[2], b});
        out.used_conn = 1;
    } else {
        for (int i = 0; i < N - 1; ++i) {
            const R* nb = states + 6 * (i + (i >= self ? 1 : 0));
            R a[3], b;
            clf_row<R>(ego, nb, a, b);
            rows.push_back({a[0], a[1], a[2], -b});
        }
    }
    DenseQP<R> qp;
    qp.n = 3;
    qp.P = Mat<R>::identity(3);
    qp.q = {-2 * desired_u[0], -2 * desired_u[1], -2 * desired_u[2]};
    const int m = static_cast<int>(rows.size());
    qp.A = Mat<R>(m, 3);
    qp.l.assign(m, -std::numeric_limits<R>::max());
    qp.u.resize(m);
    for (int i = 0; i < m; ++i) {
        for (int j = 0; j < 3; ++j) qp.A(i, j) = rows[i][j];
        qp.u[i] = rows[i][3];
    }
    out.qp = solve_dense_qp<R>(qp, feas_tol);
    out.status = out.qp.status;
    if (out.status == OPTIMAL)
        for (int j = 0; j < 3; ++j) out.u[j] = out.qp.x[j];
    return out;
}

}  // namespace oracle
