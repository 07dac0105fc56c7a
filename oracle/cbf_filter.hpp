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

// ConnectivityControl::optimize (ConnectivityControl.cpp:22-99):
//   min |u - desired_u|^2 (+ sum_i w_i s_i in slack mode)  s.t.  safety rows (:50-55), velocity CBF rows (:58-59), and the
//   connectivity row when lambda2 > 0.1 else one CLF row per neighbour (:69-84).
// Slack mode (:30-45, CBFQPGeneratorBase.cpp:9-30,59-70): N slack variables s_i >= 0 with linear cost
// w_i = slack_cost * slack_decay^i (index order, no distance ranking here); safety row i carries -s_i
// (ConnectivityQPGenerator.cpp:74-98), the connectivity row the LAST slack -s_{N-1} (:31-41), CLF row i carries -s_i, the
// same slack as safety row i (ConnectivityControl.cpp:71-82: local_slack_idx counts the neighbours); velocity rows none.
template <typename R>
FilterResult<R> cbf_filter(const mpccbf_config& cfg, const R* states, int N, int self, const R desired_u[3], R feas_tol) {
    FilterResult<R> out;
    const R gamma = R(cfg.cbf_gamma > 0 ? cfg.cbf_gamma : 5.0);
    const bool slack = cfg.slack_mode != 0;
    const int nv = 3 + (slack ? N : 0);
    std::vector<std::vector<R>> rows;  // a (nv) | ub   meaning a.x <= ub
    auto add_row = [&](R a0, R a1, R a2, R ub, int slack_idx) {
        std::vector<R> r(nv + 1, R(0));
        r[0] = a0; r[1] = a1; r[2] = a2; r[nv] = ub;
        if (slack && slack_idx >= 0) r[3 + slack_idx] = R(-1);
        rows.push_back(r);
    };
    const R* ego = states + 6 * self;
    for (int i = 0; i < N - 1; ++i) {
        const R* nb = states + 6 * (i + (i >= self ? 1 : 0));
        R a[3], b;
        safety_cbf_row<R>(ego, nb, R(cfg.d_min), gamma, a, b);
        add_row(-a[0], -a[1], -a[2], b, i);
    }
    for (int d = 0; d < 3; ++d) {  // addMinVelConstraints: Ac = e_d, Bc = v_d - vmin_d
        R a[3] = {0, 0, 0};
        a[d] = -1;
        add_row(a[0], a[1], a[2], ego[3 + d] - R(cfg.v_min[d]), -1);
    }
    for (int d = 0; d < 3; ++d) {  // addMaxVelConstraints: Ac = -e_d, Bc = vmax_d - v_d
        R a[3] = {0, 0, 0};
        a[d] = 1;
        add_row(a[0], a[1], a[2], R(cfg.v_max[d]) - ego[3 + d], -1);
    }
    Vec<R> fv;
    out.lambda2 = lambda2<R>(states, N, R(cfg.d_max), fv);
    if (out.lambda2 > R(0.1)) {
        R a[3], b;
        conn_cbf_row<R>(states, N, self, fv, out.lambda2, R(cfg.d_max), gamma, R(0.1), a, b);
        add_row(-a[0], -a[1], -a[2], b, N - 1);
        out.used_conn = 1;
    } else {
        for (int i = 0; i < N - 1; ++i) {
            const R* nb = states + 6 * (i + (i >= self ? 1 : 0));
            R a[3], b;
            clf_row<R>(ego, nb, a, b);
            add_row(a[0], a[1], a[2], -b, i);
        }
    }
    DenseQP<R> qp;
    qp.n = nv;
    qp.P = Mat<R>(nv, nv);
    for (int j = 0; j < 3; ++j) qp.P(j, j) = R(1);
    qp.q.assign(nv, R(0));
    for (int j = 0; j < 3; ++j) qp.q[j] = -2 * desired_u[j];
    if (slack) {
        qp.lb.assign(nv, R(std::numeric_limits<double>::lowest()));
        qp.ub.assign(nv, R(std::numeric_limits<double>::max()));
        for (int i = 0; i < N; ++i) {
            qp.lb[3 + i] = 0;
            qp.q[3 + i] = R(cfg.slack_cost) * std::pow(R(cfg.slack_decay), R(i));
        }
    }
    const int m = static_cast<int>(rows.size());
    qp.A = Mat<R>(m, nv);
    qp.l.assign(m, -std::numeric_limits<R>::max());
    qp.u.resize(m);
    for (int i = 0; i < m; ++i) {
        for (int j = 0; j < nv; ++j) qp.A(i, j) = rows[i][j];
        qp.u[i] = rows[i][nv];
    }
    out.qp = solve_dense_qp<R>(qp, feas_tol);
    out.status = out.qp.status;
    if (out.status == OPTIMAL)
        for (int j = 0; j < 3; ++j) out.u[j] = out.qp.x[j];
    return out;
}

}  // namespace oracle
