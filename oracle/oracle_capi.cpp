// TEST INFRASTRUCTURE — NOT PRODUCT CODE (see oracle/README.md).
// C entry points of the CPU oracle, loaded with ctypes by tests/, smoke() and bench.py's
// cpu_baseline / --impl reference legs. `prec` selects the scalar: 0 = double (timed CPU
// baseline, mirrors the reference's <double,3U>), 1 = long double (oracle accuracy).
#include <atomic>
#include <cstring>
#include <thread>

#include "cbf_filter.hpp"
#include "mpc_cbf.hpp"

using namespace oracle;

static thread_local std::string g_err;
#define ORACLE_CATCH(rc) catch (const std::exception& e) { g_err = e.what(); return rc; } catch (...) { g_err = "unknown"; return rc; }

namespace {

template <typename R>
void to_double(const Vec<R>& v, double* out) {
    for (size_t i = 0; i < v.size(); ++i) out[i] = static_cast<double>(v[i]);
}

template <typename R>
int optimize_impl(const mpccbf_config* cfg, const Controller<R>& ctl, const double* states, int N,
                  int self, const double* ref, double feas_tol, double* ctrl_pts, int32_t* status,
                  int32_t* n_curves, double* objective, double* kkt, int32_t* m_rows) {
    std::vector<R> st(6 * N);
    for (int i = 0; i < 6 * N; ++i) st[i] = R(states[i]);
    Vec<R> rf(3 * ctl.K);
    for (int i = 0; i < 3 * ctl.K; ++i) rf[i] = R(ref[i]);
    auto res = ctl.optimize(st.data(), N, self, rf, R(feas_tol));
    for (int it = 0; it < cfg->impc_iter; ++it) status[it] = res.status[it];
    if (n_curves) *n_curves = static_cast<int32_t>(res.curves.size());
    for (size_t it = 0; it < res.curves.size(); ++it) {
        if (ctrl_pts) to_double(res.curves[it], ctrl_pts + it * ctl.n);
    }
    for (size_t it = 0; it < res.qps.size(); ++it) {
        if (objective) objective[it] = static_cast<double>(res.qps[it].objective);
        if (kkt) {
            kkt[4 * it + 0] = static_cast<double>(res.qps[it].res_stationarity);
            kkt[4 * it + 1] = static_cast<double>(res.qps[it].res_primal);
            kkt[4 * it + 2] = static_cast<double>(res.qps[it].res_dual);
            kkt[4 * it + 3] = static_cast<double>(res.qps[it].res_comp);
        }
        if (m_rows) m_rows[it] = res.m_rows[it];
    }
    return 0;
}

template <typename R>
void step_batch_impl(const mpccbf_config* cfg, int S, const double* states, const double* targets,
                     double* ctrl_pts, int32_t* status, int32_t* n_curves, double* objective,
                     int threads, double feas_tol) {
    const int N = cfg->num_agents;
    Controller<R> ctl(*cfg);
    const int n = ctl.n, K = ctl.K, it_n = cfg->impc_iter;
    std::atomic<int> next{0};
    auto work = [&]() {
        std::vector<double> cp(static_cast<size_t>(it_n) * n), obj(it_n), ref(3 * K);
        std::vector<int32_t> st(it_n);
        for (;;) {
            const int id = next.fetch_add(1);
            if (id >= S * N) break;
            const int s = id / N, a = id % N;
            for (int k = 0; k < K; ++k)
                for (int d = 0; d < 3; ++d) ref[3 * k + d] = targets[(static_cast<size_t>(s) * N + a) * 3 + d];
            int32_t nc = 0;
            optimize_impl<R>(cfg, ctl, states + static_cast<size_t>(s) * N * 6, N, a, ref.data(), feas_tol,
                             cp.data(), st.data(), &nc, obj.data(), nullptr, nullptr);
            double* out = ctrl_pts + (static_cast<size_t>(s) * N + a) * n;
            if (nc > 0) std::memcpy(out, cp.data() + static_cast<size_t>(nc - 1) * n, sizeof(double) * n);
            else std::memset(out, 0, sizeof(double) * n);
            for (int it = 0; it < it_n; ++it) status[(static_cast<size_t>(s) * N + a) * it_n + it] = st[it];
            if (n_curves) n_curves[static_cast<size_t>(s) * N + a] = nc;
            if (objective) objective[static_cast<size_t>(s) * N + a] = nc > 0 ? obj[nc - 1] : 0.0;
        }
    };
    if (threads <= 1) {
        work();
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t) pool.emplace_back(work);
        for (auto& t : pool) t.join();
    }
}

template <typename R>
int64_t table_impl(const mpccbf_config* cfg, const char* name, double* out, int64_t cap) {
    Controller<R> ctl(*cfg);
    const Mat<R>* M = nullptr;
    Mat<R> tmp;
    std::string nm(name);
    if (nm == "U_basis") M = &ctl.U;
    else if (nm == "A0_pos") M = &ctl.A0pos;
    else if (nm == "A0_vel") M = &ctl.A0vel;
    else if (nm == "Lambda_pos") M = &ctl.Lpos;
    else if (nm == "Lambda_vel") M = &ctl.Lvel;
    else if (nm == "h_samples") {
        tmp = Mat<R>(1, ctl.K);
        for (int k = 0; k < ctl.K; ++k) tmp(0, k) = ctl.h_samples[k];
        M = &tmp;
    } else if (nm == "V_basis") {
        tmp = ctl.eval_sampling_basis(1);
        M = &tmp;
    } else return -1;
    const int64_t cnt = static_cast<int64_t>(M->rows) * M->cols;
    if (!out) return cnt;
    if (cap < cnt) return -2;
    for (int64_t i = 0; i < cnt; ++i) out[i] = static_cast<double>(M->a[i]);
    return cnt;
}

}  // namespace

extern "C" {

uint64_t oracle_fac(uint64_t n) { try { return fac(n); } catch (...) { return 0; } }
uint64_t oracle_comb(uint64_t n, uint64_t k) { return comb(n, k); }
uint64_t oracle_perm(uint64_t n, uint64_t k) { return perm(n, k); }

int oracle_bernstein_basis(int deg, double T, double t, int d, double* out, int prec) {
    try {
        if (prec) to_double(bernstein_basis<long double>(deg, T, t, d), out);
        else to_double(bernstein_basis<double>(deg, T, t, d), out);
    } ORACLE_CATCH(-1)
    return 0;
}

int oracle_bernstein_coefficient_matrix(int deg, double T, int d, double* out) {
    try {
        auto M = bernstein_coefficient_matrix<double>(deg, T, d);
        std::copy(M.a.begin(), M.a.end(), out);
    } ORACLE_CATCH(-1)
    return 0;
}

void oracle_double_integrator_apply(double ts, const double* x, const double* u, double* out) {
    DoubleIntegratorXYYaw<double> m(ts);
    m.apply_input(x, u, out);
}

void oracle_safety_cbf_row(const double* ego, const double* nbr, double dmin, double gamma,
                           double* a, double* b, int prec) {
    if (prec) {
        long double e[6], nb[6], aa[3], bb;
        for (int i = 0; i < 6; ++i) { e[i] = ego[i]; nb[i] = nbr[i]; }
        safety_cbf_row<long double>(e, nb, dmin, gamma, aa, bb);
        for (int i = 0; i < 3; ++i) a[i] = static_cast<double>(aa[i]);
        *b = static_cast<double>(bb);
    } else {
        safety_cbf_row<double>(ego, nbr, dmin, gamma, a, *b);
    }
}

int64_t oracle_get_table(const mpccbf_config* cfg, const char* name, double* out, int64_t cap, int prec) {
    try {
        return prec ? table_impl<long double>(cfg, name, out, cap) : table_impl<double>(cfg, name, out, cap);
    } ORACLE_CATCH(-3)
}

// Builds the QP of IMPC iteration `iter` for agent `self` exactly as the reference generator
// does. pred_states [cbf_horizon][6] is only read when iter > 0. Call with P == NULL to get
// the shapes (n_out, m_out) first. slack_weights [N-1] or NULL.
int oracle_build_qp(const mpccbf_config* cfg, const double* states, int N, int self,
                    const double* ref, int iter, const double* pred_states,
                    const double* slack_weights, int32_t* n_out, int32_t* m_out, int32_t* n_eq,
                    int32_t* n_cbf, double* P, double* q, double* A, double* l, double* u,
                    double* lb, double* ub) {
    try {
        Controller<double> ctl(*cfg);
        Vec<double> rf(ref, ref + 3 * ctl.K);
        std::vector<std::array<double, 6>> pred;
        if (iter > 0)
            for (int k = 0; k < cfg->cbf_horizon; ++k) {
                std::array<double, 6> ps;
                std::copy(pred_states + 6 * k, pred_states + 6 * k + 6, ps.begin());
                pred.push_back(ps);
            }
        Vec<double> sw(N - 1, 0.0);
        if (slack_weights) sw.assign(slack_weights, slack_weights + (N - 1));
        auto b = ctl.build_qp(states, N, self, rf, iter, pred, sw);
        *n_out = b.qp.n;
        *m_out = b.qp.m();
        if (n_eq) *n_eq = b.n_eq;
        if (n_cbf) *n_cbf = b.n_cbf;
        if (!P) return 0;
        std::copy(b.qp.P.a.begin(), b.qp.P.a.end(), P);
        std::copy(b.qp.q.begin(), b.qp.q.end(), q);
        std::copy(b.qp.A.a.begin(), b.qp.A.a.end(), A);
        std::copy(b.qp.l.begin(), b.qp.l.end(), l);
        std::copy(b.qp.u.begin(), b.qp.u.end(), u);
        if (lb) std::copy(b.qp.lb.begin(), b.qp.lb.end(), lb);
        if (ub) std::copy(b.qp.ub.begin(), b.qp.ub.end(), ub);
    } ORACLE_CATCH(-1)
    return 0;
}

// Dense QP, reference objective convention x'Px + q'x. kkt[4] = stationarity (reduced
// gradient), primal, dual, complementarity residuals. y [m] row multipliers (may be NULL).
int oracle_solve_qp(int n, int m, const double* P, const double* q, const double* A,
                    const double* l, const double* u, const double* lb, const double* ub,
                    double feas_tol, int prec, double* x, double* objective, int32_t* status,
                    int32_t* iters, double* kkt, double* y) {
    try {
        auto run = [&](auto tag) {
            using R = decltype(tag);
            DenseQP<R> qp;
            qp.n = n;
            qp.P = Mat<R>(n, n);
            for (int i = 0; i < n * n; ++i) qp.P.a[i] = R(P[i]);
            qp.q.resize(n);
            for (int i = 0; i < n; ++i) qp.q[i] = R(q[i]);
            qp.A = Mat<R>(m, n);
            for (int i = 0; i < m * n; ++i) qp.A.a[i] = R(A[i]);
            qp.l.resize(m);
            qp.u.resize(m);
            for (int i = 0; i < m; ++i) { qp.l[i] = R(l[i]); qp.u[i] = R(u[i]); }
            if (lb && ub) {
                qp.lb.resize(n);
                qp.ub.resize(n);
                for (int i = 0; i < n; ++i) { qp.lb[i] = R(lb[i]); qp.ub[i] = R(ub[i]); }
            }
            QPResult<R> r = solve_dense_qp<R>(qp, R(feas_tol));
            *status = r.status;
            if (iters) *iters = r.iterations;
            to_double(r.x, x);
            if (objective) *objective = static_cast<double>(r.objective);
            if (kkt) {
                kkt[0] = static_cast<double>(r.res_stationarity);
                kkt[1] = static_cast<double>(r.res_primal);
                kkt[2] = static_cast<double>(r.res_dual);
                kkt[3] = static_cast<double>(r.res_comp);
            }
            if (y) to_double(r.y, y);
        };
        if (prec) run((long double)0);
        else run((double)0);
    } ORACLE_CATCH(-1)
    return 0;
}

// ConnectivityIMPCCBF::optimize for one agent. ctrl_pts [impc_iter][n] (successful iterates in
// order), status [impc_iter], objective [impc_iter], kkt [impc_iter][4], m_rows [impc_iter].
int oracle_optimize(const mpccbf_config* cfg, const double* states, int N, int self,
                    const double* ref, double feas_tol, int prec, double* ctrl_pts,
                    int32_t* status, int32_t* n_curves, double* objective, double* kkt,
                    int32_t* m_rows) {
    try {
        if (prec) {
            Controller<long double> ctl(*cfg);
            return optimize_impl<long double>(cfg, ctl, states, N, self, ref, feas_tol, ctrl_pts, status,
                                              n_curves, objective, kkt, m_rows);
        }
        Controller<double> ctl(*cfg);
        return optimize_impl<double>(cfg, ctl, states, N, self, ref, feas_tol, ctrl_pts, status, n_curves,
                                     objective, kkt, m_rows);
    } ORACLE_CATCH(-1)
}

// Jacobi control step over S scenarios x N agents (same I/O contract as mpccbf_step_batch_host),
// fanned over `threads` host threads. This is the timed CPU baseline.
int oracle_step_batch(const mpccbf_config* cfg, int S, const double* states, const double* targets,
                      double* ctrl_pts, int32_t* status, int32_t* n_curves, double* objective,
                      int threads, double feas_tol, int prec) {
    try {
        if (prec) step_batch_impl<long double>(cfg, S, states, targets, ctrl_pts, status, n_curves, objective, threads, feas_tol);
        else step_batch_impl<double>(cfg, S, states, targets, ctrl_pts, status, n_curves, objective, threads, feas_tol);
    } ORACLE_CATCH(-1)
    return 0;
}

// Curve evaluation (rollout): out[3] = d-th derivative of the piecewise curve at t.
int oracle_eval_curve(const mpccbf_config* cfg, const double* ctrl_pts, double t, int d, double* out) {
    try {
        Controller<double> ctl(*cfg);
        Vec<double> x(ctrl_pts, ctrl_pts + ctl.n);
        ctl.eval_curve(x, t, d, out);
    } ORACLE_CATCH(-1)
    return 0;
}

const char* oracle_last_error() { return g_err.c_str(); }

// ---- CBF-only filter (cbf::ConnectivityControl) -------------------------------
// lambda2 + normalised Fiedler vector of one scenario's positions (states [N][6]).
double oracle_lambda2(const double* states, int N, double dmax, int prec, double* fiedler) {
    auto run = [&](auto tag) -> double {
        using R = decltype(tag);
        std::vector<R> st(6 * N);
        for (int i = 0; i < 6 * N; ++i) st[i] = R(states[i]);
        Vec<R> fv;
        const R l2 = lambda2<R>(st.data(), N, R(dmax), fv);
        if (fiedler) to_double(fv, fiedler);
        return static_cast<double>(l2);
    };
    return prec ? run((long double)0) : run((double)0);
}

// Connectivity HOCBF row of agent `self`: Ac [3], *Bc  (Ac.u + Bc >= 0), *lam2.
void oracle_conn_cbf_row(const double* states, int N, int self, double dmax, double gamma, int prec,
                         double* Ac, double* Bc, double* lam2) {
    auto run = [&](auto tag) {
        using R = decltype(tag);
        std::vector<R> st(6 * N);
        for (int i = 0; i < 6 * N; ++i) st[i] = R(states[i]);
        Vec<R> fv;
        const R l2 = lambda2<R>(st.data(), N, R(dmax), fv);
        R a[3], b;
        conn_cbf_row<R>(st.data(), N, self, fv, l2, R(dmax), R(gamma), R(0.1), a, b);
        for (int i = 0; i < 3; ++i) Ac[i] = static_cast<double>(a[i]);
        *Bc = static_cast<double>(b);
        if (lam2) *lam2 = static_cast<double>(l2);
    };
    if (prec) run((long double)0);
    else run((double)0);
}

// Distance CLF row: Ac [3], *Bc  (Ac.u <= -Bc).
void oracle_clf_row(const double* ego, const double* nbr, double* Ac, double* Bc) {
    long double e[6], n[6], a[3], b;
    for (int i = 0; i < 6; ++i) { e[i] = ego[i]; n[i] = nbr[i]; }
    clf_row<long double>(e, n, a, b);
    for (int i = 0; i < 3; ++i) Ac[i] = static_cast<double>(a[i]);
    *Bc = static_cast<double>(b);
}

// ConnectivityControl::optimize for every agent of every scenario (Jacobi: one snapshot).
// states [S][N][6], desired_u [S][N][3] -> u [S][N][3] (zeros unless OPTIMAL), status [S][N],
// lambda2 [S] (may be NULL).
int oracle_cbf_filter_batch(const mpccbf_config* cfg, int S, const double* states, const double* desired_u,
                            double feas_tol, int prec, double* u, int32_t* status, double* lam2) {
    try {
        const int N = cfg->num_agents;
        auto run = [&](auto tag) {
            using R = decltype(tag);
            std::vector<R> st(6 * N);
            for (int s = 0; s < S; ++s) {
                for (int i = 0; i < 6 * N; ++i) st[i] = R(states[static_cast<size_t>(s) * 6 * N + i]);
                for (int a = 0; a < N; ++a) {
                    const size_t g = static_cast<size_t>(s) * N + a;
                    const R du[3] = {R(desired_u[3 * g]), R(desired_u[3 * g + 1]), R(desired_u[3 * g + 2])};
                    auto r = cbf_filter<R>(*cfg, st.data(), N, a, du, R(feas_tol));
                    status[g] = r.status;
                    for (int j = 0; j < 3; ++j) u[3 * g + j] = static_cast<double>(r.u[j]);
                    if (lam2 && a == 0) lam2[s] = static_cast<double>(r.lambda2);
                }
            }
        };
        if (prec) run((long double)0);
        else run((double)0);
    } ORACLE_CATCH(-1)
    return 0;
}

int oracle_hardware_threads() { return static_cast<int>(std::thread::hardware_concurrency()); }

}  // extern "C"
