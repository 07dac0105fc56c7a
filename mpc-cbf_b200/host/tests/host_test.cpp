// Tests of the C++ host mirror (mpc-cbf_b200/host/include/mpccbf_host/*.hpp), written the way the
// reference's own gtest files read (qpcpp/tests/ProblemTest.cpp, qpcpp/tests/CPLEXTest.cpp) but
// without a gtest dependency.
//   host_test --cpu   container bookkeeping + argument/exception behaviour (no device needed)
//   host_test         + everything that computes on the B200 through the C ABI
#include <cmath>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <limits>
#include <vector>

#include "mpccbf_host/cbf_control.hpp"
#include "mpccbf_host/controller.hpp"
#include "mpccbf_host/qpcpp.hpp"

static int g_fail = 0;
#define EXPECT(cond)                                                              \
    do {                                                                          \
        if (!(cond)) {                                                            \
            std::printf("FAIL %s:%d  %s\n", __FILE__, __LINE__, #cond);           \
            ++g_fail;                                                             \
        }                                                                         \
    } while (0)
#define EXPECT_NEAR(a, b, tol) EXPECT(std::fabs((a) - (b)) <= (tol))
#define EXPECT_THROW(stmt, exc)                                                   \
    do {                                                                          \
        bool thrown = false;                                                      \
        try { stmt; } catch (const exc&) { thrown = true; } catch (...) {}        \
        if (!thrown) {                                                            \
            std::printf("FAIL %s:%d  expected %s from %s\n", __FILE__, __LINE__, #exc, #stmt); \
            ++g_fail;                                                             \
        }                                                                         \
    } while (0)

using Problem = qpcpp::Problem<double>;
constexpr unsigned int DIM = 3U;
using Controller = mpc_cbf::ConnectivityIMPCCBF<double, DIM>;
using Batched = mpc_cbf::BatchedConnectivityIMPCCBF<double, DIM>;
using State = model::State<double, DIM>;

// base_config.json (effective parameters of every shipped experiment after preprocess.py)
struct BaseConfig {
    mpc::PiecewiseBezierParams<double, DIM> bezier{3, 4, 0.5, 3};
    mpc::MPCParams<double> mpc;
    cbf::ConnectivityCBFParams<double> cbf{2.0, 4.0};
    Controller::IMPCParams impc{2, 2, 50000.0, 0.1, false};
    BaseConfig() {
        mpc.h_ = 0.1;
        mpc.Ts_ = 0.01;
        mpc.k_hor_ = 16;
        mpc.tuning_ = {10.0, 10.0, 8};
        mpc.limits_.v_min_ = {-2.0, -2.0, -2.6179938779914944};
        mpc.limits_.v_max_ = {2.0, 2.0, 2.6179938779914944};
        mpc.limits_.a_min_ = {-5.0, -5.0, -3.141592653589793};
        mpc.limits_.a_max_ = {5.0, 5.0, 3.141592653589793};
    }
};

static void test_problem_container() {
    // counts, coefficient get/set, quadratic-term symmetry, resetProblem (ProblemTest.cpp semantics)
    Problem p;
    auto* x = p.addVariable();
    auto* y = p.addVariable(-1.0, 2.0);
    EXPECT(p.numVariables() == 2 && p.hasVariable(x) && p.hasVariable(y));
    EXPECT(x->min() == std::numeric_limits<double>::lowest() && x->max() == std::numeric_limits<double>::max());
    EXPECT(y->min() == -1.0 && y->max() == 2.0);
    auto* c = p.addLinearConstraint(1.0, 5.0);
    c->setCoefficient(x, 2.0);
    c->setCoefficient(x, 3.0);  // overwrite, not accumulate
    EXPECT(c->getCoefficient(x) == 3.0 && c->getCoefficient(y) == 0.0 && c->min() == 1.0 && c->max() == 5.0);
    EXPECT(p.numLinearConstraints() == 1);
    auto* f = p.cost_function();
    f->addQuadraticTerm(x, y, 1.5);
    f->addQuadraticTerm(y, x, 0.5);  // (x,y) and (y,x) accumulate into one coefficient
    EXPECT(f->getQuadraticCoefficient(x, y) == 2.0 && f->getQuadraticCoefficient(y, x) == 2.0);
    f->addLinearTerm(y, 4.0);
    f->addLinearTerm(y, -1.0);
    EXPECT(f->getLinearCoefficient(y) == 3.0 && f->getLinearCoefficient(x) == 0.0);
    f->add_constant(7.0);
    EXPECT(f->constant() == 7.0);
    Problem other;
    auto* z = other.addVariable();
    EXPECT(!p.hasVariable(z));
    EXPECT_THROW(c->setCoefficient(z, 1.0), std::runtime_error);
    EXPECT_THROW(f->addQuadraticTerm(x, z, 1.0), std::runtime_error);
    EXPECT_THROW(f->addLinearTerm(z, 1.0), std::runtime_error);
    std::vector<double> P, q, A, l, u, lb, ub;
    p.flatten(P, q, A, l, u, lb, ub);
    EXPECT(P.size() == 4 && P[0 * 2 + 1] == 2.0 && P[1 * 2 + 0] == 0.0 && q[1] == 3.0 && A[0] == 3.0 && lb[1] == -1.0);
    p.resetProblem();
    EXPECT(p.numLinearConstraints() == 0 && p.numVariables() == 2 && f->getQuadraticCoefficient(x, y) == 0.0 &&
           f->constant() == 0.0);
    EXPECT(qpcpp::SolveStatusToStr(qpcpp::SolveStatus::INFEASIBLE) == "INFEASIBLE");
    EXPECT(static_cast<int>(qpcpp::SolveStatus::INFEASIBLEORUNBOUNDED) == MPCCBF_STATUS_INFEASIBLEORUNBOUNDED);
}

static void test_curve_eval() {
    // one cubic piece on [0, 0.5]: control points of p(t) = t^3 on axis 0 (T = 0.5 -> s = 2t)
    const double T = 0.5;
    double cp[2 * 3 * 4] = {0};
    cp[3] = T * T * T;                       // piece 0 axis 0: (0,0,0,T^3) is s^3 * T^3 = t^3
    for (int j = 0; j < 4; ++j) cp[12 + j] = T * T * T;  // piece 1 axis 0: constant
    splines::SingleParameterPiecewiseCurve<double, DIM> c(cp, 2, 4, T);
    EXPECT_NEAR(c.max_parameter(), 1.0, 0);
    EXPECT_NEAR(c.eval(0.3, 0)(0), 0.027, 1e-15);
    EXPECT_NEAR(c.eval(0.3, 1)(0), 3 * 0.09, 1e-14);
    EXPECT_NEAR(c.eval(0.3, 2)(0), 6 * 0.3, 1e-13);
    EXPECT_NEAR(c.eval(0.3, 3)(0), 6.0, 1e-12);
    EXPECT_NEAR(c.eval(0.5, 0)(0), 0.125, 1e-15);  // knot belongs to the earlier piece (lower_bound)
    EXPECT_NEAR(c.eval(0.5, 1)(0), 0.75, 1e-14);
    EXPECT_NEAR(c.eval(0.7, 1)(0), 0.0, 0);
    EXPECT_THROW(c.eval(1.01, 0), std::runtime_error);
    EXPECT_THROW(c.eval(-0.1, 0), std::runtime_error);
}

static void test_argument_errors_without_device() {
    BaseConfig b;
    Controller::MPCCBFParams mp{b.bezier, b.mpc, b.cbf};
    Controller::Params params{mp, b.impc};
    // host-only handle (device = -1): construction works, compute refuses (no CPU fallback)
    Controller ctl(params, /*num_neighbors=*/1, /*device=*/-1);
    std::vector<State> states(2);
    math::Vector<double> ref(3 * 16);
    std::vector<Controller::SingleParameterPiecewiseCurve> curves;
    EXPECT_THROW(ctl.optimize(curves, states, 0, ref), std::runtime_error);
    EXPECT_THROW(ctl.optimize(curves, states, 2, ref), std::out_of_range);
    math::Vector<double> short_ref(5);
    EXPECT_THROW(ctl.optimize(curves, states, 0, short_ref), std::invalid_argument);
    std::vector<State> three(3);
    EXPECT_THROW(ctl.optimize(curves, three, 0, ref), std::invalid_argument);
    // config validation follows common/parsing.hpp: (K-1)h must fit in num_pieces * piece_max
    b.mpc.k_hor_ = 20;
    EXPECT_THROW(Controller bad(params, 1, -1), std::invalid_argument);
}

static void test_toy_qp_on_device() {
    // qpcpp/tests/CPLEXTest.cpp:28-56 with B200Solver in place of CPLEXSolver
    Problem p;
    auto* x = p.addVariable();
    auto* y = p.addVariable();
    p.cost_function()->addQuadraticTerm(x, x, 1.0);
    p.cost_function()->addQuadraticTerm(y, y, 1.0);
    auto* c = p.addLinearConstraint(1.0, std::numeric_limits<double>::max());
    c->setCoefficient(x, 1.0);
    c->setCoefficient(y, 1.0);
    qpcpp::B200Solver<double> solver(0);
    qpcpp::Solver<double>& base = solver;  // through the abstract interface (Solver.h:27-37)
    EXPECT(base.solve(p) == qpcpp::SolveStatus::OPTIMAL);
    EXPECT_NEAR(x->solution_value(), 0.5, 1e-6);
    EXPECT_NEAR(y->solution_value(), 0.5, 1e-6);
    // infeasible: x >= 1 and x <= 0 -> INFEASIBLE, solution untouched
    Problem q;
    auto* a = q.addVariable();
    q.cost_function()->addQuadraticTerm(a, a, 1.0);
    q.addLinearConstraint(1.0, std::numeric_limits<double>::max())->setCoefficient(a, 1.0);
    q.addLinearConstraint(std::numeric_limits<double>::lowest(), 0.0)->setCoefficient(a, 1.0);
    a->set_solution_value(42.0);
    EXPECT(solver.solve(q) == qpcpp::SolveStatus::INFEASIBLE);
    EXPECT(a->solution_value() == 42.0);
    // variable bounds: min (x-3)^2 with x <= 1  -> x = 1
    Problem r;
    auto* v = r.addVariable(-10.0, 1.0);
    r.cost_function()->addQuadraticTerm(v, v, 1.0);
    r.cost_function()->addLinearTerm(v, -6.0);
    EXPECT(solver.solve(r) == qpcpp::SolveStatus::OPTIMAL);
    EXPECT_NEAR(v->solution_value(), 1.0, 1e-7);
}

static void test_two_robot_swap_on_device() {
    // BASELINE configs[0]: base_config.json + baseline/2r/circle.json, robot 0 at (2,2,0) at rest.
    // Survey-derived expectations: first acceleration -270/91 per axis,
    // objective -1378.9905, piece-0 x control points 2, 2, 1.8763736, 1.6749084.
    BaseConfig b;
    Controller::MPCCBFParams mp{b.bezier, b.mpc, b.cbf};
    Controller::Params params{mp, b.impc};
    Controller ctl(params, 1, 0);
    std::vector<State> states(2);
    states[0].pos_ = {2.0, 2.0, 0.0};
    states[1].pos_ = {-2.0, -2.0, 0.0};
    math::Vector<double> ref(3 * 16);
    for (int k = 0; k < 16; ++k) { ref(3 * k) = -2.0; ref(3 * k + 1) = -2.0; ref(3 * k + 2) = 0.0; }
    std::vector<Controller::SingleParameterPiecewiseCurve> curves;
    ctl.resetProblem();
    const bool ok = ctl.optimize(curves, states, 0, ref);
    EXPECT(ok && curves.size() == 2);
    if (curves.size() == 2) {
        const auto& c = curves.back();
        EXPECT_NEAR(c.eval(0.0, 0)(0), 2.0, 1e-9);
        EXPECT_NEAR(c.eval(0.0, 1)(1), 0.0, 1e-9);
        EXPECT_NEAR(c.eval(0.0, 2)(0), -270.0 / 91.0, 1e-6);
        EXPECT_NEAR(c.eval(0.0, 2)(1), -270.0 / 91.0, 1e-6);
        EXPECT_NEAR(c.eval(0.0, 2)(2), 0.0, 1e-9);
        EXPECT_NEAR(c.control_points()[2], 1.8763736, 1e-6);
        EXPECT_NEAR(c.control_points()[3], 1.6749084, 1e-6);
        EXPECT_NEAR(c.max_parameter(), 1.5, 0);
    }
    // infeasible start (|v0| > v_max): returns false with no curve, does not throw
    states[0].vel_ = {2.5, 0.0, 0.0};
    curves.clear();
    EXPECT(!ctl.optimize(curves, states, 0, ref) && curves.empty());
    EXPECT(ctl.last_status()[0] == MPCCBF_STATUS_INFEASIBLE && ctl.last_status()[1] == MPCCBF_STATUS_SKIPPED);
    states[0].vel_ = {0.0, 0.0, 0.0};

    // batched sibling agrees with the per-agent call (Jacobi snapshot == what optimize sees)
    Batched batch(params, 2, 4, 0);
    std::vector<std::vector<State>> S(3, states);
    std::vector<std::vector<math::VectorDIM<double, DIM>>> Tg(3);
    for (auto& t : Tg) t = {{-2.0, -2.0, 0.0}, {2.0, 2.0, 0.0}};
    S[1][0].pos_ = {1.0, 1.2, 0.1};
    auto r = batch.optimize(S, Tg);
    EXPECT(batch.success(r, 0, 0) && batch.success(r, 2, 1));
    curves.clear();
    ctl.optimize(curves, states, 0, ref);
    const auto cb = batch.curve(r, 0, 0);
    for (int j = 0; j < 36; ++j) EXPECT_NEAR(cb.control_points()[j], curves.back().control_points()[j], 1e-9);
    EXPECT_NEAR(r.objective[0], -1378.9905, 1e-3);
}

// cbf::ConnectivityControl mirror: the reference's TestInitConnectivity geometry (three robots on a
// line, 2 m apart, dmax 3) is connected (lambda2 = 0.2385 > 0.1), so the connectivity row is used.
static void test_cbf_filter_on_device() {
    using Control = cbf::ConnectivityControl<double, DIM>;
    auto cb = std::make_shared<cbf::ConnectivityCBF>(0.8, 3.0, math::Vector<double>{-3.0, -3.0, -3.0},
                                                     math::Vector<double>{3.0, 3.0, 3.0});
    Control ctl(cb, 3);
    std::vector<State> states(3);
    states[0].pos_ = {1.0, 2.0, 0.0};
    states[1].pos_ = {1.0, 4.0, 0.0};
    states[2].pos_ = {1.0, 6.0, 0.0};
    math::VectorDIM<double, DIM> u, lim{5.0, 5.0, 5.0}, nlim{-5.0, -5.0, -5.0};
    // nothing binds: the filter returns the desired input
    EXPECT(ctl.optimize(u, {0.1, 0.2, 0.05}, states, 1, nlim, lim));
    EXPECT_NEAR(ctl.last_lambda2(), 0.2385412985210327, 1e-12);
    EXPECT_NEAR(u(0), 0.1, 1e-12);
    EXPECT_NEAR(u(1), 0.2, 1e-12);
    EXPECT_NEAR(u(2), 0.05, 1e-12);
    // velocity CBF rows: v_x = 2.9, v_max = 3 -> u_x <= 0.1; yaw rate -3.2 < v_min -> u_yaw >= 0.2
    states[1].vel_ = {2.9, 0.0, -3.2};
    EXPECT(ctl.optimize(u, {2.0, 0.0, 0.0}, states, 1, nlim, lim));
    EXPECT_NEAR(u(0), 0.1, 1e-9);
    EXPECT_NEAR(u(2), 0.2, 1e-9);
    states[1].vel_ = {0.0, 0.0, 0.0};
    // the reference's golden row of agent 0 (TestInitConnectivity.cpp:103-127): g = (0, -2.703392),
    // b = 3.4635..., i.e.  -2.703392 u_y + 3.4635 >= 0  caps u_y at b / |g_y| = 1.2812
    EXPECT(ctl.optimize(u, {0.0, 4.0, 0.0}, states, 0, nlim, lim));
    EXPECT_NEAR(u(1), 3.4635324630258153 / 2.70339242, 1e-6);
    EXPECT_NEAR(u(0), 0.0, 1e-12);
    EXPECT_THROW(ctl.optimize(u, {0.0, 0.0, 0.0}, states, 3, nlim, lim), std::out_of_range);
    {   // slack mode (ConnectivityControl.cpp:30-45): the connectivity row carries the LAST slack, weight
        // slack_cost * decay^(N-1) = 0.01. min (u_y - 2.5)^2 + 0.01 s  s.t.  -2.70339 u_y + 3.46353 + s >= 0: the multiplier
        // of the row stops at its cap, u_y = 2.5 - (0.01 / 2) * 2.70339 (the hard rows would cap u_y at 1.2812)
        Control soft(cb, 3, true, 1.0, 0.1);
        math::VectorDIM<double, DIM> us;
        EXPECT(soft.optimize(us, {0.0, 2.5, 0.0}, states, 0, nlim, lim));
        EXPECT_NEAR(us(1), 2.5 - 0.005 * 2.70339242, 1e-6);
        EXPECT_NEAR(us(0), 0.0, 1e-12);
    }
    // batched sibling gives the same answer for the same snapshot
    cbf::BatchedConnectivityControl<double, DIM> batch(*cb, 3, 2);
    std::vector<double> st(2 * 3 * 6, 0.0), du(2 * 3 * 3, 0.0);
    for (int s = 0; s < 2; ++s)
        for (int i = 0; i < 3; ++i) {
            st[(s * 3 + i) * 6 + 0] = 1.0;
            st[(s * 3 + i) * 6 + 1] = 2.0 + 2.0 * i;
        }
    du[(1 * 3 + 0) * 3 + 1] = 4.0;
    auto r = batch.optimize(2, st, du);
    EXPECT(r.status[3] == MPCCBF_STATUS_OPTIMAL);
    EXPECT_NEAR(r.u[(1 * 3 + 0) * 3 + 1], u(1), 1e-12);
    EXPECT_NEAR(r.lambda2[0], r.lambda2[1], 1e-15);
}

// Multi-GPU behind the C ABI, no Python: agent-sharded closed loop on 2 devices == the same loop on 1 device,
// bit for bit; scenario-sharded step == one device. Skipped (with a note) on a box with a single GPU.
static void test_multi_gpu_on_device() {
    mpccbf_config cfg;
    mpccbf_config_default(&cfg);
    cfg.k_hor = 20;
    cfg.num_pieces = 4;
    cfg.d_min = 0.8;
    cfg.num_agents = 16;
    cfg.max_scenarios = 8;
    const int S = 8, N = 16, steps = 12;
    std::vector<double> st(S * N * 6, 0.0), tg(S * N * 3, 0.0);
    for (int s = 0; s < S; ++s)
        for (int a = 0; a < N; ++a) {   // ring of radius 2.4 + 0.2 s: neighbours ~0.94 .. 1.5 m apart, goals antipodal
            const double ang = 2.0 * 3.14159265358979323846 * a / N + 0.1 * s, R = 2.4 + 0.2 * s;
            double* x = &st[(s * N + a) * 6];
            x[0] = R * std::cos(ang); x[1] = R * std::sin(ang);
            x[3] = 0.3 * std::cos(3.0 * ang); x[4] = 0.3 * std::sin(2.0 * ang);
            tg[(s * N + a) * 3] = -x[0]; tg[(s * N + a) * 3 + 1] = -x[1];
        }
    mpccbf_multi *m1 = nullptr, *m2 = nullptr;
    EXPECT(mpccbf_multi_create(&cfg, 1, nullptr, &m1) == MPCCBF_OK);
    if (!m1) return;
    const int n = mpccbf_num_decision_vars(mpccbf_multi_handle(m1, 0));
    std::vector<double> s1 = st, c1(S * N * n), c2(S * N * n);
    std::vector<int32_t> q1(S * N * 2), q2(S * N * 2);
    double ms1 = 0, ms2 = 0;
    EXPECT(mpccbf_multi_step_batch_host(m1, S, st.data(), tg.data(), c1.data(), q1.data(), nullptr, nullptr) == MPCCBF_OK);
    EXPECT(mpccbf_multi_run_sharded(m1, S, s1.data(), tg.data(), steps, &ms1) == MPCCBF_OK);
    double moved = 0;
    for (size_t i = 0; i < s1.size(); ++i) moved = std::max(moved, std::fabs(s1[i] - st[i]));
    EXPECT(moved > 0.05 && ms1 > 0);
    if (mpccbf_multi_create(&cfg, 2, nullptr, &m2) != MPCCBF_OK) {
        std::printf("note: 2-GPU part skipped (%s)\n", mpccbf_last_error());
    } else {
        EXPECT(mpccbf_multi_num_devices(m2) == 2 && mpccbf_multi_comm_ranks(m2) == 2);
        std::vector<double> s2 = st;
        EXPECT(mpccbf_multi_step_batch_host(m2, S, st.data(), tg.data(), c2.data(), q2.data(), nullptr, nullptr) == MPCCBF_OK);
        EXPECT(c1 == c2 && q1 == q2);
        EXPECT(mpccbf_multi_run_sharded(m2, S, s2.data(), tg.data(), steps, &ms2) == MPCCBF_OK);
        EXPECT(std::memcmp(s1.data(), s2.data(), sizeof(double) * s1.size()) == 0);
        std::printf("multi-GPU: agent-sharded closed loop %.3f ms/step on 1 device, %.3f ms/step on 2 (bit-identical states)\n", ms1, ms2);
        mpccbf_multi_destroy(m2);
    }
    mpccbf_multi_destroy(m1);
}

int main(int argc, char** argv) {
    const bool cpu_only = argc > 1 && std::strcmp(argv[1], "--cpu") == 0;
    test_problem_container();
    test_curve_eval();
    test_argument_errors_without_device();
    if (!cpu_only) {
        test_toy_qp_on_device();
        test_two_robot_swap_on_device();
        test_cbf_filter_on_device();
        test_multi_gpu_on_device();
    }
    std::printf("%s: %d failure(s)%s\n", argv[0], g_fail, cpu_only ? " (cpu-only subset)" : "");
    return g_fail == 0 ? 0 : 1;
}
