// Host mirror of the reference's CBF-only controller, bound to the B200 filter kernel through the C
// ABI (mpccbf_cbf_filter_batch_host). Same names, argument meaning and return convention as
//   cbf::ConnectivityCBF(min_dist, max_dist, vmin, vmax)   cbf/include/cbf/detail/ConnectivityCBF.h
//   cbf::ConnectivityControl<T,DIM>::optimize(cbf_u, desired_u, current_states, self_idx, u_min, u_max)
//        cbf/include/cbf/controller/ConnectivityControl.h:16-60, src/controller/ConnectivityControl.cpp:22-99
// What differs, on purpose: ConnectivityCBF here is a parameter holder (the barrier expressions are
// closed forms inside the kernel, gamma = 5 and lambda2_min = 0.1 as hard-coded in the reference,
// ConnectivityCBF.cpp:62-63); slack_mode = true runs the same kernel with capped multipliers (the slack of neighbour i
// shared by its safety and CLF rows, the last slack on the connectivity row, ConnectivityControl.cpp:30-45);
// on a failed solve cbf_u is zero (what the reference's driver substitutes,
// CBFFormationControl_example.cpp:168-172). BatchedConnectivityControl filters every agent of many
// scenarios in one launch.
#pragma once
#include <memory>

#include "controller.hpp"

namespace cbf {

class ConnectivityCBF {
  public:
    ConnectivityCBF(double min_dist, double max_dist, const math::Vector<double>& vmin, const math::Vector<double>& vmax)
        : dmin(min_dist), dmax(max_dist), vmin_(vmin), vmax_(vmax) {
        if (vmin.size() != 3 || vmax.size() != 3) throw std::invalid_argument("ConnectivityCBF: vmin / vmax must have 3 entries");
    }
    double dmin, dmax;
    const math::Vector<double>& vmin() const { return vmin_; }
    const math::Vector<double>& vmax() const { return vmax_; }

  private:
    math::Vector<double> vmin_, vmax_;
};

namespace detail {
inline mpccbf_config filter_config(const ConnectivityCBF& c, int num_robots, int max_scenarios, int device) {
    mpccbf_config cfg;
    mpccbf_config_default(&cfg);
    cfg.d_min = c.dmin;
    cfg.d_max = c.dmax;
    for (int d = 0; d < 3; ++d) {
        cfg.v_min[d] = c.vmin()(d);
        cfg.v_max[d] = c.vmax()(d);
    }
    cfg.num_agents = num_robots;
    cfg.max_scenarios = max_scenarios;
    cfg.device = device;
    return cfg;
}
}  // namespace detail

template <typename T, unsigned int DIM>
class ConnectivityControl {
    static_assert(DIM == 3, "instantiated for <double, 3U> only (ConnectivityControl.cpp:103)");

  public:
    using State = model::State<T, DIM>;
    using VectorDIM = math::VectorDIM<T, DIM>;

    ConnectivityControl(std::shared_ptr<ConnectivityCBF> cbf, int num_robots, bool slack_mode = false, T slack_cost = T(0),
                        T slack_decay_rate = T(0), int device = 0)
        : cbf_(std::move(cbf)), num_robots_(num_robots) {
        if (!cbf_) throw std::invalid_argument("ConnectivityControl: cbf is null");
        cfg_ = detail::filter_config(*cbf_, num_robots, 1, device);
        cfg_.slack_mode = slack_mode ? 1 : 0;
        cfg_.slack_cost = slack_cost;
        cfg_.slack_decay = slack_decay_rate;
        mpc_cbf::detail::check(mpccbf_create(&cfg_, &h_), "ConnectivityControl");
    }
    ConnectivityControl(const ConnectivityControl&) = delete;
    ConnectivityControl& operator=(const ConnectivityControl&) = delete;
    ~ConnectivityControl() {
        if (h_) mpccbf_destroy(h_);
    }

    // true iff the QP is OPTIMAL; cbf_u receives the filtered input (zero otherwise). u_min / u_max are
    // accepted and ignored exactly like the reference (the control box is commented out, :60).
    bool optimize(VectorDIM& cbf_u, const VectorDIM& desired_u, std::vector<State> current_states, std::size_t self_idx,
                  const VectorDIM& u_min, const VectorDIM& u_max) {
        (void)u_min;
        (void)u_max;
        const std::size_t N = static_cast<std::size_t>(num_robots_);
        if (current_states.size() != N) throw std::invalid_argument("ConnectivityControl::optimize: current_states must hold num_robots states");
        if (self_idx >= N) throw std::out_of_range("ConnectivityControl::optimize: self_idx out of range");
        std::vector<double> st(N * 6), du(N * 3, 0.0), u(N * 3), l2(1);
        std::vector<int32_t> status(N);
        for (std::size_t i = 0; i < N; ++i)
            for (unsigned int d = 0; d < DIM; ++d) {
                st[i * 6 + d] = current_states[i].pos_(d);
                st[i * 6 + 3 + d] = current_states[i].vel_(d);
            }
        for (unsigned int d = 0; d < DIM; ++d) du[self_idx * 3 + d] = desired_u(d);
        mpc_cbf::detail::check(mpccbf_cbf_filter_batch_host(h_, 1, st.data(), du.data(), u.data(), status.data(), l2.data()),
                               "ConnectivityControl::optimize");
        for (unsigned int d = 0; d < DIM; ++d) cbf_u(d) = u[self_idx * 3 + d];
        last_status_ = status[self_idx];
        last_lambda2_ = l2[0];
        return last_status_ == MPCCBF_STATUS_OPTIMAL;
    }

    int32_t last_status() const { return last_status_; }
    double last_lambda2() const { return last_lambda2_; }

  private:
    std::shared_ptr<ConnectivityCBF> cbf_;
    int num_robots_;
    mpccbf_config cfg_;
    mpccbf_handle* h_ = nullptr;
    int32_t last_status_ = MPCCBF_STATUS_UNKNOWN;
    double last_lambda2_ = 0.0;
};

// Batched sibling: the filter for every agent of `num_scenarios` scenarios in one launch.
template <typename T, unsigned int DIM>
class BatchedConnectivityControl {
    static_assert(DIM == 3, "instantiated for <double, 3U> only");

  public:
    struct Result {
        std::vector<double> u;         // [S][N][3]
        std::vector<int32_t> status;   // [S][N]
        std::vector<double> lambda2;   // [S]
    };
    BatchedConnectivityControl(const ConnectivityCBF& cbf, int num_robots, int max_scenarios, int device = 0,
                               bool slack_mode = false, double slack_cost = 0.0, double slack_decay_rate = 0.0)
        : cfg_(detail::filter_config(cbf, num_robots, max_scenarios, device)) {
        cfg_.slack_mode = slack_mode ? 1 : 0;
        cfg_.slack_cost = slack_cost;
        cfg_.slack_decay = slack_decay_rate;
        mpc_cbf::detail::check(mpccbf_create(&cfg_, &h_), "BatchedConnectivityControl");
    }
    BatchedConnectivityControl(const BatchedConnectivityControl&) = delete;
    BatchedConnectivityControl& operator=(const BatchedConnectivityControl&) = delete;
    ~BatchedConnectivityControl() {
        if (h_) mpccbf_destroy(h_);
    }
    // states [S][N][6] (pos3, vel3), desired_u [S][N][3], row-major
    Result optimize(int num_scenarios, const std::vector<double>& states, const std::vector<double>& desired_u) {
        const std::size_t B = static_cast<std::size_t>(num_scenarios) * cfg_.num_agents;
        if (states.size() != B * 6 || desired_u.size() != B * 3)
            throw std::invalid_argument("BatchedConnectivityControl::optimize: states must be [S][N][6], desired_u [S][N][3]");
        Result r;
        r.u.resize(B * 3);
        r.status.resize(B);
        r.lambda2.resize(num_scenarios);
        mpc_cbf::detail::check(mpccbf_cbf_filter_batch_host(h_, num_scenarios, states.data(), desired_u.data(), r.u.data(),
                                                            r.status.data(), r.lambda2.data()),
                               "BatchedConnectivityControl::optimize");
        return r;
    }

  private:
    mpccbf_config cfg_;
    mpccbf_handle* h_ = nullptr;
};

}  // namespace cbf
