// Host mirror of the reference's controller API for the hot path, bound to the B200 kernels through
// the C ABI (include/mpccbf.h). Same names, argument meaning and error behaviour as
//   mpc::PiecewiseBezierParams / PhysicalLimits / TuningParams / MPCParams
//        mpc/include/mpc/optimization/PiecewiseBezierMPCQPOperations.h:17-41
//   cbf::ConnectivityCBFParams                       cbf/include/cbf/detail/ConnectivityCBF.h:13-16
//   splines::SingleParameterPiecewiseCurve::eval / max_parameter
//        splines/include/splines/curves/SingleParameterPiecewiseCurve.h:31-43
//   mpc_cbf::ConnectivityIMPCCBF<T,DIM>::{IMPCParams, Params, optimize, resetProblem}
//        mpc_cbf/include/mpc_cbf/controller/ConnectivityIMPCCBF.h:39-62
// What differs, on purpose: the constructor takes no model / CBF / collision-shape objects (the
// double integrator and the closed-form safety HOCBF are baked into the device tables; gamma is
// the reference's hard-coded 5.0, ConnectivityCBF.cpp:62), and there is a batched sibling,
// BatchedConnectivityIMPCCBF, that steps every agent of many scenarios in one launch.
#pragma once
#include <cmath>
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../../../include/mpccbf.h"
#include "types.hpp"

namespace mpc {

template <typename T, unsigned int DIM>
struct PiecewiseBezierParams {
    std::size_t num_pieces_, num_control_points_;
    T piece_max_parameter_;
    std::uint64_t bezier_continuity_upto_degree_;
};
template <typename T>
struct PhysicalLimits {
    math::Vector<T> p_min_, p_max_, v_min_, v_max_, a_min_, a_max_;
};
template <typename T>
struct TuningParams {
    T w_pos_err_, w_u_eff_;
    int spd_f_;
};
template <typename T>
struct MPCParams {
    T h_, Ts_;
    int k_hor_;
    TuningParams<T> tuning_;
    PhysicalLimits<T> limits_;
};

}  // namespace mpc

namespace cbf {
template <typename T>
struct ConnectivityCBFParams {
    T dmin_, dmax_;
};
}  // namespace cbf

namespace splines {

// Piecewise Bezier curve as returned by optimize(): pieces of `num_control_points` control points
// per axis, evaluated with the reference's rules: the piece of parameter t is the first one whose
// cumulative end is >= t (lower_bound; SingleParameterPiecewiseCurve.cpp:94-127), parameters
// outside [0, max_parameter()] throw.
template <typename T, unsigned int DIM>
class SingleParameterPiecewiseCurve {
  public:
    using VectorDIM = math::VectorDIM<T, DIM>;
    SingleParameterPiecewiseCurve() = default;
    // ctrl_pts: x[piece * DIM*c + axis * c + cp]  (BezierQPOperations.cpp:37)
    SingleParameterPiecewiseCurve(const T* ctrl_pts, std::size_t pieces, std::size_t c, T piece_max)
        : x_(ctrl_pts, ctrl_pts + pieces * DIM * c), pieces_(pieces), c_(c), T_(piece_max) {}
    std::size_t numPieces() const { return pieces_; }
    T max_parameter() const { return T_ * static_cast<T>(pieces_); }
    const std::vector<T>& control_points() const { return x_; }

    VectorDIM eval(T t, std::uint64_t derivative_degree) const {
        if (pieces_ == 0) throw std::runtime_error("SingleParameterPiecewiseCurve::eval: curve has no pieces");
        if (t < T(0) || t > max_parameter())
            throw std::runtime_error("SingleParameterPiecewiseCurve::eval: parameter is out of range");
        std::size_t piece = 0;
        T cum = T(0), prev = T(0);
        for (std::size_t i = 0; i < pieces_; ++i) {
            prev = cum;
            cum = (i == 0) ? T_ : cum + T_;
            piece = i;
            if (!(cum < t)) break;
        }
        T local = (piece == 0) ? t : t - prev;
        if (local > T_) local = T_;
        if (local < T(0)) local = T(0);
        // d-th derivative of a degree-(c-1) Bezier: repeated differences, then de Casteljau
        VectorDIM out;
        const std::size_t deg = c_ - 1;
        if (derivative_degree > deg) return out;
        std::vector<T> w(c_);
        for (unsigned int ax = 0; ax < DIM; ++ax) {
            const T* p = x_.data() + piece * DIM * c_ + ax * c_;
            for (std::size_t j = 0; j < c_; ++j) w[j] = p[j];
            std::size_t m = c_;
            T scale = T(1);
            for (std::uint64_t d = 0; d < derivative_degree; ++d) {
                for (std::size_t j = 0; j + 1 < m; ++j) w[j] = w[j + 1] - w[j];
                --m;
                scale *= static_cast<T>(deg - d) / T_;
            }
            const T s = local / T_;
            for (std::size_t r = 1; r < m; ++r)
                for (std::size_t j = 0; j + r < m; ++j) w[j] = (T(1) - s) * w[j] + s * w[j + 1];
            out[ax] = scale * w[0];
        }
        return out;
    }

  private:
    std::vector<T> x_;
    std::size_t pieces_ = 0, c_ = 0;
    T T_ = T(0);
};

}  // namespace splines

namespace mpc_cbf {

namespace detail {
inline void check(int rc, const char* who) {
    if (rc == MPCCBF_OK) return;
    const std::string msg = std::string(who) + ": " + mpccbf_last_error();
    if (rc == MPCCBF_ERR_INVALID) throw std::invalid_argument(msg);
    throw std::runtime_error(msg);
}
}  // namespace detail

template <typename T, unsigned int DIM>
struct MPCCBFParams {  // ConnectivityMPCCBFQPOperations<T,DIM>::Params (references, like the reference)
    mpc::PiecewiseBezierParams<T, DIM>& piecewise_bezier_params;
    mpc::MPCParams<T>& mpc_params;
    cbf::ConnectivityCBFParams<T>& connectivity_cbf_params;
};

template <typename T, unsigned int DIM>
class ConnectivityIMPCCBF {
    static_assert(DIM == 3, "the hot path is instantiated for <double, 3U> only (ConnectivityIMPCCBF.cpp:222)");

  public:
    using State = model::State<T, DIM>;
    using SingleParameterPiecewiseCurve = splines::SingleParameterPiecewiseCurve<T, DIM>;
    using VectorDIM = math::VectorDIM<T, DIM>;
    using Vector = math::Vector<T>;
    using MPCCBFParams = mpc_cbf::MPCCBFParams<T, DIM>;

    struct IMPCParams {
        int cbf_horizon_;
        int impc_iter_;
        T slack_cost_;
        T slack_decay_rate_;
        bool slack_mode_;
    };
    struct Params {
        MPCCBFParams& mpc_cbf_params;
        IMPCParams& impc_params;
    };

    static mpccbf_config to_config(const Params& p, int num_agents, int max_scenarios, int device) {
        const auto& bz = p.mpc_cbf_params.piecewise_bezier_params;
        const auto& mp = p.mpc_cbf_params.mpc_params;
        const auto& cb = p.mpc_cbf_params.connectivity_cbf_params;
        mpccbf_config c;
        mpccbf_config_default(&c);
        c.h = mp.h_;
        c.Ts = mp.Ts_;
        c.k_hor = mp.k_hor_;
        c.num_pieces = static_cast<int32_t>(bz.num_pieces_);
        c.num_ctrl_pts = static_cast<int32_t>(bz.num_control_points_);
        c.continuity_degree = static_cast<int32_t>(bz.bezier_continuity_upto_degree_);
        c.piece_max_param = bz.piece_max_parameter_;
        c.w_pos_err = mp.tuning_.w_pos_err_;
        c.w_u_eff = mp.tuning_.w_u_eff_;
        c.spd_f = mp.tuning_.spd_f_;
        if (mp.limits_.v_min_.size() != DIM || mp.limits_.v_max_.size() != DIM || mp.limits_.a_min_.size() != DIM ||
            mp.limits_.a_max_.size() != DIM)
            throw std::invalid_argument("ConnectivityIMPCCBF: physical limits must have DIM entries");
        for (unsigned int d = 0; d < DIM; ++d) {
            c.v_min[d] = mp.limits_.v_min_(d);
            c.v_max[d] = mp.limits_.v_max_(d);
            c.a_min[d] = mp.limits_.a_min_(d);
            c.a_max[d] = mp.limits_.a_max_(d);
        }
        c.d_min = cb.dmin_;
        c.d_max = cb.dmax_;
        c.cbf_horizon = p.impc_params.cbf_horizon_;
        c.impc_iter = p.impc_params.impc_iter_;
        c.slack_mode = p.impc_params.slack_mode_ ? 1 : 0;
        c.slack_cost = p.impc_params.slack_cost_;
        c.slack_decay = p.impc_params.slack_decay_rate_;
        c.num_agents = num_agents;
        c.max_scenarios = max_scenarios;
        c.device = device;
        return c;
    }

    // num_neighbors as in the reference constructor (robots - 1). device: CUDA ordinal.
    ConnectivityIMPCCBF(Params& p, int num_neighbors = 0, int device = 0) : cfg_(to_config(p, num_neighbors + 1, 1, device)) {
        detail::check(mpccbf_create(&cfg_, &h_), "ConnectivityIMPCCBF");
        n_ = mpccbf_num_decision_vars(h_);
    }
    ConnectivityIMPCCBF(const ConnectivityIMPCCBF&) = delete;
    ConnectivityIMPCCBF& operator=(const ConnectivityIMPCCBF&) = delete;
    ~ConnectivityIMPCCBF() {
        if (h_) mpccbf_destroy(h_);
    }

    // Returns true iff every IMPC iteration is OPTIMAL; result_curve receives the curves of the
    // successful iterations in order (ConnectivityIMPCCBF.cpp:203-211). Shape mismatches throw
    // std::invalid_argument; a solver failure never throws.
    bool optimize(std::vector<SingleParameterPiecewiseCurve>& result_curve, const std::vector<State>& current_states,
                  std::size_t self_idx, const Vector& ref_positions) {
        const std::size_t N = static_cast<std::size_t>(cfg_.num_agents);
        if (current_states.size() != N)
            throw std::invalid_argument("ConnectivityIMPCCBF::optimize: current_states must hold num_neighbors + 1 states");
        if (self_idx >= N) throw std::out_of_range("ConnectivityIMPCCBF::optimize: self_idx out of range");
        if (ref_positions.size() != static_cast<std::size_t>(DIM * cfg_.k_hor))
            throw std::invalid_argument("ConnectivityIMPCCBF::optimize: ref_positions must have DIM * k_hor entries");
        std::vector<double> st(N * 6), ref(ref_positions.size()), cp(static_cast<std::size_t>(cfg_.impc_iter) * n_);
        for (std::size_t i = 0; i < N; ++i)
            for (unsigned int d = 0; d < DIM; ++d) {
                st[i * 6 + d] = current_states[i].pos_(d);
                st[i * 6 + 3 + d] = current_states[i].vel_(d);
            }
        for (std::size_t i = 0; i < ref.size(); ++i) ref[i] = ref_positions(i);
        std::vector<int32_t> status(cfg_.impc_iter);
        int32_t n_curves = 0;
        detail::check(mpccbf_optimize_one(h_, st.data(), static_cast<int32_t>(self_idx), ref.data(), cp.data(), &n_curves,
                                          status.data()),
                      "ConnectivityIMPCCBF::optimize");
        last_status_ = status;
        for (int i = 0; i < n_curves; ++i) {
            std::vector<T> x(cp.begin() + static_cast<std::size_t>(i) * n_, cp.begin() + static_cast<std::size_t>(i + 1) * n_);
            result_curve.emplace_back(x.data(), static_cast<std::size_t>(cfg_.num_pieces),
                                      static_cast<std::size_t>(cfg_.num_ctrl_pts), static_cast<T>(cfg_.piece_max_param));
        }
        return n_curves == cfg_.impc_iter;
    }

    void resetProblem() {}  // the device path keeps no per-call problem state (ConnectivityIMPCCBF.h:62)

    const std::vector<int32_t>& last_status() const { return last_status_; }  // SolveStatus codes per IMPC iteration
    const mpccbf_config& config() const { return cfg_; }
    mpccbf_handle* handle() { return h_; }

  private:
    mpccbf_config cfg_;
    mpccbf_handle* h_ = nullptr;
    int n_ = 0;
    std::vector<int32_t> last_status_;
};

// Batched sibling: one control step for every agent of `num_scenarios` scenarios (Jacobi: all
// agents see the same snapshot). Host buffers in, host buffers out; device work is one fused launch.
template <typename T, unsigned int DIM>
class BatchedConnectivityIMPCCBF {
    static_assert(DIM == 3, "instantiated for <double, 3U> only");

  public:
    using Single = ConnectivityIMPCCBF<T, DIM>;
    using State = model::State<T, DIM>;
    using SingleParameterPiecewiseCurve = splines::SingleParameterPiecewiseCurve<T, DIM>;

    BatchedConnectivityIMPCCBF(typename Single::Params& p, int num_agents, int max_scenarios, int device = 0)
        : cfg_(Single::to_config(p, num_agents, max_scenarios, device)) {
        detail::check(mpccbf_create(&cfg_, &h_), "BatchedConnectivityIMPCCBF");
        n_ = mpccbf_num_decision_vars(h_);
    }
    BatchedConnectivityIMPCCBF(const BatchedConnectivityIMPCCBF&) = delete;
    BatchedConnectivityIMPCCBF& operator=(const BatchedConnectivityIMPCCBF&) = delete;
    ~BatchedConnectivityIMPCCBF() {
        if (h_) mpccbf_destroy(h_);
    }

    struct Result {
        std::vector<double> ctrl_pts;   // [S][N][n]   last successful iterate (zeros if none)
        std::vector<int32_t> status;    // [S][N][impc_iter]
        std::vector<int32_t> n_curves;  // [S][N]
        std::vector<double> objective;  // [S][N]
    };

    // states[s][i], targets[s][i]: scenario s, agent i. success[s][i] == optimize()'s return value.
    Result optimize(const std::vector<std::vector<State>>& states,
                    const std::vector<std::vector<math::VectorDIM<T, DIM>>>& targets) {
        const std::size_t S = states.size(), N = static_cast<std::size_t>(cfg_.num_agents);
        if (targets.size() != S) throw std::invalid_argument("BatchedConnectivityIMPCCBF::optimize: targets/states mismatch");
        std::vector<double> st(S * N * 6), tg(S * N * 3);
        for (std::size_t s = 0; s < S; ++s) {
            if (states[s].size() != N || targets[s].size() != N)
                throw std::invalid_argument("BatchedConnectivityIMPCCBF::optimize: every scenario must hold num_agents entries");
            for (std::size_t i = 0; i < N; ++i)
                for (unsigned int d = 0; d < DIM; ++d) {
                    st[(s * N + i) * 6 + d] = states[s][i].pos_(d);
                    st[(s * N + i) * 6 + 3 + d] = states[s][i].vel_(d);
                    tg[(s * N + i) * 3 + d] = targets[s][i](d);
                }
        }
        Result r;
        r.ctrl_pts.resize(S * N * n_);
        r.status.resize(S * N * cfg_.impc_iter);
        r.n_curves.resize(S * N);
        r.objective.resize(S * N);
        detail::check(mpccbf_step_batch_host(h_, static_cast<int32_t>(S), st.data(), tg.data(), r.ctrl_pts.data(),
                                             r.status.data(), r.n_curves.data(), r.objective.data()),
                      "BatchedConnectivityIMPCCBF::optimize");
        return r;
    }
    SingleParameterPiecewiseCurve curve(const Result& r, std::size_t scenario, std::size_t agent) const {
        const double* x = r.ctrl_pts.data() + (scenario * cfg_.num_agents + agent) * n_;
        return SingleParameterPiecewiseCurve(x, static_cast<std::size_t>(cfg_.num_pieces),
                                             static_cast<std::size_t>(cfg_.num_ctrl_pts), static_cast<T>(cfg_.piece_max_param));
    }
    bool success(const Result& r, std::size_t scenario, std::size_t agent) const {
        return r.n_curves[scenario * cfg_.num_agents + agent] == cfg_.impc_iter;
    }
    int num_decision_vars() const { return n_; }
    const mpccbf_config& config() const { return cfg_; }
    mpccbf_handle* handle() { return h_; }

  private:
    mpccbf_config cfg_;
    mpccbf_handle* h_ = nullptr;
    int n_ = 0;
};

}  // namespace mpc_cbf
