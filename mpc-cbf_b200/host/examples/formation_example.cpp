// Closed-loop formation-control episode on the B200 path — the driver that plays the role of the
// reference's lib/mpc_cbf/examples/connectivity/MPCCBFFormationControl_example.cpp:18-233: same
// flags (--config_file, --write_filename, --sim_runtime), same config JSON (base_config.json
// schema), same per-step logic (optimize per robot; follow the newest successful curve for h/Ts
// rollout sub-steps; keep following the previous curve when a solve fails; stay put with zero
// velocity when a robot never had a curve) and the same states.json output schema
// ("dt", "Ts", robots.<id>.states [t][6], robots.<id>.pred_curve [step][0][..][3]) that
// workspace/experiments/python/metrics/collision_check.py and plot_results.py consume.
// Extra flags: --mode gauss_seidel (reference order: robot i sees the already-advanced states of
// robots < i; one mpccbf_optimize_one per robot) | jacobi (one batched launch per control step);
// --noise_seed S (S = 0: noise off; otherwise a seeded N(0,pos_std)/N(0,vel_std) replaces the
// reference's std::random_device, lib/math/src/Random.cpp:7-28).
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <memory>
#include <random>
#include <sstream>
#include <string>

#include "mpccbf_host/controller.hpp"

constexpr unsigned int DIM = 3U;
using Curve = splines::SingleParameterPiecewiseCurve<double, DIM>;
using State = model::State<double, DIM>;

static void die(const std::string& msg) {
    std::fprintf(stderr, "formation_example: %s\n", msg.c_str());
    std::exit(2);
}

int main(int argc, char** argv) {
    std::string config_file, write_filename = "states.json", mode = "gauss_seidel";
    double sim_runtime = 20.0, pos_std = 0.0, vel_std = 0.0;
    unsigned long long noise_seed = 0;
    int device = 0;
    for (int i = 1; i < argc; ++i) {
        auto next = [&](const char* flag) -> std::string {
            if (i + 1 >= argc) die(std::string("missing value for ") + flag);
            return argv[++i];
        };
        if (!std::strcmp(argv[i], "--config_file")) config_file = next("--config_file");
        else if (!std::strcmp(argv[i], "--write_filename")) write_filename = next("--write_filename");
        else if (!std::strcmp(argv[i], "--sim_runtime")) sim_runtime = std::stod(next("--sim_runtime"));
        else if (!std::strcmp(argv[i], "--mode")) mode = next("--mode");
        else if (!std::strcmp(argv[i], "--noise_seed")) noise_seed = std::stoull(next("--noise_seed"));
        else if (!std::strcmp(argv[i], "--pos_std")) pos_std = std::stod(next("--pos_std"));
        else if (!std::strcmp(argv[i], "--vel_std")) vel_std = std::stod(next("--vel_std"));
        else if (!std::strcmp(argv[i], "--device")) device = std::stoi(next("--device"));
        else die(std::string("unknown flag ") + argv[i]);
    }
    if (config_file.empty()) die("--config_file is required");
    const bool jacobi = mode == "jacobi";
    if (!jacobi && mode != "gauss_seidel") die("--mode must be gauss_seidel or jacobi");

    constexpr int kCap = 1024;
    std::vector<double> so(3 * kCap), sf(3 * kCap);
    mpccbf_config cfg;
    mpccbf_config_default(&cfg);
    if (mpccbf_config_from_json(config_file.c_str(), &cfg, so.data(), sf.data(), kCap) != MPCCBF_OK) die(mpccbf_last_error());
    cfg.device = device;
    cfg.max_scenarios = 1;
    const int N = cfg.num_agents, K = cfg.k_hor, it_n = cfg.impc_iter;
    mpccbf_handle* h = nullptr;
    if (mpccbf_create(&cfg, &h) != MPCCBF_OK) die(mpccbf_last_error());
    const int n = mpccbf_num_decision_vars(h);

    std::vector<double> states(6 * N, 0.0), targets(3 * N);
    for (int i = 0; i < N; ++i)
        for (int d = 0; d < 3; ++d) {
            states[6 * i + d] = so[3 * i + d];
            targets[3 * i + d] = sf[3 * i + d];
        }
    std::vector<std::unique_ptr<Curve>> traj(N);
    std::vector<double> eval_t(N, 0.0);
    std::vector<std::ostringstream> js_states(N), js_pred(N);
    for (auto& s : js_states) s << std::setprecision(17);
    for (auto& s : js_pred) s << std::setprecision(17);
    std::vector<bool> first_state(N, true), first_pred(N, true);
    std::mt19937_64 rng(noise_seed);
    std::normal_distribution<double> gauss(0.0, 1.0);
    const double pred_horizon = cfg.num_pieces * cfg.piece_max_param;
    const int substeps = static_cast<int>(cfg.h / cfg.Ts);
    long solves = 0, failures = 0;
    double solve_seconds = 0.0;

    std::vector<double> cp(static_cast<size_t>(it_n) * n), ref(3 * K), cp_all(static_cast<size_t>(N) * n), obj(N);
    std::vector<int32_t> status(it_n), status_all(static_cast<size_t>(N) * it_n), ncv_all(N);
    int loop_idx = 0;
    for (double sim_t = 0.0; sim_t < sim_runtime; sim_t += cfg.h, ++loop_idx) {
        if (jacobi) {
            const auto t0 = std::chrono::steady_clock::now();
            if (mpccbf_step_batch_host(h, 1, states.data(), targets.data(), cp_all.data(), status_all.data(), ncv_all.data(),
                                       obj.data()) != MPCCBF_OK)
                die(mpccbf_last_error());
            solve_seconds += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        }
        for (int r = 0; r < N; ++r) {
            int32_t n_curves = 0;
            const double* newest = nullptr;
            if (jacobi) {
                n_curves = ncv_all[r];
                newest = cp_all.data() + static_cast<size_t>(r) * n;
            } else {
                for (int k = 0; k < K; ++k)
                    for (int d = 0; d < 3; ++d) ref[3 * k + d] = targets[3 * r + d];
                const auto t0 = std::chrono::steady_clock::now();
                if (mpccbf_optimize_one(h, states.data(), r, ref.data(), cp.data(), &n_curves, status.data()) != MPCCBF_OK)
                    die(mpccbf_last_error());
                solve_seconds += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
                if (n_curves > 0) newest = cp.data() + static_cast<size_t>(n_curves - 1) * n;
            }
            solves += it_n;
            if (n_curves != it_n) ++failures;
            if (n_curves > 0) {  // trajs.back()
                traj[r] = std::make_unique<Curve>(newest, static_cast<size_t>(cfg.num_pieces),
                                                  static_cast<size_t>(cfg.num_ctrl_pts), cfg.piece_max_param);
                eval_t[r] = 0.0;
            }
            // prediction log: curve sampled every 0.05 over the horizon
            js_pred[r] << (first_pred[r] ? "" : ",") << "[[";
            first_pred[r] = false;
            if (traj[r]) {
                bool first = true;
                for (double t = 0.0; t <= pred_horizon; t += 0.05) {
                    const double pt = std::min(eval_t[r] + t, traj[r]->max_parameter());
                    const auto p = traj[r]->eval(pt, 0);
                    js_pred[r] << (first ? "" : ",") << "[" << p(0) << "," << p(1) << "," << p(2) << "]";
                    first = false;
                }
            } else {
                js_pred[r] << "[" << states[6 * r] << "," << states[6 * r + 1] << "," << states[6 * r + 2] << "]";
            }
            js_pred[r] << "]]";
            // rollout: h/Ts sub-steps along the newest available curve (or stay put)
            double et = eval_t[r];
            for (int ti = 1; ti <= substeps; ++ti) {
                double next[6];
                if (traj[r]) {
                    et = std::min(eval_t[r] + cfg.Ts * ti, traj[r]->max_parameter());
                    const auto p = traj[r]->eval(et, 0), v = traj[r]->eval(et, 1);
                    for (int d = 0; d < 3; ++d) { next[d] = p(d); next[3 + d] = v(d); }
                } else {
                    for (int d = 0; d < 3; ++d) { next[d] = states[6 * r + d]; next[3 + d] = 0.0; }
                }
                if (noise_seed != 0)
                    for (int d = 0; d < 3; ++d) {
                        next[d] += pos_std * gauss(rng);
                        next[3 + d] += vel_std * gauss(rng);
                    }
                // Gauss-Seidel: robot r's new state is visible to robots > r within this control
                // step; Jacobi: every robot was solved on the snapshot taken before this loop.
                for (int d = 0; d < 6; ++d) states[6 * r + d] = next[d];
                js_states[r] << (first_state[r] ? "" : ",") << "[" << next[0] << "," << next[1] << "," << next[2] << ","
                             << next[3] << "," << next[4] << "," << next[5] << "]";
                first_state[r] = false;
            }
            if (traj[r]) eval_t[r] = et;
        }
    }
    std::ofstream o(write_filename, std::ofstream::trunc);
    o << std::setprecision(17) << "{\"dt\": " << cfg.h << ", \"Ts\": " << cfg.Ts << ", \"robots\": {";
    for (int r = 0; r < N; ++r)
        o << (r ? ", " : "") << "\"" << r << "\": {\"states\": [" << js_states[r].str() << "], \"pred_curve\": ["
          << js_pred[r].str() << "]}";
    o << "}}\n";
    o.close();
    int64_t launches = 0, qps = 0;
    mpccbf_get_counters(h, &launches, &qps);
    std::printf("{\"mode\": \"%s\", \"robots\": %d, \"control_steps\": %d, \"qps\": %ld, \"failed_optimize_calls\": %ld, "
                "\"kernel_launches\": %lld, \"solve_seconds\": %.6f, \"qps_per_sec\": %.1f, \"states_file\": \"%s\"}\n",
                mode.c_str(), N, loop_idx, solves, failures, static_cast<long long>(launches), solve_seconds,
                solve_seconds > 0 ? solves / solve_seconds : 0.0, write_filename.c_str());
    mpccbf_destroy(h);
    return 0;
}
