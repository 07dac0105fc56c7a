// Configuration boundary: defaults, validation and JSON loading.
// Mirrors common/include/common/parsing.hpp:20-214 of the reference (same keys, same
// validation rules and messages) and workspace/experiments/config/base_config.json.
#include <cmath>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>

#include "../../include/mpccbf.h"
#include "json_min.hpp"

namespace mpccbf {
void set_error(const std::string& msg);
}
using mpccbf::set_error;

extern "C" {

void mpccbf_config_default(mpccbf_config* c) {
    if (!c) return;
    std::memset(c, 0, sizeof(*c));
    c->h = 0.1;
    c->Ts = 0.01;
    c->k_hor = 16;
    c->num_pieces = 3;
    c->num_ctrl_pts = 4;
    c->continuity_degree = 3;
    c->piece_max_param = 0.5;
    c->w_pos_err = 10.0;
    c->w_u_eff = 10.0;
    c->spd_f = 8;
    c->cbf_horizon = 2;
    const double v[3] = {2.0, 2.0, 2.6179938779914944};
    const double a[3] = {5.0, 5.0, 3.141592653589793};
    for (int i = 0; i < 3; ++i) {
        c->v_min[i] = -v[i];
        c->v_max[i] = v[i];
        c->a_min[i] = -a[i];
        c->a_max[i] = a[i];
    }
    c->d_min = 2.0;
    c->d_max = 4.0;
    c->cbf_gamma = 5.0;
    c->impc_iter = 2;
    c->slack_mode = 0;
    c->slack_cost = 50000.0;
    c->slack_decay = 0.1;
    c->num_agents = 2;
    c->max_scenarios = 1;
    c->device = 0;
    c->solver = MPCCBF_SOLVER_AUTO;
    c->feas_tol = 1e-9;
}

int mpccbf_config_validate(const mpccbf_config* c) {
    if (!c) {
        set_error("config is NULL");
        return MPCCBF_ERR_INVALID;
    }
    auto bad = [](const std::string& m) {
        set_error(m);
        return MPCCBF_ERR_INVALID;
    };
    // parseMPCParams (parsing.hpp:50-75)
    if (c->Ts > c->h)
        return bad("Control timestep Ts (" + std::to_string(c->Ts) + ") must be <= MPC timestep h (" +
                   std::to_string(c->h) + ")");
    if (c->h <= 0 || c->Ts <= 0) return bad("Time parameters h and Ts must be positive");
    const double ratio = c->h / c->Ts;
    if (std::fabs(ratio - std::round(ratio)) > 1e-10)
        return bad("MPC timestep h (" + std::to_string(c->h) +
                   ") must be an integer multiple of control timestep Ts (" + std::to_string(c->Ts) + ")");
    if (c->spd_f > c->k_hor)
        return bad("Speed factor spd_f (" + std::to_string(c->spd_f) +
                   ") must be <= prediction horizon k_hor (" + std::to_string(c->k_hor) + ")");
    if (c->spd_f < 1) return bad("Speed factor spd_f must be at least 1");
    if (c->k_hor < 1) return bad("Prediction horizon k_hor must be at least 1");
    // parseIMPCParams (parsing.hpp:121-132)
    if (c->cbf_horizon < 1) return bad("CBF horizon must be at least 1");
    if (c->impc_iter < 1) return bad("IMPC iterations must be at least 1");
    if (c->conn_mode != 0 && c->conn_mode != 1) return bad("conn_mode must be 0 or 1");
    if (c->slack_mode && c->slack_cost <= 0)
        return bad("Slack cost must be positive when slack_mode is enabled");
    if (c->slack_mode && (c->slack_decay <= 0 || c->slack_decay > 1))
        return bad("Slack decay rate must be in (0,1] when slack_mode is enabled");
    // validateCrossParameterRelationships (parsing.hpp:195-213)
    if (c->cbf_horizon > c->k_hor)
        return bad("CBF horizon (" + std::to_string(c->cbf_horizon) +
                   ") must be <= MPC prediction horizon k_hor (" + std::to_string(c->k_hor) + ")");
    const double total = static_cast<double>(c->num_pieces) * c->piece_max_param;
    const double need = static_cast<double>(c->k_hor - 1) * c->h;
    if (need > total)
        return bad("MPC sampling range [0, " + std::to_string(need) + "] exceeds Bezier curve parameter range [0, " +
                   std::to_string(total) + "]. Either reduce k_hor to <= " +
                   std::to_string(static_cast<int>(total / c->h) + 1) +
                   " or increase num_pieces/piece_max_parameter in bezier_params");
    // structural requirements of this implementation (the reference would throw later, inside the
    // generator, or index out of range)
    if (c->num_pieces < 1 || c->num_ctrl_pts < 2) return bad("num_pieces must be >= 1 and num_control_points >= 2");
    if (c->piece_max_param <= 0) return bad("piece_max_parameter must be positive");
    if (c->continuity_degree < 0) return bad("bezier_continuity_upto_degree must be >= 0");
    if (c->num_agents < 1) return bad("num_agents must be at least 1");
    if (c->max_scenarios < 1) return bad("max_scenarios must be at least 1");
    if (!(c->cbf_gamma > 0)) return bad("cbf_gamma must be positive");
    for (int i = 0; i < 3; ++i)
        if (c->v_min[i] > c->v_max[i] || c->a_min[i] > c->a_max[i])
            return bad("physical_limits: min exceeds max");
    return MPCCBF_OK;
}

int mpccbf_config_from_json(const char* path, mpccbf_config* c, double* so, double* sf, int32_t cap_agents) {
    if (!path || !c) {
        set_error("mpccbf_config_from_json: NULL argument");
        return MPCCBF_ERR_INVALID;
    }
    try {
        std::ifstream f(path);
        if (!f) {
            set_error(std::string("cannot open ") + path);
            return MPCCBF_ERR_INVALID;
        }
        std::stringstream ss;
        ss << f.rdbuf();
        const std::string text = ss.str();
        mpccbf::JsonParser parser(text);
        const mpccbf::JsonValue j = parser.parse();
        const mpccbf_config keep = *c;
        mpccbf_config_default(c);
        // handle-level fields are not part of the experiment file: keep the caller's
        c->max_scenarios = keep.max_scenarios > 0 ? keep.max_scenarios : 1;
        c->device = keep.device;
        c->solver = keep.solver;
        if (keep.feas_tol > 0) c->feas_tol = keep.feas_tol;
        c->admm_max_iter = keep.admm_max_iter;
        c->admm_check_every = keep.admm_check_every;
        c->admm_eps = keep.admm_eps;
        c->admm_rho = keep.admm_rho;
        c->admm_sigma = keep.admm_sigma;

        const auto& bz = j.at("bezier_params");
        c->num_pieces = static_cast<int32_t>(bz.at("num_pieces").number());
        c->num_ctrl_pts = static_cast<int32_t>(bz.at("num_control_points").number());
        c->piece_max_param = bz.at("piece_max_parameter").number();
        c->continuity_degree = static_cast<int32_t>(bz.at("bezier_continuity_upto_degree").number());
        const auto& mp = j.at("mpc_params");
        c->h = mp.at("h").number();
        c->Ts = mp.at("Ts").number();
        c->k_hor = static_cast<int32_t>(mp.at("k_hor").number());
        c->w_pos_err = mp.at("mpc_tuning").at("w_pos_err").number();
        c->w_u_eff = mp.at("mpc_tuning").at("w_u_eff").number();
        c->spd_f = static_cast<int32_t>(mp.at("mpc_tuning").at("spd_f").number());
        const auto& pl = j.at("physical_limits");
        for (int i = 0; i < 3; ++i) {
            c->v_min[i] = pl.at("v_min").at(i).number();
            c->v_max[i] = pl.at("v_max").at(i).number();
            c->a_min[i] = pl.at("a_min").at(i).number();
            c->a_max[i] = pl.at("a_max").at(i).number();
        }
        const auto& cb = j.at("cbf_params");
        c->d_min = cb.at("d_min").number();
        c->d_max = cb.at("d_max").number();
        c->cbf_horizon = static_cast<int32_t>(cb.at("cbf_horizon").number());
        c->impc_iter = static_cast<int32_t>(cb.at("impc_iter").number());
        c->slack_mode = cb.at("slack_mode").boolean() ? 1 : 0;
        c->slack_cost = cb.at("slack_cost").number();
        c->slack_decay = cb.at("slack_decay_rate").number();
        if (j.has("tasks")) {
            const auto& so_j = j.at("tasks").at("so");
            const auto& sf_j = j.at("tasks").at("sf");
            const int32_t nr = static_cast<int32_t>(so_j.size());
            c->num_agents = nr;
            if (so && sf) {
                if (nr > cap_agents) {
                    set_error("tasks hold more agents than cap_agents");
                    return MPCCBF_ERR_INVALID;
                }
                for (int32_t i = 0; i < nr; ++i)
                    for (int d = 0; d < 3; ++d) {
                        so[3 * i + d] = so_j.at(i).at(d).number();
                        sf[3 * i + d] = sf_j.at(i).at(d).number();
                    }
            }
        } else {
            c->num_agents = keep.num_agents > 0 ? keep.num_agents : 2;
        }
    } catch (const std::exception& e) {
        set_error(e.what());
        return MPCCBF_ERR_INVALID;
    }
    return mpccbf_config_validate(c);
}


int mpccbf_episode_opts_from_json(const char* path, mpccbf_episode_opts* o) {
    if (!path || !o) {
        set_error("mpccbf_episode_opts_from_json: NULL argument");
        return MPCCBF_ERR_INVALID;
    }
    try {
        std::ifstream f(path);
        if (!f) {
            set_error(std::string("cannot open ") + path);
            return MPCCBF_ERR_INVALID;
        }
        std::stringstream ss;
        ss << f.rdbuf();
        const std::string text = ss.str();
        mpccbf::JsonParser parser(text);
        const mpccbf::JsonValue j = parser.parse();
        std::memset(o, 0, sizeof(*o));
        o->goal_radius = 1.0;  // instance_success(..., radius=1, ...) collision_check.py:104
        o->order = MPCCBF_ORDER_JACOBI;
        const auto& shape = j.at("robot_params").at("collision_shape");
        if (shape.has("aligned_box")) {  // collision_check.py:93-101: aligned_box wins over radius
            o->collision_shape = MPCCBF_SHAPE_BOX;
            o->box_half[0] = shape.at("aligned_box").at(0).number();
            o->box_half[1] = shape.at("aligned_box").at(1).number();
            if (shape.has("radius")) o->collision_radius = shape.at("radius").number();
        } else if (shape.has("radius")) {
            o->collision_shape = MPCCBF_SHAPE_CIRCLE;
            o->collision_radius = shape.at("radius").number();
        } else {
            set_error("Missing collision shape: must provide either 'aligned_box' or 'radius'");
            return MPCCBF_ERR_INVALID;
        }
    } catch (const std::exception& e) {
        set_error(std::string("mpccbf_episode_opts_from_json: ") + e.what());
        return MPCCBF_ERR_INVALID;
    }
    return MPCCBF_OK;
}

}  // extern "C"
