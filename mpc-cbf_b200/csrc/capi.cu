// C ABI of libmpccbf.so (include/mpccbf.h). Thin: argument checks, table upload, launches.
// No CPU fallback anywhere: without a usable sm_100-class device every compute entry point
// returns MPCCBF_ERR_CUDA.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/mpccbf.h"
#include "kernels.h"
#include "tables.hpp"

namespace mpccbf {

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }

}  // namespace mpccbf

using namespace mpccbf;

struct mpccbf_handle {
    mpccbf_config cfg;
    HostTables tab;
    DevTables dev;
    GenTables gen;            // general reduced kernel (r != 2 free dof per axis), see impc_general.cu
    bool general = false;
    std::vector<double> g_gt, g_vs, g_vt, g_vr, g_kr, g_yw;
    double *dg_gt = nullptr, *dg_vs = nullptr, *dg_vt = nullptr, *dg_vr = nullptr, *dg_kr = nullptr, *dg_yw = nullptr;
    RolloutConst rc;
    bool has_device = false;
    int device = -1;
    int sm_count = 0;
    int solver = MPCCBF_SOLVER_REDUCED;
    int lanes = -1;  // lanes per agent-QP: 0 = thread-per-agent kernel, 8/32 = lane-group kernel, -1 = by batch size
    cudaStream_t stream = nullptr, stream2 = nullptr;
    // device copies of the tables
    double *d_tmpl = nullptr, *d_yw = nullptr, *d_vr = nullptr, *d_kr = nullptr;
    std::vector<double> h_tmpl, h_yw, h_vr, h_kr;  // host copies (uploaded at create)
    // workspaces for the host-pointer entry points (capacity: max_scenarios * num_agents)
    double *w_states = nullptr, *w_targets = nullptr, *w_ctrl = nullptr, *w_obj = nullptr;
    int32_t *w_status = nullptr, *w_ncurves = nullptr, *w_iters = nullptr;
    // single-agent workspace (optimize_one)
    double *o_states = nullptr, *o_ref = nullptr, *o_ctrl = nullptr, *o_all = nullptr;
    int32_t* o_status = nullptr;
    // episode driver workspaces (allocated on first use)
    double *e_curve = nullptr, *e_evalt = nullptr, *e_mind2 = nullptr, *e_ctrl = nullptr, *e_sub = nullptr, *e_tg = nullptr;
    int32_t *e_has = nullptr, *e_done = nullptr, *e_finish = nullptr, *e_failed = nullptr, *e_status = nullptr, *e_nc = nullptr;
    unsigned char* e_reached = nullptr;
    // slack mode (cbf_params.slack_mode): capped multipliers inside the fused thread-per-agent kernel
    // (gi_solver.h); scratch = one leader byte per (agent, neighbour)
    unsigned char* w_lead = nullptr;
    void* w_filter = nullptr;   // workspace of the CBF-only filter for swarms of 33..256 agents (allocated on first use)
    int w_filter_blocks = 0;
    // conn_mode: lambda2 [S] and Fiedler vector [S][N] of every scenario (cbf_filter.cu), refreshed before every step
    double *w_lam2 = nullptr, *w_fied = nullptr;
    std::vector<double> hc_lam2, hc_fied;  // host-only handles (tests/emul)
    // Latency path for small batches (<= small_cap agents) and mpccbf_optimize_one: inputs are copied by the CPU into a
    // pinned, device-mapped staging buffer that the kernel reads in place, outputs are stored by the kernel straight
    // into pinned mapped memory and copied out by the CPU after the stream sync — one launch and one sync per call
    // instead of five cudaMemcpyAsync + launch + sync.
    size_t small_cap = 0;
    unsigned char *pin_in = nullptr, *pin_out = nullptr;     // host addresses
    unsigned char *pin_in_d = nullptr, *pin_out_d = nullptr;  // the same memory as the device sees it
    struct Pending {
        bool active = false;
        size_t cnt = 0, n = 0, it = 0;
        double* ctrl = nullptr;
        int32_t* status = nullptr;
        int32_t* n_curves = nullptr;
        double* objective = nullptr;
    } pending;
    int64_t launches = 0, qps = 0;
};

#define CUDA_TRY(expr)                                                                     \
    do {                                                                                   \
        cudaError_t _e = static_cast<cudaError_t>(expr);                                   \
        if (_e != cudaSuccess) {                                                           \
            set_error(std::string(#expr) + ": " + cudaGetErrorString(_e));                 \
            return MPCCBF_ERR_CUDA;                                                        \
        }                                                                                  \
    } while (0)

namespace {

int require_device(const mpccbf_handle* h) {
    if (!h) {
        set_error("handle is NULL");
        return MPCCBF_ERR_INVALID;
    }
    if (!h->has_device) {
        set_error("no CUDA device bound to this handle (host-only handle, device = -1): there is no CPU fallback");
        return MPCCBF_ERR_CUDA;
    }
    return MPCCBF_OK;
}

template <typename T>
int upload(const std::vector<T>& v, T** out) {
    CUDA_TRY(cudaMalloc(reinterpret_cast<void**>(out), sizeof(T) * std::max<size_t>(v.size(), 1)));
    CUDA_TRY(cudaMemcpy(*out, v.data(), sizeof(T) * v.size(), cudaMemcpyHostToDevice));
    return MPCCBF_OK;
}

int build_device_tables(mpccbf_handle* h) {
    const HostTables& t = h->tab;
    DevTables& d = h->dev;
    std::memset(&d, 0, sizeof(d));
    d.K = t.K;
    d.p = t.p;
    d.c = t.c;
    d.m1 = t.m1;
    d.n = t.n;
    d.N = h->cfg.num_agents;
    d.H = h->cfg.cbf_horizon;
    d.iters = h->cfg.impc_iter;
    for (int ax = 0; ax < 3; ++ax) {
        d.lim_lo[0][ax] = h->cfg.a_min[ax];
        d.lim_hi[0][ax] = h->cfg.a_max[ax];
        d.lim_lo[1][ax] = h->cfg.v_min[ax];
        d.lim_hi[1][ax] = h->cfg.v_max[ax];
    }
    d.dmin2 = h->cfg.d_min * h->cfg.d_min;
    d.gamma = h->cfg.cbf_gamma;
    d.feas_tol = h->cfg.feas_tol > 0 ? h->cfg.feas_tol : 1e-9;
    d.slack = h->cfg.slack_mode ? 1 : 0;
    d.slack_cost = h->cfg.slack_cost;
    d.slack_decay = h->cfg.slack_decay;
    d.conn_mode = h->cfg.conn_mode ? 1 : 0;
    d.dmax = h->cfg.d_max;
    // rollout constants: t = h with the reference's lower_bound piece rule
    {
        RolloutConst& rc = h->rc;
        std::memset(&rc, 0, sizeof(rc));
        rc.n = t.n;
        rc.c = t.c;
        double cum = 0.0;
        int piece = 0;
        double prev = 0.0;
        for (int i = 0; i < t.p; ++i) {
            prev = cum;
            cum = (i == 0) ? h->cfg.piece_max_param : cum + h->cfg.piece_max_param;
            piece = i;
            if (!(cum < h->cfg.h)) break;  // first cumulative end >= t
        }
        const double local = std::min(h->cfg.piece_max_param, (piece == 0) ? h->cfg.h : h->cfg.h - prev);
        rc.piece = piece;
        xreal row[8];
        if (t.c <= 8) {
            bernstein_deriv(t.c, h->cfg.piece_max_param, local, 0, row);
            for (int j = 0; j < t.c; ++j) rc.b0[j] = static_cast<double>(row[j]);
            bernstein_deriv(t.c, h->cfg.piece_max_param, local, 1, row);
            for (int j = 0; j < t.c; ++j) rc.b1[j] = static_cast<double>(row[j]);
        }
    }
    if (t.reducible && t.r1 >= 1 && t.r1 <= kGenMaxR && t.r1 != 2) {
        // ---- general reduced kernel: same whitened null-space coordinates, r free dof per axis ------------------
        GenTables& G = h->gen;
        std::memset(&G, 0, sizeof(G));
        const int K = t.K, r = t.r1;
        G.K = K; G.r = r; G.m1 = t.m1; G.n = t.n; G.c = t.c; G.p = t.p;
        G.N = h->cfg.num_agents; G.H = h->cfg.cbf_horizon; G.iters = h->cfg.impc_iter;
        G.max_gi = 50 * (6 * r + 12 * K + (d.N - 1) * d.H) + 100;
        std::memcpy(G.lim_lo, d.lim_lo, sizeof(G.lim_lo));
        std::memcpy(G.lim_hi, d.lim_hi, sizeof(G.lim_hi));
        G.dmin2 = d.dmin2; G.gamma = d.gamma; G.feas_tol = d.feas_tol;
        for (int i = 0; i < 2; ++i) {
            xreal sk = 0.0L;
            for (int k = 0; k < K; ++k) sk += t.Kr(i, k);
            G.Kt[i] = static_cast<double>(sk);
            for (int j = 0; j < 2; ++j) G.Kc[i][j] = static_cast<double>(t.Kc(i, j));
        }
        h->g_gt.assign(static_cast<size_t>(3) * K * (r + 2), 0.0);
        for (int dd = 0; dd < 3; ++dd) {
            xreal nmax = 0.0L;
            std::vector<xreal> nrm(K, 0.0L);
            for (int k = 0; k < K; ++k) {
                for (int j = 0; j < r; ++j) nrm[k] += t.Ct[dd](k, j) * t.Ct[dd](k, j);
                nrm[k] = std::sqrt(nrm[k]);
                nmax = std::max(nmax, nrm[k]);
            }
            for (int k = 0; k < K; ++k) {
                double* e = &h->g_gt[(static_cast<size_t>(dd) * K + k) * (r + 2)];
                const bool constant_row = !(nrm[k] > 1e-12L * std::max(nmax, 1e-300L));  // pinned by the equality rows
                for (int j = 0; j < r; ++j) e[j] = constant_row ? 0.0 : static_cast<double>(t.Ct[dd](k, j));
                e[r] = static_cast<double>(t.Ot[dd](k, 0));
                e[r + 1] = static_cast<double>(t.Ot[dd](k, 1));
            }
        }
        h->g_vs.assign(static_cast<size_t>(r) * 2, 0.0);
        h->g_vt.assign(r, 0.0);
        h->g_vr.assign(static_cast<size_t>(r) * K, 0.0);
        h->g_kr.assign(static_cast<size_t>(2) * K, 0.0);
        h->g_yw.assign(static_cast<size_t>(t.m1) * (2 + r), 0.0);
        for (int j = 0; j < r; ++j) {
            xreal sv = 0.0L;
            for (int k = 0; k < K; ++k) {
                h->g_vr[static_cast<size_t>(j) * K + k] = static_cast<double>(t.Vr(j, k));
                sv += t.Vr(j, k);
            }
            h->g_vt[j] = static_cast<double>(sv);
            h->g_vs[2 * j] = static_cast<double>(t.Vs(j, 0));
            h->g_vs[2 * j + 1] = static_cast<double>(t.Vs(j, 1));
        }
        for (int i = 0; i < 2; ++i)
            for (int k = 0; k < K; ++k) h->g_kr[static_cast<size_t>(i) * K + k] = static_cast<double>(t.Kr(i, k));
        for (int i = 0; i < t.m1; ++i) {
            h->g_yw[static_cast<size_t>(i) * (2 + r)] = static_cast<double>(t.Y1(i, 0));
            h->g_yw[static_cast<size_t>(i) * (2 + r) + 1] = static_cast<double>(t.Y1(i, 1));
            for (int j = 0; j < r; ++j) h->g_yw[static_cast<size_t>(i) * (2 + r) + 2 + j] = static_cast<double>(t.W1(i, j));
        }
        if (h->has_device) {
            int rc;
            if ((rc = upload(h->g_gt, &h->dg_gt)) != MPCCBF_OK) return rc;
            if ((rc = upload(h->g_vs, &h->dg_vs)) != MPCCBF_OK) return rc;
            if ((rc = upload(h->g_vt, &h->dg_vt)) != MPCCBF_OK) return rc;
            if ((rc = upload(h->g_vr, &h->dg_vr)) != MPCCBF_OK) return rc;
            if ((rc = upload(h->g_kr, &h->dg_kr)) != MPCCBF_OK) return rc;
            if ((rc = upload(h->g_yw, &h->dg_yw)) != MPCCBF_OK) return rc;
            G.gt = h->dg_gt; G.vs = h->dg_vs; G.vt = h->dg_vt; G.vr = h->dg_vr; G.kr = h->dg_kr; G.yw = h->dg_yw;
        } else {
            G.gt = h->g_gt.data(); G.vs = h->g_vs.data(); G.vt = h->g_vt.data(); G.vr = h->g_vr.data();
            G.kr = h->g_kr.data(); G.yw = h->g_yw.data();
        }
        h->general = true;
        return MPCCBF_OK;
    }
    if (!(t.reducible && t.r1 == 2)) return MPCCBF_OK;
    const int K = t.K;
    d.max_gi = 50 * (6 + 12 * K + (d.N - 1) * d.H) + 100;
    for (int i = 0; i < 2; ++i) {
        for (int j = 0; j < 2; ++j) {
            d.Vs[i][j] = static_cast<double>(t.Vs(i, j));
            d.Kc[i][j] = static_cast<double>(t.Kc(i, j));
        }
        xreal sv = 0.0L, sk = 0.0L;
        for (int k = 0; k < K; ++k) {
            sv += t.Vr(i, k);
            sk += t.Kr(i, k);
        }
        d.Vt[i] = static_cast<double>(sv);
        d.Kt[i] = static_cast<double>(sk);
    }
    std::vector<double>&tmpl = h->h_tmpl, &yw = h->h_yw, &vr = h->h_vr, &kr = h->h_kr;
    tmpl.assign(static_cast<size_t>(3) * K * kTmplStride, 0.0);
    yw.assign(static_cast<size_t>(t.m1) * 4, 0.0);
    vr.assign(2 * K, 0.0);
    kr.assign(2 * K, 0.0);
    for (int dd = 0; dd < 3; ++dd) {
        xreal nmax = 0.0L;
        for (int k = 0; k < K; ++k)
            nmax = std::max(nmax, std::sqrt(t.Ct[dd](k, 0) * t.Ct[dd](k, 0) + t.Ct[dd](k, 1) * t.Ct[dd](k, 1)));
        for (int k = 0; k < K; ++k) {
            double* e = &tmpl[(static_cast<size_t>(dd) * K + k) * kTmplStride];
            const xreal nrm = std::sqrt(t.Ct[dd](k, 0) * t.Ct[dd](k, 0) + t.Ct[dd](k, 1) * t.Ct[dd](k, 1));
            const bool constant_row = !(nrm > 1e-12L * std::max(nmax, 1e-300L));
            // rows that do not depend on the free variables (e.g. velocity at t = 0, pinned by the
            // equality rows) get an exactly zero normal and an infinite score when violated
            e[0] = constant_row ? 0.0 : static_cast<double>(t.Ct[dd](k, 0));
            e[1] = constant_row ? 0.0 : static_cast<double>(t.Ct[dd](k, 1));
            e[2] = static_cast<double>(t.Ot[dd](k, 0));
            e[3] = static_cast<double>(t.Ot[dd](k, 1));
            e[4] = constant_row ? static_cast<double>(INFINITY) : static_cast<double>(1.0L / nrm);
            e[5] = t.h_samples[k];  // sample time t_k (read by the polynomial row scan)
        }
    }
    for (int i = 0; i < t.m1; ++i) {
        yw[4 * i + 0] = static_cast<double>(t.Y1(i, 0));
        yw[4 * i + 1] = static_cast<double>(t.Y1(i, 1));
        yw[4 * i + 2] = static_cast<double>(t.W1(i, 0));
        yw[4 * i + 3] = static_cast<double>(t.W1(i, 1));
    }
    for (int i = 0; i < 2; ++i)
        for (int k = 0; k < K; ++k) {
            vr[i * K + k] = static_cast<double>(t.Vr(i, k));
            kr[i * K + k] = static_cast<double>(t.Kr(i, k));
        }
    // ---- polynomial structure of the bound rows along the horizon (see device_tables.h) ---------
    d.poly = 0;
    if (K >= 4) {
        const std::vector<double>& ts = t.h_samples;
        auto col = [&](int dd, int k, int j) -> xreal { return j < 2 ? t.Ct[dd](k, j) : t.Ot[dd](k, j - 2); };
        bool ok = true;
        xreal scale = 0.0L;
        for (int dd = 1; dd <= 2; ++dd)
            for (int k = 0; k < K; ++k)
                for (int j = 0; j < 4; ++j) scale = std::max(scale, std::fabs(col(dd, k, j)));
        const int ka = 0, kb = K / 2, kc = K - 1;
        const xreal ta = ts[ka], tb = ts[kb], tc = ts[kc];
        for (int j = 0; j < 4 && ok; ++j) {
            // acceleration: affine through the end samples
            const xreal a1 = (col(2, kc, j) - col(2, ka, j)) / (tc - ta), a0 = col(2, ka, j) - a1 * ta;
            for (int k = 0; k < K; ++k)
                if (std::fabs(a0 + a1 * ts[k] - col(2, k, j)) > 1e-12L * scale) ok = false;
            // velocity: quadratic through three samples (Newton form -> monomials about t = 0)
            const xreal f0 = col(1, ka, j);
            const xreal f01 = (col(1, kb, j) - f0) / (tb - ta);
            const xreal f12 = (col(1, kc, j) - col(1, kb, j)) / (tc - tb);
            const xreal f012 = (f12 - f01) / (tc - ta);
            const xreal c2 = f012, c1 = f01 - f012 * (ta + tb), c0 = f0 - f01 * ta + f012 * ta * tb;
            for (int k = 0; k < K; ++k) {
                const xreal tk = ts[k];
                if (std::fabs(c0 + (c1 + c2 * tk) * tk - col(1, k, j)) > 1e-12L * scale) ok = false;
            }
            d.pq[0][j] = static_cast<double>(a0);
            d.pq[1][j] = static_cast<double>(a1);
            d.pq[2][j] = static_cast<double>(c0);
            d.pq[3][j] = static_cast<double>(c1);
            d.pq[4][j] = static_cast<double>(c2);
        }
        // the key scan of the thread kernel also needs finite two-sided limits and k < 128
        for (int dd = 0; dd < 2; ++dd)
            for (int ax = 0; ax < 3; ++ax)
                if (!(std::fabs(d.lim_lo[dd][ax]) < 1e20 && std::fabs(d.lim_hi[dd][ax]) < 1e20)) ok = false;
        if (K > 127) ok = false;
        d.poly = ok ? 1 : 0;
    }
    if (const char* env = std::getenv("MPCCBF_FULL_SCAN"))
        if (std::atoi(env) != 0) d.poly = 0;
    if (h->has_device) {
        int rc;
        if ((rc = upload(tmpl, &h->d_tmpl)) != MPCCBF_OK) return rc;
        if ((rc = upload(yw, &h->d_yw)) != MPCCBF_OK) return rc;
        if ((rc = upload(vr, &h->d_vr)) != MPCCBF_OK) return rc;
        if ((rc = upload(kr, &h->d_kr)) != MPCCBF_OK) return rc;
        d.tmpl = h->d_tmpl;
        d.yw = h->d_yw;
        d.Vr = h->d_vr;
        d.Kr = h->d_kr;
    } else {  // host-only handle: tables stay on the host (table queries, tests/emul)
        d.tmpl = tmpl.data();
        d.yw = yw.data();
        d.Vr = vr.data();
        d.Kr = kr.data();
    }
    return MPCCBF_OK;
}

int ensure_episode_ws(mpccbf_handle* h) {
    if (h->e_curve) return MPCCBF_OK;
    const size_t B = static_cast<size_t>(h->cfg.max_scenarios) * h->cfg.num_agents, n = h->tab.n, it = h->cfg.impc_iter;
    const int nsub = static_cast<int>(h->cfg.h / h->cfg.Ts);
    bool ok = true;
    ok &= cudaMalloc(&h->e_curve, sizeof(double) * B * n) == cudaSuccess;
    ok &= cudaMalloc(&h->e_ctrl, sizeof(double) * B * n) == cudaSuccess;
    ok &= cudaMalloc(&h->e_evalt, sizeof(double) * B) == cudaSuccess;
    ok &= cudaMalloc(&h->e_sub, sizeof(double) * B * std::max(nsub, 1) * 2) == cudaSuccess;
    ok &= cudaMalloc(&h->e_tg, sizeof(double) * h->cfg.max_scenarios * 3) == cudaSuccess;
    ok &= cudaMalloc(&h->e_has, sizeof(int32_t) * B) == cudaSuccess;
    ok &= cudaMalloc(&h->e_nc, sizeof(int32_t) * B) == cudaSuccess;
    ok &= cudaMalloc(&h->e_status, sizeof(int32_t) * B * it) == cudaSuccess;
    ok &= cudaMalloc(&h->e_reached, B) == cudaSuccess;
    ok &= cudaMalloc(&h->e_mind2, sizeof(double) * h->cfg.max_scenarios) == cudaSuccess;
    ok &= cudaMalloc(&h->e_done, sizeof(int32_t) * h->cfg.max_scenarios) == cudaSuccess;
    ok &= cudaMalloc(&h->e_finish, sizeof(int32_t) * h->cfg.max_scenarios) == cudaSuccess;
    ok &= cudaMalloc(&h->e_failed, sizeof(int32_t) * h->cfg.max_scenarios) == cudaSuccess;
    if (!ok) {
        set_error("episode / curve-following workspace: out of device memory");
        return MPCCBF_ERR_ALLOC;
    }
    return MPCCBF_OK;
}

// the handle-constant part of EpisodeArgs
void fill_episode_consts(const mpccbf_handle* h, EpisodeArgs& a) {
    a.curve = h->e_curve; a.eval_t = h->e_evalt; a.has_curve = h->e_has; a.reached = h->e_reached;
    a.done = h->e_done; a.finish = h->e_finish; a.min_dist2 = h->e_mind2; a.failed = h->e_failed;
    a.sub = h->e_sub;
    a.N = h->cfg.num_agents; a.n = h->tab.n; a.c = h->tab.c; a.p = h->tab.p; a.iters = h->cfg.impc_iter;
    a.nsub = static_cast<int>(h->cfg.h / h->cfg.Ts);
    a.Ts = h->cfg.Ts; a.piece_T = h->cfg.piece_max_param;
    double cum = 0.0;
    for (int i = 0; i < 16; ++i) {
        if (i < h->tab.p) cum += h->cfg.piece_max_param;  // cumulative ends by repeated addition, like the reference
        a.cum[i] = cum;
    }
    a.max_param = cum;
    a.final_only = 0;
    a.box = 0; a.box_x = a.box_y = 0.0;
    a.goal_radius = a.collision_radius = 0.0;
    a.pos_std = a.vel_std = 0.0; a.seed = 0;
    a.traj = nullptr; a.traj_len = 0; a.step = 0;
}

// Which fused kernel a step of `num_qp` agent-QPs runs: 0 = thread-per-agent, 8 / 32 = lanes per agent.
int resolve_lanes(const mpccbf_handle* h, int64_t num_qp, bool force_auto = false) {
    int lanes = force_auto ? -1 : h->lanes;
    if (lanes < 0) {
        // Measured crossover on B200 (profiles/r2_kernel_ab.md section 5; round 1: profiles/r1h_kernel_crossover.md): a warp
        // per agent (32 lanes split the row scan; ~3.6 us per active-set step against ~5 us for one thread) wins while
        // its warps fit in two waves of 8 per SM (2368 agents: 0.094 vs 0.127 ms at 2048 agents, a tie at 4096); the
        // thread-per-agent kernel wins beyond. The 8-lane variant is only the large-batch fallback for configurations
        // without the polynomial row structure.
        const int wave = 16 * h->sm_count * (h->cfg.num_agents >= 128 ? 2 : 1);   // dense swarms: a scan walks hundreds of rows
        lanes = (num_qp <= wave) ? 32 : 0;
    }
    if (h->cfg.slack_mode || h->cfg.conn_mode) lanes = 0;  // capped multipliers / lifted connectivity rows: thread-per-agent kernel only
    if (lanes == 0 && !h->dev.poly)  // the thread kernel needs the polynomial row structure
        lanes = (num_qp >= 2048 && h->cfg.num_agents < 64) ? 8 : 32;
    return lanes;
}

// Workspace of the CBF-only filter for swarms of more than 32 agents: Laplacian + eigenvector matrix (and slack-mode leaders)
// of every resident CTA, 2 CTAs per SM, allocated on the first call that needs it.
int ensure_filter_ws(mpccbf_handle* h) {
    if (h->cfg.num_agents <= 32 || h->w_filter) return MPCCBF_OK;
    const int blocks = std::max(1, std::min(h->cfg.max_scenarios, 2 * h->sm_count));
    if (cudaMalloc(&h->w_filter, cbf_filter_wide_ws_bytes(h->cfg, blocks)) != cudaSuccess) {
        h->w_filter = nullptr;
        set_error("CBF-only filter: cudaMalloc of the eigen-solve workspace failed");
        return MPCCBF_ERR_ALLOC;
    }
    h->w_filter_blocks = blocks;
    return MPCCBF_OK;
}

int run_step(mpccbf_handle* h, int32_t S, int32_t agent_begin, int32_t agent_count, const double* states,
             const double* targets, int ref_mode, double* ctrl, double* ctrl_all, double* v_out, int32_t* status,
             int32_t* n_curves, double* objective, int32_t* iters, cudaStream_t stream) {
    if (h->solver != MPCCBF_SOLVER_REDUCED) {
        set_error("controller path: this configuration does not reduce to the 6-dof form (needs cubic pieces with "
                  "C^3 continuity); use mpccbf_solve_qp_dense for the assembled QP");
        return MPCCBF_ERR_UNSUPPORTED;
    }
    StepArgs a;
    a.states = states;
    a.targets = targets;
    a.ctrl_pts = ctrl;
    a.ctrl_all = ctrl_all;
    a.v_out = v_out;
    a.status = status;
    a.n_curves = n_curves;
    a.objective = objective;
    a.gi_iters = iters;
    a.num_qp = S * agent_count;
    a.agent_begin = agent_begin;
    a.agent_count = agent_count;
    a.ref_mode = ref_mode;
    a.lead_ws = h->w_lead;
    if (h->general) {
        if (v_out) {
            set_error("the compact result form needs cubic pieces with C^3 continuity (2 free dof per axis)");
            return MPCCBF_ERR_UNSUPPORTED;
        }
        CUDA_TRY(launch_impc_general(h->gen, a, h->sm_count, stream));
        h->launches += 1;
        h->qps += static_cast<int64_t>(a.num_qp) * h->cfg.impc_iter;
        return MPCCBF_OK;
    }
    if (h->cfg.conn_mode) {  // Laplacian, Jacobi eigen-solve, second smallest eigenpair of every scenario (cbf_filter.cu)
        const int wrc = ensure_filter_ws(h);
        if (wrc != MPCCBF_OK) return wrc;
        CUDA_TRY(launch_cbf_filter(h->cfg, S, states, nullptr, nullptr, nullptr, h->w_lam2, nullptr, h->sm_count, stream, h->w_fied,
                                   h->w_filter, h->w_filter_blocks));
        a.lambda2 = h->w_lam2;
        a.fiedler = h->w_fied;
        h->launches += 1;
    }
    const int lanes = resolve_lanes(h, a.num_qp);
    if (lanes == 0)
        CUDA_TRY(launch_impc_thread(h->dev, a, h->sm_count, stream));
    else
        CUDA_TRY(launch_impc_reduced(h->dev, a, lanes, h->sm_count, stream));
    h->launches += 1;
    h->qps += static_cast<int64_t>(a.num_qp) * h->cfg.impc_iter;
    return MPCCBF_OK;
}

__global__ void dfma_peak_kernel(double* out, int iters, double a, double b) {
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; ++i) {
        x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
        x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
    }
    const double s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == 123.456) out[0] = s;  // keeps the chain alive without a store in practice
}

}  // namespace

extern "C" {

const char* mpccbf_last_error(void) { return g_error.c_str(); }
const char* mpccbf_version(void) { return "mpccbf-b200 0.1 (sm_100a)"; }

int mpccbf_create(const mpccbf_config* cfg, mpccbf_handle** out) {
    if (!cfg || !out) {
        set_error("mpccbf_create: NULL argument");
        return MPCCBF_ERR_INVALID;
    }
    *out = nullptr;
    int rc = mpccbf_config_validate(cfg);
    if (rc != MPCCBF_OK) return rc;
    mpccbf_handle* h = new (std::nothrow) mpccbf_handle();
    if (!h) {
        set_error("out of host memory");
        return MPCCBF_ERR_ALLOC;
    }
    h->cfg = *cfg;
    if (h->cfg.cbf_gamma <= 0) h->cfg.cbf_gamma = 5.0;
    if (!h->tab.build(h->cfg)) {
        set_error(h->tab.error);
        delete h;
        return MPCCBF_ERR_INVALID;
    }
    const bool can_reduce = h->tab.reducible && h->tab.r1 >= 1 && h->tab.r1 <= kGenMaxR && h->cfg.cbf_horizon <= 8 && h->cfg.num_agents >= 1;
    if (cfg->solver == MPCCBF_SOLVER_REDUCED && !can_reduce) {
        set_error("MPCCBF_SOLVER_REDUCED requested but the configuration does not reduce to the 6-dof form");
        delete h;
        return MPCCBF_ERR_UNSUPPORTED;
    }
    h->solver = (cfg->solver == MPCCBF_SOLVER_DENSE || !can_reduce) ? MPCCBF_SOLVER_DENSE : MPCCBF_SOLVER_REDUCED;
    if (const char* env = std::getenv("MPCCBF_LANES")) {
        const int w = std::atoi(env);
        if (w == 32 || w == 8 || w == 1 || w == 0) h->lanes = w;
    }
    h->device = cfg->device;
    if (cfg->device >= 0) {
        int count = 0;
        cudaError_t e = cudaGetDeviceCount(&count);
        if (e != cudaSuccess || cfg->device >= count) {
            set_error(std::string("no usable CUDA device ") + std::to_string(cfg->device) + ": " +
                      (e != cudaSuccess ? cudaGetErrorString(e) : "ordinal out of range") +
                      " (there is no CPU fallback)");
            delete h;
            return MPCCBF_ERR_CUDA;
        }
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, cfg->device) != cudaSuccess || prop.major < 10) {
            set_error("device is not sm_100-class (kernels are built for sm_100a only)");
            delete h;
            return MPCCBF_ERR_CUDA;
        }
        if (cudaSetDevice(cfg->device) != cudaSuccess) {
            set_error("cudaSetDevice failed");
            delete h;
            return MPCCBF_ERR_CUDA;
        }
        h->sm_count = prop.multiProcessorCount;
        h->has_device = true;
        if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess ||
            cudaStreamCreateWithFlags(&h->stream2, cudaStreamNonBlocking) != cudaSuccess) {
            set_error("cudaStreamCreate failed");
            mpccbf_destroy(h);
            return MPCCBF_ERR_CUDA;
        }
    }
    rc = build_device_tables(h);
    if (rc != MPCCBF_OK) {
        mpccbf_destroy(h);
        return rc;
    }
    if (h->general && (h->cfg.slack_mode || h->cfg.conn_mode)) {
        set_error("slack_mode / conn_mode need cubic pieces with C^3 continuity (the fused thread-per-agent kernel); this "
                  "configuration is served by the general reduced kernel");
        mpccbf_destroy(h);
        return MPCCBF_ERR_UNSUPPORTED;
    }
    if (h->cfg.conn_mode && (h->cfg.slack_mode || h->cfg.num_agents > 256 || h->solver != MPCCBF_SOLVER_REDUCED || !h->dev.poly)) {
        set_error("conn_mode needs slack_mode = 0 (the reference indexes the slack of the connectivity row by self_idx, out of "
                  "range for the last robot: ConnectivityMPCCBFQPGenerator.cpp:56-58), num_agents <= 256 (the eigen-solve's "
                  "limit) and a configuration the thread-per-agent kernel serves");
        mpccbf_destroy(h);
        return MPCCBF_ERR_UNSUPPORTED;
    }
    if (h->cfg.slack_mode && h->solver == MPCCBF_SOLVER_REDUCED && !h->dev.poly) {
        set_error("slack mode needs the polynomial row structure of the thread-per-agent kernel (k_hor in 4..127, finite "
                  "velocity and acceleration limits)");
        mpccbf_destroy(h);
        return MPCCBF_ERR_UNSUPPORTED;
    }
    if (h->has_device) {
        const size_t B = static_cast<size_t>(cfg->max_scenarios) * cfg->num_agents;
        const size_t n = h->tab.n, it = cfg->impc_iter, K = cfg->k_hor, N = cfg->num_agents;
        bool ok = true;
        ok &= cudaMalloc(&h->w_states, sizeof(double) * B * 6) == cudaSuccess;
        ok &= cudaMalloc(&h->w_targets, sizeof(double) * B * 3) == cudaSuccess;
        ok &= cudaMalloc(&h->w_ctrl, sizeof(double) * B * n) == cudaSuccess;
        ok &= cudaMalloc(&h->w_obj, sizeof(double) * B) == cudaSuccess;
        ok &= cudaMalloc(&h->w_status, sizeof(int32_t) * B * it) == cudaSuccess;
        ok &= cudaMalloc(&h->w_ncurves, sizeof(int32_t) * B) == cudaSuccess;
        ok &= cudaMalloc(&h->w_iters, sizeof(int32_t) * B * it) == cudaSuccess;
        ok &= cudaMalloc(&h->o_states, sizeof(double) * N * 6) == cudaSuccess;
        ok &= cudaMalloc(&h->o_ref, sizeof(double) * 3 * K) == cudaSuccess;
        ok &= cudaMalloc(&h->o_ctrl, sizeof(double) * n) == cudaSuccess;
        ok &= cudaMalloc(&h->o_all, sizeof(double) * n * it) == cudaSuccess;
        ok &= cudaMalloc(&h->o_status, sizeof(int32_t) * (it + 1 + it)) == cudaSuccess;
        {
            static const bool allow_small = [] { const char* e = std::getenv("MPCCBF_SMALL_BATCH"); return !e || std::atoi(e) != 0; }();
            h->small_cap = allow_small ? std::min<size_t>(B, 1024) : 0;
            const size_t in_bytes = h->small_cap * 9 * sizeof(double) + (N * 6 + 3 * K) * sizeof(double);
            const size_t out_bytes = h->small_cap * (n * sizeof(double) + it * 4 + 4 + 8) + (n * it * sizeof(double) + (2 * it + 1) * 4) + 64;
            if (h->small_cap > 0 &&
                (cudaHostAlloc(reinterpret_cast<void**>(&h->pin_in), in_bytes, cudaHostAllocMapped) != cudaSuccess ||
                 cudaHostAlloc(reinterpret_cast<void**>(&h->pin_out), out_bytes, cudaHostAllocMapped) != cudaSuccess ||
                 cudaHostGetDevicePointer(reinterpret_cast<void**>(&h->pin_in_d), h->pin_in, 0) != cudaSuccess ||
                 cudaHostGetDevicePointer(reinterpret_cast<void**>(&h->pin_out_d), h->pin_out, 0) != cudaSuccess)) {
                (void)cudaGetLastError();
                h->small_cap = 0;  // no mapped pinned memory: the staged-copy path serves every size
            }
        }
        if (h->cfg.conn_mode) {
            ok &= cudaMalloc(&h->w_lam2, sizeof(double) * std::max<size_t>(cfg->max_scenarios, 1)) == cudaSuccess;
            ok &= cudaMalloc(&h->w_fied, sizeof(double) * std::max<size_t>(B, 1)) == cudaSuccess;
        }
        if (h->cfg.slack_mode) ok &= cudaMalloc(&h->w_lead, std::max<size_t>(B * (N > 1 ? N - 1 : 1), 1)) == cudaSuccess;
        if (!ok) {
            set_error("device workspace allocation failed");
            mpccbf_destroy(h);
            return MPCCBF_ERR_ALLOC;
        }
    }
    *out = h;
    return MPCCBF_OK;
}

int mpccbf_destroy(mpccbf_handle* h) {
    if (!h) return MPCCBF_OK;
    if (h->has_device) {
        cudaSetDevice(h->device);
        void* eptrs[] = {h->e_sub, h->e_tg, h->e_curve, h->e_ctrl, h->e_evalt, h->e_has, h->e_nc, h->e_status, h->e_reached, h->e_mind2,
                         h->e_done, h->e_finish, h->e_failed};
        for (void* q : eptrs)
            if (q) cudaFree(q);
        if (h->w_lead) cudaFree(h->w_lead);
        if (h->w_filter) cudaFree(h->w_filter);
        if (h->w_lam2) cudaFree(h->w_lam2);
        if (h->pin_in) cudaFreeHost(h->pin_in);
        if (h->pin_out) cudaFreeHost(h->pin_out);
        if (h->w_fied) cudaFree(h->w_fied);
        void* gptrs[] = {h->dg_gt, h->dg_vs, h->dg_vt, h->dg_vr, h->dg_kr, h->dg_yw};
        for (void* q : gptrs)
            if (q) cudaFree(q);
        void* ptrs[] = {h->d_tmpl, h->d_yw, h->d_vr, h->d_kr, h->w_states, h->w_targets, h->w_ctrl, h->w_obj,
                        h->w_status, h->w_ncurves, h->w_iters, h->o_states, h->o_ref, h->o_ctrl, h->o_all, h->o_status};
        for (void* p : ptrs)
            if (p) cudaFree(p);
        if (h->stream) cudaStreamDestroy(h->stream);
        if (h->stream2) cudaStreamDestroy(h->stream2);
    }
    delete h;
    return MPCCBF_OK;
}

int32_t mpccbf_num_decision_vars(const mpccbf_handle* h) { return h ? h->tab.n : 0; }

int mpccbf_set_kernel_variant(mpccbf_handle* h, int32_t lanes) {
    if (!h || !(lanes == -1 || lanes == 0 || lanes == 8 || lanes == 32)) {
        set_error("mpccbf_set_kernel_variant: lanes must be -1 (auto), 0 (thread per agent), 8 or 32");
        return MPCCBF_ERR_INVALID;
    }
    h->lanes = lanes;
    return MPCCBF_OK;
}
int32_t mpccbf_auto_kernel_variant(const mpccbf_handle* h, int64_t num_agent_qps) {
    if (!h) return -1;
    return resolve_lanes(h, num_agent_qps, true);
}
int32_t mpccbf_active_solver(const mpccbf_handle* h) { return h ? h->solver : 0; }

void* mpccbf_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
        set_error("cudaHostAlloc failed");
        return nullptr;
    }
    return p;
}
void mpccbf_host_free(void* p) {
    if (p) cudaFreeHost(p);
}

int mpccbf_step_agents(mpccbf_handle* h, int32_t S, int32_t agent_begin, int32_t agent_count,
                       const double* states, const double* targets, double* ctrl_pts, int32_t* status,
                       int32_t* n_curves, double* objective, void* cuda_stream) {
    int rc = require_device(h);
    if (rc != MPCCBF_OK) return rc;
    if (S < 0 || agent_begin < 0 || agent_count < 0 || agent_begin + agent_count > h->cfg.num_agents ||
        !states || !targets || !ctrl_pts || !status) {
        set_error("mpccbf_step_agents: invalid argument");
        return MPCCBF_ERR_INVALID;
    }
    if (static_cast<int64_t>(S) * agent_count > static_cast<int64_t>(h->cfg.max_scenarios) * h->cfg.num_agents) {
        set_error("num_scenarios exceeds the handle's max_scenarios");
        return MPCCBF_ERR_INVALID;
    }
    CUDA_TRY(cudaSetDevice(h->device));
    return run_step(h, S, agent_begin, agent_count, states, targets, 0, ctrl_pts, nullptr, nullptr, status, n_curves,
                    objective, h->w_iters, static_cast<cudaStream_t>(cuda_stream));
}

int mpccbf_step_batch(mpccbf_handle* h, int32_t S, const double* states, const double* targets,
                      double* ctrl_pts, int32_t* status, int32_t* n_curves, double* objective,
                      void* cuda_stream) {
    if (!h) {
        set_error("handle is NULL");
        return MPCCBF_ERR_INVALID;
    }
    return mpccbf_step_agents(h, S, 0, h->cfg.num_agents, states, targets, ctrl_pts, status, n_curves, objective,
                              cuda_stream);
}

const int32_t* mpccbf_last_iteration_counts(const mpccbf_handle* h) { return h ? h->w_iters : nullptr; }

int mpccbf_copy_iteration_counts(mpccbf_handle* h, int32_t* out_host, int64_t count) {
    int rc = require_device(h);
    if (rc != MPCCBF_OK) return rc;
    const int64_t cap = static_cast<int64_t>(h->cfg.max_scenarios) * h->cfg.num_agents * h->cfg.impc_iter;
    if (!out_host || count < 0 || count > cap) {
        set_error("mpccbf_copy_iteration_counts: invalid argument");
        return MPCCBF_ERR_INVALID;
    }
    CUDA_TRY(cudaSetDevice(h->device));
    // the handle's own streams are non-blocking (not ordered against the legacy stream cudaMemcpy uses): finish
    // them first. A step enqueued on a CALLER's stream is the caller's to synchronise (documented in mpccbf.h).
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream2));
    CUDA_TRY(cudaMemcpy(out_host, h->w_iters, sizeof(int32_t) * count, cudaMemcpyDeviceToHost));
    return MPCCBF_OK;
}

int mpccbf_synchronize(mpccbf_handle* h) {
    int rc = require_device(h);
    if (rc != MPCCBF_OK) return rc;
    CUDA_TRY(cudaSetDevice(h->device));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream2));
    if (h->pending.active) {  // small-batch latency path: results sit in the pinned staging buffer
        const mpccbf_handle::Pending& p = h->pending;
        const unsigned char* o = h->pin_out;
        std::memcpy(p.ctrl, o, sizeof(double) * p.cnt * p.n);
        o += sizeof(double) * p.cnt * p.n;
        if (p.objective) std::memcpy(p.objective, o, sizeof(double) * p.cnt);
        o += sizeof(double) * p.cnt;
        std::memcpy(p.status, o, sizeof(int32_t) * p.cnt * p.it);
        o += sizeof(int32_t) * p.cnt * p.it;
        if (p.n_curves) std::memcpy(p.n_curves, o, sizeof(int32_t) * p.cnt);
        h->pending.active = false;
    }
    return MPCCBF_OK;
}

static int step_host_enqueue(mpccbf_handle* h, int32_t S, const double* states, const double* targets, double* ctrl_pts,
                             int32_t* status, int32_t* n_curves, double* objective, bool zero_copy, bool compact);

int mpccbf_step_batch_host(mpccbf_handle* h, int32_t S, const double* states, const double* targets,
                           double* ctrl_pts, int32_t* status, int32_t* n_curves, double* objective) {
    const int rc = step_host_enqueue(h, S, states, targets, ctrl_pts, status, n_curves, objective, true, false);
    if (rc != MPCCBF_OK || S == 0) return rc;
    return mpccbf_synchronize(h);
}

// The pipelined form always stages through device memory: measured on B200, kernels of two handles that
// both store control points straight into pinned host memory collapse to ~11 ms per batch (PCIe reads of
// one kernel queue behind the posted writes of the other), while copy-engine transfers overlap cleanly
// with the other handle's solve (0.55 ms per batch, the PCIe bound for 31 MB).
int mpccbf_step_batch_host_async(mpccbf_handle* h, int32_t S, const double* states, const double* targets,
                                 double* ctrl_pts, int32_t* status, int32_t* n_curves, double* objective) {
    return step_host_enqueue(h, S, states, targets, ctrl_pts, status, n_curves, objective, false, false);
}

int mpccbf_step_batch_compact(mpccbf_handle* h, int32_t S, const double* states, const double* targets, double* v_free,
                              int32_t* status, int32_t* n_curves, double* objective, void* cuda_stream) {
    int rc = require_device(h);
    if (rc != MPCCBF_OK) return rc;
    if (S < 0 || S > h->cfg.max_scenarios || !states || !targets || !v_free || !status) {
        set_error("mpccbf_step_batch_compact: invalid argument (num_scenarios must be <= max_scenarios)");
        return MPCCBF_ERR_INVALID;
    }
    CUDA_TRY(cudaSetDevice(h->device));
    return run_step(h, S, 0, h->cfg.num_agents, states, targets, 0, nullptr, nullptr, v_free, status, n_curves, objective,
                    h->w_iters, static_cast<cudaStream_t>(cuda_stream));
}

int mpccbf_step_batch_host_compact(mpccbf_handle* h, int32_t S, const double* states, const double* targets, double* v_free,
                                   int32_t* status, int32_t* n_curves, double* objective) {
    const int rc = step_host_enqueue(h, S, states, targets, v_free, status, n_curves, objective, false, true);
    if (rc != MPCCBF_OK || S == 0) return rc;
    return mpccbf_synchronize(h);
}

int mpccbf_step_batch_host_compact_async(mpccbf_handle* h, int32_t S, const double* states, const double* targets,
                                         double* v_free, int32_t* status, int32_t* n_curves, double* objective) {
    return step_host_enqueue(h, S, states, targets, v_free, status, n_curves, objective, false, true);
}

// x[piece, axis, cp] = Y1 [pos; vel] + W1 [va; vb] per axis: the exact linear map the kernels apply on the device
int mpccbf_expand_ctrl_pts(const mpccbf_handle* h, int64_t count, const double* states, const double* v_free,
                           const int32_t* n_curves, double* ctrl_pts) {
    if (!h || count < 0 || !states || !v_free || !ctrl_pts) {
        set_error("mpccbf_expand_ctrl_pts: invalid argument");
        return MPCCBF_ERR_INVALID;
    }
    if (h->solver != MPCCBF_SOLVER_REDUCED || h->general) {
        set_error("mpccbf_expand_ctrl_pts: the compact form needs cubic pieces with C^3 continuity");
        return MPCCBF_ERR_UNSUPPORTED;
    }
    const int n = h->tab.n, c = h->tab.c, m1 = h->tab.m1;
    const double* yw = h->h_yw.data();
    for (int64_t g = 0; g < count; ++g) {
        const double* s = states + g * 6;
        const double* v = v_free + g * 6;
        double* out = ctrl_pts + g * n;
        if (n_curves && n_curves[g] <= 0) {
            for (int i = 0; i < n; ++i) out[i] = 0.0;
            continue;
        }
        for (int j = 0; j < m1; ++j) {
            const int piece = j / c, cp = j - piece * c;
            const double* y = yw + 4 * j;
            for (int ax = 0; ax < 3; ++ax)
                out[piece * 3 * c + ax * c + cp] = std::fma(y[0], s[ax], std::fma(y[1], s[3 + ax], std::fma(y[2], v[2 * ax], y[3] * v[2 * ax + 1])));
        }
    }
    return MPCCBF_OK;
}

// curve^(deriv)(t) straight from the compact record: piece lookup and clamping of
// SingleParameterPiecewiseCurve::eval (splines/src/curves/SingleParameterPiecewiseCurve.cpp:94-127)
int mpccbf_eval_compact(const mpccbf_handle* h, int64_t count, const double* states, const double* v_free, double t,
                        int32_t deriv, double* out) {
    if (!h || count < 0 || !states || !v_free || !out || deriv < 0 || t < 0) {
        set_error("mpccbf_eval_compact: invalid argument");
        return MPCCBF_ERR_INVALID;
    }
    if (h->solver != MPCCBF_SOLVER_REDUCED || h->general || h->tab.c > 8) {
        set_error("mpccbf_eval_compact: the compact form needs cubic pieces with C^3 continuity");
        return MPCCBF_ERR_UNSUPPORTED;
    }
    const int c = h->tab.c, p = h->tab.p;
    const double T = h->cfg.piece_max_param;
    double cum = 0.0, prev = 0.0;
    int piece = p - 1;
    for (int i = 0; i < p; ++i) {  // first piece whose cumulative end >= t (lower_bound)
        prev = cum;
        cum += T;
        if (!(cum < t)) {
            piece = i;
            break;
        }
    }
    if (t > cum) {
        set_error("mpccbf_eval_compact: parameter beyond the curve's max_parameter");
        return MPCCBF_ERR_INVALID;
    }
    const double local = std::min(T, std::max(0.0, t - prev));
    xreal basis[8];
    bernstein_deriv(c, T, local, deriv, basis);
    double r[4] = {0, 0, 0, 0};  // basis' [Y1 | W1] of this piece
    for (int cp = 0; cp < c; ++cp)
        for (int k = 0; k < 4; ++k) r[k] += static_cast<double>(basis[cp]) * h->h_yw[4 * (piece * c + cp) + k];
    for (int64_t g = 0; g < count; ++g) {
        const double* s = states + g * 6;
        const double* v = v_free + g * 6;
        for (int ax = 0; ax < 3; ++ax) out[g * 3 + ax] = r[0] * s[ax] + r[1] * s[3 + ax] + r[2] * v[2 * ax] + r[3] * v[2 * ax + 1];
    }
    return MPCCBF_OK;
}

// compact: `ctrl_pts` receives [S][N][6] whitened free dof instead of [S][N][n] control points (48 instead of 8 n
// bytes per agent over PCIe)
static int step_host_enqueue(mpccbf_handle* h, int32_t S, const double* states, const double* targets, double* ctrl_pts,
                             int32_t* status, int32_t* n_curves, double* objective, bool zero_copy, bool compact) {
    int rc = require_device(h);
    if (rc != MPCCBF_OK) return rc;
    if (S < 0 || S > h->cfg.max_scenarios || !states || !targets || !ctrl_pts || !status) {
        set_error("mpccbf_step_batch_host: invalid argument (num_scenarios must be <= max_scenarios)");
        return MPCCBF_ERR_INVALID;
    }
    if (S == 0) return MPCCBF_OK;
    CUDA_TRY(cudaSetDevice(h->device));
    const size_t N = h->cfg.num_agents, n = compact ? 6 : h->tab.n, it = h->cfg.impc_iter;
    // Control points are 384 of the 404 output bytes per agent. When the caller's ctrl_pts buffer is
    // pinned host memory (mpccbf_host_alloc, cudaHostAlloc / cudaHostRegister: mapped under UVA) the
    // kernel stores them straight into it — coalesced posted writes over PCIe that overlap the solve —
    // instead of writing HBM and copying afterwards. The small per-agent outputs (status, n_curves,
    // objective) are single 4/8-byte stores and stay on the copy path. MPCCBF_ZEROCOPY=0 disables it.
    if (static_cast<size_t>(S) * N <= h->small_cap) {
        if (h->pending.active) {  // an earlier small batch has not been collected yet: finish it first
            rc = mpccbf_synchronize(h);
            if (rc != MPCCBF_OK) return rc;
        }
        const size_t cnt = static_cast<size_t>(S) * N;
        double* in = reinterpret_cast<double*>(h->pin_in);
        std::memcpy(in, states, sizeof(double) * cnt * 6);
        std::memcpy(in + cnt * 6, targets, sizeof(double) * cnt * 3);
        const double* in_d = reinterpret_cast<const double*>(h->pin_in_d);
        unsigned char* o = h->pin_out_d;
        double* ctrl_d = reinterpret_cast<double*>(o);
        double* obj_d = reinterpret_cast<double*>(o + sizeof(double) * cnt * n);
        int32_t* status_d = reinterpret_cast<int32_t*>(o + sizeof(double) * cnt * (n + 1));
        int32_t* nc_d = status_d + cnt * it;
        rc = run_step(h, S, 0, h->cfg.num_agents, in_d, in_d + cnt * 6, 0, compact ? nullptr : ctrl_d, nullptr, compact ? ctrl_d : nullptr,
                      status_d, nc_d, obj_d, h->w_iters, h->stream);
        if (rc != MPCCBF_OK) return rc;
        h->pending.active = true;
        h->pending.cnt = cnt; h->pending.n = n; h->pending.it = it;
        h->pending.ctrl = ctrl_pts; h->pending.status = status; h->pending.n_curves = n_curves; h->pending.objective = objective;
        return MPCCBF_OK;
    }
    double* ctrl_dev = nullptr;
    const double *states_dev = nullptr, *targets_dev = nullptr;  // pinned inputs are read in place as well
    {
        static const bool allow = [] { const char* e = std::getenv("MPCCBF_ZEROCOPY"); return !e || std::atoi(e) != 0; }();
        static const bool allow_in = [] { const char* e = std::getenv("MPCCBF_ZEROCOPY_IN"); return !e || std::atoi(e) != 0; }();
        auto mapped = [](const void* p) -> void* {
            cudaPointerAttributes pa;
            if (cudaPointerGetAttributes(&pa, p) == cudaSuccess && pa.type == cudaMemoryTypeHost && pa.devicePointer) return pa.devicePointer;
            (void)cudaGetLastError();
            return nullptr;
        };
        if (allow && zero_copy) ctrl_dev = static_cast<double*>(mapped(ctrl_pts));
        if (allow && allow_in && zero_copy) {
            states_dev = static_cast<const double*>(mapped(states));
            targets_dev = static_cast<const double*>(mapped(targets));
            if (!states_dev || !targets_dev) states_dev = targets_dev = nullptr;
        }
    }
    // The batch can also be cut into scenario chunks that alternate between two streams so that
    // copies and kernels overlap (MPCCBF_HOST_CHUNKS). Measured on B200 (profiles/r1d_host_chunks.md)
    // this does NOT pay at 65 536 agents: the fused kernel is latency-bound, a quarter-size launch
    // takes almost as long as the full one, so the default is one chunk.
    int nchunks = 1;
    if (const char* env = std::getenv("MPCCBF_HOST_CHUNKS")) nchunks = std::max(1, std::min(std::atoi(env), S));
    cudaStream_t streams[2] = {h->stream, h->stream2};
    for (int c = 0; c < nchunks; ++c) {
        const int s0 = static_cast<int>(static_cast<int64_t>(S) * c / nchunks);
        const int s1 = static_cast<int>(static_cast<int64_t>(S) * (c + 1) / nchunks);
        const size_t a0 = s0 * N, cnt = (s1 - s0) * N;
        if (cnt == 0) continue;
        cudaStream_t st = streams[c & 1];
        if (!states_dev) {
            CUDA_TRY(cudaMemcpyAsync(h->w_states + a0 * 6, states + a0 * 6, sizeof(double) * cnt * 6, cudaMemcpyHostToDevice, st));
            CUDA_TRY(cudaMemcpyAsync(h->w_targets + a0 * 3, targets + a0 * 3, sizeof(double) * cnt * 3, cudaMemcpyHostToDevice, st));
        }
        double* ctrl_out = ctrl_dev ? ctrl_dev + a0 * n : h->w_ctrl + a0 * n;
        rc = run_step(h, s1 - s0, 0, h->cfg.num_agents, states_dev ? states_dev + a0 * 6 : h->w_states + a0 * 6,
                      states_dev ? targets_dev + a0 * 3 : h->w_targets + a0 * 3, 0, compact ? nullptr : ctrl_out,
                      nullptr, compact ? ctrl_out : nullptr, h->w_status + a0 * it, h->w_ncurves + a0, h->w_obj + a0, h->w_iters + a0 * it, st);
        if (rc != MPCCBF_OK) return rc;
        if (!ctrl_dev)
            CUDA_TRY(cudaMemcpyAsync(ctrl_pts + a0 * n, h->w_ctrl + a0 * n, sizeof(double) * cnt * n, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaMemcpyAsync(status + a0 * it, h->w_status + a0 * it, sizeof(int32_t) * cnt * it, cudaMemcpyDeviceToHost, st));
        if (n_curves) CUDA_TRY(cudaMemcpyAsync(n_curves + a0, h->w_ncurves + a0, sizeof(int32_t) * cnt, cudaMemcpyDeviceToHost, st));
        if (objective) CUDA_TRY(cudaMemcpyAsync(objective + a0, h->w_obj + a0, sizeof(double) * cnt, cudaMemcpyDeviceToHost, st));
    }
    return MPCCBF_OK;
}

int mpccbf_optimize_one(mpccbf_handle* h, const double* states, int32_t self_idx, const double* ref,
                        double* ctrl_pts, int32_t* n_curves, int32_t* status) {
    int rc = require_device(h);
    if (rc != MPCCBF_OK) return rc;
    const int N = h->cfg.num_agents, K = h->cfg.k_hor, n = h->tab.n, it = h->cfg.impc_iter;
    if (!states || !ref || !ctrl_pts || !n_curves || !status || self_idx < 0 || self_idx >= N) {
        set_error("mpccbf_optimize_one: invalid argument");
        return MPCCBF_ERR_INVALID;
    }
    CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    if (h->small_cap > 0) {
        // latency path: pinned mapped staging, one launch + one sync (see mpccbf_handle::pin_in)
        if (h->pending.active) {
            rc = mpccbf_synchronize(h);
            if (rc != MPCCBF_OK) return rc;
        }
        double* in = reinterpret_cast<double*>(h->pin_in + h->small_cap * 9 * sizeof(double));
        std::memcpy(in, states, sizeof(double) * N * 6);
        std::memcpy(in + N * 6, ref, sizeof(double) * 3 * K);
        const double* in_d = reinterpret_cast<const double*>(h->pin_in_d + h->small_cap * 9 * sizeof(double));
        const size_t off = ((h->small_cap * (static_cast<size_t>(n) * sizeof(double) + it * 4 + 4 + 8) + 63) / 64) * 64;
        double* all_d = reinterpret_cast<double*>(h->pin_out_d + off);
        int32_t* st_d = reinterpret_cast<int32_t*>(h->pin_out_d + off + sizeof(double) * n * it);
        const double* all_h = reinterpret_cast<const double*>(h->pin_out + off);
        const int32_t* st_h = reinterpret_cast<const int32_t*>(h->pin_out + off + sizeof(double) * n * it);
        rc = run_step(h, 1, self_idx, 1, in_d, in_d + N * 6, 1, nullptr, all_d, nullptr, st_d, st_d + it, nullptr, st_d + it + 1, st);
        if (rc != MPCCBF_OK) return rc;
        CUDA_TRY(cudaStreamSynchronize(st));
        for (int i = 0; i < it; ++i) status[i] = st_h[i];
        *n_curves = st_h[it];
        if (st_h[it] > 0) std::memcpy(ctrl_pts, all_h, sizeof(double) * n * st_h[it]);
        return MPCCBF_OK;
    }
    CUDA_TRY(cudaMemcpyAsync(h->o_states, states, sizeof(double) * N * 6, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(h->o_ref, ref, sizeof(double) * 3 * K, cudaMemcpyHostToDevice, st));
    int32_t* d_status = h->o_status;
    int32_t* d_nc = h->o_status + it;
    int32_t* d_iters = h->o_status + it + 1;
    rc = run_step(h, 1, self_idx, 1, h->o_states, h->o_ref, 1, h->o_ctrl, h->o_all, nullptr, d_status, d_nc, nullptr, d_iters, st);
    if (rc != MPCCBF_OK) return rc;
    std::vector<int32_t> hs(it + 1);
    CUDA_TRY(cudaMemcpyAsync(hs.data(), d_status, sizeof(int32_t) * (it + 1), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    for (int i = 0; i < it; ++i) status[i] = hs[i];
    *n_curves = hs[it];
    if (hs[it] > 0) CUDA_TRY(cudaMemcpy(ctrl_pts, h->o_all, sizeof(double) * n * hs[it], cudaMemcpyDeviceToHost));
    return MPCCBF_OK;
}

int mpccbf_solve_qp_dense(mpccbf_handle* h, int32_t batch, int32_t n, int32_t m, const double* P,
                          const double* q, const double* A, const double* l, const double* u,
                          const double* lb, const double* ub, double* x, double* objective,
                          int32_t* status, int32_t* iters) {
    int rc = require_device(h);
    if (rc != MPCCBF_OK) return rc;
    if (batch < 0 || n < 1 || m < 0 || !P || !q || (m > 0 && (!A || !l || !u)) || !x || !status) {
        set_error("mpccbf_solve_qp_dense: invalid argument");
        return MPCCBF_ERR_INVALID;
    }
    CUDA_TRY(cudaSetDevice(h->device));
    DenseOpts o;
    o.max_iter = h->cfg.admm_max_iter;
    o.check_every = h->cfg.admm_check_every;
    o.eps = h->cfg.admm_eps;
    o.rho = h->cfg.admm_rho;
    o.sigma = h->cfg.admm_sigma;
    o.sm_count = h->sm_count;
    o.stream = h->stream;
    rc = dense_qp_solve_host(o, batch, n, m, P, q, A, l, u, lb, ub, x, objective, status, iters);
    if (rc == MPCCBF_OK) {
        h->launches += 1;
        h->qps += batch;
    }
    return rc;
}

int mpccbf_cbf_rows(mpccbf_handle* h, int32_t S, const double* states, double* rows, void* cuda_stream) {
    int rc = require_device(h);
    if (rc != MPCCBF_OK) return rc;
    if (S < 0 || !states || !rows) {
        set_error("mpccbf_cbf_rows: invalid argument");
        return MPCCBF_ERR_INVALID;
    }
    CUDA_TRY(cudaSetDevice(h->device));
    CUDA_TRY(launch_cbf_rows(states, rows, S, h->cfg.num_agents, h->cfg.d_min * h->cfg.d_min, h->cfg.cbf_gamma,
                             h->sm_count, static_cast<cudaStream_t>(cuda_stream)));
    h->launches += 1;
    return MPCCBF_OK;
}

int mpccbf_cbf_filter_batch(mpccbf_handle* h, int32_t S, const double* states, const double* desired_u, double* u,
                            int32_t* status, double* lambda2, void* cuda_stream) {
    int rc = require_device(h);
    if (rc != MPCCBF_OK) return rc;
    if (S < 0 || !states || !desired_u || !u || !status) {
        set_error("mpccbf_cbf_filter_batch: invalid argument");
        return MPCCBF_ERR_INVALID;
    }
    if (h->cfg.num_agents > 256) {
        set_error("mpccbf_cbf_filter_batch: num_agents must be <= 256 (one thread per agent of a scenario's CTA)");
        return MPCCBF_ERR_UNSUPPORTED;
    }
    CUDA_TRY(cudaSetDevice(h->device));
    rc = ensure_filter_ws(h);
    if (rc != MPCCBF_OK) return rc;
    CUDA_TRY(launch_cbf_filter(h->cfg, S, states, desired_u, u, status, lambda2, nullptr, h->sm_count,
                               static_cast<cudaStream_t>(cuda_stream), nullptr, h->w_filter, h->w_filter_blocks));
    h->launches += 1;
    h->qps += static_cast<int64_t>(S) * h->cfg.num_agents;
    return MPCCBF_OK;
}

int mpccbf_cbf_filter_batch_host(mpccbf_handle* h, int32_t S, const double* states, const double* desired_u, double* u,
                                 int32_t* status, double* lambda2) {
    int rc = require_device(h);
    if (rc != MPCCBF_OK) return rc;
    if (S < 0 || S > h->cfg.max_scenarios || !states || !desired_u || !u || !status) {
        set_error("mpccbf_cbf_filter_batch_host: invalid argument (num_scenarios must be <= max_scenarios)");
        return MPCCBF_ERR_INVALID;
    }
    if (S == 0) return MPCCBF_OK;
    CUDA_TRY(cudaSetDevice(h->device));
    const size_t B = static_cast<size_t>(S) * h->cfg.num_agents;
    cudaStream_t st = h->stream;
    // workspaces of the controller path are reused: w_targets [B][3] <- desired_u, w_ctrl (>= [B][3]) <- u,
    // w_obj [B] (>= [S]) <- lambda2
    CUDA_TRY(cudaMemcpyAsync(h->w_states, states, sizeof(double) * B * 6, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(h->w_targets, desired_u, sizeof(double) * B * 3, cudaMemcpyHostToDevice, st));
    rc = mpccbf_cbf_filter_batch(h, S, h->w_states, h->w_targets, h->w_ctrl, h->w_status, h->w_obj, st);
    if (rc != MPCCBF_OK) return rc;
    CUDA_TRY(cudaMemcpyAsync(u, h->w_ctrl, sizeof(double) * B * 3, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(status, h->w_status, sizeof(int32_t) * B, cudaMemcpyDeviceToHost, st));
    if (lambda2) CUDA_TRY(cudaMemcpyAsync(lambda2, h->w_obj, sizeof(double) * S, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return MPCCBF_OK;
}

int mpccbf_conn_rows(mpccbf_handle* h, int32_t S, const double* states, double* rows, double* lambda2, void* cuda_stream) {
    int rc = require_device(h);
    if (rc != MPCCBF_OK) return rc;
    if (S < 0 || !states || !rows) {
        set_error("mpccbf_conn_rows: invalid argument");
        return MPCCBF_ERR_INVALID;
    }
    if (h->cfg.num_agents > 256) {
        set_error("mpccbf_conn_rows: num_agents must be <= 256");
        return MPCCBF_ERR_UNSUPPORTED;
    }
    CUDA_TRY(cudaSetDevice(h->device));
    rc = ensure_filter_ws(h);
    if (rc != MPCCBF_OK) return rc;
    CUDA_TRY(launch_cbf_filter(h->cfg, S, states, nullptr, nullptr, nullptr, lambda2, rows, h->sm_count,
                               static_cast<cudaStream_t>(cuda_stream), nullptr, h->w_filter, h->w_filter_blocks));
    h->launches += 1;
    return MPCCBF_OK;
}

int mpccbf_run_episode_ex(mpccbf_handle* h, int32_t S, double* states, const double* targets, int32_t steps,
                          const mpccbf_episode_opts* opts, int32_t* success, int32_t* finish_step, double* min_dist,
                          int32_t* failed_calls, double* traj, void* cuda_stream) {
    int rc = require_device(h);
    if (rc != MPCCBF_OK) return rc;
    if (S < 0 || S > h->cfg.max_scenarios || steps < 0 || !states || !targets || !opts ||
        (opts->order != MPCCBF_ORDER_JACOBI && opts->order != MPCCBF_ORDER_GAUSS_SEIDEL) || opts->pos_std < 0 || opts->vel_std < 0 ||
        (opts->collision_shape != MPCCBF_SHAPE_CIRCLE && opts->collision_shape != MPCCBF_SHAPE_BOX)) {
        set_error("mpccbf_run_episode: invalid argument (num_scenarios must be <= max_scenarios)");
        return MPCCBF_ERR_INVALID;
    }
    if (h->tab.c > 8 || h->tab.p > 16) {
        set_error("mpccbf_run_episode: curves with more than 8 control points or 16 pieces are not supported");
        return MPCCBF_ERR_UNSUPPORTED;
    }
    if (S == 0) return MPCCBF_OK;
    CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    const int nsub = static_cast<int>(h->cfg.h / h->cfg.Ts);  // int(h / Ts), MPCCBFFormationControl_example.cpp:193
    rc = ensure_episode_ws(h);
    if (rc != MPCCBF_OK) return rc;
    const int N = h->cfg.num_agents;
    CUDA_TRY(launch_episode_init(S, N, h->e_evalt, h->e_has, h->e_reached, h->e_done, h->e_finish, h->e_mind2, h->e_failed, st));
    EpisodeArgs a;
    fill_episode_consts(h, a);
    a.states = states; a.targets = targets; a.ctrl_new = h->e_ctrl; a.n_curves = h->e_nc;
    a.traj = traj;
    a.S = S;
    a.traj_len = steps * a.nsub;
    a.goal_radius = opts->goal_radius; a.collision_radius = opts->collision_radius;
    a.box = opts->collision_shape == MPCCBF_SHAPE_BOX ? 1 : 0;
    a.box_x = opts->box_half[0]; a.box_y = opts->box_half[1];
    a.pos_std = opts->pos_std; a.vel_std = opts->vel_std; a.seed = opts->seed;
    for (int step = 0; step < steps; ++step) {
        a.step = step;
        if (opts->order == MPCCBF_ORDER_JACOBI) {
            rc = run_step(h, S, 0, N, states, targets, 0, h->e_ctrl, nullptr, nullptr, h->e_status, h->e_nc, nullptr, h->w_iters, st);
            if (rc != MPCCBF_OK) return rc;
            a.agent_begin = 0; a.agent_count = N; a.do_metrics = 1;
            CUDA_TRY(launch_episode_advance(a, h->sm_count, st));
            h->launches += 1;
        } else {
            for (int ag = 0; ag < N; ++ag) {
                CUDA_TRY(launch_gather_agent_targets(S, N, ag, targets, h->e_tg, st));
                rc = run_step(h, S, ag, 1, states, h->e_tg, 0, h->e_ctrl, nullptr, nullptr, h->e_status, h->e_nc, nullptr, h->w_iters, st);
                if (rc != MPCCBF_OK) return rc;
                a.agent_begin = ag; a.agent_count = 1; a.do_metrics = (ag == N - 1) ? 1 : 0;
                CUDA_TRY(launch_episode_advance(a, h->sm_count, st));
                h->launches += 2;
            }
        }
    }
    CUDA_TRY(launch_episode_final(S, steps * a.nsub, h->e_done, h->e_finish, h->e_mind2, success, finish_step, min_dist, st));
    if (failed_calls) CUDA_TRY(cudaMemcpyAsync(failed_calls, h->e_failed, sizeof(int32_t) * S, cudaMemcpyDeviceToDevice, st));
    return MPCCBF_OK;
}

int mpccbf_run_episode(mpccbf_handle* h, int32_t S, double* states, const double* targets, int32_t steps, double goal_radius,
                       double collision_radius, int32_t* success, int32_t* finish_step, double* min_dist, int32_t* failed_calls,
                       double* traj, void* cuda_stream) {
    mpccbf_episode_opts o;
    o.goal_radius = goal_radius;
    o.collision_radius = collision_radius;
    o.order = MPCCBF_ORDER_JACOBI;
    o.pos_std = o.vel_std = 0.0;
    o.seed = 0;
    o.collision_shape = MPCCBF_SHAPE_CIRCLE;
    o.box_half[0] = o.box_half[1] = 0.0;
    return mpccbf_run_episode_ex(h, S, states, targets, steps, &o, success, finish_step, min_dist, failed_calls, traj, cuda_stream);
}

int mpccbf_follow_reset(mpccbf_handle* h, int32_t S, void* cuda_stream) {
    int rc = require_device(h);
    if (rc != MPCCBF_OK) return rc;
    if (S < 0 || S > h->cfg.max_scenarios) {
        set_error("mpccbf_follow_reset: invalid argument (num_scenarios must be <= max_scenarios)");
        return MPCCBF_ERR_INVALID;
    }
    CUDA_TRY(cudaSetDevice(h->device));
    rc = ensure_episode_ws(h);
    if (rc != MPCCBF_OK) return rc;
    CUDA_TRY(launch_episode_init(S, h->cfg.num_agents, h->e_evalt, h->e_has, h->e_reached, h->e_done, h->e_finish, h->e_mind2,
                                 h->e_failed, static_cast<cudaStream_t>(cuda_stream)));
    return MPCCBF_OK;
}

int mpccbf_follow_step(mpccbf_handle* h, int32_t S, int32_t agent_begin, int32_t agent_count, double* states,
                       const double* ctrl_new, const int32_t* n_curves, void* cuda_stream) {
    int rc = require_device(h);
    if (rc != MPCCBF_OK) return rc;
    if (S < 0 || S > h->cfg.max_scenarios || agent_begin < 0 || agent_count < 0 || agent_begin + agent_count > h->cfg.num_agents ||
        !states || !ctrl_new || !n_curves) {
        set_error("mpccbf_follow_step: invalid argument");
        return MPCCBF_ERR_INVALID;
    }
    if (h->tab.c > 8 || h->tab.p > 16) {
        set_error("mpccbf_follow_step: curves with more than 8 control points or 16 pieces are not supported");
        return MPCCBF_ERR_UNSUPPORTED;
    }
    if (S == 0 || agent_count == 0) return MPCCBF_OK;
    CUDA_TRY(cudaSetDevice(h->device));
    rc = ensure_episode_ws(h);
    if (rc != MPCCBF_OK) return rc;
    EpisodeArgs a;
    fill_episode_consts(h, a);
    a.states = states; a.targets = nullptr; a.ctrl_new = ctrl_new; a.n_curves = n_curves;
    a.S = S;
    a.agent_begin = agent_begin; a.agent_count = agent_count; a.do_metrics = 0;
    a.final_only = 1;
    CUDA_TRY(launch_episode_advance(a, h->sm_count, static_cast<cudaStream_t>(cuda_stream)));
    h->launches += 1;
    return MPCCBF_OK;
}

int mpccbf_rollout(mpccbf_handle* h, int32_t S, const double* states, const double* ctrl_pts,
                   const int32_t* n_curves, double* next_states, void* cuda_stream) {
    if (!h) {
        set_error("handle is NULL");
        return MPCCBF_ERR_INVALID;
    }
    return mpccbf_rollout_agents(h, S, h->cfg.num_agents, states, ctrl_pts, n_curves, next_states, cuda_stream);
}

int mpccbf_rollout_agents(mpccbf_handle* h, int32_t S, int32_t agent_count, const double* states,
                          const double* ctrl_pts, const int32_t* n_curves, double* next_states,
                          void* cuda_stream) {
    int rc = require_device(h);
    if (rc != MPCCBF_OK) return rc;
    if (S < 0 || agent_count < 0 || !states || !ctrl_pts || !next_states) {
        set_error("mpccbf_rollout: invalid argument");
        return MPCCBF_ERR_INVALID;
    }
    if (h->tab.c > 8) {
        set_error("mpccbf_rollout: more than 8 control points per piece is not supported");
        return MPCCBF_ERR_UNSUPPORTED;
    }
    CUDA_TRY(cudaSetDevice(h->device));
    CUDA_TRY(launch_rollout(states, ctrl_pts, n_curves, next_states, S * agent_count, h->rc,
                            static_cast<cudaStream_t>(cuda_stream)));
    h->launches += 1;
    return MPCCBF_OK;
}

int64_t mpccbf_get_table(const mpccbf_handle* h, const char* name, double* out, int64_t cap) {
    if (!h || !name) {
        set_error("mpccbf_get_table: NULL argument");
        return MPCCBF_ERR_INVALID;
    }
    const std::string nm(name);
    std::vector<double> v;
    const HostTables& t = h->tab;
    if (nm == "U_basis") v = t.full_U();
    else if (nm == "V_basis") v = t.full_basis(1);
    else if (nm == "P") v = t.full_P();
    else if (nm == "E") v = t.full_E();
    else if (nm == "A_fix") v = t.full_Afix();
    else if (nm == "h_samples") v = t.h_samples;
    else if (nm == "Z" && t.reducible) v = t.full_Z();
    else if (nm == "Y" && t.reducible) v = t.full_Y();
    else if (nm == "H_red" && t.reducible) v = t.full_Hred();
    else {
        set_error("mpccbf_get_table: unknown table '" + nm + "'");
        return MPCCBF_ERR_INVALID;
    }
    const int64_t cnt = static_cast<int64_t>(v.size());
    if (!out) return cnt;
    if (cap < cnt) {
        set_error("mpccbf_get_table: buffer too small");
        return MPCCBF_ERR_INVALID;
    }
    std::memcpy(out, v.data(), sizeof(double) * v.size());
    return cnt;
}

int mpccbf_linear_cost(const mpccbf_handle* h, const double* x0, const double* ref, double* q) {
    if (!h || !x0 || !ref || !q) {
        set_error("mpccbf_linear_cost: NULL argument");
        return MPCCBF_ERR_INVALID;
    }
    h->tab.linear_cost(x0, ref, q);
    return MPCCBF_OK;
}

// Round 1 had a second (ADMM) tier for slack-mode QPs; the caps now live in the active-set kernel and
// nothing falls back. Kept so that callers built against the round-1 header keep linking.
int64_t mpccbf_slack_fallback_count(const mpccbf_handle* h) { return h ? 0 : -1; }

int mpccbf_get_counters(const mpccbf_handle* h, int64_t* kernel_launches, int64_t* qps_solved) {
    if (!h) return MPCCBF_ERR_INVALID;
    if (kernel_launches) *kernel_launches = h->launches;
    if (qps_solved) *qps_solved = h->qps;
    return MPCCBF_OK;
}

#ifdef MPCCBF_HOST_EMUL
// Test-only entry points (tests/emul/libmpccbf_emul.so): the per-agent device function run on the
// host for CPU-only CI. NOT part of include/mpccbf.h and NOT compiled into libmpccbf.so.
int mpccbf_emul_step(mpccbf_handle* h, int32_t S, int32_t agent_begin, int32_t agent_count, int32_t ref_mode,
                     const double* states, const double* targets, double* ctrl_pts, double* ctrl_all,
                     int32_t* status, int32_t* n_curves, double* objective, int32_t* iters, double* v_out) {
    if (!h || h->has_device || h->solver != MPCCBF_SOLVER_REDUCED) {
        set_error("mpccbf_emul_step: needs a host-only (device = -1) reducible handle");
        return MPCCBF_ERR_INVALID;
    }
    StepArgs a;
    a.states = states;
    a.targets = targets;
    a.ctrl_pts = ctrl_pts;
    a.ctrl_all = ctrl_all;
    a.status = status;
    a.n_curves = n_curves;
    a.objective = objective;
    a.gi_iters = iters;
    a.v_out = v_out;
    a.num_qp = S * agent_count;
    a.agent_begin = agent_begin;
    a.agent_count = agent_count;
    a.ref_mode = ref_mode;
    if (h->general) {
        emul_impc_general_host(h->gen, a);
        return MPCCBF_OK;
    }
    if (h->cfg.conn_mode) {  // host stand-in for cbf_filter.cu's warp-level eigen-solve (cyclic Jacobi, same weights)
        const int N = h->cfg.num_agents;
        h->hc_lam2.assign(S, 0.0);
        h->hc_fied.assign(static_cast<size_t>(S) * N, 0.0);
        const double R2 = h->cfg.d_max * h->cfg.d_max, sigma = R2 * R2 / std::log(2.0);
        std::vector<double> Lm(static_cast<size_t>(N) * N), Vm(static_cast<size_t>(N) * N);
        for (int s = 0; s < S && N > 1; ++s) {
            const double* st = states + static_cast<size_t>(s) * N * 6;
            std::fill(Lm.begin(), Lm.end(), 0.0);
            std::fill(Vm.begin(), Vm.end(), 0.0);
            for (int i = 0; i < N; ++i) {
                Vm[i * N + i] = 1.0;
                for (int j = 0; j < N; ++j) {
                    if (i == j) continue;
                    const double dx = st[i * 6] - st[j * 6], dy = st[i * 6 + 1] - st[j * 6 + 1], d2 = dx * dx + dy * dy;
                    if (d2 <= R2) {
                        const double w = std::exp((R2 - d2) * (R2 - d2) / sigma) - 1.0;
                        Lm[i * N + j] = -w;
                        Lm[i * N + i] += w;
                    }
                }
            }
            for (int sweep = 0; sweep < 60; ++sweep) {
                double off = 0.0, dg = 0.0;
                for (int i = 0; i < N; ++i)
                    for (int j = 0; j < N; ++j) (i == j ? dg : off) += Lm[i * N + j] * Lm[i * N + j];
                if (!(off > 4.930380657631324e-32 * dg)) break;
                for (int p = 0; p < N - 1; ++p)
                    for (int q = p + 1; q < N; ++q) {
                        const double apq = Lm[p * N + q];
                        if (apq == 0.0) continue;
                        const double app = Lm[p * N + p], aqq = Lm[q * N + q];
                        const double theta = (aqq - app) / (2.0 * apq);
                        const double t = std::copysign(1.0, theta) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
                        const double c = 1.0 / std::sqrt(t * t + 1.0), sn = t * c;
                        for (int k = 0; k < N; ++k) {
                            if (k != p && k != q) {
                                const double akp = Lm[k * N + p], akq = Lm[k * N + q];
                                Lm[k * N + p] = Lm[p * N + k] = c * akp - sn * akq;
                                Lm[k * N + q] = Lm[q * N + k] = sn * akp + c * akq;
                            }
                            const double vkp = Vm[k * N + p], vkq = Vm[k * N + q];
                            Vm[k * N + p] = c * vkp - sn * vkq;
                            Vm[k * N + q] = sn * vkp + c * vkq;
                        }
                        Lm[p * N + p] = app - t * apq;
                        Lm[q * N + q] = aqq + t * apq;
                        Lm[p * N + q] = Lm[q * N + p] = 0.0;
                    }
            }
            int i1 = 0, i2 = -1;
            for (int i = 1; i < N; ++i)
                if (Lm[i * N + i] < Lm[i1 * N + i1]) i1 = i;
            for (int i = 0; i < N; ++i)
                if (i != i1 && (i2 < 0 || Lm[i * N + i] < Lm[i2 * N + i2])) i2 = i;
            h->hc_lam2[s] = Lm[i2 * N + i2];
            double nn = 0.0;
            for (int i = 0; i < N; ++i) nn += Vm[i * N + i2] * Vm[i * N + i2];
            for (int i = 0; i < N; ++i) h->hc_fied[static_cast<size_t>(s) * N + i] = Vm[i * N + i2] / std::sqrt(nn);
        }
        a.lambda2 = h->hc_lam2.data();
        a.fiedler = h->hc_fied.data();
    }
    if (h->lanes > 0 || !h->dev.poly)
        emul_impc_reduced_host(h->dev, a);
    else
        emul_impc_thread_host(h->dev, a);
    return MPCCBF_OK;
}
#endif

double mpccbf_measure_fp64_tflops(mpccbf_handle* h) {
    if (require_device(h) != MPCCBF_OK) return MPCCBF_ERR_CUDA;
    if (cudaSetDevice(h->device) != cudaSuccess) return MPCCBF_ERR_CUDA;
    double* d = nullptr;
    if (cudaMalloc(&d, 8) != cudaSuccess) return MPCCBF_ERR_CUDA;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = h->sm_count * 8, threads = 256, iters = 20000;
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0, h->stream);
        dfma_peak_kernel<<<blocks, threads, 0, h->stream>>>(d, iters, 1.0000001, 1e-9);
        cudaEventRecord(e1, h->stream);
        if (cudaEventSynchronize(e1) != cudaSuccess) break;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flops = 2.0 * 8.0 * iters * static_cast<double>(blocks) * threads;
        if (rep > 0 && ms > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    return best;
}

}  // extern "C"
