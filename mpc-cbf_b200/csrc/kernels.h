// Launch interfaces shared by the .cu translation units of libmpccbf.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/mpccbf.h"
#include "device_tables.h"

struct mpccbf_handle;

namespace mpccbf {

struct StepArgs {
    const double* states;    // [S][N][6]  full snapshot of every scenario
    const double* targets;   // [B][3]  (ref_mode 0)  or  [B][3K] (ref_mode 1)
    double* ctrl_pts;        // [B][n]            last successful iterate (zeros if none); NULL: not wanted
    double* ctrl_all;        // [B][iters][n]     every successful iterate, may be NULL
    double* v_out = nullptr; // [B][6]            compact result: whitened free dof of the last successful iterate
                             //                   (zeros if none); x = Y [pos; vel] + W v. May be NULL
    int32_t* status;         // [B][iters]
    int32_t* n_curves;       // [B]   may be NULL
    double* objective;       // [B]   may be NULL
    int32_t* gi_iters;       // [B][iters] may be NULL
    const double* lambda2 = nullptr;   // [S]    conn_mode: algebraic connectivity of every scenario (cbf_filter.cu)
    const double* fiedler = nullptr;   // [S][N] conn_mode: its normalised eigenvector
    unsigned char* lead_ws = nullptr;  // [B][N-1] slack mode: group leaders (kernel scratch)
    int32_t num_qp;          // B = scenarios * agent_count
    int32_t agent_begin;     // first agent of each scenario handled by this launch
    int32_t agent_count;     // agents per scenario handled by this launch
    int32_t ref_mode;
};

// K3 constants: position / velocity Bernstein rows at t = h inside piece `piece`.
struct RolloutConst {
    int n, c, piece;
    double b0[8], b1[8];
};

// Tables of the general reduced kernel (impc_general.cu): r free dof per axis, any continuity degree / control-point count.
constexpr int kGenMaxR = 8;
struct GenTables {
    int32_t K, r, m1, n, c, p, N, H, iters, max_gi;
    double lim_lo[2][3], lim_hi[2][3];  // [0] acceleration rows, [1] velocity rows
    double dmin2, gamma, feas_tol;
    double Kc[2][2], Kt[2];
    const double* gt;   // [3][K][r + 2]  d = 0 pos, 1 vel, 2 acc: {Ct (r), Ot (2)}; rows independent of v have Ct = 0 exactly
    const double* vs;   // [r][2]
    const double* vt;   // [r]     row sums of vr (constant reference)
    const double* vr;   // [r][K]
    const double* kr;   // [2][K]
    const double* yw;   // [m1][2 + r]  {Y1 (2), W1 (r)}
};
int launch_impc_general(const GenTables& T, const StepArgs& a, int sm_count, cudaStream_t stream);

int launch_impc_reduced(const DevTables& T, const StepArgs& a, int width, int sm_count, cudaStream_t stream);
int launch_impc_thread(const DevTables& T, const StepArgs& a, int sm_count, cudaStream_t stream);
int launch_cbf_rows(const double* states, double* rows, int S, int N, double dmin2, double gamma,
                    int sm_count, cudaStream_t stream);
int launch_rollout(const double* states, const double* ctrl, const int32_t* n_curves, double* next,
                   int B, const RolloutConst& rc, cudaStream_t stream);
int launch_cbf_filter(const mpccbf_config& cfg, int S, const double* states, const double* desired_u, double* u,
                      int32_t* status, double* lambda2, double* conn_rows, int sm_count, cudaStream_t stream,
                      double* fiedler = nullptr, void* wide_ws = nullptr, int wide_blocks = 0);
size_t cbf_filter_wide_ws_bytes(const mpccbf_config& cfg, int blocks);  // workspace of the 33..256-agent kernel
// Episode driver (episode.cu): one control step's worth of curve following + metric update.
struct EpisodeArgs {
    double* states;            // [S][N][6] in: state the step was solved at; out: state after h
    const double* targets;     // [S][N][3]
    const double* ctrl_new;    // [S][N][n] result of this step's optimize
    const int32_t* n_curves;   // [S][N]
    double* curve;             // [S][N][n] curve being followed (persistent)
    double* eval_t;            // [S][N]    its evaluation time (persistent)
    int32_t* has_curve;        // [S][N]
    unsigned char* reached;    // [S][N] goal-area flags
    int32_t* done;             // [S] 0 running, 1 success decided, 2 collision
    int32_t* finish;           // [S]
    double* min_dist2;         // [S]
    int32_t* failed;           // [S] optimize calls that did not return all IMPC iterates (may be NULL)
    double* traj;              // [S][N][traj_len][6] or NULL
    double* sub;               // [S][N][nsub][2] x/y of this control step's sub-steps (scratch)
    int S, N, n, c, p, iters, nsub, step, traj_len;
    int agent_begin, agent_count;  // phase A range; ctrl_new / n_curves are [S][agent_count][..]
    int do_metrics;            // phase B
    int final_only;            // phase A evaluates only the last sub-step (no metrics, trajectory dump or noise wanted)
    double Ts, piece_T, max_param, goal_radius, collision_radius;
    int box;                   // collision_check.py box shape instead of the circle
    double box_x, box_y;
    double pos_std, vel_std;   // Gaussian state noise per sub-step (0: none)
    unsigned long long seed;
    double cum[16];            // cumulative piece ends
};
int launch_episode_init(int S, int N, double* eval_t, int32_t* has_curve, unsigned char* reached, int32_t* done, int32_t* finish,
                        double* min_dist2, int32_t* failed, cudaStream_t stream);
int launch_episode_advance(const EpisodeArgs& a, int sm_count, cudaStream_t stream);
int launch_gather_agent_targets(int S, int N, int agent, const double* targets, double* out, cudaStream_t stream);
int launch_episode_final(int S, int total_substeps, const int32_t* done, const int32_t* finish, const double* min_dist2,
                         int32_t* success, int32_t* finish_step, double* min_dist, cudaStream_t stream);
struct DenseOpts {
    int max_iter, check_every;  // <= 0: defaults (20000, 10)
    double eps, rho, sigma;     // <= 0: defaults (1e-7, 0.1, 1e-6)
    int sm_count;
    cudaStream_t stream;
};
int dense_qp_solve_host(const DenseOpts& opts, int32_t batch, int32_t n, int32_t m, const double* P,
                        const double* q, const double* A, const double* l, const double* u,
                        const double* lb, const double* ub, double* x, double* objective,
                        int32_t* status, int32_t* iters);
bool dense_qp_needs_wsA(int n, int m);
int dense_qp_solve_device(const DenseOpts& opts, int32_t batch, int32_t n, int32_t m, const double* P, const double* q,
                          const double* A, const double* l, const double* u, double* wsP, double* wsA, double* x,
                          double* objective, int32_t* status, int32_t* iters);

void set_error(const std::string& msg);
#ifdef MPCCBF_HOST_EMUL
void emul_impc_reduced_host(const DevTables& T, const StepArgs& a);
void emul_impc_thread_host(const DevTables& T, const StepArgs& a);
void emul_impc_general_host(const GenTables& T, const StepArgs& a);
#endif

}  // namespace mpccbf
