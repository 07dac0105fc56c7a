/* mpccbf.h — C ABI of the B200-native batched MPC-CBF QP path.
 *
 * Drop-in boundary for the reference's per-agent, per-control-step QP build+solve
 * (paths relative to /root/reference/workspace/lib):
 *   S-ctl  bool mpc_cbf::ConnectivityIMPCCBF<double,3>::optimize(result_curves, current_states,
 *          self_idx, ref_positions)        mpc_cbf/include/mpc_cbf/controller/ConnectivityIMPCCBF.h:58-60
 *   S-qp   qpcpp::SolveStatus qpcpp::Solver<double>::solve(qpcpp::Problem<double>&)
 *                                           qpcpp/include/qpcpp/solvers/Solver.h:27-37
 * The reference has no FFI layer; these entry points are what a C++ shim with the two signatures
 * above binds (see INTEGRATION.md and mpc-cbf_b200/host/). Plain pointers and sizes only; no
 * exceptions cross the ABI; solver outcomes are data (status[]), not errors.
 *
 * All arithmetic is FP64 (the reference instantiates <double,3U> only:
 * mpc_cbf/src/controller/ConnectivityIMPCCBF.cpp:222).
 *
 * Error convention: 0 = OK; negative = invalid argument / CUDA error, message available from
 * mpccbf_last_error() (thread-local). There is NO CPU fallback: every compute entry point
 * returns MPCCBF_ERR_CUDA when no sm_100-class device is usable.
 */
#ifndef MPCCBF_H
#define MPCCBF_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPCCBF_OK 0
#define MPCCBF_ERR_INVALID (-1)   /* bad argument / failed config validation            */
#define MPCCBF_ERR_CUDA (-2)      /* CUDA runtime error or no usable device             */
#define MPCCBF_ERR_UNSUPPORTED (-3) /* configuration outside what the kernels implement */
#define MPCCBF_ERR_ALLOC (-4)

/* qpcpp::SolveStatus codes (qpcpp/include/qpcpp/solvers/Solver.h:13-21), same order. */
#define MPCCBF_STATUS_OPTIMAL 0
#define MPCCBF_STATUS_FEASIBLE 1
#define MPCCBF_STATUS_UNBOUNDED 2
#define MPCCBF_STATUS_INFEASIBLE 3
#define MPCCBF_STATUS_ERROR 4
#define MPCCBF_STATUS_UNKNOWN 5
#define MPCCBF_STATUS_INFEASIBLEORUNBOUNDED 6
/* "not attempted": IMPC iteration skipped because an earlier one was not OPTIMAL
 * (ConnectivityIMPCCBF.cpp:203-211 breaks out of the loop). */
#define MPCCBF_STATUS_SKIPPED 7

/* Which solver the controller path uses. */
#define MPCCBF_SOLVER_AUTO 0     /* reduced when the structure allows, else dense           */
#define MPCCBF_SOLVER_REDUCED 1  /* fused null-space active-set kernels: cubic C^3 pieces   */
                                 /* (2 free dof per axis; slack_mode, conn_mode, compact    */
                                 /* results) or any configuration with 1..8 free dof per    */
                                 /* axis (general kernel)                                   */
#define MPCCBF_SOLVER_DENSE 2    /* the equality rows are rank-deficient or leave more than */
                                 /* 8 free dof per axis: only the S-qp seam                 */
                                 /* (mpccbf_solve_qp_dense, ADMM/OSQP kernel) is available  */

/* Configuration: union of the reference's parameter structs
 *   mpc::PiecewiseBezierParams / MPCParams / TuningParams / PhysicalLimits
 *        mpc/include/mpc/optimization/PiecewiseBezierMPCQPOperations.h:17-41
 *   ConnectivityIMPCCBF::IMPCParams   mpc_cbf/include/mpc_cbf/controller/ConnectivityIMPCCBF.h:39-45
 *   cbf::ConnectivityCBFParams {dmin,dmax}; gamma is hard-coded 5.0 at
 *        cbf/src/detail/ConnectivityCBF.cpp:62
 * JSON keys: workspace/experiments/config/config.md:5-55. */
typedef struct mpccbf_config {
    double h;                 /* mpc_params.h   : MPC step                              */
    double Ts;                /* mpc_params.Ts  : control (rollout) step                */
    int32_t k_hor;            /* mpc_params.k_hor                                       */
    int32_t num_pieces;       /* bezier_params.num_pieces                               */
    int32_t num_ctrl_pts;     /* bezier_params.num_control_points (4 = cubic)           */
    int32_t continuity_degree;/* bezier_params.bezier_continuity_upto_degree            */
    double piece_max_param;   /* bezier_params.piece_max_parameter                      */
    double w_pos_err;         /* mpc_tuning.w_pos_err                                   */
    double w_u_eff;           /* mpc_tuning.w_u_eff                                     */
    int32_t spd_f;            /* mpc_tuning.spd_f                                       */
    int32_t cbf_horizon;      /* cbf_params.cbf_horizon                                 */
    double v_min[3], v_max[3], a_min[3], a_max[3]; /* physical_limits                   */
    double d_min, d_max;      /* cbf_params.d_min / d_max                               */
    double cbf_gamma;         /* 5.0 in the reference                                   */
    int32_t impc_iter;        /* cbf_params.impc_iter                                   */
    int32_t slack_mode;       /* cbf_params.slack_mode: one slack >= 0 per neighbour on its
                               * safety rows (ConnectivityIMPCCBF.cpp:76-100), solved inside
                               * the fused active-set kernel with per-neighbour multiplier
                               * caps (no limit on num_agents), see DESIGN.md section 8       */
    double slack_cost;        /* cbf_params.slack_cost                                  */
    double slack_decay;       /* cbf_params.slack_decay_rate                            */
    int32_t num_agents;       /* robots per scenario (tasks.so length)                  */
    int32_t max_scenarios;    /* batch capacity of the handle                           */
    int32_t device;           /* CUDA device ordinal; -1 = host-only handle (tables and
                               * validation only, every compute call returns ERR_CUDA)   */
    int32_t solver;           /* MPCCBF_SOLVER_*                                        */
    double feas_tol;          /* row feasibility tolerance (<=0: default 1e-9)          */
    /* dense ADMM kernel knobs (<=0: defaults) */
    int32_t admm_max_iter;    /* default 4000                                           */
    int32_t admm_check_every; /* default 10                                             */
    double admm_eps;          /* abs = rel termination tolerance, default 1e-7          */
    double admm_rho;          /* initial rho, default 0.1                               */
    double admm_sigma;        /* default 1e-6                                           */
    int32_t conn_mode;        /* 0 (the reference as shipped: the block is commented out,
                               * ConnectivityIMPCCBF.cpp:143-157,177-192) or 1: lambda2 > 0.1 ->
                               * the MPC-lifted connectivity HOCBF row at every CBF horizon step,
                               * else one distance-CLF row per neighbour and step
                               * (ConnectivityMPCCBFQPOperations.cpp:45-86,111-173). Needs
                               * num_agents <= 256 and slack_mode = 0                       */
} mpccbf_config;

typedef struct mpccbf_handle mpccbf_handle;

/* Fills *cfg with workspace/experiments/config/base_config.json (the effective parameters
 * after preprocess.py: K=16, 3 cubic pieces, C^3, d_min 2.0, H=2, 2 IMPC iterations). */
void mpccbf_config_default(mpccbf_config* cfg);

/* Same validation rules (and messages) as common/include/common/parsing.hpp:50-75,121-132,
 * 195-213. Returns MPCCBF_OK or MPCCBF_ERR_INVALID (message in mpccbf_last_error()). */
int mpccbf_config_validate(const mpccbf_config* cfg);

/* Reads a reference experiment JSON (base_config.json schema) into *cfg and, when `so`/`sf` are
 * non-NULL, the tasks' start/goal poses ([num_agents][3], up to cap_agents). Host only. */
int mpccbf_config_from_json(const char* path, mpccbf_config* cfg, double* so, double* sf,
                            int32_t cap_agents);

/* Builds the constant tables (ConnectivityIMPCCBF ctor + PiecewiseBezierMPCQPOperations ctor:
 * ConnectivityIMPCCBF.cpp:9-44, PiecewiseBezierMPCQPOperations.cpp:9-38), uploads them and
 * allocates device workspaces for max_scenarios x num_agents. */
int mpccbf_create(const mpccbf_config* cfg, mpccbf_handle** out);
int mpccbf_destroy(mpccbf_handle* h);

const char* mpccbf_last_error(void);
const char* mpccbf_version(void);

/* Decision-vector length n = num_pieces*3*num_ctrl_pts, layout x[piece*3c + axis*c + cp]
 * (BezierQPOperations.cpp:37, PiecewiseBezierMPCQPGenerator.cpp:9-34). */
int32_t mpccbf_num_decision_vars(const mpccbf_handle* h);
/* Which solver the controller path resolved to: MPCCBF_SOLVER_REDUCED (fused active-set kernels, also in
 * slack mode) or MPCCBF_SOLVER_DENSE (the configuration does not reduce: only mpccbf_solve_qp_dense is
 * available and the controller entries return MPCCBF_ERR_UNSUPPORTED). */
int32_t mpccbf_active_solver(const mpccbf_handle* h);

/* Batched control step == ConnectivityIMPCCBF::optimize for every agent of every scenario
 * (Jacobi: all agents see the same state snapshot).
 *   states   [S][N][6]  (pos3, vel3)            device pointer
 *   targets  [S][N][3]  goal, replicated over the horizon like
 *                       MPCCBFFormationControl_example.cpp:144        device pointer
 *   ctrl_pts [S][N][n]  control points of the LAST successful IMPC iterate (what the caller of
 *                       optimize uses: trajs.back()); zero-filled if none   device pointer
 *   status   [S][N][impc_iter]  SolveStatus per IMPC iteration (SKIPPED after a failure)
 *   n_curves [S][N]     number of curves optimize would have pushed (may be NULL)
 *   objective[S][N]     x'Px+q'x of the last successful iterate (may be NULL)
 * Asynchronous on `cuda_stream` (a cudaStream_t, NULL = default stream). */
int mpccbf_step_batch(mpccbf_handle* h, int32_t num_scenarios, const double* states,
                      const double* targets, double* ctrl_pts, int32_t* status, int32_t* n_curves,
                      double* objective, void* cuda_stream);

/* Same, HOST pointers: copies in, runs, copies out, synchronises (this is the call the
 * reference-facing plugin makes and what bench.py times as `e2e`). */
int mpccbf_step_batch_host(mpccbf_handle* h, int32_t num_scenarios, const double* states,
                           const double* targets, double* ctrl_pts, int32_t* status,
                           int32_t* n_curves, double* objective);

/* Split form of the call above for callers that pipeline batches: _async returns once the copies and
 * the launch are enqueued on the handle's own streams (pageable buffers may make it block), the
 * buffers must stay untouched until mpccbf_synchronize(h) returns. Two handles used alternately
 * overlap the PCIe phase of one batch with the solve of the next (bench.py's e2e leg does). */
int mpccbf_step_batch_host_async(mpccbf_handle* h, int32_t num_scenarios, const double* states,
                                 const double* targets, double* ctrl_pts, int32_t* status,
                                 int32_t* n_curves, double* objective);
int mpccbf_synchronize(mpccbf_handle* h);

/* ---- Compact result form ------------------------------------------------------------------------------
 * With C^3 cubic pieces the n control points of a curve are an exact linear image of the agent's own
 * state and 6 numbers: x[piece, axis, cp] = Y [pos; vel] + W v, v = the whitened free dof (2 per axis).
 * These entries return v_free [S][N][6] (zeros where n_curves == 0) instead of ctrl_pts [S][N][n]: 48
 * instead of 8 n bytes per agent cross PCIe, which is what bounds the host-pointer calls. The caller of
 * ConnectivityIMPCCBF::optimize only evaluates trajs.back() at h / Ts sub-steps
 * (MPCCBFFormationControl_example.cpp:191-205): mpccbf_eval_compact does that straight from (state, v),
 * mpccbf_expand_ctrl_pts rebuilds the control points (bit-identical to mpccbf_step_batch's) on the host.
 * Only for handles whose solver is MPCCBF_SOLVER_REDUCED. */
int mpccbf_step_batch_compact(mpccbf_handle* h, int32_t num_scenarios, const double* states,
                              const double* targets, double* v_free, int32_t* status, int32_t* n_curves,
                              double* objective, void* cuda_stream);                 /* device pointers */
int mpccbf_step_batch_host_compact(mpccbf_handle* h, int32_t num_scenarios, const double* states,
                                   const double* targets, double* v_free, int32_t* status,
                                   int32_t* n_curves, double* objective);            /* host pointers   */
int mpccbf_step_batch_host_compact_async(mpccbf_handle* h, int32_t num_scenarios, const double* states,
                                         const double* targets, double* v_free, int32_t* status,
                                         int32_t* n_curves, double* objective);
/* Host only. states [count][6], v_free [count][6], n_curves [count] or NULL, ctrl_pts [count][n]. */
int mpccbf_expand_ctrl_pts(const mpccbf_handle* h, int64_t count, const double* states,
                           const double* v_free, const int32_t* n_curves, double* ctrl_pts);
/* Host only. out [count][3] = deriv-th derivative of every agent's curve at time t, with the piece lookup
 * and clamping of SingleParameterPiecewiseCurve::eval (splines/src/curves/SingleParameterPiecewiseCurve.cpp:94-127). */
int mpccbf_eval_compact(const mpccbf_handle* h, int64_t count, const double* states, const double* v_free,
                        double t, int32_t deriv, double* out);

/* Single-agent drop-in for S-ctl (host pointers): states [N][6], ref [3*k_hor] (only used as a
 * per-step reference: ref[3k..3k+2] is the target at horizon step k), ctrl_pts [impc_iter][n]
 * receives EVERY successful iterate in order (result_curves), *n_curves how many,
 * status [impc_iter]. Returns MPCCBF_OK even when the QP is infeasible (see status). */
int mpccbf_optimize_one(mpccbf_handle* h, const double* states, int32_t self_idx,
                        const double* ref, double* ctrl_pts, int32_t* n_curves, int32_t* status);

/* S-qp: batch of `batch` dense QPs of identical shape, reference objective convention
 *      min x'Px + q'x   s.t.  l <= A x <= u ,  lb <= x <= ub        (|bound| >= 1e20: none)
 * P [batch][n][n], q [batch][n], A [batch][m][n], l,u [batch][m], lb,ub [batch][n] or NULL,
 * x [batch][n], objective [batch] (may be NULL), status [batch], iters [batch] (may be NULL).
 * HOST pointers; one CTA per QP on the device (ADMM/OSQP iteration with a shared-memory
 * Cholesky of P~+sigma I+rho A'A, followed by an active-set polish). */
int mpccbf_solve_qp_dense(mpccbf_handle* h, int32_t batch, int32_t n, int32_t m, const double* P,
                          const double* q, const double* A, const double* l, const double* u,
                          const double* lb, const double* ub, double* x, double* objective,
                          int32_t* status, int32_t* iters);

/* Allocation-free variant of mpccbf_solve_qp_dense. It has no counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it). Same kernel, same results; what differs:
 *   - P, q, A, l, u, x, objective, status, iters are DEVICE pointers on the CURRENT device (all required except
 *     A, l, u when m == 0); variable bounds are passed as rows of A; no handle: cfg (may be NULL = defaults)
 *     only supplies admm_max_iter / admm_check_every / admm_eps / admm_rho / admm_sigma;
 *   - asynchronous on cuda_stream (a cudaStream_t, NULL = default stream): one memset + one kernel launch, no
 *     cudaMalloc / cudaFree / host synchronisation;
 *   - the kernel's scratch (scaled P~ [batch][n*n]; scaled A [batch][m][n|1] when it does not fit in shared
 *     memory) is a file-scope device array of MPCCBF_NOALLOC_SCRATCH_BYTES that every GPU which loads the library
 *     holds until exit. A call needing more returns MPCCBF_ERR_UNSUPPORTED (message in mpccbf_last_error()).
 *     Every call on a device shares the array: calls that overlap on the GPU (different streams, or threads) race. */
#define MPCCBF_NOALLOC_SCRATCH_BYTES (8u << 20)
int mpccbf_solve_qp_dense_noalloc(const mpccbf_config* cfg, int32_t batch, int32_t n, int32_t m, const double* P,
                                  const double* q, const double* A, const double* l, const double* u, double* x,
                                  double* objective, int32_t* status, int32_t* iters, void* cuda_stream);

/* Sub-kernel K1: pairwise safety HOCBF rows for every ordered (ego, neighbour) pair at the
 * CURRENT states (iter-0 rows, ConnectivityIMPCCBF.cpp:135-141; closed form of
 * ConnectivityCBF.cpp:152-198). rows [S][N][N-1][4] = (ax, ay, az=0, b) meaning a.u + b >= 0,
 * neighbour order = ascending agent index skipping ego. Device pointers, async on stream. */
int mpccbf_cbf_rows(mpccbf_handle* h, int32_t num_scenarios, const double* states, double* rows,
                    void* cuda_stream);

/* ---- CBF-only safety / connectivity filter ------------------------------
 * == cbf::ConnectivityControl<double,3>::optimize(cbf_u, desired_u, current_states, self_idx, u_min,
 * u_max)  cbf/include/cbf/controller/ConnectivityControl.h:16-60, cbf/src/controller/
 * ConnectivityControl.cpp:22-99 (slack_mode = false) for EVERY agent of every scenario (Jacobi):
 *      min |u - desired_u|^2   s.t.  pairwise safety HOCBF rows, velocity CBF rows (cfg v_min/v_max),
 *      the lambda2-connectivity HOCBF row when lambda2 > 0.1, else one distance-CLF row per neighbour
 * (ConnectivityQPGenerator.cpp:13-150; the u_min/u_max box is commented out in the reference, :60).
 *   states    [S][N][6]   desired_u [S][N][3]
 *   u         [S][N][3]   filtered input; zeros where the QP is not OPTIMAL (the reference example
 *                         falls back to zero control: CBFFormationControl_example.cpp:168-172)
 *   status    [S][N]      SolveStatus;  lambda2 [S] algebraic connectivity (may be NULL)
 * num_agents <= 32: one warp per scenario, matrices of the eigen-solve in shared memory; 33 .. 256: one CTA per scenario,
 * matrices in a handle-owned workspace allocated on the first such call; beyond: MPCCBF_ERR_UNSUPPORTED. Uses cfg d_min, d_max,
 * cbf_gamma, v_min, v_max, feas_tol, and slack_mode / slack_cost / slack_decay: in slack mode
 * (ConnectivityControl.cpp:30-45) slack i (cost slack_cost * slack_decay^i) relaxes safety row i and CLF row i,
 * the last slack the connectivity row; the velocity rows stay hard. Device pointers, async on `cuda_stream`. */
int mpccbf_cbf_filter_batch(mpccbf_handle* h, int32_t num_scenarios, const double* states,
                            const double* desired_u, double* u, int32_t* status, double* lambda2,
                            void* cuda_stream);
/* Same with HOST pointers (copies in, runs, copies out, synchronises). */
int mpccbf_cbf_filter_batch_host(mpccbf_handle* h, int32_t num_scenarios, const double* states,
                                 const double* desired_u, double* u, int32_t* status, double* lambda2);

/* Sub-kernel: lambda2 and the connectivity HOCBF row of every agent ==
 * ConnectivityCBF::getLambda2 / initConnCBF / getConnConstraints / getConnBound
 * (cbf/src/detail/ConnectivityCBF.cpp:388-424,472-552). rows [S][N][4] = (ax, ay, az = 0, b) meaning
 * a.u + b >= 0; lambda2 [S]. Device pointers, async on stream. num_agents <= 256. */
int mpccbf_conn_rows(mpccbf_handle* h, int32_t num_scenarios, const double* states, double* rows,
                     double* lambda2, void* cuda_stream);

/* ---- Device-resident closed-loop episodes ----------------------------------------
 * Runs `num_control_steps` control steps of every scenario without leaving the device: batched
 * optimize (== mpccbf_step_batch), then each agent follows its newest successful curve for h / Ts
 * sub-steps exactly like the reference's experiment driver (noise-free, Jacobi order)
 *      mpc_cbf/examples/connectivity/MPCCBFFormationControl_example.cpp:135-231
 * (an agent whose optimize fails keeps following its previous curve, the parameter clamped to
 * max_parameter; one that never had a curve stays put with zero velocity), and the success metric of
 *      workspace/experiments/python/metrics/collision_check.py:48-79  (instance_success; circle shape
 *      here, the box shape through mpccbf_run_episode_ex)
 * is updated at every sub-step.
 *   states       [S][N][6]  in: start states, out: final states            device pointer
 *   targets      [S][N][3]  goals (tasks.sf)                               device pointer
 *   goal_radius             reach_goal_area radius (1.0 in the reference)
 *   collision_radius        robot_params.collision_shape.radius: overlap when dist <= 2 r
 *   success      [S]  1 unless two agents overlapped before every agent had reached its goal area
 *   finish_step  [S]  the sub-step index instance_success returns (number of sub-steps when the goals
 *                     were not all reached; -1 for its float('inf'))
 *   min_dist     [S]  smallest pairwise centre distance over all sub-steps
 *   failed_calls [S]  optimize calls that returned fewer than impc_iter curves
 *   traj         NULL or [S][N][num_control_steps * (h/Ts)][6]: every sub-step state (states.json)
 * Output pointers may be NULL individually. Device pointers; the call is asynchronous on
 * `cuda_stream` except for its first use on a handle (workspace allocation). */
int mpccbf_run_episode(mpccbf_handle* h, int32_t num_scenarios, double* states, const double* targets,
                       int32_t num_control_steps, double goal_radius, double collision_radius,
                       int32_t* success, int32_t* finish_step, double* min_dist, int32_t* failed_calls,
                       double* traj, void* cuda_stream);

/* Options of mpccbf_run_episode_ex. */
#define MPCCBF_ORDER_JACOBI 0        /* every agent optimises on the same snapshot (one launch per step)   */
#define MPCCBF_ORDER_GAUSS_SEIDEL 1  /* the reference's loop order (MPCCBFFormationControl_example.cpp:140):
                                        agent i optimises and advances before agent i + 1 optimises        */
#define MPCCBF_SHAPE_CIRCLE 0
#define MPCCBF_SHAPE_BOX 1
typedef struct mpccbf_episode_opts {
    double goal_radius;       /* reach_goal_area radius                                               */
    double collision_radius;  /* robot_params.collision_shape.radius                                   */
    int32_t order;            /* MPCCBF_ORDER_*                                                        */
    double pos_std, vel_std;  /* mpc_params-level state noise of the driver (math::addRandomNoise,
                                 math/src/Random.cpp:7-29: N(0, std) on every position / velocity
                                 component at every sub-step); 0 = noise-free                          */
    uint64_t seed;            /* the reference draws from random_device; here the noise is a
                                 counter-based stream of (seed, scenario, agent, sub-step)             */
    int32_t collision_shape;  /* MPCCBF_SHAPE_CIRCLE (dist <= 2 r) or MPCCBF_SHAPE_BOX: the axis-aligned
                                 rectangle test of collision_check.py:10-39 with its offsets (corner at
                                 centre - half/2, width 2*half)                                        */
    double box_half[2];       /* robot_params.collision_shape.aligned_box[0..1]                        */
} mpccbf_episode_opts;

/* Fills goal_radius (1, the metric's default), collision_radius / collision_shape / box_half from
 * robot_params.collision_shape of a reference experiment JSON with the metric's own precedence
 * (collision_check.py:93-101: "aligned_box" first, else "radius"); order Jacobi, no noise. Host only. */
int mpccbf_episode_opts_from_json(const char* path, mpccbf_episode_opts* opts);

/* Same as mpccbf_run_episode with the loop order, the state noise and the collision shape of the
 * reference driver / metric. */
int mpccbf_run_episode_ex(mpccbf_handle* h, int32_t num_scenarios, double* states, const double* targets,
                          int32_t num_control_steps, const mpccbf_episode_opts* opts, int32_t* success,
                          int32_t* finish_step, double* min_dist, int32_t* failed_calls, double* traj,
                          void* cuda_stream);

/* Sub-kernel K3: advance every agent along its curve by one MPC step, i.e. the state after the
 * h/Ts rollout sub-steps of MPCCBFFormationControl_example.cpp:191-205 without noise:
 * next[S][N][6] = (curve(h), curve'(h)); agents whose n_curves == 0 keep their position with zero
 * velocity (:208-221). n_curves may be NULL (all agents have a curve). next may alias states.
 * Device pointers, async on stream. */
int mpccbf_rollout(mpccbf_handle* h, int32_t num_scenarios, const double* states,
                   const double* ctrl_pts, const int32_t* n_curves, double* next_states,
                   void* cuda_stream);

/* Same for an agent shard: states / ctrl_pts / n_curves / next_states are indexed
 * [S][agent_count][...] (the shard's own agents only). */
int mpccbf_rollout_agents(mpccbf_handle* h, int32_t num_scenarios, int32_t agent_count,
                          const double* states, const double* ctrl_pts, const int32_t* n_curves,
                          double* next_states, void* cuda_stream);

/* Agent-sharded control step (one scenario's agents split across GPUs):
 * solves only agents [agent_begin, agent_begin + agent_count) of every scenario. states is the
 * FULL snapshot [S][N][6] (after the all-gather); targets/ctrl_pts/status/n_curves/objective are
 * indexed [S][agent_count][...]. Device pointers, async on stream. */
int mpccbf_step_agents(mpccbf_handle* h, int32_t num_scenarios, int32_t agent_begin,
                       int32_t agent_count, const double* states, const double* targets,
                       double* ctrl_pts, int32_t* status, int32_t* n_curves, double* objective,
                       void* cuda_stream);

/* Curve following of the reference driver for an agent shard, stateful per handle (curve being followed,
 * its evaluation time, "has a curve" flag per agent): agents [agent_begin, agent_begin + agent_count) adopt
 * their new curve where n_curves > 0 (evaluation time back to 0), then advance h / Ts sub-steps of Ts along
 * the curve they follow, the parameter clamped to max_parameter; an agent that never had a curve keeps its
 * position with zero velocity (MPCCBFFormationControl_example.cpp:161-221). Their rows of `states`
 * ([S][N][6], the FULL snapshot) are updated in place; ctrl_new / n_curves are [S][agent_count][..].
 * mpccbf_follow_reset forgets every curve. Device pointers, async on stream. */
int mpccbf_follow_reset(mpccbf_handle* h, int32_t num_scenarios, void* cuda_stream);
int mpccbf_follow_step(mpccbf_handle* h, int32_t num_scenarios, int32_t agent_begin, int32_t agent_count,
                       double* states, const double* ctrl_new, const int32_t* n_curves, void* cuda_stream);

/* Which fused kernel the controller entries run: -1 = by batch size (default), 0 = one thread per agent,
 * 8 / 32 = that many lanes per agent. mpccbf_auto_kernel_variant returns what -1 resolves to for a step of
 * `num_agent_qps` agents. (One agent's arithmetic does not depend on the batch it is solved in, only on the
 * variant: pinning it makes sharded and unsharded runs agree bit for bit.) */
int mpccbf_set_kernel_variant(mpccbf_handle* h, int32_t lanes);
int32_t mpccbf_auto_kernel_variant(const mpccbf_handle* h, int64_t num_agent_qps);

/* ---- Multi-GPU behind the ABI: one process, one host thread + stream per device ------------
 * mpccbf_multi_create builds one handle per device (devices == NULL: ordinals 0..num_devices-1; cfg->device is
 * ignored, cfg->max_scenarios is PER DEVICE) and, for more than one device, an NCCL communicator per device
 * (ncclCommInitAll; libnccl.so.2 is loaded on demand). The reference's only fan-out is a shell loop over
 * processes (experiments/bash/ScaleScript.sh:17-34). */
typedef struct mpccbf_multi mpccbf_multi;
int mpccbf_multi_create(const mpccbf_config* cfg, int32_t num_devices, const int32_t* devices, mpccbf_multi** out);
int mpccbf_multi_destroy(mpccbf_multi* m);
int32_t mpccbf_multi_num_devices(const mpccbf_multi* m);
int32_t mpccbf_multi_comm_ranks(const mpccbf_multi* m);      /* ranks the NCCL communicator reports (1 without NCCL) */
mpccbf_handle* mpccbf_multi_handle(mpccbf_multi* m, int32_t index);  /* the per-device handle (owned by m)   */
/* Scenario sharding: same contract as mpccbf_step_batch_host / _host_compact (HOST pointers, pinned for
 * overlap), scenarios split contiguously across the devices, no collective. */
int mpccbf_multi_step_batch_host(mpccbf_multi* m, int32_t num_scenarios, const double* states, const double* targets,
                                 double* ctrl_pts, int32_t* status, int32_t* n_curves, double* objective);
int mpccbf_multi_step_batch_host_compact(mpccbf_multi* m, int32_t num_scenarios, const double* states,
                                         const double* targets, double* v_free, int32_t* status, int32_t* n_curves,
                                         double* objective);
/* Agent sharding (BASELINE configs[3]): `num_control_steps` closed-loop control steps of num_scenarios swarms whose
 * num_agents agents are split evenly across the devices: per step and device mpccbf_step_agents on the full
 * snapshot -> mpccbf_follow_step -> ONE ncclAllGather of the [S][N/G][6] new states. states [S][N][6] HOST, in:
 * start, out: final; targets [S][N][3] HOST. *ms_per_step (may be NULL): device time per control step, max over
 * the devices. The result does not depend on the number of devices (bit for bit). */
int mpccbf_multi_run_sharded(mpccbf_multi* m, int32_t num_scenarios, double* states, const double* targets,
                             int32_t num_control_steps, double* ms_per_step);

/* Active-set iteration counts of the last mpccbf_step_batch / _agents call on this handle:
 * iters [S][agents][impc_iter] (device pointer, valid until the next call), for flop accounting. */
const int32_t* mpccbf_last_iteration_counts(const mpccbf_handle* h);
/* Copies the first `count` of them to a HOST buffer (synchronises the handle's stream and the
 * given one is the caller's business: call after the step has completed). */
int mpccbf_copy_iteration_counts(mpccbf_handle* h, int32_t* out_host, int64_t count);

/* Measures the FP64 FMA throughput of the device with a register-resident DFMA loop (the
 * roofline denominator of the solve kernel). Returns TFLOP/s, or a negative error code. */
double mpccbf_measure_fp64_tflops(mpccbf_handle* h);

/* Constant tables for tests / integrators (host copies, row-major). name is one of
 * "U_basis" [3K][n], "V_basis" [3K][n], "P" [n][n], "E" [me][n] (equality rows: eval pos/vel at 0, continuity),
 * "A_fix" [6K][n] (acc rows then vel rows), "h_samples" [K], "Z" [n][r], "Y" [n][6],
 * "H_red" [r][r]. Returns the number of doubles written, or the required count when
 * out == NULL, negative on error. */
int64_t mpccbf_get_table(const mpccbf_handle* h, const char* name, double* out, int64_t cap);

/* Linear cost q [n] of one agent, q = Phi'Q(2 A0 x0 - 2 ref) (positionErrorPenaltyCost,
 * mpc/src/optimization/PiecewiseBezierMPCQPOperations.cpp:64-91): x0 [6], ref [3K]. Host only;
 * with table "P" this is everything an integrator needs to assemble the reference's QP. */
int mpccbf_linear_cost(const mpccbf_handle* h, const double* x0, const double* ref, double* q);

/* Pinned host memory for the host-pointer entry points (cudaHostAlloc / cudaFreeHost), so a
 * C++ caller without the CUDA runtime can give mpccbf_step_batch_host DMA-able buffers. */
void* mpccbf_host_alloc(size_t bytes);
void mpccbf_host_free(void* p);

/* Round 1 solved slack-mode QPs in two tiers and reported here how many agent-steps took the second (ADMM)
 * one. The slack caps now live in the active-set kernel: always 0 (-1 for a NULL handle). Kept so that
 * callers built against the round-1 header keep linking. */
int64_t mpccbf_slack_fallback_count(const mpccbf_handle* h);

/* Counters since create (may be NULL individually). */
int mpccbf_get_counters(const mpccbf_handle* h, int64_t* kernel_launches, int64_t* qps_solved);

#ifdef __cplusplus
}
#endif
#endif /* MPCCBF_H */
