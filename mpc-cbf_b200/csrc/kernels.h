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
    double* ctrl_pts;        // [B][n]            last successful iterate (zeros if none)
    double* ctrl_all;        // [B][iters][n]     every successful iterate, may be NULL
    int32_t* status;         // [B][iters]
    int32_t* n_curves;       // [B]   may be NULL
    double* objective;       // [B]   may be NULL
    int32_t* gi_iters;       // [B][iters] may be NULL
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

int launch_impc_reduced(const DevTables& T, const StepArgs& a, int width, int sm_count, cudaStream_t stream);
int launch_impc_thread(const DevTables& T, const StepArgs& a, int sm_count, cudaStream_t stream);
int launch_cbf_rows(const double* states, double* rows, int S, int N, double dmin2, double gamma,
                    int sm_count, cudaStream_t stream);
int launch_rollout(const double* states, const double* ctrl, const int32_t* n_curves, double* next,
                   int B, const RolloutConst& rc, cudaStream_t stream);
int launch_cbf_filter(const mpccbf_config& cfg, int S, const double* states, const double* desired_u, double* u,
                      int32_t* status, double* lambda2, double* conn_rows, int sm_count, cudaStream_t stream);
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
void set_error(const std::string& msg);
#ifdef MPCCBF_HOST_EMUL
void emul_impc_reduced_host(const DevTables& T, const StepArgs& a);
void emul_impc_thread_host(const DevTables& T, const StepArgs& a);
#endif

}  // namespace mpccbf
