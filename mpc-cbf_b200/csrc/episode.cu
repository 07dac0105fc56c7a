// Device-resident closed-loop episode driver — sm_100a.
//
// Replaces the per-control-step bookkeeping of the reference's experiment driver
// (mpc_cbf/examples/connectivity/MPCCBFFormationControl_example.cpp:135-231, noise-free, Jacobi
// order) and the success metric of workspace/experiments/python/metrics/collision_check.py:48-79
// for S scenarios at once: after every batched optimize, each agent
//   * adopts the newest successful curve and resets its evaluation time (:161-165), or keeps
//     following the previous one (or stays put with zero velocity when it never had one, :208-221),
//   * advances h / Ts sub-steps along the curve, the parameter clamped to max_parameter (:191-205),
// and the scenario's metric is updated at every sub-step exactly as instance_success does: goal-area
// flags first, then pairwise circle overlap; the episode is a success at the first sub-step at which
// every agent has already reached its goal area, a failure at the first overlap before that.
// One warp per scenario; positions of the sub-step are exchanged through shared memory.
#include <cuda_runtime.h>

#include "../../include/mpccbf.h"
#include "kernels.h"

namespace mpccbf {

namespace {

// pos / vel of one axis of a piecewise Bezier at parameter t (splines/src/curves/
// SingleParameterPiecewiseCurve.cpp:94-127 piece lookup, Bezier.cpp:61-75 evaluation): de Casteljau
// on the control points and on their differences.
__device__ __forceinline__ void eval_axis(const double* cp, int c, double s, double inv_T, double& pos, double& vel) {
    double b[8], d[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        b[i] = (i < c) ? cp[i] : 0.0;
        d[i] = 0.0;
    }
#pragma unroll
    for (int i = 0; i < 7; ++i)
        if (i < c - 1) d[i] = b[i + 1] - b[i];
#pragma unroll
    for (int r = 1; r < 8; ++r)
        if (r < c) {
#pragma unroll
            for (int i = 0; i < 7; ++i)
                if (i < c - r) b[i] = fma(s, b[i + 1] - b[i], b[i]);
        }
#pragma unroll
    for (int r = 1; r < 7; ++r)
        if (r < c - 1) {
#pragma unroll
            for (int i = 0; i < 6; ++i)
                if (i < c - 1 - r) d[i] = fma(s, d[i + 1] - d[i], d[i]);
        }
    pos = b[0];
    vel = (c > 1) ? d[0] * (c - 1) * inv_T : 0.0;
}

__global__ void __launch_bounds__(128) episode_advance_kernel(const EpisodeArgs a) {
    extern __shared__ __align__(16) unsigned char esm[];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int N = a.N, n = a.n, c = a.c;
    const size_t per_warp = static_cast<size_t>(N) * 16 + ((N + 15) / 16) * 16;
    double* spos = reinterpret_cast<double*>(esm + wid * per_warp);
    unsigned char* sreach = reinterpret_cast<unsigned char*>(spos + 2 * N);
    const int warps = (gridDim.x * blockDim.x) >> 5;
    const double inv_T = 1.0 / a.piece_T;
    const double coll2 = 4.0 * a.collision_radius * a.collision_radius;

    for (int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < a.S; s += warps) {
        const size_t b0 = static_cast<size_t>(s) * N;
        // adopt new curves
        for (int ag = lane; ag < N; ag += 32) {
            const size_t g = b0 + ag;
            sreach[ag] = a.reached[g];
            if (a.n_curves[g] > 0) {
                for (int i = 0; i < n; ++i) a.curve[g * n + i] = a.ctrl_new[g * n + i];
                a.eval_t[g] = 0.0;
                a.has_curve[g] = 1;
            } else if (a.failed) {
                // counted once per scenario below
            }
        }
        int nfail = 0;
        for (int ag = lane; ag < N; ag += 32) nfail += (a.n_curves[b0 + ag] < a.iters) ? 1 : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) nfail += __shfl_xor_sync(full, nfail, o);
        if (lane == 0 && a.failed) a.failed[s] += nfail;
        int done = a.done[s];          // 0 running, 1 success, 2 collision
        int finish = a.finish[s];
        double mind2 = a.min_dist2[s];
        __syncwarp();
        for (int ti = 1; ti <= a.nsub; ++ti) {
            const int tglob = a.step * a.nsub + (ti - 1);
            // instance_success: "if all(reach_goals): return True, max(0, t - 1)" precedes the updates
            if (!done) {
                int all = 1;
                for (int ag = lane; ag < N; ag += 32) all &= sreach[ag];
                all = __all_sync(full, all);
                if (all) {
                    done = 1;
                    finish = max(0, tglob - 1);
                }
            }
            for (int ag = lane; ag < N; ag += 32) {
                const size_t g = b0 + ag;
                double st6[6];
                if (a.has_curve[g]) {
                    double t = a.eval_t[g] + a.Ts * ti;
                    if (t > a.max_param) t = a.max_param;
                    int piece = 0;
                    while (piece < a.p - 1 && a.cum[piece] < t) ++piece;  // first cumulative end >= t
                    const double local = fmin(a.piece_T, t - (piece ? a.cum[piece - 1] : 0.0));
                    const double* cp = a.curve + g * n + piece * 3 * c;
#pragma unroll
                    for (int ax = 0; ax < 3; ++ax) eval_axis(cp + ax * c, c, local * inv_T, inv_T, st6[ax], st6[3 + ax]);
                    if (ti == a.nsub) a.eval_t[g] = t;
                } else {
#pragma unroll
                    for (int ax = 0; ax < 3; ++ax) {
                        st6[ax] = a.states[g * 6 + ax];
                        st6[3 + ax] = 0.0;
                    }
                }
                spos[2 * ag] = st6[0];
                spos[2 * ag + 1] = st6[1];
                if (ti == a.nsub)
#pragma unroll
                    for (int i = 0; i < 6; ++i) a.states[g * 6 + i] = st6[i];
                if (a.traj)
#pragma unroll
                    for (int i = 0; i < 6; ++i) a.traj[(g * a.traj_len + tglob) * 6 + i] = st6[i];
                const double gx = st6[0] - a.targets[g * 3], gy = st6[1] - a.targets[g * 3 + 1];
                if (gx * gx + gy * gy <= a.goal_radius * a.goal_radius) sreach[ag] = 1;
            }
            __syncwarp();
            int hit = 0;
            for (int i = lane; i < N; i += 32)
                for (int j = i + 1; j < N; ++j) {
                    const double dx = spos[2 * i] - spos[2 * j], dy = spos[2 * i + 1] - spos[2 * j + 1];
                    const double d2 = dx * dx + dy * dy;
                    mind2 = fmin(mind2, d2);
                    hit |= (d2 <= coll2) ? 1 : 0;
                }
            hit = __any_sync(full, hit);
            if (!done && hit) {
                done = 2;
                finish = -1;
            }
            __syncwarp();
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mind2 = fmin(mind2, __shfl_xor_sync(full, mind2, o));
        for (int ag = lane; ag < N; ag += 32) a.reached[b0 + ag] = sreach[ag];
        if (lane == 0) {
            a.done[s] = done;
            a.finish[s] = finish;
            a.min_dist2[s] = mind2;
        }
        __syncwarp();
    }
}

__global__ void episode_init_kernel(int S, int N, double* eval_t, int32_t* has_curve, unsigned char* reached, int32_t* done,
                                    int32_t* finish, double* min_dist2, int32_t* failed) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < S * N) {
        eval_t[i] = 0.0;
        has_curve[i] = 0;
        reached[i] = 0;
    }
    if (i < S) {
        done[i] = 0;
        finish[i] = 0;
        min_dist2[i] = HUGE_VAL;
        failed[i] = 0;
    }
}

// success [S] (1 / 0), finish_step [S] (sub-step index as instance_success returns it; -1 = inf),
// min_dist [S] from the running state
__global__ void episode_final_kernel(int S, int total_substeps, const int32_t* done, const int32_t* finish, const double* min_dist2,
                                     int32_t* success, int32_t* finish_step, double* min_dist) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= S) return;
    const int d = done[i];
    if (success) success[i] = (d != 2) ? 1 : 0;
    if (finish_step) finish_step[i] = (d == 0) ? total_substeps : finish[i];
    if (min_dist) min_dist[i] = sqrt(min_dist2[i]);
}

}  // namespace

int launch_episode_init(int S, int N, double* eval_t, int32_t* has_curve, unsigned char* reached, int32_t* done, int32_t* finish,
                        double* min_dist2, int32_t* failed, cudaStream_t stream) {
    const int tot = S * N;
    if (tot <= 0) return cudaSuccess;
    episode_init_kernel<<<(tot + 255) / 256, 256, 0, stream>>>(S, N, eval_t, has_curve, reached, done, finish, min_dist2, failed);
    return cudaGetLastError();
}

int launch_episode_advance(const EpisodeArgs& a, int sm_count, cudaStream_t stream) {
    if (a.S <= 0) return cudaSuccess;
    if (a.c > 8 || a.p > 16) return cudaErrorInvalidValue;
    const int threads = 128;
    const size_t per_warp = static_cast<size_t>(a.N) * 16 + ((a.N + 15) / 16) * 16;
    const size_t smem = per_warp * (threads / 32);
    if (smem > 200 * 1024) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(episode_advance_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    const int need = (a.S + threads / 32 - 1) / (threads / 32);
    int grid = sm_count * 4;
    if (need < grid) grid = need;
    episode_advance_kernel<<<grid, threads, smem, stream>>>(a);
    return cudaGetLastError();
}

int launch_episode_final(int S, int total_substeps, const int32_t* done, const int32_t* finish, const double* min_dist2,
                         int32_t* success, int32_t* finish_step, double* min_dist, cudaStream_t stream) {
    if (S <= 0) return cudaSuccess;
    episode_final_kernel<<<(S + 255) / 256, 256, 0, stream>>>(S, total_substeps, done, finish, min_dist2, success, finish_step, min_dist);
    return cudaGetLastError();
}

}  // namespace mpccbf
