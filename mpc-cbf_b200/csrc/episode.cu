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
// One warp per scenario; the sub-step positions of a control step pass through a small global scratch
// (L2-resident) so that the Gauss-Seidel order of the reference (one agent optimised and advanced at a
// time) and the Jacobi order share the kernel. Optional seeded Gaussian state noise replaces the
// reference's random_device noise (math/src/Random.cpp:7-29).
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

// counter-based normal deviates: splitmix64 of (seed, stream index) -> two uniforms -> Box-Muller
__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__device__ __forceinline__ void normal_pair(unsigned long long seed, unsigned long long idx, double& n0, double& n1) {
    const unsigned long long a = splitmix64(seed ^ (idx * 0xD1342543DE82EF95ull)), b = splitmix64(a + idx);
    const double u0 = (static_cast<double>(a >> 11) + 1.0) * (1.0 / 9007199254740993.0);  // (0, 1)
    const double u1 = static_cast<double>(b >> 11) * (1.0 / 9007199254740992.0);          // [0, 1)
    const double r = sqrt(-2.0 * log(u0));
    double sn, cs;
    sincospi(2.0 * u1, &sn, &cs);
    n0 = r * cs;
    n1 = r * sn;
}

// One warp per scenario. Phase A (a.agent_count > 0): the agents [agent_begin, agent_begin + agent_count)
// adopt their new curves and advance a.nsub sub-steps; the sub-step positions go to a.sub. Phase B
// (a.do_metrics): goal flags / overlap / minimum distance of the whole scenario from a.sub. Jacobi order
// runs A (all agents) + B in one launch; Gauss-Seidel order runs A once per agent and B at the end of
// the control step.
__global__ void __launch_bounds__(128) episode_advance_kernel(const EpisodeArgs a) {
    extern __shared__ __align__(16) unsigned char esm[];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int N = a.N, n = a.n, c = a.c;
    unsigned char* sreach = esm + static_cast<size_t>(wid) * (((N + 15) / 16) * 16);
    const int warps = (gridDim.x * blockDim.x) >> 5;
    const double inv_T = 1.0 / a.piece_T;
    const double coll2 = 4.0 * a.collision_radius * a.collision_radius;

    for (int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < a.S; s += warps) {
        const size_t b0 = static_cast<size_t>(s) * N;
        // ---- phase A: curve following ---------------------------------------------------------------
        int nfail = 0;
        for (int k = lane; k < a.agent_count; k += 32) {
            const int ag = a.agent_begin + k;
            const size_t g = b0 + ag;                                   // index into the [S][N] arrays
            const size_t gc = static_cast<size_t>(s) * a.agent_count + k;  // index into this launch's results
            const int ncv = a.n_curves[gc];
            nfail += (ncv < a.iters) ? 1 : 0;
            if (ncv > 0) {
                for (int i = 0; i < n; ++i) a.curve[g * n + i] = a.ctrl_new[gc * n + i];
                a.eval_t[g] = 0.0;
                a.has_curve[g] = 1;
            }
            const bool has = a.has_curve[g] != 0;
            const double t0 = a.eval_t[g];
            double hold[3] = {a.states[g * 6], a.states[g * 6 + 1], a.states[g * 6 + 2]};
            for (int ti = a.final_only ? a.nsub : 1; ti <= a.nsub; ++ti) {
                const int tglob = a.step * a.nsub + (ti - 1);
                double st6[6];
                if (has) {
                    double t = t0 + a.Ts * ti;
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
                        st6[ax] = hold[ax];
                        st6[3 + ax] = 0.0;
                    }
                }
                if (a.pos_std > 0.0 || a.vel_std > 0.0) {  // math::addRandomNoise (math/src/Random.cpp:7-29), seeded
                    const unsigned long long base = (static_cast<unsigned long long>(g) * 0x100000ull + tglob) * 4ull;
#pragma unroll
                    for (int pr = 0; pr < 3; ++pr) {
                        double n0, n1;
                        normal_pair(a.seed, base + pr, n0, n1);
                        const int c0 = 2 * pr, c1 = 2 * pr + 1;
                        st6[c0] += (c0 < 3 ? a.pos_std : a.vel_std) * n0;
                        st6[c1] += (c1 < 3 ? a.pos_std : a.vel_std) * n1;
                    }
                    if (!has) {
#pragma unroll
                        for (int ax = 0; ax < 3; ++ax) hold[ax] = st6[ax];  // the stationary robot drifts with its noise
                    }
                }
                a.sub[(g * a.nsub + (ti - 1)) * 2] = st6[0];
                a.sub[(g * a.nsub + (ti - 1)) * 2 + 1] = st6[1];
                if (ti == a.nsub)
#pragma unroll
                    for (int i = 0; i < 6; ++i) a.states[g * 6 + i] = st6[i];
                if (a.traj)
#pragma unroll
                    for (int i = 0; i < 6; ++i) a.traj[(g * a.traj_len + tglob) * 6 + i] = st6[i];
            }
        }
        if (a.failed && a.agent_count > 0) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) nfail += __shfl_xor_sync(full, nfail, o);
            if (lane == 0) a.failed[s] += nfail;
        }
        if (!a.do_metrics) continue;
        __syncwarp();
        // ---- phase B: instance_success bookkeeping per sub-step ----------------------------------------
        for (int ag = lane; ag < N; ag += 32) sreach[ag] = a.reached[b0 + ag];
        int done = a.done[s];          // 0 running, 1 success decided, 2 collision
        int finish = a.finish[s];
        double mind2 = a.min_dist2[s];
        __syncwarp();
        for (int ti = 1; ti <= a.nsub; ++ti) {
            const int tglob = a.step * a.nsub + (ti - 1);
            // "if all(reach_goals): return True, max(0, t - 1)" precedes the updates of sub-step t
            if (!done) {
                int all = 1;
                for (int ag = lane; ag < N; ag += 32) all &= sreach[ag];
                all = __all_sync(full, all);
                if (all) {
                    done = 1;
                    finish = max(0, tglob - 1);
                }
            }
            int hit = 0;
            for (int i = lane; i < N; i += 32) {
                const double* pi = a.sub + ((b0 + i) * a.nsub + (ti - 1)) * 2;
                const double gx = pi[0] - a.targets[(b0 + i) * 3], gy = pi[1] - a.targets[(b0 + i) * 3 + 1];
                if (gx * gx + gy * gy <= a.goal_radius * a.goal_radius) sreach[i] = 1;
                for (int j = i + 1; j < N; ++j) {
                    const double* pj = a.sub + ((b0 + j) * a.nsub + (ti - 1)) * 2;
                    const double dx = pi[0] - pj[0], dy = pi[1] - pj[1];
                    const double d2 = dx * dx + dy * dy;
                    mind2 = fmin(mind2, d2);
                    if (a.box) {
                        // collision_check.py:10-39: corner = centre - half / 2, width = 2 * half, strict inequalities
                        const double xA = pi[0] - a.box_x / 2, yA = pi[1] - a.box_y / 2;
                        const double xB = pj[0] - a.box_x / 2, yB = pj[1] - a.box_y / 2;
                        const double w = 2 * a.box_x, hh = 2 * a.box_y;
                        hit |= (xA < xB + w && xA + w > xB && yA < yB + hh && yA + hh > yB) ? 1 : 0;
                    } else {
                        hit |= (d2 <= coll2) ? 1 : 0;
                    }
                }
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

// targets of one agent of every scenario, compact: out [S][3] = targets[s][agent][:]
__global__ void gather_agent_targets_kernel(int S, int N, int agent, const double* targets, double* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < S * 3) out[i] = targets[(static_cast<size_t>(i / 3) * N + agent) * 3 + (i % 3)];
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

int launch_gather_agent_targets(int S, int N, int agent, const double* targets, double* out, cudaStream_t stream) {
    if (S <= 0) return cudaSuccess;
    gather_agent_targets_kernel<<<(S * 3 + 255) / 256, 256, 0, stream>>>(S, N, agent, targets, out);
    return cudaGetLastError();
}

int launch_episode_advance(const EpisodeArgs& a, int sm_count, cudaStream_t stream) {
    if (a.S <= 0) return cudaSuccess;
    if (a.c > 8 || a.p > 16) return cudaErrorInvalidValue;
    const int threads = 128;
    const size_t per_warp = ((a.N + 15) / 16) * 16;
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
