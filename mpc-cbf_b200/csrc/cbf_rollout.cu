// K1: pairwise safety-HOCBF row construction (materialised), K3: rollout — sm_100a.
//
// K1 replaces the GiNaC substitution chain of ConnectivityCBF::getSafetyConstraints /
// getSafetyBound (cbf/src/detail/ConnectivityCBF.cpp:292-304,560-566; expressions :152-198) for
// every ordered (ego, neighbour) pair of every scenario. It is a streaming kernel: per row it
// reads two cached 48-byte states and writes 32 bytes, so it is bound by the HBM write stream.
// K3 evaluates every agent's curve one MPC step ahead (SingleParameterPiecewiseCurve::eval,
// splines/src/curves/SingleParameterPiecewiseCurve.cpp:94-127, as driven by
// MPCCBFFormationControl_example.cpp:191-221).
#include <cuda_runtime.h>

#include <string>

#include "kernels.h"

namespace mpccbf {

// rows [S][N][N-1][4] = (ax, ay, 0, b). One CTA walks whole scenarios; the scenario's states sit in
// shared memory as four N-long arrays (x, y, vx, vy); warps take 32 consecutive rows of the
// scenario's flat N(N-1) row list and store them as fully coalesced 16-byte words.
__global__ void __launch_bounds__(256) cbf_rows_kernel(const double* __restrict__ states,
                                                       double* __restrict__ rows, int S, int N,
                                                       double dmin2, double gamma) {
    extern __shared__ double sm[];
    double* sx = sm;
    double* sy = sx + N;
    double* svx = sy + N;
    double* svy = svx + N;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int nm1 = N - 1;
    const int rows_per_scen = N * nm1;
    for (int s = blockIdx.x; s < S; s += gridDim.x) {
        __syncthreads();
        {
            const double2* src = reinterpret_cast<const double2*>(states + static_cast<size_t>(s) * N * 6);
            for (int i = threadIdx.x; i < N * 3; i += blockDim.x) {
                const double2 w = __ldg(src + i);
                const int agent = i / 3, part = i - agent * 3;  // (x,y) (yaw,vx) (vy,vyaw)
                if (part == 0) {
                    sx[agent] = w.x;
                    sy[agent] = w.y;
                } else if (part == 1) {
                    svx[agent] = w.y;
                } else {
                    svy[agent] = w.x;
                }
            }
        }
        __syncthreads();
        double2* out = reinterpret_cast<double2*>(rows + static_cast<size_t>(s) * rows_per_scen * 4);
        for (int base = warp * 32; base < rows_per_scen; base += nwarps * 32) {
            const int r = base + lane;
            double ax = 0.0, ay = 0.0, b = 0.0;
            if (r < rows_per_scen) {
                const int ego = r / nm1, j = r - ego * nm1;
                const int other = j + (j >= ego ? 1 : 0);
                const double px = sx[ego], py = sy[ego], vx = svx[ego], vy = svy[ego];
                const double dx = px - sx[other], dy = py - sy[other];
                const double dvx = vx - svx[other], dvy = vy - svy[other];
                const double h = dx * dx + dy * dy - dmin2;
                const double lfh = 2.0 * (dx * dvx + dy * dvy);
                const double lf2h = 2.0 * (dvx * dvx + dvy * dvy);
                const double lfa = 3.0 * gamma * h * h * (2.0 * dx * vx + 2.0 * dy * vy);
                const double psi1 = lfh + gamma * h * h * h;
                ax = 2.0 * dx;
                ay = 2.0 * dy;
                b = lf2h + lfa + gamma * psi1 * psi1 * psi1;
            }
            // transpose in registers: 16-byte word w of the warp's 1 KiB block is half (w & 1)
            // of row (w >> 1); lane L stores words L and L + 32.
            const int src0 = lane >> 1, src1 = 16 + (lane >> 1);
            const double a0 = __shfl_sync(0xffffffffu, ax, src0), a1 = __shfl_sync(0xffffffffu, ax, src1);
            const double y0 = __shfl_sync(0xffffffffu, ay, src0), y1 = __shfl_sync(0xffffffffu, ay, src1);
            const double b0 = __shfl_sync(0xffffffffu, b, src0), b1 = __shfl_sync(0xffffffffu, b, src1);
            const bool hi = lane & 1;
            const double2 w0 = hi ? make_double2(0.0, b0) : make_double2(a0, y0);
            const double2 w1 = hi ? make_double2(0.0, b1) : make_double2(a1, y1);
            if (base + src0 < rows_per_scen) out[static_cast<size_t>(base) * 2 + lane] = w0;
            if (base + src1 < rows_per_scen) out[static_cast<size_t>(base) * 2 + 32 + lane] = w1;
        }
    }
}

int launch_cbf_rows(const double* states, double* rows, int S, int N, double dmin2, double gamma,
                    int sm_count, cudaStream_t stream) {
    if (S <= 0 || N < 2) return cudaSuccess;
    const size_t smem = sizeof(double) * 4 * static_cast<size_t>(N);
    if (smem > 200 * 1024) return cudaErrorInvalidConfiguration;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(cbf_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return e;
    }
    int grid = sm_count * 8;
    if (grid > S) grid = S;
    cbf_rows_kernel<<<grid, 256, smem, stream>>>(states, rows, S, N, dmin2, gamma);
    return cudaGetLastError();
}

// K3.
__global__ void __launch_bounds__(256) rollout_kernel(const double* __restrict__ states,
                                                      const double* __restrict__ ctrl,
                                                      const int32_t* __restrict__ n_curves,
                                                      double* __restrict__ next, int B, RolloutConst rc) {
    // one thread per (agent, axis): strided reads of `c` control points stay inside one 128-byte line
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= B * 3) return;
    const int g = idx / 3, ax = idx - 3 * g;
    const bool ok = n_curves ? (n_curves[g] > 0) : true;
    double pos, vel;
    if (ok) {
        const double* x = ctrl + static_cast<size_t>(g) * rc.n + rc.piece * 3 * rc.c + ax * rc.c;
        pos = 0.0;
        vel = 0.0;
        for (int j = 0; j < rc.c; ++j) {
            const double xj = __ldg(x + j);
            pos = fma(rc.b0[j], xj, pos);
            vel = fma(rc.b1[j], xj, vel);
        }
    } else {  // MPCCBFFormationControl_example.cpp:208-221 — stay in place, zero velocity
        pos = states[static_cast<size_t>(g) * 6 + ax];
        vel = 0.0;
    }
    next[static_cast<size_t>(g) * 6 + ax] = pos;
    next[static_cast<size_t>(g) * 6 + 3 + ax] = vel;
}

int launch_rollout(const double* states, const double* ctrl, const int32_t* n_curves, double* next,
                   int B, const RolloutConst& rc, cudaStream_t stream) {
    if (B <= 0) return cudaSuccess;
    const int threads = 256;
    const int grid = (B * 3 + threads - 1) / threads;
    rollout_kernel<<<grid, threads, 0, stream>>>(states, ctrl, n_curves, next, B, rc);
    return cudaGetLastError();
}

}  // namespace mpccbf
