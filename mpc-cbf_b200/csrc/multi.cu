// Multi-GPU behind the C ABI: one process, one host thread + stream per device, handle-owned NCCL
// communicators. Two ways the path shards:
//   * scenario sharding (mpccbf_multi_step_batch_host[_compact]): independent scenarios are split
//     contiguously across the devices, constant tables replicated, NO collective;
//   * agent sharding (mpccbf_multi_run_sharded): the agents of every scenario are split across the devices
//     (BASELINE configs[3]); per control step each device solves its own agents against the full state
//     snapshot (mpccbf_step_agents), advances them along their curves like the reference driver
//     (mpccbf_follow_step; MPCCBFFormationControl_example.cpp:161-221) and ONE ncclAllGather of the
//     [S][N/G][6] new states rebuilds the snapshot everywhere. The IMPC iter >= 1 rows use the CURRENT
//     neighbour states (ConnectivityIMPCCBF.cpp:174), so one exchange per control step suffices.
// NCCL is loaded with dlopen (libnccl.so.2) when a multi-device handle with more than one device is
// created: single-GPU use of libmpccbf.so has no NCCL dependency.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/mpccbf.h"
#include "kernels.h"

namespace {

// The few NCCL declarations this file needs (ABI of NCCL 2.x, nccl.h): resolved at run time.
typedef struct ncclComm* ncclComm_t;
typedef int ncclResult_t;
constexpr int kNcclFloat64 = 8;  // ncclDataType_t::ncclFloat64 / ncclDouble
struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommCount)(const ncclComm_t, int*) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool load(std::string& err) {
        if (lib) return true;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {
            lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (lib) break;
        }
        if (!lib) {
            err = std::string("cannot load libnccl.so.2: ") + dlerror();
            return false;
        }
        CommInitAll = reinterpret_cast<decltype(CommInitAll)>(dlsym(lib, "ncclCommInitAll"));
        CommDestroy = reinterpret_cast<decltype(CommDestroy)>(dlsym(lib, "ncclCommDestroy"));
        AllGather = reinterpret_cast<decltype(AllGather)>(dlsym(lib, "ncclAllGather"));
        CommCount = reinterpret_cast<decltype(CommCount)>(dlsym(lib, "ncclCommCount"));
        GetErrorString = reinterpret_cast<decltype(GetErrorString)>(dlsym(lib, "ncclGetErrorString"));
        if (!CommInitAll || !CommDestroy || !AllGather || !GetErrorString) {
            err = "libnccl.so.2 lacks ncclCommInitAll / ncclAllGather";
            return false;
        }
        return true;
    }
};
NcclApi g_nccl;

// send[s][k][:] = states[s][begin + k][:]
__global__ void pack_agents_kernel(const double* states, double* send, int S, int N, int begin, int cnt) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= S * cnt * 6) return;
    const int s = i / (cnt * 6), r = i - s * cnt * 6;
    send[i] = states[(static_cast<size_t>(s) * N + begin) * 6 + r];
}
// states[s][g * cnt + k][:] = recv[g][s][k][:]
__global__ void unpack_agents_kernel(const double* recv, double* states, int S, int N, int cnt, int G) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= S * N * 6) return;
    const int s = i / (N * 6), r = i - s * N * 6;
    const int g = r / (cnt * 6), k = r - g * cnt * 6;
    states[i] = recv[(static_cast<size_t>(g) * S + s) * cnt * 6 + k];
}

struct DevSlot {
    int device = -1;
    mpccbf_handle* h = nullptr;
    cudaStream_t stream = nullptr;
    ncclComm_t comm = nullptr;
    // agent-sharded run
    double *states = nullptr, *targets = nullptr, *ctrl = nullptr, *send = nullptr, *recv = nullptr;
    int32_t *status = nullptr, *ncv = nullptr, *failed = nullptr;
    size_t cap_S = 0;
    std::string error;
    int rc = MPCCBF_OK;
    float ms = 0.f;
};

}  // namespace

struct mpccbf_multi {
    mpccbf_config cfg;
    int G = 0;
    std::vector<DevSlot> d;
    bool have_comms = false;
    int nranks_seen = 0;
};

using mpccbf::set_error;

namespace {

void fail(DevSlot& s, int rc, const std::string& msg) {
    if (s.rc == MPCCBF_OK) {
        s.rc = rc;
        s.error = msg;
    }
}
#define SLOT_CUDA(slot, expr)                                                        \
    do {                                                                             \
        cudaError_t _e = static_cast<cudaError_t>(expr);                             \
        if (_e != cudaSuccess) {                                                     \
            fail(slot, MPCCBF_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
            return;                                                                  \
        }                                                                            \
    } while (0)
#define SLOT_RC(slot, expr)                                                          \
    do {                                                                             \
        const int _rc = (expr);                                                      \
        if (_rc != MPCCBF_OK) {                                                      \
            fail(slot, _rc, std::string(#expr) + ": " + mpccbf_last_error());        \
            return;                                                                  \
        }                                                                            \
    } while (0)

// runs fn(slot index) on one host thread per device and joins; returns the first failure
template <class F>
int on_all_devices(mpccbf_multi* m, F fn) {
    for (auto& s : m->d) {
        s.rc = MPCCBF_OK;
        s.error.clear();
    }
    if (m->G == 1) {
        fn(0);
    } else {
        std::vector<std::thread> th;
        th.reserve(m->G);
        for (int g = 0; g < m->G; ++g) th.emplace_back(fn, g);
        for (auto& t : th) t.join();
    }
    for (auto& s : m->d)
        if (s.rc != MPCCBF_OK) {
            set_error("device " + std::to_string(s.device) + ": " + s.error);
            return s.rc;
        }
    return MPCCBF_OK;
}

void scenario_range(int S, int g, int G, int& begin, int& count) {
    const int base = S / G, extra = S % G;
    begin = g * base + std::min(g, extra);
    count = base + (g < extra ? 1 : 0);
}

void free_sharded(DevSlot& s) {
    void* ptrs[] = {s.states, s.targets, s.ctrl, s.send, s.recv, s.status, s.ncv, s.failed};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    s.states = s.targets = s.ctrl = s.send = s.recv = nullptr;
    s.status = s.ncv = s.failed = nullptr;
    s.cap_S = 0;
}

}  // namespace

extern "C" {

int mpccbf_multi_create(const mpccbf_config* cfg, int32_t num_devices, const int32_t* devices, mpccbf_multi** out) {
    if (!cfg || !out || num_devices < 1) {
        set_error("mpccbf_multi_create: invalid argument");
        return MPCCBF_ERR_INVALID;
    }
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count < 1) {
        set_error("mpccbf_multi_create: no CUDA device (there is no CPU fallback)");
        return MPCCBF_ERR_CUDA;
    }
    mpccbf_multi* m = new (std::nothrow) mpccbf_multi();
    if (!m) {
        set_error("out of host memory");
        return MPCCBF_ERR_ALLOC;
    }
    m->cfg = *cfg;
    m->G = num_devices;
    m->d.resize(num_devices);
    std::vector<int> devs(num_devices);
    for (int g = 0; g < num_devices; ++g) {
        devs[g] = devices ? devices[g] : g;
        if (devs[g] < 0 || devs[g] >= count || std::count(devs.begin(), devs.begin() + g, devs[g])) {
            set_error("mpccbf_multi_create: device ordinal " + std::to_string(devs[g]) + " is out of range or listed twice (" +
                      std::to_string(count) + " devices visible)");
            delete m;
            return MPCCBF_ERR_INVALID;
        }
        m->d[g].device = devs[g];
    }
    for (int g = 0; g < num_devices; ++g) {
        mpccbf_config c = *cfg;
        c.device = devs[g];
        int rc = mpccbf_create(&c, &m->d[g].h);
        if (rc == MPCCBF_OK && (cudaSetDevice(devs[g]) != cudaSuccess ||
                                cudaStreamCreateWithFlags(&m->d[g].stream, cudaStreamNonBlocking) != cudaSuccess)) {
            set_error("mpccbf_multi_create: cudaStreamCreate failed");
            rc = MPCCBF_ERR_CUDA;
        }
        if (rc != MPCCBF_OK) {
            mpccbf_multi_destroy(m);
            return rc;
        }
    }
    if (num_devices > 1) {
        std::string err;
        if (!g_nccl.load(err)) {
            set_error("mpccbf_multi_create: " + err);
            mpccbf_multi_destroy(m);
            return MPCCBF_ERR_UNSUPPORTED;
        }
        std::vector<ncclComm_t> comms(num_devices, nullptr);
        const ncclResult_t r = g_nccl.CommInitAll(comms.data(), num_devices, devs.data());
        if (r != 0) {
            set_error(std::string("ncclCommInitAll: ") + g_nccl.GetErrorString(r));
            mpccbf_multi_destroy(m);
            return MPCCBF_ERR_CUDA;
        }
        for (int g = 0; g < num_devices; ++g) m->d[g].comm = comms[g];
        m->have_comms = true;
        if (g_nccl.CommCount) g_nccl.CommCount(comms[0], &m->nranks_seen);
    } else {
        m->nranks_seen = 1;
    }
    *out = m;
    return MPCCBF_OK;
}

int mpccbf_multi_destroy(mpccbf_multi* m) {
    if (!m) return MPCCBF_OK;
    for (auto& s : m->d) {
        if (s.device >= 0) cudaSetDevice(s.device);
        if (s.comm && g_nccl.CommDestroy) g_nccl.CommDestroy(s.comm);
        free_sharded(s);
        if (s.stream) cudaStreamDestroy(s.stream);
        if (s.h) mpccbf_destroy(s.h);
    }
    delete m;
    return MPCCBF_OK;
}

int32_t mpccbf_multi_num_devices(const mpccbf_multi* m) { return m ? m->G : 0; }
int32_t mpccbf_multi_comm_ranks(const mpccbf_multi* m) { return m ? m->nranks_seen : 0; }
mpccbf_handle* mpccbf_multi_handle(mpccbf_multi* m, int32_t index) {
    return (m && index >= 0 && index < m->G) ? m->d[index].h : nullptr;
}

static int multi_step_host(mpccbf_multi* m, int32_t S, const double* states, const double* targets, double* result,
                           int32_t* status, int32_t* n_curves, double* objective, bool compact) {
    if (!m || S < 0 || !states || !targets || !result || !status) {
        set_error("mpccbf_multi_step_batch_host: invalid argument");
        return MPCCBF_ERR_INVALID;
    }
    const int64_t cap = static_cast<int64_t>(m->cfg.max_scenarios) * m->G;
    if (S > cap) {
        set_error("mpccbf_multi_step_batch_host: num_scenarios exceeds max_scenarios x num_devices");
        return MPCCBF_ERR_INVALID;
    }
    const size_t N = m->cfg.num_agents, it = m->cfg.impc_iter;
    const size_t n = compact ? 6 : static_cast<size_t>(mpccbf_num_decision_vars(m->d[0].h));
    // enqueue on every device (the per-handle streams run concurrently), then wait for all
    return on_all_devices(m, [&](int g) {
        DevSlot& s = m->d[g];
        int b, c;
        scenario_range(S, g, m->G, b, c);
        if (c > m->cfg.max_scenarios) {
            fail(s, MPCCBF_ERR_INVALID, "scenario shard exceeds max_scenarios");
            return;
        }
        if (c == 0) return;
        const size_t a0 = static_cast<size_t>(b) * N;
        auto fn = compact ? mpccbf_step_batch_host_compact_async : mpccbf_step_batch_host_async;
        SLOT_RC(s, fn(s.h, c, states + a0 * 6, targets + a0 * 3, result + a0 * n, status + a0 * it,
                      n_curves ? n_curves + a0 : nullptr, objective ? objective + a0 : nullptr));
        SLOT_RC(s, mpccbf_synchronize(s.h));
    });
}

int mpccbf_multi_step_batch_host(mpccbf_multi* m, int32_t S, const double* states, const double* targets, double* ctrl_pts,
                                 int32_t* status, int32_t* n_curves, double* objective) {
    return multi_step_host(m, S, states, targets, ctrl_pts, status, n_curves, objective, false);
}

int mpccbf_multi_step_batch_host_compact(mpccbf_multi* m, int32_t S, const double* states, const double* targets,
                                         double* v_free, int32_t* status, int32_t* n_curves, double* objective) {
    return multi_step_host(m, S, states, targets, v_free, status, n_curves, objective, true);
}

int mpccbf_multi_run_sharded(mpccbf_multi* m, int32_t S, double* states, const double* targets, int32_t steps,
                             double* ms_per_step) {
    if (!m || S < 1 || S > m->cfg.max_scenarios || !states || !targets || steps < 0) {
        set_error("mpccbf_multi_run_sharded: invalid argument (num_scenarios must be in 1..max_scenarios)");
        return MPCCBF_ERR_INVALID;
    }
    const int N = m->cfg.num_agents, G = m->G;
    if (N % G != 0) {
        set_error("mpccbf_multi_run_sharded: num_agents (" + std::to_string(N) + ") must be divisible by the number of devices (" +
                  std::to_string(G) + "): one all-gather, no padding");
        return MPCCBF_ERR_INVALID;
    }
    const int cnt = N / G;
    const size_t n = static_cast<size_t>(mpccbf_num_decision_vars(m->d[0].h)), it = m->cfg.impc_iter;
    // every shard runs the kernel variant the UNSHARDED step would run, so that the result does not depend on
    // the number of devices (an agent's arithmetic is the same in every tile composition of one variant)
    const int32_t variant = mpccbf_auto_kernel_variant(m->d[0].h, static_cast<int64_t>(S) * N);
    for (auto& s : m->d) mpccbf_set_kernel_variant(s.h, variant);
    const int rc = on_all_devices(m, [&](int g) {
        DevSlot& s = m->d[g];
        SLOT_CUDA(s, cudaSetDevice(s.device));
        if (s.cap_S < static_cast<size_t>(S)) {
            free_sharded(s);
            const size_t B = static_cast<size_t>(S) * cnt;
            SLOT_CUDA(s, cudaMalloc(&s.states, sizeof(double) * S * N * 6));
            SLOT_CUDA(s, cudaMalloc(&s.targets, sizeof(double) * B * 3));
            SLOT_CUDA(s, cudaMalloc(&s.ctrl, sizeof(double) * B * n));
            SLOT_CUDA(s, cudaMalloc(&s.send, sizeof(double) * B * 6));
            SLOT_CUDA(s, cudaMalloc(&s.recv, sizeof(double) * S * N * 6));
            SLOT_CUDA(s, cudaMalloc(&s.status, sizeof(int32_t) * B * it));
            SLOT_CUDA(s, cudaMalloc(&s.ncv, sizeof(int32_t) * B));
            SLOT_CUDA(s, cudaMalloc(&s.failed, sizeof(int32_t) * S));
            s.cap_S = S;
        }
        const int begin = g * cnt;
        cudaStream_t st = s.stream;
        SLOT_CUDA(s, cudaMemcpyAsync(s.states, states, sizeof(double) * S * N * 6, cudaMemcpyHostToDevice, st));
        SLOT_CUDA(s, cudaMemcpy2DAsync(s.targets, sizeof(double) * cnt * 3, targets + static_cast<size_t>(begin) * 3,
                                       sizeof(double) * N * 3, sizeof(double) * cnt * 3, S, cudaMemcpyHostToDevice, st));
        SLOT_RC(s, mpccbf_follow_reset(s.h, S, st));
        cudaEvent_t e0, e1;
        SLOT_CUDA(s, cudaEventCreate(&e0));
        SLOT_CUDA(s, cudaEventCreate(&e1));
        SLOT_CUDA(s, cudaEventRecord(e0, st));
        const int tot_local = S * cnt * 6, tot_full = S * N * 6;
        for (int k = 0; k < steps; ++k) {
            SLOT_RC(s, mpccbf_step_agents(s.h, S, begin, cnt, s.states, s.targets, s.ctrl, s.status, s.ncv, nullptr, st));
            SLOT_RC(s, mpccbf_follow_step(s.h, S, begin, cnt, s.states, s.ctrl, s.ncv, st));
            if (G == 1) continue;
            if (S == 1) {  // the shard is contiguous inside the snapshot: in-place all-gather
                const ncclResult_t r = g_nccl.AllGather(s.states + static_cast<size_t>(begin) * 6, s.states, static_cast<size_t>(cnt) * 6,
                                                        kNcclFloat64, s.comm, st);
                if (r != 0) {
                    fail(s, MPCCBF_ERR_CUDA, std::string("ncclAllGather: ") + g_nccl.GetErrorString(r));
                    return;
                }
            } else {
                pack_agents_kernel<<<(tot_local + 255) / 256, 256, 0, st>>>(s.states, s.send, S, N, begin, cnt);
                const ncclResult_t r = g_nccl.AllGather(s.send, s.recv, static_cast<size_t>(tot_local), kNcclFloat64, s.comm, st);
                if (r != 0) {
                    fail(s, MPCCBF_ERR_CUDA, std::string("ncclAllGather: ") + g_nccl.GetErrorString(r));
                    return;
                }
                unpack_agents_kernel<<<(tot_full + 255) / 256, 256, 0, st>>>(s.recv, s.states, S, N, cnt, G);
                SLOT_CUDA(s, cudaGetLastError());
            }
        }
        SLOT_CUDA(s, cudaEventRecord(e1, st));
        if (g == 0) SLOT_CUDA(s, cudaMemcpyAsync(states, s.states, sizeof(double) * S * N * 6, cudaMemcpyDeviceToHost, st));
        SLOT_CUDA(s, cudaStreamSynchronize(st));
        SLOT_CUDA(s, cudaEventElapsedTime(&s.ms, e0, e1));
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
    });
    for (auto& s : m->d) mpccbf_set_kernel_variant(s.h, -1);
    if (rc != MPCCBF_OK) return rc;
    if (ms_per_step) {
        float worst = 0.f;
        for (auto& s : m->d) worst = std::max(worst, s.ms);
        *ms_per_step = steps > 0 ? worst / steps : 0.0;
    }
    return MPCCBF_OK;
}

}  // extern "C"
