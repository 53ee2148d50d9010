// Internal definitions shared by the translation units of libnsdg_gpu.so.
// Nothing here is part of the C ABI (include/nsdg_gpu.h).
#pragma once
#include "../../include/nsdg_gpu.h"
#include <cuda_runtime.h>
#include <cstdio>
#include <string>
#include <vector>

struct DynState; // dyn.cu
struct CommState; // comm.cu

struct ThermoState {
    int64_t n = 0;
    int32_t nlayers = 1;
    double* fields[NSDG_TF_COUNT] = {};
    double* diag = nullptr; // [NSDG_THERMO_NDIAG][n], allocated on first diag step
    double* stage = nullptr; // device staging for tice upload/download [n*nlayers]
    nsdg_thermo_params params;
    bool have_prog = false, have_forcing = false;
    double* level[2] = {}; // forcing time levels, [8][n] each, allocated on first use
    bool have_level[2] = { false, false };
    bool slab_ocean = false;
};

// kernel classes for the per-launch event profiler (nsdg_profile_*)
enum NsdgKernelId {
    NSDG_K_THERMO = 0,
    NSDG_K_TRANSPORT,
    NSDG_K_TRANSPORT_MARCH,
    NSDG_K_LIMIT,
    NSDG_K_PRESSURE,
    NSDG_K_NODAL,
    NSDG_K_STRESS,
    NSDG_K_VELOCITY,
    NSDG_K_MEVP_FUSED,
    NSDG_K_MEVP_PAIR,
    NSDG_K_MEVP_MARCH,
    NSDG_K_PACK,
    NSDG_K_UNPACK,
    NSDG_K_BLEND,
    NSDG_K_OTHER,
    NSDG_K_COUNT
};

struct NsdgProfiler {
    bool on = false;
    std::vector<cudaEvent_t> ev; // pairs: 2*i start, 2*i+1 stop
    std::vector<int> kid;
    size_t used = 0; // pairs in flight
    double total_ms[NSDG_K_COUNT] = {};
    int64_t count[NSDG_K_COUNT] = {};
};

struct nsdg_ctx {
    int device = 0;
    cudaStream_t stream = nullptr; // compute
    cudaStream_t comm_stream = nullptr; // halo exchange
    cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr; // nsdg_dyn_step_host copies
    cudaEvent_t ev_pipe[4] = {};
    cudaEvent_t ev_halo[2] = {}; // boundary tiles done -> exchange; exchange done -> next launch
    cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
    int64_t launches = 0;
    int sm_count = 148;
    bool lring_attr_set = false;
    bool ring_attr_set = false; // dynamic shared-memory limit of the ring kernels raised on this device
    std::string error;
    bool cuda_failed = false; // sticky
    NsdgProfiler prof;
    ThermoState* thermo = nullptr;
    DynState* dyn = nullptr;
    CommState* comm = nullptr;
};

int nsdg_fail(nsdg_ctx* ctx, int code, const char* fmt, ...);

#define NSDG_CUDA(ctx, expr)                                                                      \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess) {                                                                  \
            (ctx)->cuda_failed = true;                                                            \
            return nsdg_fail((ctx), NSDG_ECUDA, "%s failed at %s:%d: %s", #expr, __FILE__,        \
                __LINE__, cudaGetErrorString(_e));                                                \
        }                                                                                         \
    } while (0)

#define NSDG_REQUIRE(ctx, cond, code, ...)                                                        \
    do {                                                                                          \
        if (!(cond))                                                                              \
            return nsdg_fail((ctx), (code), __VA_ARGS__);                                         \
    } while (0)

#define NSDG_ENTER(ctx)                                                                           \
    do {                                                                                          \
        if (!(ctx))                                                                               \
            return NSDG_EINVAL;                                                                   \
        if ((ctx)->cuda_failed)                                                                   \
            return NSDG_ECUDA;                                                                    \
        cudaError_t _e = cudaSetDevice((ctx)->device);                                            \
        if (_e != cudaSuccess) {                                                                  \
            (ctx)->cuda_failed = true;                                                            \
            return nsdg_fail((ctx), NSDG_ECUDA, "cudaSetDevice(%d): %s", (ctx)->device,           \
                cudaGetErrorString(_e));                                                          \
        }                                                                                         \
    } while (0)

// after a kernel launch
#define NSDG_LAUNCHED(ctx)                                                                        \
    do {                                                                                          \
        (ctx)->launches++;                                                                        \
        NSDG_CUDA((ctx), cudaGetLastError());                                                     \
    } while (0)

// event pair around one launch when profiling is on (ctx.cu)
void nsdg_prof_begin(nsdg_ctx* ctx, int kid);
void nsdg_prof_end(nsdg_ctx* ctx);

// NSDG_KERNEL(ctx, NSDG_K_X, kernel<<<grid, block, smem, ctx->stream>>>(args));
#define NSDG_KERNEL(ctx, kid, ...)                                                                \
    do {                                                                                          \
        nsdg_prof_begin((ctx), (kid));                                                            \
        __VA_ARGS__;                                                                              \
        nsdg_prof_end((ctx));                                                                     \
        NSDG_LAUNCHED(ctx);                                                                       \
    } while (0)

// teardown hooks implemented by the owning translation unit
void nsdg_thermo_free(nsdg_ctx* ctx);
void nsdg_dyn_free(nsdg_ctx* ctx);
void nsdg_comm_free(nsdg_ctx* ctx);
