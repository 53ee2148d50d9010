// Context life cycle, error reporting, stream timing.  See include/nsdg_gpu.h.
#include "ctx.h"
#include <cstdarg>

int nsdg_fail(nsdg_ctx* ctx, int code, const char* fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx)
        ctx->error = buf;
    return code;
}

static void prof_flush(nsdg_ctx* ctx)
{
    NsdgProfiler& p = ctx->prof;
    if (!p.used)
        return;
    cudaEventSynchronize(p.ev[2 * (p.used - 1) + 1]);
    for (size_t i = 0; i < p.used; ++i) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, p.ev[2 * i], p.ev[2 * i + 1]) == cudaSuccess) {
            p.total_ms[p.kid[i]] += ms;
            p.count[p.kid[i]]++;
        }
    }
    p.used = 0;
}

void nsdg_prof_begin(nsdg_ctx* ctx, int kid)
{
    NsdgProfiler& p = ctx->prof;
    if (!p.on)
        return;
    if (p.used >= 16384)
        prof_flush(ctx);
    if (p.ev.size() < 2 * (p.used + 1)) {
        cudaEvent_t a, b;
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        p.ev.push_back(a);
        p.ev.push_back(b);
        p.kid.push_back(kid);
    }
    p.kid[p.used] = kid;
    cudaEventRecord(p.ev[2 * p.used], ctx->stream);
}

void nsdg_prof_end(nsdg_ctx* ctx)
{
    NsdgProfiler& p = ctx->prof;
    if (!p.on)
        return;
    cudaEventRecord(p.ev[2 * p.used + 1], ctx->stream);
    p.used++;
}

static const char* const KERNEL_NAMES[NSDG_K_COUNT] = { "thermo_kernel", "transport_stage_kernel",
    "transport_march_kernel", "limit_kernel", "pressure_kernel", "nodal_kernel", "stress_kernel", "velocity_kernel",
    "mevp_fused_kernel", "mevp_pair_kernel", "mevp_march_kernel", "pack_rows", "unpack_rows", "blend_kernel", "other" };

extern "C" {

int nsdg_profile_begin(nsdg_ctx* ctx)
{
    NSDG_ENTER(ctx);
    NsdgProfiler& p = ctx->prof;
    prof_flush(ctx);
    for (int k = 0; k < NSDG_K_COUNT; ++k) {
        p.total_ms[k] = 0.;
        p.count[k] = 0;
    }
    p.on = true;
    return NSDG_OK;
}

int nsdg_profile_end(nsdg_ctx* ctx)
{
    NSDG_ENTER(ctx);
    prof_flush(ctx);
    ctx->prof.on = false;
    return NSDG_OK;
}

int nsdg_profile_query(const nsdg_ctx* ctx, int32_t kernel, double* total_ms, int64_t* launches)
{
    if (!ctx || kernel < 0 || kernel >= NSDG_K_COUNT)
        return NSDG_EINVAL;
    if (total_ms)
        *total_ms = ctx->prof.total_ms[kernel];
    if (launches)
        *launches = ctx->prof.count[kernel];
    return NSDG_OK;
}

const char* nsdg_profile_kernel_name(int32_t kernel)
{
    return (kernel >= 0 && kernel < NSDG_K_COUNT) ? KERNEL_NAMES[kernel] : nullptr;
}

int nsdg_abi_version(void) { return NSDG_ABI_VERSION; }

int nsdg_create(nsdg_ctx** out, int device)
{
    if (!out)
        return NSDG_EINVAL;
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || device < 0 || device >= ndev) {
        // there is deliberately no CPU path behind this ABI
        fprintf(stderr, "nsdg_create: no usable CUDA device %d (%s)\n", device,
            e != cudaSuccess ? cudaGetErrorString(e) : "index out of range");
        return NSDG_ECUDA;
    }
    nsdg_ctx* ctx = new nsdg_ctx;
    ctx->device = device;
    // boundary tiles and halo messages run at the highest priority so that they are scheduled
    // ahead of the interior tiles they overlap with
    int prio_lo = 0, prio_hi = 0;
    cudaSetDevice(device);
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    if (cudaSetDevice(device) != cudaSuccess
        || cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess
        || cudaStreamCreateWithPriority(&ctx->comm_stream, cudaStreamNonBlocking, prio_hi) != cudaSuccess
        || cudaStreamCreateWithFlags(&ctx->h2d_stream, cudaStreamNonBlocking) != cudaSuccess
        || cudaStreamCreateWithFlags(&ctx->d2h_stream, cudaStreamNonBlocking) != cudaSuccess
        || cudaEventCreateWithFlags(&ctx->ev_pipe[0], cudaEventDisableTiming) != cudaSuccess
        || cudaEventCreateWithFlags(&ctx->ev_pipe[1], cudaEventDisableTiming) != cudaSuccess
        || cudaEventCreateWithFlags(&ctx->ev_pipe[2], cudaEventDisableTiming) != cudaSuccess
        || cudaEventCreateWithFlags(&ctx->ev_pipe[3], cudaEventDisableTiming) != cudaSuccess
        || cudaEventCreateWithFlags(&ctx->ev_halo[0], cudaEventDisableTiming) != cudaSuccess
        || cudaEventCreateWithFlags(&ctx->ev_halo[1], cudaEventDisableTiming) != cudaSuccess
        || cudaEventCreate(&ctx->ev_start) != cudaSuccess
        || cudaEventCreate(&ctx->ev_stop) != cudaSuccess) {
        fprintf(stderr, "nsdg_create: %s\n", cudaGetErrorString(cudaGetLastError()));
        delete ctx;
        return NSDG_ECUDA;
    }
    cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
    *out = ctx;
    return NSDG_OK;
}

int nsdg_destroy(nsdg_ctx* ctx)
{
    if (!ctx)
        return NSDG_EINVAL;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    nsdg_comm_free(ctx);
    nsdg_dyn_free(ctx);
    nsdg_thermo_free(ctx);
    for (cudaEvent_t e : ctx->prof.ev)
        cudaEventDestroy(e);
    if (ctx->ev_start)
        cudaEventDestroy(ctx->ev_start);
    if (ctx->ev_stop)
        cudaEventDestroy(ctx->ev_stop);
    if (ctx->stream)
        cudaStreamDestroy(ctx->stream);
    if (ctx->comm_stream)
        cudaStreamDestroy(ctx->comm_stream);
    if (ctx->h2d_stream)
        cudaStreamDestroy(ctx->h2d_stream);
    if (ctx->d2h_stream)
        cudaStreamDestroy(ctx->d2h_stream);
    for (cudaEvent_t e : ctx->ev_pipe)
        if (e)
            cudaEventDestroy(e);
    for (cudaEvent_t e : ctx->ev_halo)
        if (e)
            cudaEventDestroy(e);
    delete ctx;
    return NSDG_OK;
}

const char* nsdg_last_error(const nsdg_ctx* ctx) { return ctx ? ctx->error.c_str() : "null context"; }

int nsdg_synchronize(nsdg_ctx* ctx)
{
    NSDG_ENTER(ctx);
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->comm_stream));
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return NSDG_OK;
}

int64_t nsdg_launch_count(const nsdg_ctx* ctx) { return ctx ? ctx->launches : -1; }

int nsdg_timer_start(nsdg_ctx* ctx)
{
    NSDG_ENTER(ctx);
    NSDG_CUDA(ctx, cudaEventRecord(ctx->ev_start, ctx->stream));
    return NSDG_OK;
}

int nsdg_timer_stop(nsdg_ctx* ctx, double* elapsed_ms)
{
    NSDG_ENTER(ctx);
    NSDG_REQUIRE(ctx, elapsed_ms, NSDG_EINVAL, "elapsed_ms is null");
    NSDG_CUDA(ctx, cudaEventRecord(ctx->ev_stop, ctx->stream));
    NSDG_CUDA(ctx, cudaEventSynchronize(ctx->ev_stop));
    float ms = 0.f;
    NSDG_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_start, ctx->ev_stop));
    *elapsed_ms = ms;
    return NSDG_OK;
}

} // extern "C"
