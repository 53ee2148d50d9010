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

extern "C" {

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
    if (cudaSetDevice(device) != cudaSuccess
        || cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess
        || cudaStreamCreateWithFlags(&ctx->comm_stream, cudaStreamNonBlocking) != cudaSuccess
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
    if (ctx->ev_start)
        cudaEventDestroy(ctx->ev_start);
    if (ctx->ev_stop)
        cudaEventDestroy(ctx->ev_stop);
    if (ctx->stream)
        cudaStreamDestroy(ctx->stream);
    if (ctx->comm_stream)
        cudaStreamDestroy(ctx->comm_stream);
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
