// nsdg_dyn_mass_noalloc: the mass diagnostic (nsdg_dyn_mass, dyn.cu) for callers that cannot
// allocate -- code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// No counterpart in the reference.
//
// nsdg_dyn_mass accumulates into a cudaMalloc'ed scratch word of the dynamics state.  This variant
// keeps its scratch (one partial sum per row) in a file-scope __device__ array of fixed size
// instead, so the entry point calls no cudaMalloc / cudaFree / allocating library routine:
//   mass_rows_kernel   one CTA per row (grid-stride over rows): g_row_sums[j] = sum_i mean[j][i]
//   mass_final_kernel  one CTA: out[0] = hx hy sum_j g_row_sums[j], summed in a fixed order
// Both sums have a fixed association order, so the result is reproducible bit for bit (the atomicAdd
// of mass_kernel is not), and every entry the second kernel reads has been written by the first in
// the same call: nothing has to start at zero and stale contents of earlier calls are never read.
//
// The array exists once per GPU that loads the library (128 KiB, held until the process exits) and
// is shared by every call: two calls that overlap on one GPU (different streams / contexts) race.
#include "ctx.h"
#include <mutex>

// scratch: row sums.  In HOST code this name is only a handle for cudaGetSymbolAddress, never a
// device address.
static __device__ __align__(256) double g_row_sums[NSDG_NOALLOC_MAX_ROWS];

namespace {

constexpr int MAX_DEVICES = 64;
std::mutex g_addr_lock;
double* g_row_sums_addr[MAX_DEVICES] = {}; // device address of g_row_sums on each GPU

__device__ __forceinline__ double cta_sum(double s)
{
    __shared__ double part[32];
    for (int off = 16; off > 0; off >>= 1)
        s += __shfl_down_sync(0xffffffffu, s, off);
    if ((threadIdx.x & 31) == 0)
        part[threadIdx.x >> 5] = s;
    __syncthreads();
    s = (threadIdx.x < (blockDim.x >> 5)) ? part[threadIdx.x] : 0.;
    if (threadIdx.x < 32)
        for (int off = 16; off > 0; off >>= 1)
            s += __shfl_down_sync(0xffffffffu, s, off);
    __syncthreads(); // part[] is reused by the next row
    return s; // valid in thread 0
}

__global__ void __launch_bounds__(256) mass_rows_kernel(const double* __restrict__ mean, int64_t nx, int64_t pitch, int64_t ny, double* __restrict__ rows)
{
    for (int64_t j = blockIdx.x; j < ny; j += gridDim.x) {
        const double* row = mean + j * pitch;
        double s = 0.;
        for (int64_t i = threadIdx.x; i < nx; i += blockDim.x)
            s += row[i];
        s = cta_sum(s);
        if (threadIdx.x == 0)
            rows[j] = s;
    }
}

__global__ void __launch_bounds__(1024) mass_final_kernel(const double* __restrict__ rows, int64_t ny, double hx, double hy, double* __restrict__ out)
{
    double s = 0.;
    for (int64_t j = threadIdx.x; j < ny; j += blockDim.x)
        s += rows[j];
    s = cta_sum(s);
    if (threadIdx.x == 0)
        out[0] = s * hx * hy; // same association as nsdg_dyn_mass
}

} // namespace

extern "C" int nsdg_dyn_mass_noalloc(nsdg_ctx* ctx, const double* mean, int64_t nx, int64_t ny, int64_t pitch, double hx, double hy, double* mass_out)
{
    NSDG_ENTER(ctx); // makes ctx->device current
    NSDG_REQUIRE(ctx, mean && mass_out, NSDG_EINVAL, "nsdg_dyn_mass_noalloc: mean and mass_out must be device pointers, not NULL");
    NSDG_REQUIRE(ctx, nx > 0 && ny > 0 && pitch >= nx, NSDG_EINVAL, "nsdg_dyn_mass_noalloc: need nx > 0, ny > 0, pitch >= nx (got nx %lld, ny %lld, pitch %lld)",
        (long long)nx, (long long)ny, (long long)pitch);
    NSDG_REQUIRE(ctx, ny <= NSDG_NOALLOC_MAX_ROWS, NSDG_EINVAL,
        "nsdg_dyn_mass_noalloc: %lld rows exceed the fixed scratch of NSDG_NOALLOC_MAX_ROWS = %d rows; use nsdg_dyn_mass or split the rows", (long long)ny,
        (int)NSDG_NOALLOC_MAX_ROWS);
    NSDG_REQUIRE(ctx, ctx->device >= 0 && ctx->device < MAX_DEVICES, NSDG_EINVAL, "nsdg_dyn_mass_noalloc: device %d out of range", ctx->device);
    double* rows;
    {
        // once per device, with that device current: the device address of the scratch array
        std::lock_guard<std::mutex> g(g_addr_lock);
        if (!g_row_sums_addr[ctx->device]) {
            void* p = nullptr;
            NSDG_CUDA(ctx, cudaGetSymbolAddress(&p, g_row_sums));
            g_row_sums_addr[ctx->device] = static_cast<double*>(p);
        }
        rows = g_row_sums_addr[ctx->device];
    }
    const unsigned grid = (unsigned)(ny < (int64_t)ctx->sm_count * 8 ? ny : (int64_t)ctx->sm_count * 8);
    mass_rows_kernel<<<grid, 256, 0, ctx->stream>>>(mean, nx, pitch, ny, rows);
    NSDG_LAUNCHED(ctx);
    mass_final_kernel<<<1, 1024, 0, ctx->stream>>>(rows, ny, hx, hy, mass_out);
    NSDG_LAUNCHED(ctx);
    return NSDG_OK; // asynchronous: mass_out is valid once the context's stream has reached this point
}
