// Scratch microbenchmark: achievable HBM bandwidth of the mEVP access pattern (23 read streams,
// 11 write streams, 32x8 tiles) in the planar layout vs an AoSoA layout with 32-cell blocks.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o stream_probe stream_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

struct Ptrs { const double* in[23]; double* out[11]; };

// planar: one thread per cell, CTA = 32x8 tile
__global__ void __launch_bounds__(256) planar(const __grid_constant__ Ptrs P, int nx, int ny)
{
    const int i = blockIdx.x * 32 + (threadIdx.x & 31), j = blockIdx.y * 8 + (threadIdx.x >> 5);
    const size_t k = (size_t)j * nx + i;
    double v[23];
#pragma unroll
    for (int q = 0; q < 23; ++q) v[q] = __ldg(P.in[q] + k);
    double s = 0;
#pragma unroll
    for (int q = 0; q < 23; ++q) s += v[q];
#pragma unroll
    for (int q = 0; q < 11; ++q) P.out[q][k] = s + q;
}

// blocked: in0 = [cells/32][9][32], in1 = [..][4][32], in2 = [..][8][32], u,v planar;
// out0 = [..][9][32], u,v planar out. Block index = j*(nx/32) + i/32.
struct BPtrs { const double *s, *pg, *nd, *u, *v; double *so, *uo, *vo; };
template <int ROWS>
__global__ void __launch_bounds__(32 * ROWS) blocked(const __grid_constant__ BPtrs P, int nx, int ny)
{
    const int lane = threadIdx.x & 31, xb = blockIdx.x, j = blockIdx.y * ROWS + (threadIdx.x >> 5);
    const size_t blk = (size_t)j * (nx / 32) + xb;
    const size_t k = (size_t)j * nx + xb * 32 + lane;
    const double* ps = P.s + blk * 9 * 32 + lane;
    const double* pp = P.pg + blk * 4 * 32 + lane;
    const double* pn = P.nd + blk * 8 * 32 + lane;
    double v[23];
#pragma unroll
    for (int q = 0; q < 9; ++q) v[q] = __ldg(ps + q * 32);
#pragma unroll
    for (int q = 0; q < 4; ++q) v[9 + q] = __ldg(pp + q * 32);
#pragma unroll
    for (int q = 0; q < 8; ++q) v[13 + q] = __ldg(pn + q * 32);
    v[21] = __ldg(P.u + k);
    v[22] = __ldg(P.v + k);
    double s = 0;
#pragma unroll
    for (int q = 0; q < 23; ++q) s += v[q];
    double* po = P.so + blk * 9 * 32 + lane;
#pragma unroll
    for (int q = 0; q < 9; ++q) po[q * 32] = s + q;
    P.uo[k] = s;
    P.vo[k] = -s;
}

__global__ void copyk(const double2* __restrict__ a, double2* __restrict__ b, size_t n)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) b[i] = a[i];
}

int main(int argc, char** argv)
{
    const int nx = 4096, ny = 4096;
    const size_t n = (size_t)nx * ny;
    double *in, *out;
    CK(cudaMalloc(&in, 23 * n * 8));
    CK(cudaMalloc(&out, 11 * n * 8));
    CK(cudaMemset(in, 0, 23 * n * 8));
    CK(cudaMemset(out, 0, 11 * n * 8));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const double gb = 34.0 * n * 8 / 1e9;
    auto time = [&](const char* name, auto&& f) {
        for (int w = 0; w < 3; ++w) f();
        CK(cudaEventRecord(e0));
        for (int r = 0; r < 10; ++r) f();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 10;
        printf("%-28s %.4f ms  %.0f GB/s\n", name, ms, gb / ms * 1e3);
    };
    Ptrs P;
    for (int q = 0; q < 23; ++q) P.in[q] = in + q * n;
    for (int q = 0; q < 11; ++q) P.out[q] = out + q * n;
    time("planar 23r+11w tile 32x8", [&] { planar<<<dim3(nx / 32, ny / 8), 256>>>(P, nx, ny); });
    BPtrs B = { in, in + 9 * n, in + 13 * n, in + 21 * n, in + 22 * n, out, out + 9 * n, out + 10 * n };
    time("blocked AoSoA tile 32x8", [&] { blocked<8><<<dim3(nx / 32, ny / 8), 256>>>(B, nx, ny); });
    time("blocked AoSoA tile 32x4", [&] { blocked<4><<<dim3(nx / 32, ny / 4), 128>>>(B, nx, ny); });
    time("blocked AoSoA tile 32x16", [&] { blocked<16><<<dim3(nx / 32, ny / 16), 512>>>(B, nx, ny); });
    // plain copy of the same volume (17 n doubles in, 17 n out)
    time("copy double2 17n->17n", [&] { copyk<<<148 * 16, 512>>>((const double2*)in, (double2*)out, 11 * n / 2); });
    printf("(copy line moves 22n doubles = %.2f GB, scale: x%.3f)\n", 22.0 * n * 8 / 1e9, 22.0 / 34.0);
    CK(cudaGetLastError());
    return 0;
}
