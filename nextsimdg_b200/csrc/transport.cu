// DG tracer transport: fused Runge-Kutta stage kernels.
//
// NOT IN THE REFERENCE SNAPSHOT (its CMakeLists.txt:43-46 comments "dynamics" out); the scheme
// is the one stated in oracle/dg_transport.c and DESIGN.md, parity is against that oracle only.
//
// Three forms of the same arithmetic (dg1_* / dg2_* device functions), bit-identical to one another:
//   transport_stage_kernel        one thread per cell, one stage per launch (dG0; small grids)
//   transport_march_kernel        a warp marches up a band of rows, one stage per launch
//   transport_pair_march_kernel   the same march with TWO stages per launch (the whole Heun step of
//                                 dG1, stages 1 + 2 of SSP-RK3 for dG2), stage 1 kept in registers
//
// transport_stage_kernel: one thread owns one cell and gathers everything it needs (no atomics,
// deterministic):
//   volume term + the four edge upwind fluxes + inverse mass + the RK stage combination
//     dst = a*base + b*(src + dt*L(src))
// for one tracer (blockIdx.z selects it).  Each thread computes both fluxes of its own
// edges (the neighbour recomputes the shared one): 2x redundant flops, zero extra DRAM bytes.
// Coefficients are component-major planes, so every load below is a fully coalesced stream;
// the +-1 neighbours in x and y hit L1/L2 (counting rule of SURVEY.md 8d: neighbour re-reads
// are free).  Algorithmic bytes per cell per stage: (NC src + NC base + NC dst)*NF*8 + 2*8
// (the velocity reads of the second tracer's CTAs hit L2).
#include "dyn.h"
#include <algorithm>
#include <cstdlib>

namespace {

template <int NQ> struct Gauss;
template <> struct Gauss<1> {
    static __device__ __forceinline__ double x(int) { return 0.5; }
    static __device__ __forceinline__ double w(int) { return 1.0; }
};
template <> struct Gauss<2> {
    static __device__ __forceinline__ double x(int k)
    {
        return k == 0 ? 0.5 - 0.28867513459481287 : 0.5 + 0.28867513459481287;
    }
    static __device__ __forceinline__ double w(int) { return 0.5; }
};
template <> struct Gauss<3> {
    static __device__ __forceinline__ double x(int k)
    {
        return k == 0 ? 0.5 - 0.3872983346207417 : (k == 1 ? 0.5 : 0.5 + 0.3872983346207417);
    }
    static __device__ __forceinline__ double w(int k) { return k == 1 ? 8. / 18. : 5. / 18.; }
};

// ---- floating-point reproducibility --------------------------------------------------------
// This translation unit is compiled with -fmad=false (build.py): the compiler never contracts a
// multiplication and an addition on its own, every fused multiply-add below is written out.
// Which products get fused is therefore the same in every kernel, in every unrolled copy of a
// loop body and for every tile / band layout, and a cell's result does not depend on which of
// them computed it: the marching kernel, the one-thread-per-cell kernel and any decomposition
// of the grid give the same bits (with automatic contraction the three unrolled copies of the
// marching step differed in the last bit, profiles/r2g_transport_tile_probe.txt).
template <int NC> __device__ __forceinline__ double eval(const double (&c)[NC], double xi, double eta)
{
    double a = xi - 0.5, b = eta - 0.5;
    double s = c[0];
    if (NC >= 3)
        s = fma(c[2], b, fma(c[1], a, s));
    if (NC >= 6)
        s = fma(c[5], a * b, fma(c[4], b * b - 1. / 12., fma(c[3], a * a - 1. / 12., s)));
    return s;
}

// r[i] += F * psi_i(xi,eta)
template <int NC> __device__ __forceinline__ void add_psi(double (&r)[NC], double F, double xi, double eta)
{
    double a = xi - 0.5, b = eta - 0.5;
    r[0] += F;
    if (NC >= 3) {
        r[1] = fma(F, a, r[1]);
        r[2] = fma(F, b, r[2]);
    }
    if (NC >= 6) {
        r[3] = fma(F, a * a - 1. / 12., r[3]);
        r[4] = fma(F, b * b - 1. / 12., r[4]);
        r[5] = fma(F, a * b, r[5]);
    }
}

template <int NC> __device__ __forceinline__ void load_cell(double (&c)[NC], const double* p, int plane, int k)
{
#pragma unroll
    for (int q = 0; q < NC; ++q)
        c[q] = __ldg(p + q * plane + k);
}

// ---- dG2 cell arithmetic shared by the tiled and the marching stage kernels ------------------
// Three-point Gauss rule, written out in tensor-product form (same integrals as the generic
// quadrature loops, ~45 % fewer FP64 instructions): with a = xi - 1/2, b = eta - 1/2
//   phi(a, b) = (c0 + c2 b + c4 (b^2 - 1/12)) + (c1 + c5 b) a + c3 (a^2 - 1/12)
// is evaluated row by row, the velocity at a volume point is a lerp of the edge point
// velocities, an edge trace is a 1-D quadratic trace(t) = T0 + T1 t + T2 (t^2 - 1/12) along the
// edge, and the moments against psi are taken from row / column sums.  Gauss abscissae -g, 0, +g
// with g^2 = 3/20: g^2 - 1/12 = 1/15.
struct EdgeVel { // edge-normal velocities times weight / cell size at the edge Gauss points
    double uL[3], uR[3], vB[3], vT[3];
};
struct Traces { // a cell's own traces on its four edges
    double l[3], r[3], b[3], t[3]; // xi = 0, xi = 1, eta = 0, eta = 1
};
__device__ __forceinline__ void dg2_traces(const double (&c)[6], Traces& T)
{
    const double m0 = fma(c[3], 1. / 6., c[0]), n0 = fma(c[4], 1. / 6., c[0]);
    T.l[0] = fma(-0.5, c[1], m0);
    T.l[1] = fma(-0.5, c[5], c[2]);
    T.l[2] = c[4];
    T.r[0] = fma(0.5, c[1], m0);
    T.r[1] = fma(0.5, c[5], c[2]);
    T.r[2] = c[4];
    T.b[0] = fma(-0.5, c[2], n0);
    T.b[1] = fma(-0.5, c[5], c[1]);
    T.b[2] = c[3];
    T.t[0] = fma(0.5, c[2], n0);
    T.t[1] = fma(0.5, c[5], c[1]);
    T.t[2] = c[3];
}
// flux moments of one edge: M0 = sum F, M1 = sum F t, M2 = sum F (t^2 - 1/12); `minus` is the
// trace on the side the normal velocity points away from
__device__ __forceinline__ void dg2_edge(const double (&vn)[3], const double (&minus)[3], const double (&plus)[3],
    double& M0, double& M1, double& M2)
{
    constexpr double g = 0.3872983346207417;
    const double em = fma(minus[2], 1. / 15., minus[0]), om = minus[1] * g;
    const double ep = fma(plus[2], 1. / 15., plus[0]), op = plus[1] * g;
    const double F0 = vn[0] * (vn[0] > 0. ? em - om : ep - op);
    const double F1 = vn[1] * (vn[1] > 0. ? fma(minus[2], -1. / 12., minus[0]) : fma(plus[2], -1. / 12., plus[0]));
    const double F2 = vn[2] * (vn[2] > 0. ? em + om : ep + op);
    const double s02 = F0 + F2;
    M0 = s02 + F1;
    M1 = g * (F2 - F0);
    M2 = fma(F1, -1. / 12., s02 * (1. / 15.));
}
// rr = int_K phi (v . grad psi) - sum_e int_e (v.n) phi^ psi for one cell from
// lm / rm / bm / tm: the flux moments (M0, M1, M2 of dg2_edge) of the left, right, bottom and top
// edge.  The flux through an edge is the same number whichever of its two cells computes it.
__device__ __forceinline__ void dg2_cell_m(const double (&c)[6], const EdgeVel& V, const double (&lm)[3],
    const double (&rm)[3], const double (&bm)[3], const double (&tm)[3], double (&rr)[6])
{
    constexpr double g = 0.3872983346207417, W0 = 5. / 18., W1 = 8. / 18.;
    constexpr double XI[3] = { 0.5 - g, 0.5, 0.5 + g };
    constexpr double WQ[3] = { W0, W1, W0 };
    double r0 = 0., r1 = 0., r2 = 0., r3 = 0., r4 = 0., r5 = 0.;
    { // ---- volume term: int phi (u/hx dpsi/dxi + v/hy dpsi/deta)
        // fx[b][a] = w_a w_b phi u/hx, fy[b][a] = w_a w_b phi v/hy (uL.. carry w_b/hx, vB.. w_a/hy)
        double sx0 = 0., sx2 = 0., sxa = 0.; // sum_a fx of row b = 0, 2; sum over all of fx a/g
        double sy_a0 = 0., sy_a2 = 0., syb = 0.; // sum_b fy of column a = 0, 2; sum of fy b/g
        double dv[3];
#pragma unroll
        for (int a = 0; a < 3; ++a)
            dv[a] = V.vT[a] - V.vB[a];
#pragma unroll
        for (int b = 0; b < 3; ++b) {
            const double bb = XI[b] - 0.5;
            const double p0 = (b == 1) ? fma(c[4], -1. / 12., c[0]) : fma(c[4], 1. / 15., fma(c[2], bb, c[0]));
            const double p1 = (b == 1) ? c[1] : fma(c[5], bb, c[1]);
            const double e = fma(c[3], 1. / 15., p0), o = p1 * g;
            const double ph[3] = { e - o, fma(c[3], -1. / 12., p0), e + o };
            const double du = V.uR[b] - V.uL[b];
            double fxr[3], fyr[3], uu[3], vv[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                uu[a] = fma(du, XI[a], V.uL[b]); // w_b u / hx
                vv[a] = fma(dv[a], XI[b], V.vB[a]); // w_a v / hy
            }
            fxr[0] = (ph[0] * WQ[0]) * uu[0];
            fxr[2] = (ph[2] * WQ[2]) * uu[2];
            fyr[0] = (ph[0] * WQ[b]) * vv[0];
            fyr[2] = (ph[2] * WQ[b]) * vv[2];
            // the middle products are only ever summed: fused into the row sums
            const double sx = fma(ph[1] * WQ[1], uu[1], fxr[0] + fxr[2]);
            const double sy = fma(ph[1] * WQ[b], vv[1], fyr[0] + fyr[2]);
            r1 += sx;
            r2 += sy;
            sxa += fxr[2] - fxr[0];
            if (b == 0) {
                sx0 = sx;
                sy_a0 = fyr[0];
                sy_a2 = fyr[2];
                syb -= sy;
            } else if (b == 2) {
                sx2 = sx;
                sy_a0 += fyr[0];
                sy_a2 += fyr[2];
                syb += sy;
            } else {
                sy_a0 += fyr[0];
                sy_a2 += fyr[2];
            }
        }
        r3 = (2. * g) * sxa; // sum fx 2a
        r4 = (2. * g) * syb; // sum fy 2b
        r5 = g * ((sx2 - sx0) + (sy_a2 - sy_a0)); // sum fx b + fy a
    }
    // psi on xi = 0 / 1: (1, -+1/2, t, 1/6, t^2 - 1/12, -+t/2); inflow through the left edge
    // adds, outflow through the right edge subtracts
    {
        const double d0 = lm[0] - rm[0], s0 = lm[0] + rm[0], d1 = lm[1] - rm[1], s1 = lm[1] + rm[1];
        r0 += d0;
        r1 = fma(-0.5, s0, r1);
        r2 += d1;
        r3 = fma(d0, 1. / 6., r3);
        r4 += lm[2] - rm[2];
        r5 = fma(-0.5, s1, r5);
    }
    // psi on eta = 0 / 1: (1, t, -+1/2, t^2 - 1/12, 1/6, -+t/2)
    {
        const double d0 = bm[0] - tm[0], s0 = bm[0] + tm[0], d1 = bm[1] - tm[1], s1 = bm[1] + tm[1];
        r0 += d0;
        r1 += d1;
        r2 = fma(-0.5, s0, r2);
        r3 += bm[2] - tm[2];
        r4 = fma(d0, 1. / 6., r4);
        r5 = fma(-0.5, s1, r5);
    }
    rr[0] = r0;
    rr[1] = r1;
    rr[2] = r2;
    rr[3] = r3;
    rr[4] = r4;
    rr[5] = r5;
}
// the same from the NEIGHBOURS' traces on the shared edges (the left neighbour's xi = 1 trace, ...)
__device__ __forceinline__ void dg2_cell(const double (&c)[6], const EdgeVel& V, const Traces& T,
    const double (&nbL)[3], const double (&nbR)[3], const double (&nbB)[3], const double (&nbT)[3], double (&rr)[6])
{
    double lm[3], rm[3], bm[3], tm[3];
    dg2_edge(V.uL, nbL, T.l, lm[0], lm[1], lm[2]);
    dg2_edge(V.uR, T.r, nbR, rm[0], rm[1], rm[2]);
    dg2_edge(V.vB, nbB, T.b, bm[0], bm[1], bm[2]);
    dg2_edge(V.vT, T.t, nbT, tm[0], tm[1], tm[2]);
    dg2_cell_m(c, V, lm, rm, bm, tm, rr);
}
// Normal velocity of the bilinear field at an edge Gauss point, times weight / cell size:
// (a (1 - s) + b s) w / h = a k0 + b k1 with k0 = (1 - s) w / h, k1 = s w / h taken on the host
// (StageArgs::ku0 ...: FP64 instructions read them straight from the constant bank).
__device__ __forceinline__ double edge_lerp(double a, double b, double k0, double k1)
{
    return fma(b, k1, a * k0);
}
// dst = a*base + b*(c + dt*IM*rr); dtim = dt * IM taken on the host
template <bool HAS_BASE>
__device__ __forceinline__ double stage_value(double c, double rr, double dtim, double a, double b, double base)
{
    const double val = fma(dtim, rr, c);
    return HAS_BASE ? fma(a, base, b * val) : val;
}

// ---- dG1 cell arithmetic, written out like dg2_cell (two-point Gauss rule, abscissae -+g) -------
//   phi(a, b) = c0 + c1 a + c2 b; an edge trace is T0 + T1 t; flux moments M0 = sum F, M1 = sum F t
__device__ __forceinline__ void dg1_edge(const double (&vn)[3], double m0, double m1, double p0, double p1, double& M0,
    double& M1)
{
    constexpr double g = 0.28867513459481287;
    const double om = m1 * g, op = p1 * g;
    const double F0 = vn[0] * (vn[0] > 0. ? m0 - om : p0 - op);
    const double F1 = vn[1] * (vn[1] > 0. ? m0 + om : p0 + op);
    M0 = F0 + F1;
    M1 = g * (F1 - F0);
}
// rr from the flux moments (M0, M1) of the four edges
__device__ __forceinline__ void dg1_cell_m(const double (&c)[3], const double (&uL)[3], const double (&uR)[3],
    const double (&vB)[3], const double (&vT)[3], const double (&lm)[2], const double (&rm)[2], const double (&bm)[2],
    const double (&tm)[2], double (&rr)[3])
{
    constexpr double g = 0.28867513459481287;
    constexpr double XI[2] = { 0.5 - g, 0.5 + g };
    double r0 = 0., r1 = 0., r2 = 0.;
    { // volume term; the weights 1/2 of the other direction are folded into the polynomial
        const double ch0 = 0.5 * c[0], oh = c[1] * (0.5 * g);
        const double dv[2] = { vT[0] - vB[0], vT[1] - vB[1] };
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const double p0h = fma(c[2], b ? 0.5 * g : -0.5 * g, ch0);
            const double ph[2] = { p0h - oh, p0h + oh };
            const double du = uR[b] - uL[b];
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                r1 = fma(ph[a], fma(du, XI[a], uL[b]), r1); // w_b u / hx at (xi_a, eta_b)
                r2 = fma(ph[a], fma(dv[a], XI[b], vB[a]), r2); // w_a v / hy
            }
        }
    }
    r0 += lm[0] - rm[0];
    r1 = fma(-0.5, lm[0] + rm[0], r1);
    r2 += lm[1] - rm[1];
    r0 += bm[0] - tm[0];
    r1 += bm[1] - tm[1];
    r2 = fma(-0.5, bm[0] + tm[0], r2);
    rr[0] = r0;
    rr[1] = r1;
    rr[2] = r2;
}
// own traces of a dG1 cell: x-normal edges (c0 -+ c1/2, c2), y-normal edges (c0 -+ c2/2, c1)
struct Traces1 {
    double l[2], r[2], b[2], t[2];
};
__device__ __forceinline__ void dg1_traces(const double (&c)[3], Traces1& T)
{
    T.l[0] = fma(-0.5, c[1], c[0]);
    T.l[1] = c[2];
    T.r[0] = fma(0.5, c[1], c[0]);
    T.r[1] = c[2];
    T.b[0] = fma(-0.5, c[2], c[0]);
    T.b[1] = c[1];
    T.t[0] = fma(0.5, c[2], c[0]);
    T.t[1] = c[1];
}
// c: own coefficients; cl / cr / cb / ct: the neighbours' (left, right, below, above)
__device__ __forceinline__ void dg1_cell(const double (&c)[3], const double (&cl)[3], const double (&cr)[3],
    const double (&cb)[3], const double (&ct)[3], const double (&uL)[3], const double (&uR)[3], const double (&vB)[3],
    const double (&vT)[3], double (&rr)[3])
{
    Traces1 T, L, R, B, U;
    dg1_traces(c, T);
    dg1_traces(cl, L);
    dg1_traces(cr, R);
    dg1_traces(cb, B);
    dg1_traces(ct, U);
    double lm[2], rm[2], bm[2], tm[2];
    dg1_edge(uL, L.r[0], L.r[1], T.l[0], T.l[1], lm[0], lm[1]);
    dg1_edge(uR, T.r[0], T.r[1], R.l[0], R.l[1], rm[0], rm[1]);
    dg1_edge(vB, B.t[0], B.t[1], T.b[0], T.b[1], bm[0], bm[1]);
    dg1_edge(vT, T.t[0], T.t[1], U.b[0], U.b[1], tm[0], tm[1]);
    dg1_cell_m(c, uL, uR, vB, vT, lm, rm, bm, tm, rr);
}

struct StageArgs {
    const double* src[2];
    const double* base[2];
    double* dst[2];
    const double* u;
    const double* v;
    double a, b, dt, ihx, ihy;
    double ku0[3], ku1[3], kv0[3], kv1[3]; // edge_lerp constants per Gauss point (x- / y-normal edges)
    double dtim[6]; // dt times the inverse mass of each mode
    int nx, nyl, plane, r0, r1;
    int wrap, prefetch;
    int band; // rows per band of the marching kernel
};

template <int NQ>
__device__ __forceinline__ void edge_velocities(const StageArgs& A, double u00, double u10, double u01, double u11,
    double v00, double v10, double v01, double v11, EdgeVel& V)
{
    if (NQ < 3)
        V.uL[2] = V.uR[2] = V.vB[2] = V.vT[2] = 0.;
#pragma unroll
    for (int g = 0; g < NQ; ++g) {
        V.uL[g] = edge_lerp(u00, u01, A.ku0[g], A.ku1[g]);
        V.uR[g] = edge_lerp(u10, u11, A.ku0[g], A.ku1[g]);
        V.vB[g] = edge_lerp(v00, v10, A.kv0[g], A.kv1[g]);
        V.vT[g] = edge_lerp(v01, v11, A.kv0[g], A.kv1[g]);
    }
}

__device__ __forceinline__ void prefetch_line(const double* p)
{
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

// one tracer per blockIdx.z (the velocity reads of the second tracer hit L2): half the registers
// of a two-tracer thread, so three CTAs instead of one fit an SM
template <int NC, bool HAS_BASE, int MINB>
__global__ void __launch_bounds__(256, MINB) transport_stage_kernel(const __grid_constant__ StageArgs A)
{
    constexpr int NQ = (NC == 1) ? 1 : (NC == 3 ? 2 : 3);
    using G = Gauss<NQ>;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int r = A.r0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= A.nx || r >= A.r1)
        return;
    int rm = r - 1, rp = r + 1;
    if (A.wrap) { // interior rows are [NSDG_GHOST, NSDG_GHOST + nyl)
        if (rm < NSDG_GHOST)
            rm += A.nyl;
        if (rp >= NSDG_GHOST + A.nyl)
            rp -= A.nyl;
    }
    const int im = (i == 0) ? A.nx - 1 : i - 1;
    const int ip = (i == A.nx - 1) ? 0 : i + 1;
    const int k = r * A.nx + i;
    const int f = blockIdx.z;
    if (NC == 6 && A.prefetch) {
        // the kernel is latency bound with 24 warps per SM and cannot keep all 36 loads of a
        // thread in flight in 80 registers: pull the streams that are used last (the RK base, the
        // rows above and below the CTA) towards L1/L2 now, without holding registers
        if (HAS_BASE)
#pragma unroll
            for (int q = 0; q < NC; ++q)
                prefetch_line(A.base[f] + q * A.plane + k);
        if (threadIdx.y == 0)
#pragma unroll
            for (int q = 0; q < NC; ++q)
                prefetch_line(A.src[f] + q * A.plane + rm * A.nx + i);
        if (threadIdx.y == blockDim.y - 1)
#pragma unroll
            for (int q = 0; q < NC; ++q)
                prefetch_line(A.src[f] + q * A.plane + rp * A.nx + i);
    }

    const double u00 = __ldg(A.u + r * A.nx + i), u10 = __ldg(A.u + r * A.nx + ip);
    const double u01 = __ldg(A.u + rp * A.nx + i), u11 = __ldg(A.u + rp * A.nx + ip);
    const double v00 = __ldg(A.v + r * A.nx + i), v10 = __ldg(A.v + r * A.nx + ip);
    const double v01 = __ldg(A.v + rp * A.nx + i), v11 = __ldg(A.v + rp * A.nx + ip);

    // edge-normal velocities (times weight / cell size) at the edge Gauss points; the upwind
    // trace is selected by their sign: max(vn,0)*minus + min(vn,0)*plus == vn*(vn > 0 ? minus : plus)
    double uL[NQ], uR[NQ], vB[NQ], vT[NQ];
#pragma unroll
    for (int g = 0; g < NQ; ++g) {
        uL[g] = edge_lerp(u00, u01, A.ku0[g], A.ku1[g]);
        uR[g] = edge_lerp(u10, u11, A.ku0[g], A.ku1[g]);
        vB[g] = edge_lerp(v00, v10, A.kv0[g], A.kv1[g]);
        vT[g] = edge_lerp(v01, v11, A.kv0[g], A.kv1[g]);
    }

    if constexpr (NC == 6) {
        // dG2: written-out tensor-product form (dg2_cell); the neighbours' traces are evaluated
        // from their coefficients with the expressions of dg2_traces
        double c[NC], cn[NC], rr[NC];
        load_cell<NC>(c, A.src[f], A.plane, k);
        EdgeVel V;
#pragma unroll
        for (int g = 0; g < 3; ++g) {
            V.uL[g] = uL[g];
            V.uR[g] = uR[g];
            V.vB[g] = vB[g];
            V.vT[g] = vT[g];
        }
        Traces T, N;
        dg2_traces(c, T);
        double nbL[3], nbR[3], nbB[3], nbT[3];
        load_cell<NC>(cn, A.src[f], A.plane, r * A.nx + im); // left neighbour: its xi = 1 trace
        dg2_traces(cn, N);
#pragma unroll
        for (int q = 0; q < 3; ++q)
            nbL[q] = N.r[q];
        load_cell<NC>(cn, A.src[f], A.plane, r * A.nx + ip); // right neighbour: its xi = 0 trace
        dg2_traces(cn, N);
#pragma unroll
        for (int q = 0; q < 3; ++q)
            nbR[q] = N.l[q];
        load_cell<NC>(cn, A.src[f], A.plane, rm * A.nx + i); // lower neighbour: its eta = 1 trace
        dg2_traces(cn, N);
#pragma unroll
        for (int q = 0; q < 3; ++q)
            nbB[q] = N.t[q];
        load_cell<NC>(cn, A.src[f], A.plane, rp * A.nx + i); // upper neighbour: its eta = 0 trace
        dg2_traces(cn, N);
#pragma unroll
        for (int q = 0; q < 3; ++q)
            nbT[q] = N.b[q];
        dg2_cell(c, V, T, nbL, nbR, nbB, nbT, rr);
#pragma unroll
        for (int q = 0; q < NC; ++q)
            A.dst[f][q * A.plane + k] = stage_value<HAS_BASE>(c[q], rr[q], A.dtim[q], A.a, A.b,
                HAS_BASE ? __ldg(A.base[f] + q * A.plane + k) : 0.);
    } else if constexpr (NC == 3) {
        // dG1: written out (dg1_cell), ~117 FP64 instructions against ~147 for the generic loops
        double c[3], cl[3], cr[3], cb[3], ct[3], rr[3];
        load_cell<3>(c, A.src[f], A.plane, k);
        load_cell<3>(cl, A.src[f], A.plane, r * A.nx + im);
        load_cell<3>(cr, A.src[f], A.plane, r * A.nx + ip);
        load_cell<3>(cb, A.src[f], A.plane, rm * A.nx + i);
        load_cell<3>(ct, A.src[f], A.plane, rp * A.nx + i);
        const double ul[3] = { uL[0], uL[1], 0. }, ur[3] = { uR[0], uR[1], 0. };
        const double vb[3] = { vB[0], vB[1], 0. }, vt[3] = { vT[0], vT[1], 0. };
        dg1_cell(c, cl, cr, cb, ct, ul, ur, vb, vt, rr);
#pragma unroll
        for (int q = 0; q < 3; ++q)
            A.dst[f][q * A.plane + k] = stage_value<HAS_BASE>(c[q], rr[q], A.dtim[q], A.a, A.b,
                HAS_BASE ? __ldg(A.base[f] + q * A.plane + k) : 0.);
    } else {
        double c[NC], cn[NC], rr[NC];
        load_cell<NC>(c, A.src[f], A.plane, k);
#pragma unroll
        for (int q = 0; q < NC; ++q)
            rr[q] = 0.;
        // volume term: int phi (u/hx dpsi/dxi + v/hy dpsi/deta)
        if (NC > 1) {
#pragma unroll
            for (int ga = 0; ga < NQ; ++ga)
#pragma unroll
                for (int gb = 0; gb < NQ; ++gb) {
                    const double xi = G::x(ga), eta = G::x(gb), w = G::w(ga) * G::w(gb);
                    const double uu = fma(fma(u11, xi, u01 * (1 - xi)), eta, fma(u10, xi, u00 * (1 - xi)) * (1 - eta));
                    const double vv = fma(fma(v11, xi, v01 * (1 - xi)), eta, fma(v10, xi, v00 * (1 - xi)) * (1 - eta));
                    const double p = w * eval<NC>(c, xi, eta);
                    const double fx = (p * uu) * A.ihx, fy = (p * vv) * A.ihy;
                    rr[1] += fx;
                    rr[2] += fy;
                    if (NC >= 6) {
                        rr[3] = fma(fx, 2. * (xi - 0.5), rr[3]);
                        rr[4] = fma(fy, 2. * (eta - 0.5), rr[4]);
                        rr[5] = fma(fy, xi - 0.5, fma(fx, eta - 0.5, rr[5]));
                    }
                }
        }
        // left edge (xi = 0): inflow from the left neighbour's right trace
        load_cell<NC>(cn, A.src[f], A.plane, r * A.nx + im);
#pragma unroll
        for (int g = 0; g < NQ; ++g) {
            const double s = G::x(g);
            double F = uL[g] * (uL[g] > 0. ? eval<NC>(cn, 1., s) : eval<NC>(c, 0., s));
            add_psi<NC>(rr, F, 0., s);
        }
        // right edge (xi = 1)
        load_cell<NC>(cn, A.src[f], A.plane, r * A.nx + ip);
#pragma unroll
        for (int g = 0; g < NQ; ++g) {
            const double s = G::x(g);
            double F = uR[g] * (uR[g] > 0. ? eval<NC>(c, 1., s) : eval<NC>(cn, 0., s));
            add_psi<NC>(rr, -F, 1., s);
        }
        // bottom edge (eta = 0)
        load_cell<NC>(cn, A.src[f], A.plane, rm * A.nx + i);
#pragma unroll
        for (int g = 0; g < NQ; ++g) {
            const double s = G::x(g);
            double F = vB[g] * (vB[g] > 0. ? eval<NC>(cn, s, 1.) : eval<NC>(c, s, 0.));
            add_psi<NC>(rr, F, s, 0.);
        }
        // top edge (eta = 1)
        load_cell<NC>(cn, A.src[f], A.plane, rp * A.nx + i);
#pragma unroll
        for (int g = 0; g < NQ; ++g) {
            const double s = G::x(g);
            double F = vT[g] * (vT[g] > 0. ? eval<NC>(c, s, 1.) : eval<NC>(cn, s, 0.));
            add_psi<NC>(rr, -F, s, 1.);
        }
#pragma unroll
        for (int q = 0; q < NC; ++q)
            A.dst[f][q * A.plane + k] = stage_value<HAS_BASE>(c[q], rr[q], A.dtim[q], A.a, A.b,
                HAS_BASE ? __ldg(A.base[f] + q * A.plane + k) : 0.);
    }
}

// bounds-preserving rescaling of the higher moments about the cell mean (oracle/dg_transport.c)
template <int NC>
__global__ void limit_kernel(double* phi, int64_t nx, int64_t nyl, int64_t plane, double lo, double hi)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t r = NSDG_GHOST + (int64_t)blockIdx.y;
    if (i >= nx || (int64_t)blockIdx.y >= nyl)
        return;
    const int64_t k = r * nx + i;
    double c[NC];
#pragma unroll
    for (int q = 0; q < NC; ++q)
        c[q] = phi[q * plane + k];
    double mean = c[0], mn = mean, mx = mean;
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) {
            double p = eval<NC>(c, 0.5 * a, 0.5 * b);
            mn = fmin(mn, p);
            mx = fmax(mx, p);
        }
    double theta = 1.;
    if (mn < lo)
        theta = fmin(theta, fmax(0., (mean - lo) / (mean - mn)));
    if (mx > hi)
        theta = fmin(theta, fmax(0., (hi - mean) / (mx - mean)));
    if (theta < 1.) {
#pragma unroll
        for (int q = 1; q < NC; ++q)
            phi[q * plane + k] = c[q] * theta;
    }
}

// ---------------------------------------------------------------------------------------------
// Marching stage kernel (dG2): a WARP owns 28 columns and marches up a band of rows; a lane is a
// column.  The neighbours' edge traces come from the lanes next door (6 shuffles per row) and
// from the lane's own registers (the row below: kept from the previous step; the row above:
// already loaded, because it is the next row to be processed), so every coefficient is loaded
// ONCE per stage (plus a 2-column / 2-row halo per warp and band) instead of five times through
// L1/L2, and the loads of row r+2 are issued at the top of step r: a register software pipeline
// two rows deep.  Three row records rotate through the roles current / next / being loaded (loop
// unrolled by three), so a register that is the target of a load in flight is never copied.
// The per-cell arithmetic is dg2_cell, shared with the tiled kernel: results are bit-identical.
constexpr int TM_OWN = 28, TM_OFF = 2, TM_WPB = 4;

template <int NC> struct TRow {
    double c[NC]; // src coefficients of cell row r
    double u, v; // velocity at node row r, own column
    double base[NC]; // RK base of cell row r
    int k; // plane index of (row r, own column)
};

__device__ __forceinline__ int t_wrap(int i, int n)
{
    if (i < 0)
        i += n;
    else if (i >= n)
        i -= n;
    if (i < 0 || i >= n) { // grids narrower than the halo
        i %= n;
        if (i < 0)
            i += n;
    }
    return i;
}
__device__ __forceinline__ int t_plane_row(const StageArgs& A, int r)
{
    if (A.wrap)
        return NSDG_GHOST + t_wrap(r - NSDG_GHOST, A.nyl);
    return max(0, min(r, A.nyl + 2 * NSDG_GHOST - 1));
}
template <int NC> __device__ __forceinline__ void tm_load_src(const StageArgs& A, int f, int r, int gi, TRow<NC>& R)
{
    const int k = t_plane_row(A, r) * A.nx + gi;
    R.k = k;
#pragma unroll
    for (int q = 0; q < NC; ++q)
        R.c[q] = __ldg(A.src[f] + q * A.plane + k);
    R.u = __ldg(A.u + k);
    R.v = __ldg(A.v + k);
}
template <int NC> __device__ __forceinline__ void tm_load_base(const StageArgs& A, int f, TRow<NC>& R)
{
#pragma unroll
    for (int q = 0; q < NC; ++q)
        R.base[q] = __ldg(A.base[f] + q * A.plane + R.k);
}

// one step of the march: X = row r (coefficients, velocity, base all arrived), Y = row r+1
// (coefficients and velocity arrived, base requested here), Z = target of the loads of row r+2.
// Every edge flux is computed ONCE: a lane computes the flux through its left edge and takes the
// one through its right edge from the lane on its right; the flux through the top edge is kept
// (bm) as the next row's bottom flux, and so are the top edge's normal velocities (vb).
template <int NC, bool HAS_BASE>
__device__ __forceinline__ void tm_step(const StageArgs& A, int f, int r, int gi, bool own, TRow<NC>& X, TRow<NC>& Y,
    TRow<NC>& Z, double (&bm)[3], double (&vb)[3])
{
    constexpr int NT = NC == 6 ? 3 : 2; // numbers per edge trace / flux moments per edge / Gauss points
    tm_load_src<NC>(A, f, r + 2, gi, Z);
    if (HAS_BASE)
        tm_load_base<NC>(A, f, Y);
    EdgeVel V;
    {
        const double u10 = __shfl_down_sync(0xffffffffu, X.u, 1), u11 = __shfl_down_sync(0xffffffffu, Y.u, 1);
        const double v10 = __shfl_down_sync(0xffffffffu, X.v, 1), v11 = __shfl_down_sync(0xffffffffu, Y.v, 1);
        edge_velocities<NT>(A, X.u, u10, Y.u, u11, X.v, v10, Y.v, v11, V);
#pragma unroll
        for (int q = 0; q < 3; ++q)
            V.vB[q] = vb[q]; // == the previous row's vT (same nodes, same expression)
    }
    double lm[3], rm[3], tm[3], rr[NC];
    if constexpr (NC == 6) {
        Traces T, TY;
        dg2_traces(X.c, T);
        dg2_traces(Y.c, TY);
        double nbL[3];
#pragma unroll
        for (int q = 0; q < 3; ++q)
            nbL[q] = __shfl_up_sync(0xffffffffu, T.r[q], 1); // left neighbour's xi = 1 trace
        dg2_edge(V.uL, nbL, T.l, lm[0], lm[1], lm[2]);
#pragma unroll
        for (int q = 0; q < 3; ++q)
            rm[q] = __shfl_down_sync(0xffffffffu, lm[q], 1); // right edge = the right neighbour's left edge
        dg2_edge(V.vT, T.t, TY.b, tm[0], tm[1], tm[2]);
        dg2_cell_m(X.c, V, lm, rm, bm, tm, rr);
    } else {
        Traces1 T, TY;
        dg1_traces(X.c, T);
        dg1_traces(Y.c, TY);
        const double nb0 = __shfl_up_sync(0xffffffffu, T.r[0], 1), nb1 = __shfl_up_sync(0xffffffffu, T.r[1], 1);
        dg1_edge(V.uL, nb0, nb1, T.l[0], T.l[1], lm[0], lm[1]);
        rm[0] = __shfl_down_sync(0xffffffffu, lm[0], 1);
        rm[1] = __shfl_down_sync(0xffffffffu, lm[1], 1);
        dg1_edge(V.vT, T.t[0], T.t[1], TY.b[0], TY.b[1], tm[0], tm[1]);
        const double l2[2] = { lm[0], lm[1] }, r2[2] = { rm[0], rm[1] }, b2[2] = { bm[0], bm[1] }, t2[2] = { tm[0], tm[1] };
        dg1_cell_m(X.c, V.uL, V.uR, V.vB, V.vT, l2, r2, b2, t2, rr);
        tm[2] = 0.;
    }
    if (own) {
#pragma unroll
        for (int q = 0; q < NC; ++q)
            A.dst[f][q * A.plane + X.k] = stage_value<HAS_BASE>(X.c[q], rr[q], A.dtim[q], A.a, A.b, HAS_BASE ? X.base[q] : 0.);
    }
#pragma unroll
    for (int q = 0; q < NT; ++q)
        bm[q] = tm[q];
#pragma unroll
    for (int q = 0; q < 3; ++q)
        vb[q] = V.vT[q];
}

template <int NC, bool HAS_BASE>
__global__ void __launch_bounds__(TM_WPB * 32, NC == 6 ? 3 : 5) transport_march_kernel(const __grid_constant__ StageArgs A)
{
    const int lane = threadIdx.x & 31;
    const int cb = (blockIdx.x * TM_WPB + (threadIdx.x >> 5)) * TM_OWN; // first owned column
    if (cb >= A.nx)
        return; // whole warp
    const int cu = cb - TM_OFF + lane; // unwrapped column of this lane
    const int gi = t_wrap(cu, A.nx);
    const bool own = lane >= TM_OFF && lane < TM_OFF + TM_OWN && cu < A.nx;
    const int f = blockIdx.z;
    const int b0 = A.r0 + blockIdx.y * A.band; // owned rows [b0, b1)
    const int b1 = min(b0 + A.band, A.r1);
    // The march starts one row below the band: that step computes the flux through the bottom
    // edge of the band's first row (its own top edge) with the SAME instruction sequence every
    // other row goes through, so a cell's result does not depend on where the band boundaries
    // fall (no separately compiled prologue arithmetic); nothing is stored for it.
    double bm[3] = { 0., 0., 0. }, vb[3] = { 0., 0., 0. };
    TRow<NC> X, Y, Z;
    tm_load_src<NC>(A, f, b0 - 1, gi, X);
    tm_load_src<NC>(A, f, b0, gi, Y);
#pragma unroll
    for (int q = 0; q < NC; ++q)
        X.base[q] = 0.;
    // (the not-stored first step is the loop's first iteration with a run-time predicate, not a
    // separately specialised copy of the step)
    int r = b0 - 1;
#pragma unroll 1
    for (; r + 2 < b1; r += 3) {
        tm_step<NC, HAS_BASE>(A, f, r, gi, own && r >= b0, X, Y, Z, bm, vb);
        tm_step<NC, HAS_BASE>(A, f, r + 1, gi, own, Y, Z, X, bm, vb);
        tm_step<NC, HAS_BASE>(A, f, r + 2, gi, own, Z, X, Y, bm, vb);
    }
    if (r < b1) {
        tm_step<NC, HAS_BASE>(A, f, r, gi, own && r >= b0, X, Y, Z, bm, vb);
        if (r + 1 < b1)
            tm_step<NC, HAS_BASE>(A, f, r + 1, gi, own, Y, Z, X, bm, vb);
    }
}

// ---------------------------------------------------------------------------------------------
// TWO Runge-Kutta stages in one launch (temporal blocking, like the sub-cycle's pair kernel):
//   k1  = phi + dt L(phi)                       stage 1, kept in registers, never stored
//   dst = a phi + b (k1 + dt L(k1))             stage 2, the RK base is the source itself
// = the whole Heun step of dG1 (a = b = 1/2) and stages 1 + 2 of the SSP-RK3 step of dG2
// (a = 3/4, b = 1/4; stage 3 follows as an ordinary stage launch).  Step r of the march runs
// stage 1 on cell row r in lanes 1..30 and stage 2 on row r-1 in the owned lanes 2..29, whose
// neighbours are k1 of rows r-2 (through the kept bottom flux), r-1 and r and of the lanes next
// door.  The velocity does not change between the stages, so the edge velocities of a row are
// computed once and kept for its second stage.  Heun: phi crosses HBM / L2 once per STEP (1 read
// + 1 write of every coefficient instead of 3 + 2), one launch instead of two; RK3: 12 + 18
// doubles per cell and tracer for the step instead of 12 + 18 + 18.  The price is stage 1 on a
// 1-column / 1-row halo (30/28 in x, 3 extra march steps per band).  Same device functions, same
// operand order as the stage kernels: bit-identical to separate transport_march / transport_stage
// launches.  dst must not be the source (neighbouring warps and bands still read it): Heun
// writes to the stage buffer and the host swaps the two pointers.  Limiter off.  Strips: two
// ghost rows of the source and of the velocity on either side must be current.
template <int NC> struct HRow {
    double c[NC]; // phi of cell row r
    double u, v; // velocity at node row r, own column
    int k;
};
template <int NC> __device__ __forceinline__ void th_load(const StageArgs& A, int f, int r, int gi, HRow<NC>& R)
{
    const int k = t_plane_row(A, r) * A.nx + gi;
    R.k = k;
#pragma unroll
    for (int q = 0; q < NC; ++q)
        R.c[q] = __ldg(A.src[f] + q * A.plane + k);
    R.u = __ldg(A.u + k);
    R.v = __ldg(A.v + k);
}
template <int NC> struct HKeep { // what stage 2 of row r-1 needs from the steps before
    double k0[NC]; // k1 of row r-1
    double pc[NC]; // phi of row r-1 (the RK base)
    EdgeVel V; // edge velocities of row r-1
    double bm1[3], bm2[3]; // flux through the bottom edge of row r (stage 1) / row r-1 (stage 2)
    double vb[3]; // stage 1: the previous row's vT
    int pk; // plane index of (row r-1, own column)
};
// rr of the cells of one row from their coefficients c, the row above's cy, the edge velocities V
// and the kept flux through the bottom edge bm (replaced by the flux through the top edge):
// exactly the arithmetic of tm_step
template <int NC>
__device__ __forceinline__ void th_row(const double (&c)[NC], const double (&cy)[NC], const EdgeVel& V, double (&bm)[3], double (&rr)[NC])
{
    double lm[3], rm[3], tm[3];
    if constexpr (NC == 6) {
        Traces T, TY;
        dg2_traces(c, T);
        dg2_traces(cy, TY);
        double nbL[3];
#pragma unroll
        for (int q = 0; q < 3; ++q)
            nbL[q] = __shfl_up_sync(0xffffffffu, T.r[q], 1);
        dg2_edge(V.uL, nbL, T.l, lm[0], lm[1], lm[2]);
#pragma unroll
        for (int q = 0; q < 3; ++q)
            rm[q] = __shfl_down_sync(0xffffffffu, lm[q], 1);
        dg2_edge(V.vT, T.t, TY.b, tm[0], tm[1], tm[2]);
        dg2_cell_m(c, V, lm, rm, bm, tm, rr);
    } else {
        Traces1 T, TY;
        dg1_traces(c, T);
        dg1_traces(cy, TY);
        const double nb0 = __shfl_up_sync(0xffffffffu, T.r[0], 1), nb1 = __shfl_up_sync(0xffffffffu, T.r[1], 1);
        dg1_edge(V.uL, nb0, nb1, T.l[0], T.l[1], lm[0], lm[1]);
        rm[0] = __shfl_down_sync(0xffffffffu, lm[0], 1);
        rm[1] = __shfl_down_sync(0xffffffffu, lm[1], 1);
        dg1_edge(V.vT, T.t[0], T.t[1], TY.b[0], TY.b[1], tm[0], tm[1]);
        const double l2[2] = { lm[0], lm[1] }, r2[2] = { rm[0], rm[1] }, b2[2] = { bm[0], bm[1] }, t2[2] = { tm[0], tm[1] };
        dg1_cell_m(c, V.uL, V.uR, V.vB, V.vT, l2, r2, b2, t2, rr);
        tm[2] = 0.;
    }
#pragma unroll
    for (int q = 0; q < 3; ++q)
        bm[q] = tm[q];
}
template <int NC>
__device__ __forceinline__ void th_step(const StageArgs& A, int f, int r, int gi, bool store, HRow<NC>& X, HRow<NC>& Y, HRow<NC>& Z,
    HKeep<NC>& K)
{
    constexpr int NT = NC == 6 ? 3 : 2;
    th_load<NC>(A, f, r + 2, gi, Z);
    EdgeVel V;
    {
        const double u10 = __shfl_down_sync(0xffffffffu, X.u, 1), u11 = __shfl_down_sync(0xffffffffu, Y.u, 1);
        const double v10 = __shfl_down_sync(0xffffffffu, X.v, 1), v11 = __shfl_down_sync(0xffffffffu, Y.v, 1);
        edge_velocities<NT>(A, X.u, u10, Y.u, u11, X.v, v10, Y.v, v11, V);
#pragma unroll
        for (int q = 0; q < 3; ++q)
            V.vB[q] = K.vb[q];
    }
    double k1[NC], rr[NC];
    // ---- stage 1 on row r (as tm_step<NC, false>)
    th_row<NC>(X.c, Y.c, V, K.bm1, rr);
#pragma unroll
    for (int q = 0; q < NC; ++q)
        k1[q] = stage_value<false>(X.c[q], rr[q], A.dtim[q], 0., 1., 0.);
    // ---- stage 2 on row r-1 (as tm_step<NC, true> with src = k1, base = phi)
    th_row<NC>(K.k0, k1, K.V, K.bm2, rr);
    if (store) {
#pragma unroll
        for (int q = 0; q < NC; ++q)
            A.dst[f][q * A.plane + K.pk] = stage_value<true>(K.k0[q], rr[q], A.dtim[q], A.a, A.b, K.pc[q]);
    }
#pragma unroll
    for (int q = 0; q < NC; ++q) {
        K.k0[q] = k1[q];
        K.pc[q] = X.c[q];
    }
    K.V = V;
#pragma unroll
    for (int q = 0; q < 3; ++q)
        K.vb[q] = V.vT[q];
    K.pk = X.k;
}

// dG1: 128 registers at 4 CTAs per SM (3 CTAs: no faster, 5: spills, 26.4 against 24.6 us on the
// 1024^2 cone, profiles/r2ab_heun_sweep.txt); dG2: 2 CTAs per SM
template <int NC>
__global__ void __launch_bounds__(TM_WPB * 32, NC == 6 ? 2 : 4) transport_pair_march_kernel(const __grid_constant__ StageArgs A)
{
    const int lane = threadIdx.x & 31;
    const int cb = (blockIdx.x * TM_WPB + (threadIdx.x >> 5)) * TM_OWN; // first owned column
    if (cb >= A.nx)
        return; // whole warp
    const int cu = cb - TM_OFF + lane;
    const int gi = t_wrap(cu, A.nx);
    const bool own = lane >= TM_OFF && lane < TM_OFF + TM_OWN && cu < A.nx;
    const int f = blockIdx.z;
    const int b0 = A.r0 + blockIdx.y * A.band; // owned rows [b0, b1)
    const int b1 = min(b0 + A.band, A.r1);
    // The march starts two rows below the band and ends one row above it; the first steps are
    // ordinary loop iterations on zero-initialised kept values (never stored), so a cell's result
    // does not depend on where the band boundaries fall.
    HKeep<NC> K = {};
    HRow<NC> X, Y, Z;
    th_load<NC>(A, f, b0 - 2, gi, X);
    th_load<NC>(A, f, b0 - 1, gi, Y);
    int r = b0 - 2;
#pragma unroll 1
    for (; r + 2 <= b1; r += 3) {
        th_step<NC>(A, f, r, gi, own && r > b0, X, Y, Z, K);
        th_step<NC>(A, f, r + 1, gi, own && r + 1 > b0, Y, Z, X, K);
        th_step<NC>(A, f, r + 2, gi, own && r + 2 > b0, Z, X, Y, K);
    }
    if (r <= b1) {
        th_step<NC>(A, f, r, gi, own && r > b0, X, Y, Z, K);
        if (r + 1 <= b1)
            th_step<NC>(A, f, r + 1, gi, own && r + 1 > b0, Y, Z, X, K);
    }
}

// rows per band: all CTAs of a launch cost rows + 3 steps and `slots` are resident at a time, so
// the launch takes whole waves; taller bands have fewer halo rows (cf. march_auto_rows, mevp.cu)
int transport_band_rows(int64_t nrows, int64_t ctas_per_band, int64_t slots)
{
    int best = (int)std::min<int64_t>(nrows, 64);
    double best_t = 1e300;
    for (int64_t nb = 1; nb <= nrows; ++nb) {
        const int64_t ry = (nrows + nb - 1) / nb;
        if (ry > 2048)
            continue;
        if (ry < 12)
            break;
        const int64_t nbands = (nrows + ry - 1) / ry;
        const int64_t waves = (nbands * ctas_per_band + slots - 1) / slots;
        const double t = (double)waves * (double)(ry + 3);
        if (t < best_t) {
            best_t = t;
            best = (int)ry;
        }
    }
    return std::max(best, 1);
}

template <int NC, int MINB>
void launch_stage(const StageArgs& A, bool has_base, dim3 grid, dim3 block, cudaStream_t s)
{
    if (has_base)
        transport_stage_kernel<NC, true, MINB><<<grid, block, 0, s>>>(A);
    else
        transport_stage_kernel<NC, false, MINB><<<grid, block, 0, s>>>(A);
}

} // namespace

static StageArgs stage_args(DynState* d, int nfields, double* const* src, double* const* base, double* const* dst,
    double a, double b, double dt, int64_t r0, int64_t r1)
{
    StageArgs A;
    for (int f = 0; f < 2; ++f) {
        A.src[f] = f < nfields ? src[f] : nullptr;
        A.base[f] = (f < nfields && base) ? base[f] : nullptr;
        A.dst[f] = f < nfields ? dst[f] : nullptr;
    }
    A.u = d->f[NSDG_DF_U];
    A.v = d->f[NSDG_DF_V];
    A.a = a;
    A.b = b;
    A.dt = dt;
    A.ihx = 1. / d->cfg.hx;
    A.ihy = 1. / d->cfg.hy;
    A.nx = (int)d->nx;
    A.nyl = (int)d->nyl;
    A.plane = (int)d->plane;
    A.r0 = (int)r0;
    A.r1 = (int)r1;
    A.wrap = d->wrap ? 1 : 0;
    A.prefetch = 0;
    A.band = 0;
    const int nq = d->nc == 1 ? 1 : (d->nc == 3 ? 2 : 3);
    const double g2 = 0.28867513459481287, g3 = 0.3872983346207417;
    const double xs[3][3] = { { 0.5, 0., 0. }, { 0.5 - g2, 0.5 + g2, 0. }, { 0.5 - g3, 0.5, 0.5 + g3 } };
    const double ws[3][3] = { { 1., 0., 0. }, { 0.5, 0.5, 0. }, { 5. / 18., 8. / 18., 5. / 18. } };
    for (int g = 0; g < 3; ++g) {
        const double sx = xs[nq - 1][g], w = ws[nq - 1][g];
        A.ku0[g] = (1. - sx) * w * A.ihx;
        A.ku1[g] = sx * w * A.ihx;
        A.kv0[g] = (1. - sx) * w * A.ihy;
        A.kv1[g] = sx * w * A.ihy;
    }
    const double im[6] = { 1., 12., 12., 180., 180., 144. };
    for (int q = 0; q < 6; ++q)
        A.dtim[q] = dt * im[q];
    return A;
}

// the launcher's choice between the marching and the one-thread-per-cell stage kernel
// (nsdg_dyn_tuning.transport_march: 1 marching on any grid, -1 never, 0 from 2^20 cells)
bool nsdg_transport_marching(const nsdg_ctx* ctx)
{
    const DynState* d = ctx->dyn;
    const int tm = d->tune.transport_march;
    return d->nc >= 3 && tm >= 0 && (tm > 0 || d->nx * d->nyl >= ((int64_t)1 << 20));
}

int nsdg_transport_pair(nsdg_ctx* ctx, int nfields, double* const* src, double* const* dst, double a, double b, double dt,
    int64_t r0, int64_t r1)
{
    DynState* d = ctx->dyn;
    if (r1 <= r0)
        return NSDG_OK;
    NSDG_REQUIRE(ctx, d->nc >= 3, NSDG_ESTATE, "the two-stage launch is for dG1 / dG2");
    StageArgs A = stage_args(d, nfields, src, nullptr, dst, a, b, dt, r0, r1);
    const int64_t nwarps = (d->nx + TM_OWN - 1) / TM_OWN, gx = (nwarps + TM_WPB - 1) / TM_WPB;
    A.band = transport_band_rows(r1 - r0, gx * nfields, (d->nc == 6 ? 2 : 4) * (int64_t)ctx->sm_count);
    dim3 block(TM_WPB * 32), grid((unsigned)gx, (unsigned)((r1 - r0 + A.band - 1) / A.band), (unsigned)nfields);
    if (d->nc == 6)
        NSDG_KERNEL(ctx, NSDG_K_TRANSPORT_MARCH, (transport_pair_march_kernel<6><<<grid, block, 0, ctx->stream>>>(A)));
    else
        NSDG_KERNEL(ctx, NSDG_K_TRANSPORT_MARCH, (transport_pair_march_kernel<3><<<grid, block, 0, ctx->stream>>>(A)));
    return NSDG_OK;
}

int nsdg_transport_stage(nsdg_ctx* ctx, int nfields, double* const* src, double* const* base,
    double* const* dst, double a, double b, double dt, int64_t r0, int64_t r1)
{
    DynState* d = ctx->dyn;
    if (r1 <= r0)
        return NSDG_OK;
    StageArgs A = stage_args(d, nfields, src, base, dst, a, b, dt, r0, r1);
    // tiled kernel only: L2 prefetch of the streams a thread uses last.  (For the marching kernel an
    // L2 prefetch 1-8 rows ahead of the register pipeline was measured SLOWER, 3.04-3.72 ms against
    // 2.77-2.85 ms per H+A step at 4096^2: profiles/r2m_transport_prefetch.txt.)
    A.prefetch = 1;
    const bool hb = base != nullptr;
    // dG2 on grids that do not live in L2: the marching kernel (nsdg_dyn_tuning.transport_march:
    // 1 forces it on any grid, -1 selects the one-thread-per-cell kernel)
    if (nsdg_transport_marching(ctx)) {
        const int64_t nwarps = (d->nx + TM_OWN - 1) / TM_OWN, gx = (nwarps + TM_WPB - 1) / TM_WPB;
        // resident CTAs per SM: 3 (dG2, 168 registers) / 5 (dG1)
        A.band = transport_band_rows(r1 - r0, gx * nfields, (d->nc == 6 ? 3 : 5) * (int64_t)ctx->sm_count);
        dim3 mblock(TM_WPB * 32), mgrid((unsigned)gx, (unsigned)((r1 - r0 + A.band - 1) / A.band), (unsigned)nfields);
        nsdg_prof_begin(ctx, NSDG_K_TRANSPORT_MARCH);
        if (d->nc == 6) {
            if (hb)
                transport_march_kernel<6, true><<<mgrid, mblock, 0, ctx->stream>>>(A);
            else
                transport_march_kernel<6, false><<<mgrid, mblock, 0, ctx->stream>>>(A);
        } else if (hb)
            transport_march_kernel<3, true><<<mgrid, mblock, 0, ctx->stream>>>(A);
        else
            transport_march_kernel<3, false><<<mgrid, mblock, 0, ctx->stream>>>(A);
        nsdg_prof_end(ctx);
        NSDG_LAUNCHED(ctx);
        return NSDG_OK;
    }
    dim3 block(64, 4);
    dim3 grid((unsigned)((d->nx + 63) / 64), (unsigned)((r1 - r0 + 3) / 4), (unsigned)nfields);
    nsdg_prof_begin(ctx, NSDG_K_TRANSPORT);
    if (d->nc == 1)
        launch_stage<1, 4>(A, hb, grid, block, ctx->stream);
    else if (d->nc == 3)
        launch_stage<3, 4>(A, hb, grid, block, ctx->stream);
    else
        launch_stage<6, 4>(A, hb, grid, block, ctx->stream);
    nsdg_prof_end(ctx);
    NSDG_LAUNCHED(ctx);
    return NSDG_OK;
}

int nsdg_transport_limit(nsdg_ctx* ctx, double* phi, double lo, double hi)
{
    DynState* d = ctx->dyn;
    if (d->nc == 1)
        return NSDG_OK;
    dim3 block(128), grid((unsigned)((d->nx + 127) / 128), (unsigned)d->nyl);
    if (d->nc == 3)
        NSDG_KERNEL(ctx, NSDG_K_LIMIT, limit_kernel<3><<<grid, block, 0, ctx->stream>>>(phi, d->nx, d->nyl, d->plane, lo, hi));
    else
        NSDG_KERNEL(ctx, NSDG_K_LIMIT, limit_kernel<6><<<grid, block, 0, ctx->stream>>>(phi, d->nx, d->nyl, d->plane, lo, hi));
    return NSDG_OK;
}
