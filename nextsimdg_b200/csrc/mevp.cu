// mEVP stress / velocity sub-cycle kernels.
//
// NOT IN THE REFERENCE SNAPSHOT (CMakeLists.txt:43-46 comments "dynamics" out; the only trace
// is the comment "Ice momentum fluxes are handled by the dynamics" at
// physics/src/modules/NextsimPhysics.cpp:124,129).  Scheme: oracle/mevp.c and DESIGN.md.
//
// Once per momentum step ("prepare"):
//   pressure  ice strength P_g at the 2x2 Gauss points of every cell (4 planes)
//   nodal     per-node constants of the sub-cycle (4 planes):
//               mdt = m/dt, acd = A_n C_o rho_o,
//               cx0 = mdt u^n + A_n tau_ax - m f v_o,  cy0 = mdt v^n + A_n tau_ay + m f u_o
//             so that one sub-iteration of the momentum equation is
//               cd = acd |u_o - u|,  u' = (mdt beta u + cx0 + cd u_o + mdt dt f v + D_x)/(mdt (1+beta) + cd)
//             (the oracle's formula with the p-independent terms summed once)
// Per sub-iteration, four interchangeable, mutually bit-identical implementations (all HBM
// streams over component-major planes, coalesced):
//   split   stress_kernel + velocity_kernel                          24 + 17 = 41 doubles/cell
//   fused   mevp_fused_kernel: stress + velocity in one launch, the new stress goes to the
//           velocity phase through shared memory                     32 doubles/cell
//   pair    mevp_pair_kernel: TWO sub-iterations per launch (temporal blocking on a 2-cell
//           halo; periodic wrap or the strip's ghost rows)           32 doubles/cell per PAIR
//   march   mevp_march_kernel (default): the same pair, but a warp marches up a band of rows
//           with a register software pipeline, x-neighbours by shuffle   32 doubles/cell per PAIR
// All index arithmetic is 32-bit (validated by the host: every array of the sub-cycle has fewer
// than 2^31 elements) and every quotient that does not depend on the cell is taken on the host.
#include "dyn.h"
#include "fp64_fast.cuh"
#include <cuda.h> // CUtensorMap (types only: cuTensorMapEncodeTiled is looked up at run time)
#include <algorithm>
#include <cstdlib>
#include <string>

namespace {

constexpr double GQ = 0.28867513459481287; // 1/(2 sqrt 3)

struct MevpArgs {
    int nx, nyl, plane, r0, r1, j0;
    int wrap, nc, bc, ny_global;
    int band; // rows per band of the marching kernel
    double ihx, ihy, ie2, dmin2, ia; // stress: 1/hx, 1/hy, 1/e^2, Delta_min^2, 1/alpha
    double hx2, hy2, hx4, hy4, ome, e2x2, oma, ia4, ia3g; // products of the above (stress_relax)
    double c1, c2, c3, c4; // divergence: 0.5/hx, 1/(12 hx), 0.5/hy, 1/(12 hy)
    double beta, beta1, dtf; // beta, 1+beta, dt*f
    double p_star, c_star, rho_ice, h_min, idt, car, cor, fcor; // prepare kernels
    const double* H;
    const double* A;
    const double* u;
    const double* v;
    const double* uo;
    const double* vo;
    const double* ua;
    const double* va;
    const double* pg;
    const double* nd; // cx0, cy0, mdt, acd planes
    const double* S11;
    const double* S12;
    const double* S22;
    double* S11n; // new stress (ping-pong partner in the fused kernels)
    double* S12n;
    double* S22n;
    double* out0;
    double* out1;
    // peer halo (strips, marching kernel): the neighbours' copies of S11n, S12n, S22n, out0, out1
    // and their flag arrays, mapped with CUDA IPC (comm.cu); flag layout [2][nflagw]
    double* up[5];
    double* dn[5];
    int plane_up, plane_dn, nyl_dn, nflagw;
    const unsigned* flag_mine;
    unsigned* flag_up;
    unsigned* flag_dn;
    unsigned seq;
    int pf_dist, pf_align; // RING == 3: rows ahead of the L2 prefetch; boxes on 256-byte boundaries
};

__device__ __forceinline__ int row_below(const MevpArgs& A, int r)
{
    int rm = r - 1;
    if (A.wrap && rm < NSDG_GHOST)
        rm += A.nyl;
    return rm;
}
__device__ __forceinline__ int row_above(const MevpArgs& A, int r)
{
    int rp = r + 1;
    if (A.wrap && rp >= NSDG_GHOST + A.nyl)
        rp -= A.nyl;
    return rp;
}

__device__ __forceinline__ void prefetch_l1(const double* p)
{
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
}
// pull the six node streams of node_update towards L1 ahead of their use (no registers held)
template <typename Args> __device__ __forceinline__ void prefetch_node(const Args& A, int n)
{
    prefetch_l1(A.uo + n);
    prefetch_l1(A.vo + n);
    prefetch_l1(A.nd + n);
    prefetch_l1(A.nd + A.plane + n);
    prefetch_l1(A.nd + 2 * A.plane + n);
    prefetch_l1(A.nd + 3 * A.plane + n);
}

__device__ __forceinline__ double eval_dg(const double* f, int plane, int k, int nc, double a, double b)
{
    double s = __ldg(f + k);
    if (nc >= 3)
        s += __ldg(f + plane + k) * a + __ldg(f + 2 * plane + k) * b;
    if (nc >= 6)
        s += __ldg(f + 3 * plane + k) * (a * a - 1. / 12.) + __ldg(f + 4 * plane + k) * (b * b - 1. / 12.)
            + __ldg(f + 5 * plane + k) * a * b;
    return s;
}

// HALF the ice strength at the Gauss points (the VP law only ever uses P/2): out0 = pg (4 planes),
// g = 2*b + a
__global__ void __launch_bounds__(256) pressure_kernel(const __grid_constant__ MevpArgs A)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int r = A.r0 + blockIdx.y;
    if (i >= A.nx || r >= A.r1)
        return;
    const int k = r * A.nx + i;
#pragma unroll
    for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int a = 0; a < 2; ++a) {
            const double xa = a ? GQ : -GQ, yb = b ? GQ : -GQ;
            double h = eval_dg(A.H, A.plane, k, A.nc, xa, yb);
            double c = eval_dg(A.A, A.plane, k, A.nc, xa, yb);
            A.out0[(2 * b + a) * A.plane + k] = (0.5 * A.p_star) * h * exp(-A.c_star * (1. - c));
        }
}

// per-node constants of the sub-cycle: out0 = nd (cx0, cy0, mdt, acd); nodal H, A are the mean
// of the 4 adjacent cell means; u, v are the velocity at the start of the momentum step
__global__ void __launch_bounds__(256) nodal_kernel(const __grid_constant__ MevpArgs A)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int r = A.r0 + blockIdx.y;
    if (i >= A.nx || r >= A.r1)
        return;
    const int rm = row_below(A, r);
    const int im = (i == 0) ? A.nx - 1 : i - 1;
    const int n = r * A.nx + i;
    const double hn = 0.25 * (__ldg(A.H + rm * A.nx + im) + __ldg(A.H + rm * A.nx + i) + __ldg(A.H + r * A.nx + im) + __ldg(A.H + n));
    const double an = 0.25 * (__ldg(A.A + rm * A.nx + im) + __ldg(A.A + rm * A.nx + i) + __ldg(A.A + r * A.nx + im) + __ldg(A.A + n));
    const double ua = __ldg(A.ua + n), va = __ldg(A.va + n);
    const double wa = sqrt(ua * ua + va * va);
    const double tax = A.car * wa * ua, tay = A.car * wa * va;
    const double m = A.rho_ice * fmax(hn, A.h_min);
    const double mdt = m * A.idt;
    const double mf = m * A.fcor;
    A.out0[n] = mdt * __ldg(A.u + n) + an * tax - mf * __ldg(A.vo + n);
    A.out0[A.plane + n] = mdt * __ldg(A.v + n) + an * tay + mf * __ldg(A.uo + n);
    A.out0[2 * A.plane + n] = mdt;
    A.out0[3 * A.plane + n] = an * A.cor;
}

// ---- arithmetic shared by all sub-iteration kernels ------------------------------------------

// S <- S + (proj - S)/alpha, where proj is the L2 projection to dG1 of the viscous-plastic stress
// of the bilinear velocity (u00..v11 at the cell's corner nodes) evaluated at the 2x2 Gauss points
// with HALF strengths hp[g] = P_g/2, g = 2*b + a.  Every product of constants is taken on the
// host (hx2 = 1/(2hx), hx4 = 1/(4hx), ome = 1 - 1/e^2, e2x2 = 2/e^2, oma = 1 - 1/alpha,
// ia4 = 1/(4 alpha), ia3g = 3 GQ/alpha), which removes 29 of ~190 FP64 instructions per call.
__device__ __forceinline__ void stress_relax(const MevpArgs& A, double u00, double u10, double u01, double u11,
    double v00, double v10, double v01, double v11, const double (&hp)[4], double (&so11)[3],
    double (&so12)[3], double (&so22)[3])
{
    // dG1 coefficients (mean, xi, eta) of the strain rates, exact for bilinear velocity
    const double ax = u10 - u00, bx = u11 - u01, cyy = u01 - u00, dyy = u11 - u10;
    const double avx = v10 - v00, bvx = v11 - v01, cvy = v01 - v00, dvy = v11 - v10;
    const double e11_0 = (ax + bx) * A.hx2, e11_2 = (bx - ax) * A.ihx;
    const double e22_0 = (cvy + dvy) * A.hy2, e22_1 = (dvy - cvy) * A.ihy;
    const double e12_0 = (cyy + dyy) * A.hy4 + (avx + bvx) * A.hx4;
    const double e12_1 = (dyy - cyy) * A.hy2, e12_2 = (bvx - avx) * A.hx2;
    double s11[4], s12[4], s22[4];
#pragma unroll
    for (int b = 0; b < 2; ++b) {
        const double yb = b ? GQ : -GQ;
        const double g11 = e11_0 + e11_2 * yb; // e11 varies with eta only
        const double g12b = e12_0 + e12_2 * yb;
#pragma unroll
        for (int a = 0; a < 2; ++a) {
            const double xa = a ? GQ : -GQ;
            const double g22 = e22_0 + e22_1 * xa; // e22 varies with xi only
            const double g12 = g12b + e12_1 * xa;
            const double tr = g11 + g22, df = g11 - g22;
            // zeta = P/(2 Delta) with one reciprocal square root instead of sqrt + divide
            const double h = hp[2 * b + a];
            const double zeta = h * rsqrt_pos(A.dmin2 + tr * tr + A.ie2 * (df * df + 4. * g12 * g12));
            const double iso = (zeta * tr) * A.ome - h; // (zeta - eta) tr - P/2, eta = zeta/e^2
            const double eta2 = zeta * A.e2x2;
            s11[2 * b + a] = eta2 * g11 + iso;
            s22[2 * b + a] = eta2 * g22 + iso;
            s12[2 * b + a] = eta2 * g12;
        }
    }
    // (s, psi_c)/(psi_c, psi_c): weights 1/4, +-3 GQ, +-3 GQ at the points (-,-) (+,-) (-,+) (+,+);
    // relaxation folded in: S_c <- (1 - 1/alpha) S_c + proj_c/alpha
#define NSDG_RELAX(so, s)                                                                         \
    do {                                                                                          \
        const double lo = s[0] + s[1], hi = s[2] + s[3];                                          \
        so[0] = (lo + hi) * A.ia4 + so[0] * A.oma;                                                \
        so[1] = ((s[1] - s[0]) + (s[3] - s[2])) * A.ia3g + so[1] * A.oma;                         \
        so[2] = (hi - lo) * A.ia3g + so[2] * A.oma;                                               \
    } while (0)
    NSDG_RELAX(so11, s11);
    NSDG_RELAX(so12, s12);
    NSDG_RELAX(so22, s22);
#undef NSDG_RELAX
}

// -(S, grad phi_n)/(hx hy) at a node from the 7 stress coefficients (S11_0, S11_2, S12_0, S12_1,
// S12_2, S22_0, S22_1) of its 4 adjacent cells x00 (lower left), x10, x01, x11.  The analytic
// dG1 x Q1 integrals collapse to signed sums times four constants; they are taken cell column by
// cell column: with lo / hi the cells below / above the node row in one column,
//   P = c1 (lo0 + hi0) + c2 (lo1 - hi1) + c4 (lo3 - hi3),   Q = c3 (lo2 - hi2)
// the column gives P + Q to the node on its right and Q - P to the node on its left (same for y
// with the coefficients 2, 4, 6 and 5), so that a node needs TWO numbers per direction from its
// neighbour column (the marching kernel shuffles those instead of 14 coefficients).
struct ColPart {
    double lx, ly; // this column is the LEFT column of the node
    double rx, ry; // this column is the RIGHT column of the node
};
__device__ __forceinline__ ColPart col_part(const MevpArgs& A, const double (&lo)[7], const double (&hi)[7])
{
    const double px = A.c1 * (lo[0] + hi[0]) + A.c2 * (lo[1] - hi[1]) + A.c4 * (lo[3] - hi[3]);
    const double qx = A.c3 * (lo[2] - hi[2]);
    const double py = A.c1 * (lo[2] + hi[2]) + A.c2 * (lo[4] - hi[4]) + A.c4 * (lo[6] - hi[6]);
    const double qy = A.c3 * (lo[5] - hi[5]);
    ColPart c;
    c.lx = px + qx;
    c.ly = py + qy;
    c.rx = qx - px;
    c.ry = qy - py;
    return c;
}
struct Div7 {
    double x00[7], x10[7], x01[7], x11[7];
};
__device__ __forceinline__ void stress_div(const MevpArgs& A, const Div7& q, double& dx, double& dy)
{
    const ColPart l = col_part(A, q.x00, q.x01), r = col_part(A, q.x10, q.x11);
    dx = -(l.lx + r.rx);
    dy = -(l.ly + r.ry);
}

// one sub-iteration of the momentum equation at a node (see the header of this file)
__device__ __forceinline__ void node_update_vals(const MevpArgs& A, double uo, double vo, double cx0, double cy0,
    double mdt, double acd, double dx, double dy, double u, double v, double& unew, double& vnew)
{
    const double du = uo - u, dv = vo - v;
    const double cd = acd * sqrt(du * du + dv * dv);
    const double mb = mdt * A.beta, mf = mdt * A.dtf;
    const double inv = rcp_pos(mdt * A.beta1 + cd);
    unew = (mb * u + cx0 + cd * uo + mf * v + dx) * inv;
    vnew = (mb * v + cy0 + cd * vo - mf * u + dy) * inv;
}
__device__ __forceinline__ void node_update(const MevpArgs& A, int n, double dx, double dy, double u, double v,
    double& unew, double& vnew)
{
    const double uo = __ldg(A.uo + n), vo = __ldg(A.vo + n);
    const double cx0 = __ldg(A.nd + n), cy0 = __ldg(A.nd + A.plane + n);
    const double mdt = __ldg(A.nd + 2 * A.plane + n), acd = __ldg(A.nd + 3 * A.plane + n);
    node_update_vals(A, uo, vo, cx0, cy0, mdt, acd, dx, dy, u, v, unew, vnew);
}

// closed box: the node lines i == 0 and global j == 0 (jrow = plane row - NSDG_GHOST; a strip's
// upper ghost rows on the last rank are the ring's row 0 again)
__device__ __forceinline__ bool dirichlet(const MevpArgs& A, int gi, int jrow)
{
    const int j = A.j0 + jrow;
    return A.bc == NSDG_BC_BOX && (gi == 0 || j == 0 || j == A.ny_global);
}

// ---- split variant --------------------------------------------------------------------------
__global__ void __launch_bounds__(256) stress_kernel(const __grid_constant__ MevpArgs A)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int r = A.r0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= A.nx || r >= A.r1)
        return;
    const int rp = row_above(A, r);
    const int ip = (i == A.nx - 1) ? 0 : i + 1;
    const int k = r * A.nx + i;
    double so11[3], so12[3], so22[3], pr[4];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        so11[c] = A.S11[c * A.plane + k];
        so12[c] = A.S12[c * A.plane + k];
        so22[c] = A.S22[c * A.plane + k];
    }
#pragma unroll
    for (int g = 0; g < 4; ++g)
        pr[g] = __ldg(A.pg + g * A.plane + k);
    stress_relax(A, __ldg(A.u + k), __ldg(A.u + r * A.nx + ip), __ldg(A.u + rp * A.nx + i), __ldg(A.u + rp * A.nx + ip),
        __ldg(A.v + k), __ldg(A.v + r * A.nx + ip), __ldg(A.v + rp * A.nx + i), __ldg(A.v + rp * A.nx + ip), pr, so11, so12, so22);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        A.S11n[c * A.plane + k] = so11[c];
        A.S12n[c * A.plane + k] = so12[c];
        A.S22n[c * A.plane + k] = so22[c];
    }
}

// out0 = unew, out1 = vnew
__global__ void __launch_bounds__(256) velocity_kernel(const __grid_constant__ MevpArgs A)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int r = A.r0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= A.nx || r >= A.r1)
        return;
    const int n = r * A.nx + i;
    double un_ = 0., vn_ = 0.;
    if (!dirichlet(A, i, r - NSDG_GHOST)) {
        const int rm = row_below(A, r);
        const int im = (i == 0) ? A.nx - 1 : i - 1;
        Div7 q;
#define NSDG_LOAD7(dst, k)                                                                        \
    do {                                                                                          \
        dst[0] = __ldg(A.S11 + (k));                                                              \
        dst[1] = __ldg(A.S11 + 2 * A.plane + (k));                                                \
        dst[2] = __ldg(A.S12 + (k));                                                              \
        dst[3] = __ldg(A.S12 + A.plane + (k));                                                    \
        dst[4] = __ldg(A.S12 + 2 * A.plane + (k));                                                \
        dst[5] = __ldg(A.S22 + (k));                                                              \
        dst[6] = __ldg(A.S22 + A.plane + (k));                                                    \
    } while (0)
        NSDG_LOAD7(q.x00, rm * A.nx + im);
        NSDG_LOAD7(q.x10, rm * A.nx + i);
        NSDG_LOAD7(q.x01, r * A.nx + im);
        NSDG_LOAD7(q.x11, n);
#undef NSDG_LOAD7
        double dx, dy;
        stress_div(A, q, dx, dy);
        node_update(A, n, dx, dy, __ldg(A.u + n), __ldg(A.v + n), un_, vn_);
    }
    A.out0[n] = un_;
    A.out1[n] = vn_;
}

// gather the divergence coefficients of the cells at slots t00, t00+1, t00+w, t00+w+1 of ss[7][stride]
__device__ __forceinline__ void div_from_smem(const MevpArgs& A, const double* ss, int stride, int t00, int w,
    double& dx, double& dy)
{
    Div7 q;
#pragma unroll
    for (int c = 0; c < 7; ++c) {
        q.x00[c] = ss[c * stride + t00];
        q.x10[c] = ss[c * stride + t00 + 1];
        q.x01[c] = ss[c * stride + t00 + w];
        q.x11[c] = ss[c * stride + t00 + w + 1];
    }
    stress_div(A, q, dx, dy);
}
__device__ __forceinline__ void div_to_smem(double* ss, int stride, int t, const double (&s11)[3],
    const double (&s12)[3], const double (&s22)[3])
{
    ss[0 * stride + t] = s11[0];
    ss[1 * stride + t] = s11[2];
    ss[2 * stride + t] = s12[0];
    ss[3 * stride + t] = s12[1];
    ss[4 * stride + t] = s12[2];
    ss[5 * stride + t] = s22[0];
    ss[6 * stride + t] = s22[1];
}

// ---------------------------------------------------------------------------------------------
// Fused sub-iteration: stress update + velocity update in ONE kernel.
//
// A CTA owns a TX x TY tile of nodes (and the cells of the same indices).  One thread per cell
// of the (TX+1) x (TY+1) cells that touch its nodes computes the new stress (the left column and
// bottom row are recomputed redundantly: ~16 % extra flops, their old-stress reads hit L2; all
// 21 loads of a cell are issued before the first use), keeps the 7 stress coefficients the
// divergence needs in shared memory, stores all 9 to HBM for the owned cells only; after one
// barrier the first TX*TY threads update their nodes from shared memory.  The velocity kernel's
// re-read of S from HBM disappears.  Stress is ping-ponged (old -> new) so the redundant halo
// cells always see the OLD stress.
template <int TX, int TY> struct FusedCfg {
    static constexpr int CXS = TX + 1, CYS = TY + 1;
    static constexpr int NCELL = CXS * CYS;
    static constexpr int NT = (NCELL + 31) / 32 * 32; // one thread per (tile + halo) cell
    static constexpr int MINB = NT <= 352 ? 2 : 1; // resident CTAs per SM the register budget targets
};

template <int TX, int TY>
__global__ void __launch_bounds__(FusedCfg<TX, TY>::NT, FusedCfg<TX, TY>::MINB) mevp_fused_kernel(const __grid_constant__ MevpArgs A)
{
    using C = FusedCfg<TX, TY>;
    constexpr int CXS = C::CXS, NCELL = C::NCELL;
    __shared__ double ss[7 * NCELL];
    const int tid = threadIdx.x;
    const int i0 = blockIdx.x * TX;
    const int r0 = A.r0 + blockIdx.y * TY;
    {
        const int ni = i0 + tid % TX, nr = r0 + tid / TX;
        if (tid < TX * TY && ni < A.nx && nr < A.r1)
            prefetch_node(A, nr * A.nx + ni);
    }

    // ---- new stress of the (TX+1) x (TY+1) cells touching this tile's nodes
    if (tid < NCELL) {
        const int cy = tid / CXS - 1, cx = tid % CXS - 1;
        int gi = i0 + cx, gr = r0 + cy;
        if (gi < A.nx && gr < A.r1) { // else beyond the grid (partial tile): never read by a valid node
            if (gi < 0)
                gi += A.nx;
            const int gip = (gi == A.nx - 1) ? 0 : gi + 1;
            bool ghost_row = false;
            if (gr < NSDG_GHOST) {
                if (A.wrap)
                    gr += A.nyl;
                else
                    ghost_row = true; // strips: the first lower ghost row is kept current redundantly
            }
            const int k = gr * A.nx + gi;
            const int grp = row_above(A, gr);
            // issue every global load of this cell up front
            double so11[3], so12[3], so22[3], pr[4];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                so11[c] = __ldg(A.S11 + c * A.plane + k);
                so12[c] = __ldg(A.S12 + c * A.plane + k);
                so22[c] = __ldg(A.S22 + c * A.plane + k);
            }
#pragma unroll
            for (int g = 0; g < 4; ++g)
                pr[g] = __ldg(A.pg + g * A.plane + k);
            const double u00 = __ldg(A.u + k), u10 = __ldg(A.u + gr * A.nx + gip);
            const double u01 = __ldg(A.u + grp * A.nx + gi), u11 = __ldg(A.u + grp * A.nx + gip);
            const double v00 = __ldg(A.v + k), v10 = __ldg(A.v + gr * A.nx + gip);
            const double v01 = __ldg(A.v + grp * A.nx + gi), v11 = __ldg(A.v + grp * A.nx + gip);
            stress_relax(A, u00, u10, u01, u11, v00, v10, v01, v11, pr, so11, so12, so22);
            div_to_smem(ss, NCELL, tid, so11, so12, so22);
            if ((cx >= 0 && cy >= 0) || (ghost_row && cx >= 0)) {
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    A.S11n[c * A.plane + k] = so11[c];
                    A.S12n[c * A.plane + k] = so12[c];
                    A.S22n[c * A.plane + k] = so22[c];
                }
            }
        }
    }
    __syncthreads();

    // ---- velocity of the owned nodes
    const int tx = tid % TX, ty = tid / TX;
    const int ni = i0 + tx, nr = r0 + ty;
    if (tid >= TX * TY || ni >= A.nx || nr >= A.r1)
        return;
    const int n = nr * A.nx + ni;
    double un_ = 0., vn_ = 0.;
    if (!dirichlet(A, ni, nr - NSDG_GHOST)) {
        double dx, dy;
        div_from_smem(A, ss, NCELL, ty * CXS + tx, CXS, dx, dy);
        node_update(A, n, dx, dy, __ldg(A.u + n), __ldg(A.v + n), un_, vn_);
    }
    A.out0[n] = un_;
    A.out1[n] = vn_;
}

// ---------------------------------------------------------------------------------------------
// Pair-fused sub-iterations: TWO mEVP sub-iterations (p -> p+2) in one kernel (temporal blocking).
//
// The stress S (18 of the 32 doubles a sub-iteration moves per cell) and every per-step constant
// stream then cross HBM once per TWO sub-iterations: 32 doubles = 256 B per cell per PAIR.
// The price is redundant FP64 work on a two-cell-wide halo.  A CTA owns TX x TY nodes/cells:
//   A  stress p+1   on the (TX+3) x (TY+3) cells around the tile (global S^p, P_g, u^p)
//                   -> registers (9 coefficients + P_g of the thread's own cell) and the 7
//                      coefficients the divergence needs in shared memory
//   B  velocity p+1 on the (TX+2) x (TY+2) nodes  -> shared memory only
//   C  stress p+2   on the (TX+1) x (TY+1) cells, by the SAME thread that holds the cell's
//                   S^{p+1} and P_g in registers, u^{p+1} from shared memory; owned cells -> HBM
//   D  velocity p+2 on the TX x TY owned nodes -> HBM
// Every value is computed by the same formulas as in the single sub-iteration kernels, so a
// redundantly computed halo value is bit-identical to the owner's.  Single rank: periodic wrap
// in x and y by index arithmetic.  Strips: the two ghost rows on either side carry S^p, P_g,
// u^p and the node constants of the neighbours, refreshed once per pair (dyn.cu); the launch can
// be restricted to a range of tile rows so that the tiles next to the ghost rows run first and
// their halo messages travel while the interior tiles compute.
__device__ __forceinline__ int wrap_idx(int i, int n)
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

template <int TX, int TY> struct PairCfg {
    static constexpr int AX = TX + 3, AY = TY + 3; // cells of phase A
    static constexpr int BX = TX + 2, BY = TY + 2; // nodes of phase B
    static constexpr int NA = AX * AY, NB = BX * BY;
    static constexpr int NT = (NA + 31) / 32 * 32;
};

template <int TX, int TY, int MINB>
__global__ void __launch_bounds__(PairCfg<TX, TY>::NT, MINB) mevp_pair_kernel(const __grid_constant__ MevpArgs A)
{
    using C = PairCfg<TX, TY>;
    constexpr int AX = C::AX, BX = C::BX, NA = C::NA, NB = C::NB;
    __shared__ double ss[7 * NA]; // divergence coefficients of S^{p+1}, then of S^{p+2}
    __shared__ double su1[NB], sv1[NB]; // u^{p+1}, v^{p+1}
    const int tid = threadIdx.x;
    const int i0 = blockIdx.x * TX;
    const int r0 = A.r0 + blockIdx.y * TY; // plane row of the first owned node row
    // (an L1 prefetch of the node streams of phases B and D, as in mevp_fused_kernel, was measured
    // to cost 5 % at 16x16: profiles/r1p_prefetch_occupancy_sweep.txt)

    // ---- A: stress p+1 of cell (ax, ay) = global column i0-2+ax, interior row index r0-3+ay
    const int ax = tid % AX, ay = tid / AX;
    const bool in_a = tid < NA;
    double so11[3], so12[3], so22[3], pr[4];
    int k = 0;
    if (in_a) {
        const int gi = wrap_idx(i0 - 2 + ax, A.nx);
        const int gip = (gi == A.nx - 1) ? 0 : gi + 1;
        // plane row: periodic wrap on a single rank, ghost rows of the strip otherwise (rows past
        // a partial last tile row are clamped to valid memory and never used)
        int gr = r0 - 2 + ay, grp;
        if (A.wrap) {
            gr = NSDG_GHOST + wrap_idx(gr - NSDG_GHOST, A.nyl);
            grp = (gr == NSDG_GHOST + A.nyl - 1) ? NSDG_GHOST : gr + 1;
        } else {
            gr = min(gr, A.nyl + 2 * NSDG_GHOST - 1);
            grp = min(gr + 1, A.nyl + 2 * NSDG_GHOST - 1);
        }
        k = gr * A.nx + gi;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            so11[c] = __ldg(A.S11 + c * A.plane + k);
            so12[c] = __ldg(A.S12 + c * A.plane + k);
            so22[c] = __ldg(A.S22 + c * A.plane + k);
        }
#pragma unroll
        for (int g = 0; g < 4; ++g)
            pr[g] = __ldg(A.pg + g * A.plane + k);
        const double u00 = __ldg(A.u + k), u10 = __ldg(A.u + gr * A.nx + gip);
        const double u01 = __ldg(A.u + grp * A.nx + gi), u11 = __ldg(A.u + grp * A.nx + gip);
        const double v00 = __ldg(A.v + k), v10 = __ldg(A.v + gr * A.nx + gip);
        const double v01 = __ldg(A.v + grp * A.nx + gi), v11 = __ldg(A.v + grp * A.nx + gip);
        stress_relax(A, u00, u10, u01, u11, v00, v10, v01, v11, pr, so11, so12, so22);
        div_to_smem(ss, NA, tid, so11, so12, so22);
    }
    // The node streams of phase B are requested BEFORE the barrier (nothing of phase A's arithmetic
    // is live any more): their round trip to L2 overlaps the wait for the slowest warp and the
    // shared-memory gather below instead of starting after them (profiles/r1x_pair_ablation.txt).
    // ---- B: velocity p+1 of node (bx, by) = global column i0-1+bx, interior row r0-2+by
    const int bx = tid % BX, by = tid / BX;
    bool b_active = false;
    double nuo = 0., nvo = 0., ncx = 0., ncy = 0., nmdt = 0., nacd = 0., nub = 0., nvb = 0.;
    if (tid < NB) {
        const int ni = wrap_idx(i0 - 1 + bx, A.nx);
        int nrow = r0 - 1 + by; // plane row
        if (A.wrap)
            nrow = NSDG_GHOST + wrap_idx(nrow - NSDG_GHOST, A.nyl);
        else
            nrow = min(nrow, A.nyl + 2 * NSDG_GHOST - 1);
        b_active = !dirichlet(A, ni, nrow - NSDG_GHOST);
        if (b_active) {
            const int n = nrow * A.nx + ni;
            nuo = __ldg(A.uo + n);
            nvo = __ldg(A.vo + n);
            ncx = __ldg(A.nd + n);
            ncy = __ldg(A.nd + A.plane + n);
            nmdt = __ldg(A.nd + 2 * A.plane + n);
            nacd = __ldg(A.nd + 3 * A.plane + n);
            nub = __ldg(A.u + n);
            nvb = __ldg(A.v + n);
        }
    }
    __syncthreads();
    if (tid < NB) {
        double un_ = 0., vn_ = 0.;
        if (b_active) {
            double dx, dy;
            div_from_smem(A, ss, NA, by * AX + bx, AX, dx, dy);
            node_update_vals(A, nuo, nvo, ncx, ncy, nmdt, nacd, dx, dy, nub, nvb, un_, vn_);
        }
        su1[tid] = un_;
        sv1[tid] = vn_;
    }
    __syncthreads();

    // ---- C: stress p+2 of the cells 1 <= ax <= TX+1, 1 <= ay <= TY+1 (same thread, S^{p+1} and
    // P_g still in registers); ss is dead after B and is reused for S^{p+2}
    if (in_a && ax >= 1 && ax <= TX + 1 && ay >= 1 && ay <= TY + 1) {
        const int b00 = (ay - 1) * BX + (ax - 1);
        stress_relax(A, su1[b00], su1[b00 + 1], su1[b00 + BX], su1[b00 + BX + 1], sv1[b00], sv1[b00 + 1],
            sv1[b00 + BX], sv1[b00 + BX + 1], pr, so11, so12, so22);
        div_to_smem(ss, NA, tid, so11, so12, so22);
        // owned: unwrapped column i0+ax-2 in [i0, min(i0+TX, nx)), row r0+ay-2 in [r0, min(r0+TY, r1))
        if (ax >= 2 && ay >= 2 && i0 + ax - 2 < A.nx && r0 + ay - 2 < A.r1) {
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                A.S11n[c * A.plane + k] = so11[c];
                A.S12n[c * A.plane + k] = so12[c];
                A.S22n[c * A.plane + k] = so22[c];
            }
        }
    }
    // ---- D: velocity p+2 of the owned nodes; its node streams are requested before the barrier too
    const int tx = tid % TX, ty = tid / TX;
    const int dni = i0 + tx, dnr = r0 + ty;
    const bool d_owned = tid < TX * TY && dni < A.nx && dnr < A.r1;
    const bool d_active = d_owned && !dirichlet(A, dni, dnr - NSDG_GHOST);
    const int dn = dnr * A.nx + dni;
    if (d_active) {
        nuo = __ldg(A.uo + dn);
        nvo = __ldg(A.vo + dn);
        ncx = __ldg(A.nd + dn);
        ncy = __ldg(A.nd + A.plane + dn);
        nmdt = __ldg(A.nd + 2 * A.plane + dn);
        nacd = __ldg(A.nd + 3 * A.plane + dn);
    }
    __syncthreads();
    if (d_owned) {
        double un_ = 0., vn_ = 0.;
        if (d_active) {
            const int b = (ty + 1) * BX + tx + 1;
            double dx, dy;
            div_from_smem(A, ss, NA, (ty + 1) * AX + tx + 1, AX, dx, dy);
            node_update_vals(A, nuo, nvo, ncx, ncy, nmdt, nacd, dx, dy, su1[b], sv1[b], un_, vn_);
        }
        A.out0[dn] = un_;
        A.out1[dn] = vn_;
    }
}

// ---------------------------------------------------------------------------------------------
// Marching pair (default): the same two sub-iterations per launch, but a WARP owns 28 columns
// and marches up a band of rows; a lane is a column, neighbours in x come through warp shuffles
// and neighbours in y are the lane's own registers of the previous row.  No shared memory, no
// barriers, warps free-run.  Step s of the march works on row ra = band start - 2 + s:
//   A  stress p+1   of cell row ra      (32 lanes valid; u^p rows ra, ra+1)
//   B  velocity p+1 of node row ra      (lanes 1..31;   S^{p+1} rows ra-1, ra)
//   C  stress p+2   of cell row ra-1    (lanes 1..30;   u^{p+1} rows ra-1, ra)   -> HBM (owned)
//   D  velocity p+2 of node row ra-1    (lanes 2..30;   S^{p+2} rows ra-2, ra-1) -> HBM (owned)
// Lanes 2..29 are owned (28 columns = 7 whole 32-byte sectors per stored row).  Redundant work:
// 32/28 in x and (rows+3)/rows in y instead of the (19/16)^2, (18/16)^2, (17/16)^2 of the tiled
// kernel; a node needs two numbers per direction from the lane on its left (col_part).
// The 21 loads of row ra+1 (+2 on lane 31) are issued at the top of step ra and consumed one
// step later: a register software pipeline.  Three row records rotate through the roles
// "previous / current / being loaded" (the loop is unrolled by three), so a register that is
// the target of a load in flight is never copied.  ~240 registers, 8 warps per SM: measured
// faster than every variant with more resident warps (L2 prefetch instead of the register
// pipeline, idle values parked in shared memory: profiles/r1z_march_sweep.txt) -- a load that
// is not a full step ahead costs more than the occupancy buys.
// Same formulas, same operand order as every other sub-iteration kernel: bit-identical results.
__device__ __forceinline__ int plane_row(const MevpArgs& A, int r)
{
    if (A.wrap)
        return NSDG_GHOST + wrap_idx(r - NSDG_GHOST, A.nyl);
    return min(r, A.nyl + 2 * NSDG_GHOST - 1);
}

struct MarchRow {
    double s11[3], s12[3], s22[3], pg[4]; // S^p (A: -> S^{p+1}, C: -> S^{p+2}) and P_g/2 of cell row r
    double un, vn, unr, vnr; // u^p of node row r+1: own column, right neighbour
    double uo, vo, cx0, cy0, mdt, acd; // node constants of node row r
    double u1, v1; // u^{p+1} of node row r (phase B)
    int k; // r * nx + column
    bool act; // node (column, r) is not a Dirichlet node
};

// PEER: the ghost rows of S, u, v are written by the neighbour GPU while this kernel runs, so the
// loads must not go through the (non-coherent) L1 / read-only path: ld.global.cg reads at L2, the
// point of coherence for peer stores arriving over NVLink.  Every value is loaded once anyway.
template <bool CG> __device__ __forceinline__ double ldm(const double* p)
{
    return CG ? __ldcg(p) : __ldg(p);
}

// the plane row above plane row g: what plane_row(A, r + 1) is when g = plane_row(A, r), without
// the general modulo (whose never-taken division path would sit in every copy of the march step:
// the loop body is at the size of the 32 KB instruction-cache level)
__device__ __forceinline__ int plane_row_above(const MevpArgs& A, int g)
{
    if (A.wrap)
        return g + 1 == NSDG_GHOST + A.nyl ? NSDG_GHOST : g + 1;
    return min(g + 1, A.nyl + 2 * NSDG_GHOST - 1);
}
// gr = plane_row(A, r) of the row to load, grp = plane_row(A, r + 1)
template <bool CG>
__device__ __forceinline__ void march_load(const MevpArgs& A, int gr, int grp, int gi, int gip, int lane, MarchRow& R)
{
    const int k = gr * A.nx + gi;
    R.k = k;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        R.s11[c] = ldm<CG>(A.S11 + c * A.plane + k);
        R.s12[c] = ldm<CG>(A.S12 + c * A.plane + k);
        R.s22[c] = ldm<CG>(A.S22 + c * A.plane + k);
    }
#pragma unroll
    for (int g = 0; g < 4; ++g)
        R.pg[g] = __ldg(A.pg + g * A.plane + k);
    R.un = ldm<CG>(A.u + grp * A.nx + gi);
    R.vn = ldm<CG>(A.v + grp * A.nx + gi);
    R.unr = R.vnr = 0.;
    if (lane == 31) { // the other lanes take it from their right neighbour (from_right)
        R.unr = ldm<CG>(A.u + grp * A.nx + gip);
        R.vnr = ldm<CG>(A.v + grp * A.nx + gip);
    }
    R.uo = __ldg(A.uo + k);
    R.vo = __ldg(A.vo + k);
    R.cx0 = __ldg(A.nd + k);
    R.cy0 = __ldg(A.nd + A.plane + k);
    R.mdt = __ldg(A.nd + 2 * A.plane + k);
    R.acd = __ldg(A.nd + 3 * A.plane + k);
    R.act = !dirichlet(A, gi, gr - NSDG_GHOST);
}

// ---- peer halo: flags ----------------------------------------------------------------------
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p)
{
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// wait until the warps w-1, w, w+1 (ring in x) of the neighbour's boundary band have published
// `need`: they produced the ghost columns this warp reads, and they have finished reading the
// ghost columns this warp is about to overwrite in the other buffer set
// (all lanes poll and the loop exits on a warp vote: control flow stays warp-uniform, so the
// compiler keeps treating the warp as converged for the shuffles of the march)
__device__ __forceinline__ void halo_wait(const unsigned* f, int w, int nw, unsigned need, int lane)
{
    int q = w - 1 + min(lane, 2);
    q = q < 0 ? q + nw : (q >= nw ? q - nw : q);
    while (!__all_sync(0xffffffffu, (int)(ld_acquire_sys(f + q) - need) >= 0))
        __nanosleep(64);
}

__device__ __forceinline__ double from_right(double x, double edge, int lane)
{
    const double y = __shfl_down_sync(0xffffffffu, x, 1);
    return lane == 31 ? edge : y;
}
// the 7 divergence coefficients of a cell's stress, in the order Div7 / div_to_smem use
__device__ __forceinline__ void div7_of(double (&q)[7], const double (&s11)[3], const double (&s12)[3],
    const double (&s22)[3])
{
    q[0] = s11[0];
    q[1] = s11[2];
    q[2] = s12[0];
    q[3] = s12[1];
    q[4] = s12[2];
    q[5] = s22[0];
    q[6] = s22[1];
}
// divergence at the lane's node from its own column (below: lo, at the node row: hi) and the
// column part of the lane to the left
__device__ __forceinline__ void march_div(const MevpArgs& A, const double (&lo)[7], const double (&hi)[7],
    double& dx, double& dy)
{
    const ColPart c = col_part(A, lo, hi);
    dx = -(__shfl_up_sync(0xffffffffu, c.lx, 1) + c.rx);
    dy = -(__shfl_up_sync(0xffffffffu, c.ly, 1) + c.ry);
}

constexpr int MARCH_OWN = 28, MARCH_OFF = 2, MARCH_WPB = 4;

__device__ __forceinline__ void peer_store(const MevpArgs& A, int rs, int gi, bool own, bool up, bool dn,
    const double (&s11)[3], const double (&s12)[3], const double (&s22)[3], double u2, double v2)
{
    if (!own)
        return;
    if (up) {
        const int q = (rs - A.nyl) * A.nx + gi;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            A.up[0][c * A.plane_up + q] = s11[c];
            A.up[1][c * A.plane_up + q] = s12[c];
            A.up[2][c * A.plane_up + q] = s22[c];
        }
        A.up[3][q] = u2;
        A.up[4][q] = v2;
    }
    if (dn) {
        const int q = (rs + A.nyl_dn) * A.nx + gi;
        if (rs == NSDG_GHOST) {
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                A.dn[0][c * A.plane_dn + q] = s11[c];
                A.dn[1][c * A.plane_dn + q] = s12[c];
                A.dn[2][c * A.plane_dn + q] = s22[c];
            }
        }
        A.dn[3][q] = u2;
        A.dn[4][q] = v2;
    }
}

// ---- shared-memory ring of the marching kernel (RING variant) --------------------------------
// The register pipeline keeps one row (21 values per lane) in flight per warp: 8 warps x 5.4 KB per
// SM, which is what bounds the kernel.  Here the rows ahead are brought into a
// shared-memory ring by the bulk-copy engine instead: a CTA's four warps own 112 adjacent columns,
// so ONE cp.async.bulk per array and row moves the CTA's whole row segment (120 doubles: 112 + 2
// halo columns on either side + the right neighbour of the last lane, padded to 16 bytes), 21
// copies per row issued by 21 lanes of warp 0 in one instruction, completing on an mbarrier per
// stage.  RING_NS - 1 rows (3 x 20 KB per CTA) are in flight; a lane reads its 21 values of the
// next row from shared memory at the top of a step, exactly where the global loads were, so the
// arithmetic and its order are untouched (bit-identical).  Segments that wrap around in x (first
// and last CTA) are copied in pieces.  Needs an even nx (16-byte alignment of the pieces).
constexpr int RING_NS = 4, RING_ARR = 21;
template <int WPB> struct RingSmem {
    // the CTA's owned columns plus 16 on either side: the segment starts on a 128-byte boundary
    // of the row (c0 is a multiple of 112 = 7 x 16 columns) and is a whole number of 128-byte lines
    static constexpr int W = WPB * 28 + 32;
    static constexpr int LEFT = 16; // columns in front of the first owned one
    double data[RING_NS][RING_ARR][W];
    unsigned long long full[RING_NS], empty[RING_NS];
    const double* base[RING_ARR];
};

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity)
{
    unsigned ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok)
                     : "r"(smem_u32(bar)), "r"(parity)
                     : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(__cvta_generic_to_global(src)), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// warp 0 (all lanes enter): bring row r of the 21 arrays into stage st.  c0 = first owned column
// of the CTA; the segment starts two columns to its left.
template <int WPB>
__device__ __forceinline__ void ring_fill(const MevpArgs& A, RingSmem<WPB>& R, int r, int st, int c0, int lane)
{
    constexpr int RING_W = RingSmem<WPB>::W;
    if (lane == 0)
        mbar_arrive_expect_tx(&R.full[st], RING_ARR * RING_W * (unsigned)sizeof(double));
    __syncwarp();
    if (lane < RING_ARR) {
        const int row = plane_row(A, (lane == 13 || lane == 14) ? r + 1 : r); // u, v: node row r + 1
        const double* src = R.base[lane] + row * A.nx;
        double* dst = &R.data[st][lane][0];
        int pos = c0 - RingSmem<WPB>::LEFT, rem = RING_W, off = 0;
        while (rem > 0) {
            const int g = wrap_idx(pos, A.nx);
            const int n = min(rem, A.nx - g);
            bulk_g2s(dst + off, src + g, (unsigned)n * 8u, &R.full[st]);
            pos += n;
            off += n;
            rem -= n;
        }
    }
}

// the lane's 21 values of row r (the k-th row of the band's march) from the ring
template <int WPB>
__device__ __forceinline__ void ring_load(const MevpArgs& A, RingSmem<WPB>& R, int r, int k, int gi, int col, int lane, MarchRow& Z)
{
    constexpr int RING_W = RingSmem<WPB>::W;
    const int st = k % RING_NS;
    mbar_wait(&R.full[st], (unsigned)(k / RING_NS) & 1u);
    const int gr = plane_row(A, r);
    Z.k = gr * A.nx + gi;
    const double(*D)[RING_W] = R.data[st];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        Z.s11[c] = D[c][col];
        Z.s12[c] = D[3 + c][col];
        Z.s22[c] = D[6 + c][col];
    }
#pragma unroll
    for (int g = 0; g < 4; ++g)
        Z.pg[g] = D[9 + g][col];
    Z.un = D[13][col];
    Z.vn = D[14][col];
    Z.unr = Z.vnr = 0.;
    if (lane == 31) {
        Z.unr = D[13][col + 1];
        Z.vnr = D[14][col + 1];
    }
    Z.uo = D[15][col];
    Z.vo = D[16][col];
    Z.cx0 = D[17][col];
    Z.cy0 = D[18][col];
    Z.mdt = D[19][col];
    Z.acd = D[20][col];
    Z.act = !dirichlet(A, gi, gr - NSDG_GHOST);
    // every lane's reads of the stage are done before it is handed back to the copy engine
    __syncwarp();
    if (lane == 0)
        mbar_arrive(&R.empty[st]);
}

// ---- per-lane asynchronous copies (RING == 2) ------------------------------------------------
// The same idea without the copy engine and without any coupling between warps: every lane
// brings ITS OWN 21 values of the rows ahead into a private shared-memory slot with 8-byte
// cp.async (LDGSTS: no destination register, so the depth is not bounded by the register file),
// LR_D rows deep, and reads them back into the row record one step before their use.  A lane
// only ever reads what it copied itself: cp.async.wait_group orders that, no barrier is needed.
constexpr int LR_NS = 4, LR_D = LR_NS - 1;
struct LaneRing {
    double d[MARCH_WPB][LR_NS][RING_ARR][33]; // [warp][stage][array][lane (+1: right neighbour of lane 31)]
};
__device__ __forceinline__ void cpa8(double* dst, const double* src)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(__cvta_generic_to_global(src)) : "memory");
}
__device__ __forceinline__ void lane_ring_fill(const MevpArgs& A, double (*D)[33], int r, int gi, int gip, int lane, bool valid)
{
    if (valid) {
        const int gr = plane_row(A, r), grp = plane_row(A, r + 1);
        const int k = gr * A.nx + gi, kp = grp * A.nx + gi;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            cpa8(&D[c][lane], A.S11 + c * A.plane + k);
            cpa8(&D[3 + c][lane], A.S12 + c * A.plane + k);
            cpa8(&D[6 + c][lane], A.S22 + c * A.plane + k);
        }
#pragma unroll
        for (int g = 0; g < 4; ++g)
            cpa8(&D[9 + g][lane], A.pg + g * A.plane + k);
        cpa8(&D[13][lane], A.u + kp);
        cpa8(&D[14][lane], A.v + kp);
        if (lane == 31) {
            cpa8(&D[13][32], A.u + grp * A.nx + gip);
            cpa8(&D[14][32], A.v + grp * A.nx + gip);
        }
        cpa8(&D[15][lane], A.uo + k);
        cpa8(&D[16][lane], A.vo + k);
#pragma unroll
        for (int q = 0; q < 4; ++q)
            cpa8(&D[17 + q][lane], A.nd + q * A.plane + k);
    }
    asm volatile("cp.async.commit_group;" ::: "memory"); // (an empty group keeps the count uniform)
}
__device__ __forceinline__ void lane_ring_load(const MevpArgs& A, const double (*D)[33], int r, int gi, int lane, MarchRow& Z)
{
    asm volatile("cp.async.wait_group %0;" ::"n"(LR_D) : "memory");
    const int gr = plane_row(A, r);
    Z.k = gr * A.nx + gi;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        Z.s11[c] = D[c][lane];
        Z.s12[c] = D[3 + c][lane];
        Z.s22[c] = D[6 + c][lane];
    }
#pragma unroll
    for (int g = 0; g < 4; ++g)
        Z.pg[g] = D[9 + g][lane];
    Z.un = D[13][lane];
    Z.vn = D[14][lane];
    Z.unr = Z.vnr = 0.;
    if (lane == 31) {
        Z.unr = D[13][32];
        Z.vnr = D[14][32];
    }
    Z.uo = D[15][lane];
    Z.vo = D[16][lane];
    Z.cx0 = D[17][lane];
    Z.cy0 = D[18][lane];
    Z.mdt = D[19][lane];
    Z.acd = D[20][lane];
    Z.act = !dirichlet(A, gi, gr - NSDG_GHOST);
}

// One step of the march at row ra.  X: row ra-1 (S^{p+1}, P_g, node constants, u^{p+1}; its
// un/vn/unr/vnr are u^p of node row ra), Y: the loaded data of row ra, Z: target of the loads of
// row ra+1.  d2p: divergence coefficients of S^{p+2}, cell row ra-2.  own: this lane stores row
// ra-1.
// the q-th of the 21 arrays a march row is read from (order of the ring stages)
__device__ __forceinline__ const double* march_array(const MevpArgs& A, int q)
{
    return q < 3 ? A.S11 + q * A.plane : q < 6 ? A.S12 + (q - 3) * A.plane
        : q < 9  ? A.S22 + (q - 6) * A.plane
        : q < 13 ? A.pg + (q - 9) * A.plane
        : q == 13 ? A.u
        : q == 14 ? A.v
        : q == 15 ? A.uo
        : q == 16 ? A.vo
                  : A.nd + (q - 17) * A.plane;
}

// ---- L2 prefetch by the bulk-copy engine (RING == 3) -------------------------------------------
// The register pipeline stays as it is; in addition the bulk-copy engine is asked to bring rows
// ahead of the march into L2 (no destination, no barrier, nobody waits for it), so that the
// register loads one step ahead find their lines in L2.  What the experiments on the way here
// showed (profiles/r2z_march_prefetch.txt):
//  * cp.async.bulk.prefetch.L2 with per-lane addresses (21 lanes, one per array) removes the wait
//    (long-scoreboard stall 1.18 -> 0.11 per issue, +6 % DRAM bytes) but the instruction takes its
//    operands from UNIFORM registers: the compiler wraps it in a loop of ~11 instructions per
//    active lane, +40 % instructions, 1.04 against 0.84 ms;
//  * CTA-wide segments from one warp, or dealt out over the warps, save instructions but the warps
//    of a CTA free-run and drift apart: a warp prefetches relative to ITS position what the others
//    needed earlier or later (read hit rate 34-68 %, DRAM bytes up to x1.7);
//  * so every warp prefetches its own columns at its own pace, and to make that cheap the 21
//    arrays go as 9 TENSOR boxes (3 rows x the 3 or 4 planes of S11, S12, S22, P_g, the node
//    constants; u, v, u_o, v_o one plane each) every third step from one elected lane:
//    cp.async.bulk.prefetch.tensor.3d.L2, 9 instructions per three steps (+3.8 % instructions);
//  * boxes that start at the warp's first column (16-byte aligned) straddle three 128-byte lines
//    and overlap the neighbours': DRAM reads x1.5.  Boxes on 256-byte boundaries (warp w of the
//    CTA takes columns c0 + 32 w ...; nx % 16 == 0) bring that down to x1.13.
// What is left (this variant): 0.92 against 0.80 ms under ncu when the loop body was 33.1 KB
// (34.7 KB with the prefetch: past the 32 KB instruction-cache level, instruction-fetch stalls 0.14
// -> 0.36 per issue), and still 0.98 against 0.83 ms now that the body is 29.7 KB (31.1 KB with
// the prefetch, hit rate 99.6 %): DRAM reads x1.24 and the wait for loaded rows is where it was
// (1.66 against 1.74 per issue) -- box prefetches of 256-byte rows cost bandwidth and do not arrive
// ahead of the demand loads.  Kept as an opt-in arm (march_ring = 3, single rank); bit-identical.
// Needs an even nx (16-byte strides of the tensor maps).
constexpr int PF_MAPS = 9; // S11, S12, S22, pg, nd, u, v, uo, vo
struct alignas(64) PfMaps {
    CUtensorMap m[PF_MAPS];
};
__device__ __forceinline__ bool elect_one()
{
    unsigned pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tensor_prefetch_l2(const CUtensorMap* map, int x, int y)
{
    asm volatile("cp.async.bulk.prefetch.tensor.3d.L2.global.tile [%0, {%1, %2, %3}];" ::"l"(map), "r"(x), "r"(y), "r"(0) : "memory");
}
// rows [row, row + 3) of the plane (u, v: one row higher), columns [x, x + 32); rows past the plane's
// end and columns outside the row are skipped by the copy engine.  The whole warp arrives converged.
__device__ __forceinline__ void pf_rows(const PfMaps& M, int x, int row)
{
    if (elect_one()) {
#pragma unroll
        for (int q = 0; q < PF_MAPS; ++q)
            tensor_prefetch_l2(&M.m[q], x, (q == 5 || q == 6) ? row + 1 : row);
    }
}

template <bool PEER, bool CG, int RING, int WPB>
__device__ __forceinline__ void march_step(const MevpArgs& A, int ra, int gi, int gip, int lane, bool own,
    MarchRow& X, MarchRow& Y, MarchRow& Z, double (&d2p)[7], RingSmem<WPB>* R, int s, int nsteps, int c0, int col, int& grn)
{
    if (RING == 2) {
        // row s + 1 + LR_D of the march goes out to shared memory, row s + 1 comes in from it
        LaneRing* LR = reinterpret_cast<LaneRing*>(R);
        const int w = threadIdx.x >> 5;
        lane_ring_fill(A, LR->d[w][(s + 1 + LR_D) % LR_NS], ra + 1 + LR_D, gi, gip, lane, s + 1 + LR_D <= nsteps);
        lane_ring_load(A, LR->d[w][(s + 1) % LR_NS], ra + 1, gi, lane, Z);
    } else if (RING == 1) {
        // warp 0 refills the stage every warp has read in the previous step with the row RING_NS
        // ahead of it, then every warp takes the next row (march row s + 1) from the ring
        if (threadIdx.x < 32 && s + RING_NS <= nsteps) { // warp 0 of the CTA
            mbar_wait(&R->empty[s % RING_NS], (unsigned)(s / RING_NS) & 1u);
            ring_fill(A, *R, ra + RING_NS, s % RING_NS, c0, lane);
        }
        ring_load(A, *R, ra + 1, s + 1, gi, col, lane, Z);
    } else {
        // grn = plane row of row ra + 1, carried from step to step
        const int grp = plane_row_above(A, grn);
        march_load<CG>(A, grn, grp, gi, gip, lane, Z); // consumed by the next step (wasted on the last one)
        grn = grp;
    }

    // ---- A: stress p+1 of cell row ra
    Y.unr = from_right(Y.un, Y.unr, lane);
    Y.vnr = from_right(Y.vn, Y.vnr, lane);
    stress_relax(A, X.un, X.unr, Y.un, Y.unr, X.vn, X.vnr, Y.vn, Y.vnr, Y.pg, Y.s11, Y.s12, Y.s22);

    // ---- B: velocity p+1 of node row ra
    double lo[7], hi[7], dx, dy;
    div7_of(lo, X.s11, X.s12, X.s22);
    div7_of(hi, Y.s11, Y.s12, Y.s22);
    march_div(A, lo, hi, dx, dy);
    node_update_vals(A, Y.uo, Y.vo, Y.cx0, Y.cy0, Y.mdt, Y.acd, dx, dy, X.un, X.vn, Y.u1, Y.v1);
    if (!Y.act)
        Y.u1 = Y.v1 = 0.;

    // ---- C: stress p+2 of cell row ra-1
    {
        const double u1pr = __shfl_down_sync(0xffffffffu, X.u1, 1), v1pr = __shfl_down_sync(0xffffffffu, X.v1, 1);
        const double u1cr = __shfl_down_sync(0xffffffffu, Y.u1, 1), v1cr = __shfl_down_sync(0xffffffffu, Y.v1, 1);
        stress_relax(A, X.u1, u1pr, Y.u1, u1cr, X.v1, v1pr, Y.v1, v1cr, X.pg, X.s11, X.s12, X.s22);
    }
    if (own) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            A.S11n[c * A.plane + X.k] = X.s11[c];
            A.S12n[c * A.plane + X.k] = X.s12[c];
            A.S22n[c * A.plane + X.k] = X.s22[c];
        }
    }

    // ---- D: velocity p+2 of node row ra-1
    double u2, v2;
    div7_of(hi, X.s11, X.s12, X.s22);
    march_div(A, d2p, hi, dx, dy);
    node_update_vals(A, X.uo, X.vo, X.cx0, X.cy0, X.mdt, X.acd, dx, dy, X.u1, X.v1, u2, v2);
    if (!X.act)
        u2 = v2 = 0.;
    if (own) {
        A.out0[X.k] = u2;
        A.out1[X.k] = v2;
    }
    if (PEER) {
        // rows the ring neighbours need as ghost rows go straight into their planes over NVLink:
        // the two top rows (S, u, v) to the rank above, the bottom row (S) / two rows (u, v) to
        // the rank below.  Strips never wrap, so the stored row ra-1 is a plane row.  The row
        // tests are warp-uniform and almost never true: one uniform branch keeps the stores out
        // of the march's instruction stream.
        const int rs = ra - 1;
        const bool up = rs >= NSDG_GHOST + A.nyl - 2, dn = rs < NSDG_GHOST + 2;
        if (up || dn)
            peer_store(A, rs, gi, own, up, dn, X.s11, X.s12, X.s22, u2, v2);
    }
#pragma unroll
    for (int c = 0; c < 7; ++c)
        d2p[c] = hi[c];
}

template <bool PEER, bool CG, int RING, int WPB>
__device__ __forceinline__ void march_body(const MevpArgs& A, const PfMaps* M)
{
    constexpr int MARCH_WPB = WPB; // warps (28-column blocks) per CTA: 4, or 8 for the ring variant
    extern __shared__ __align__(128) unsigned char ring_raw[];
    RingSmem<WPB>* R = reinterpret_cast<RingSmem<WPB>*>(ring_raw);
    const int lane = threadIdx.x & 31;
    const int wg = blockIdx.x * MARCH_WPB + (threadIdx.x >> 5); // warp's column block
    const int cb = wg * MARCH_OWN; // first owned column
    const int c0 = blockIdx.x * MARCH_WPB * MARCH_OWN; // first owned column of the CTA
    // this lane's column in the CTA's row segment (ring variant)
    const int col = RingSmem<WPB>::LEFT - MARCH_OFF + (threadIdx.x >> 5) * MARCH_OWN + lane;
    if (RING == 1) {
        if (threadIdx.x == 0) {
            const int nw = (A.nx + MARCH_OWN - 1) / MARCH_OWN; // warps with columns, over the grid
            const int active = min(MARCH_WPB, nw - (int)blockIdx.x * MARCH_WPB);
            for (int q = 0; q < RING_NS; ++q) {
                mbar_init(&R->full[q], 1);
                mbar_init(&R->empty[q], (unsigned)active);
            }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        if (threadIdx.x < RING_ARR) {
            R->base[threadIdx.x] = march_array(A, threadIdx.x);
        }
        __syncthreads();
    }
    if (cb >= A.nx)
        return; // whole warp
    const int cu = cb - MARCH_OFF + lane; // unwrapped column of this lane
    const int gi = wrap_idx(cu, A.nx);
    const int gip = (gi == A.nx - 1) ? 0 : gi + 1;
    const bool own_col = lane >= MARCH_OFF && lane < MARCH_OFF + MARCH_OWN && cu < A.nx;
    // PEER: the two boundary bands go first (blockIdx.y 0 and 1), their rows travel to the
    // neighbours while the interior bands compute
    int band = blockIdx.y;
    if (PEER && band > 0)
        band = (band == 1) ? (int)gridDim.y - 1 : band - 1;
    const int b0 = A.r0 + band * A.band; // owned node/cell rows [b0, b1)
    // PEER: the last band takes a one-row remainder along (the host leaves that band out of the
    // grid), so that both rows the rank above needs are stored by the band that publishes them
    const int b1 = (PEER && band == (int)gridDim.y - 1) ? A.r1 : min(b0 + A.band, A.r1);
    const int nsteps = b1 - b0 + 3;
    if (PEER) {
        // ghost rows below / above this band: produced by the neighbours' previous pair
        if (b0 == A.r0)
            halo_wait(A.flag_mine, wg, A.nflagw, A.seq - 1, lane);
        if (b1 == A.r1)
            halo_wait(A.flag_mine + A.nflagw, wg, A.nflagw, A.seq - 1, lane);
    }

    MarchRow X = {}, Y, Z;
    X.mdt = 1.;
    int grn = plane_row(A, b0 - 1); // plane row of the row the next march_load takes
    double d2p[7] = { 0., 0., 0., 0., 0., 0., 0. };
    {
        const int g0 = plane_row(A, b0 - 2);
        X.un = ldm<CG>(A.u + g0 * A.nx + gi);
        X.vn = ldm<CG>(A.v + g0 * A.nx + gi);
        X.unr = from_right(X.un, ldm<CG>(A.u + g0 * A.nx + gip), lane);
        X.vnr = from_right(X.vn, ldm<CG>(A.v + g0 * A.nx + gip), lane);
    }
    if (RING == 2) {
        LaneRing* LR = reinterpret_cast<LaneRing*>(R);
        const int w = threadIdx.x >> 5;
        for (int k = 0; k <= LR_D; ++k)
            lane_ring_fill(A, LR->d[w][k], b0 - 2 + k, gi, gip, lane, k <= nsteps);
        lane_ring_load(A, LR->d[w][0], b0 - 2, gi, lane, Y);
    } else if (RING == 1) {
        if (PEER)
            asm volatile("fence.proxy.async;" ::: "memory"); // peer stores seen above -> bulk-copy reads
        if (threadIdx.x < 32) // warp 0: the first RING_NS rows of the march
            for (int k = 0; k < RING_NS && k <= nsteps; ++k)
                ring_fill(A, *R, b0 - 2 + k, k, c0, lane);
        ring_load(A, *R, b0 - 2, 0, gi, col, lane, Y);
    } else {
        march_load<CG>(A, plane_row(A, b0 - 2), grn, gi, gip, lane, Y);
    }
    // step s works on row ra = b0 - 2 + s and stores the owned row ra - 1 when 3 <= s; the first
    // steps run phases C / D on the zero-initialised records (finite values, never stored)
#define NSDG_STEP(s, X, Y, Z) \
    march_step<PEER, CG, RING, WPB>(A, b0 - 2 + (s), gi, gip, lane, own_col && (s) >= 3, X, Y, Z, d2p, R, (s), nsteps, c0, col, grn)
    int s = 0;
#pragma unroll 1
    for (; s + 2 < nsteps; s += 3) {
        if (RING == 3 && s + A.pf_dist < nsteps) // the rows steps s + pf_dist ... + 2 will load
            pf_rows(*M, A.pf_align ? c0 + 32 * (int)(threadIdx.x >> 5) : cb - MARCH_OFF, plane_row(A, b0 - 1 + s + A.pf_dist));
        NSDG_STEP(s, X, Y, Z);
        NSDG_STEP(s + 1, Y, Z, X);
        NSDG_STEP(s + 2, Z, X, Y);
    }
    if (s < nsteps) {
        NSDG_STEP(s, X, Y, Z);
        if (s + 1 < nsteps)
            NSDG_STEP(s + 1, Y, Z, X);
    }
#undef NSDG_STEP
    if (PEER && (b0 == A.r0 || b1 == A.r1)) {
        // publish: every lane's peer stores are ordered before the flag (fence by each storing
        // lane, warp barrier, release store by lane 0)
        __threadfence_system();
        __syncwarp();
        if (lane == 0) {
            if (b1 == A.r1)
                st_release_sys(A.flag_up + wg, A.seq); // slot [0] of the rank above: "from below"
            if (b0 == A.r0)
                st_release_sys(A.flag_dn + A.nflagw + wg, A.seq); // slot [1] of the rank below
        }
    }
}

template <bool PEER, bool CG, int RING, int WPB>
__global__ void __launch_bounds__(WPB * 32, 8 / WPB) mevp_march_kernel(const __grid_constant__ MevpArgs A)
{
    march_body<PEER, CG, RING, WPB>(A, nullptr);
}
// RING == 3: the tensor maps of the prefetch travel as a second kernel argument
template <bool PEER, bool CG, int WPB>
__global__ void __launch_bounds__(WPB * 32, 8 / WPB) mevp_march_pf_kernel(const __grid_constant__ MevpArgs A, const __grid_constant__ PfMaps M)
{
    march_body<PEER, CG, 3, WPB>(A, &M);
}

// tensor maps of the prefetch arm: one per array group, planes of (nyl + 4) x nx doubles, boxes of
// 32 columns x 3 rows x all planes of the group.  cuTensorMapEncodeTiled is a pure host function
// of the driver library, looked up once (the library does not link against libcuda).
typedef CUresult (*TensorMapEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
TensorMapEncodeFn tensor_map_encoder()
{
    static const TensorMapEncodeFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<TensorMapEncodeFn>(p);
    }();
    return fn;
}
bool pf_maps(const DynState* d, const MevpArgs& A, PfMaps& M)
{
    const TensorMapEncodeFn enc = tensor_map_encoder();
    if (!enc)
        return false;
    const double* base[PF_MAPS] = { A.S11, A.S12, A.S22, A.pg, A.nd, A.u, A.v, A.uo, A.vo };
    const int planes[PF_MAPS] = { 3, 3, 3, 4, 4, 1, 1, 1, 1 };
    for (int q = 0; q < PF_MAPS; ++q) {
        const cuuint64_t dim[3] = { (cuuint64_t)d->nx, (cuuint64_t)d->rows, (cuuint64_t)planes[q] };
        const cuuint64_t stride[2] = { (cuuint64_t)d->nx * 8u, (cuuint64_t)d->plane * 8u };
        const cuuint32_t box[3] = { 32u, 3u, (cuuint32_t)planes[q] }, estr[3] = { 1u, 1u, 1u };
        if (enc(&M.m[q], CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<double*>(base[q]), dim, stride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)
            != CUDA_SUCCESS)
            return false;
    }
    return true;
}

MevpArgs base_args(DynState* d, double dt)
{
    MevpArgs A = {};
    const nsdg_mevp_params& P = d->mp;
    A.nx = (int)d->nx;
    A.nyl = (int)d->nyl;
    A.plane = (int)d->plane;
    A.j0 = (int)d->cfg.j0;
    A.ny_global = (int)d->cfg.ny_global;
    A.wrap = d->wrap ? 1 : 0;
    A.nc = d->nc;
    A.bc = P.bc;
    A.ihx = 1. / d->cfg.hx;
    A.ihy = 1. / d->cfg.hy;
    A.ie2 = 1. / (P.ellipse_e * P.ellipse_e);
    A.dmin2 = P.delta_min * P.delta_min;
    A.ia = 1. / P.alpha;
    A.hx2 = 0.5 * A.ihx;
    A.hy2 = 0.5 * A.ihy;
    A.hx4 = 0.25 * A.ihx;
    A.hy4 = 0.25 * A.ihy;
    A.ome = 1. - A.ie2;
    A.e2x2 = 2. * A.ie2;
    A.oma = 1. - A.ia;
    A.ia4 = 0.25 * A.ia;
    A.ia3g = 3. * GQ * A.ia;
    A.c1 = 0.5 * A.ihx;
    A.c2 = A.ihx * (1. / 12.);
    A.c3 = 0.5 * A.ihy;
    A.c4 = A.ihy * (1. / 12.);
    A.beta = P.beta;
    A.beta1 = 1. + P.beta;
    A.dtf = dt * P.f_coriolis;
    A.p_star = P.p_star;
    A.c_star = P.c_star;
    A.rho_ice = P.rho_ice;
    A.h_min = P.h_min;
    A.idt = dt > 0 ? 1. / dt : 0.;
    A.car = P.c_atm * P.rho_atm;
    A.cor = P.c_ocean * P.rho_ocean;
    A.fcor = P.f_coriolis;
    A.H = d->f[NSDG_DF_H];
    A.A = d->f[NSDG_DF_A];
    A.u = d->f[NSDG_DF_U];
    A.v = d->f[NSDG_DF_V];
    A.uo = d->f[NSDG_DF_UOCN];
    A.vo = d->f[NSDG_DF_VOCN];
    A.ua = d->f[NSDG_DF_UATM];
    A.va = d->f[NSDG_DF_VATM];
    A.pg = d->pg;
    A.nd = d->nd;
    A.S11 = d->f[NSDG_DF_S11];
    A.S12 = d->f[NSDG_DF_S12];
    A.S22 = d->f[NSDG_DF_S22];
    return A;
}

void set_stress(MevpArgs& A, double* const* Sold, double* const* Snew)
{
    A.S11 = Sold[0];
    A.S12 = Sold[1];
    A.S22 = Sold[2];
    A.S11n = Snew[0];
    A.S12n = Snew[1];
    A.S22n = Snew[2];
}

} // namespace

// Requires (strips): two lower and one upper ghost row of H, A current.  Single rank: kernels wrap.
int nsdg_mevp_prepare(nsdg_ctx* ctx, double dt)
{
    DynState* d = ctx->dyn;
    MevpArgs A = base_args(d, dt);
    const int G = NSDG_GHOST;
    // strips: strength also on the ghost cell rows the (pair-)fused kernels recompute
    A.r0 = d->wrap ? G : 0;
    A.r1 = G + (int)d->nyl + (d->wrap ? 0 : 1);
    A.out0 = d->pg;
    dim3 block(256), grid((unsigned)((d->nx + 255) / 256), (unsigned)(A.r1 - A.r0));
    NSDG_KERNEL(ctx, NSDG_K_PRESSURE, pressure_kernel<<<grid, block, 0, ctx->stream>>>(A));
    // node constants on the owned node rows, plus (strips) the first ghost node row on either
    // side, which phase B of the pair kernel updates redundantly
    A.r0 = d->wrap ? G : G - 1;
    A.r1 = G + (int)d->nyl + (d->wrap ? 0 : 1);
    A.out0 = d->nd;
    grid = dim3((unsigned)((d->nx + 255) / 256), (unsigned)(A.r1 - A.r0));
    NSDG_KERNEL(ctx, NSDG_K_NODAL, nodal_kernel<<<grid, block, 0, ctx->stream>>>(A));
    return NSDG_OK;
}

int nsdg_mevp_stress(nsdg_ctx* ctx, double dt, const double* u, const double* v, double* const* S,
    int64_t r0, int64_t r1)
{
    DynState* d = ctx->dyn;
    if (r1 <= r0)
        return NSDG_OK;
    MevpArgs A = base_args(d, dt);
    A.u = u;
    A.v = v;
    set_stress(A, S, S); // in place: a cell reads and writes only its own coefficients
    A.r0 = (int)r0;
    A.r1 = (int)r1;
    dim3 block(64, 4), grid((unsigned)((d->nx + 63) / 64), (unsigned)((r1 - r0 + 3) / 4));
    NSDG_KERNEL(ctx, NSDG_K_STRESS, stress_kernel<<<grid, block, 0, ctx->stream>>>(A));
    return NSDG_OK;
}

int nsdg_mevp_velocity(nsdg_ctx* ctx, double dt, const double* u, const double* v, double* unew,
    double* vnew, double* const* S, int64_t r0, int64_t r1)
{
    DynState* d = ctx->dyn;
    if (r1 <= r0)
        return NSDG_OK;
    MevpArgs A = base_args(d, dt);
    A.u = u;
    A.v = v;
    set_stress(A, S, S);
    A.out0 = unew;
    A.out1 = vnew;
    A.r0 = (int)r0;
    A.r1 = (int)r1;
    dim3 block(64, 4), grid((unsigned)((d->nx + 63) / 64), (unsigned)((r1 - r0 + 3) / 4));
    NSDG_KERNEL(ctx, NSDG_K_VELOCITY, velocity_kernel<<<grid, block, 0, ctx->stream>>>(A));
    return NSDG_OK;
}

// one fused sub-iteration: (u,v,S) -> (unew,vnew,Snew) on node rows [r0,r1)
int nsdg_mevp_fused(nsdg_ctx* ctx, double dt, const double* u, const double* v, double* unew,
    double* vnew, double* const* Sold, double* const* Snew, int64_t r0, int64_t r1)
{
    DynState* d = ctx->dyn;
    if (r1 <= r0)
        return NSDG_OK;
    MevpArgs A = base_args(d, dt);
    A.u = u;
    A.v = v;
    A.out0 = unew;
    A.out1 = vnew;
    set_stress(A, Sold, Snew);
    A.r0 = (int)r0;
    A.r1 = (int)r1;
    // tile shape: 32x8 by default; nsdg_dyn_tuning.fused_tile selects another compiled variant
    const int tile = d->tune.fused_tile;
#define NSDG_FUSED_LAUNCH(TX, TY)                                                                 \
    do {                                                                                          \
        dim3 block(FusedCfg<TX, TY>::NT),                                                         \
            grid((unsigned)((d->nx + TX - 1) / TX), (unsigned)((r1 - r0 + TY - 1) / TY));         \
        NSDG_KERNEL(ctx, NSDG_K_MEVP_FUSED,                                                       \
            (mevp_fused_kernel<TX, TY><<<grid, block, 0, ctx->stream>>>(A)));                     \
    } while (0)
    switch (tile) {
    case 1:
        NSDG_FUSED_LAUNCH(64, 4);
        break;
    case 2:
        NSDG_FUSED_LAUNCH(32, 16);
        break;
    case 3:
        NSDG_FUSED_LAUNCH(64, 8);
        break;
    case 4:
        NSDG_FUSED_LAUNCH(16, 16);
        break;
    default:
        NSDG_FUSED_LAUNCH(32, 8);
    }
#undef NSDG_FUSED_LAUNCH
    return NSDG_OK;
}

// two fused sub-iterations: (u,v,S) at p -> (unew,vnew,Snew) at p+2; single rank (wrap) only
namespace {

// Band height of the marching kernel for a strip of nrows x nx cells.  All CTAs of a launch cost
// the same (rows + 3 march steps) except those of the last, shorter band, and 2 CTAs are resident
// per SM, so the launch takes whole waves: the cost model is greedy list scheduling of the CTAs in
// launch order on the CTA slots.  Taller bands have fewer redundant rows; the best height puts the
// CTAs into full waves (4096^2 on 148 SMs: 8 bands of 512 rows = 296 CTAs = one wave, 0.844 ms
// against 0.878 ms for 64 bands of 64 rows, profiles/r1z_band_rows.txt).  Strips (multi): the
// halo exchange starts when the boundary bands are done, so a band may take at most a quarter of
// the launch (4096^2 per rank: 32 bands of 128 rows = four waves; leaving CTA slots out of the
// model for the exchange kernels was measured slower).
int march_auto_rows(int64_t nrows, int64_t nx, int sm_count, bool multi)
{
    const int64_t gx = ((nx + MARCH_OWN - 1) / MARCH_OWN + MARCH_WPB - 1) / MARCH_WPB;
    const int64_t slots = std::max<int64_t>(2 * sm_count, 1);
    int best = (int)std::min<int64_t>(nrows, 64);
    double best_t = 1e300;
    for (int64_t nb = 1; nb <= nrows; ++nb) {
        const int64_t ry = (nrows + nb - 1) / nb;
        if (ry > 1024)
            continue;
        if (ry < 32)
            break;
        const int64_t full = nrows / ry, rem = nrows - full * ry;
        const double c = (double)(ry + 3), cr = (double)(rem + 3);
        const int64_t nfull = full * gx, q = nfull / slots, r = nfull % slots;
        double t = (double)(q + (r > 0)) * c;
        if (rem > 0)
            t = (gx <= slots - r) ? std::max(t, (double)q * c + cr) : (double)(q + 1) * c + cr;
        if (multi && (t < 4. * c || full + (rem > 0) < 4))
            continue;
        if (t < best_t) {
            best_t = t;
            best = (int)ry;
        }
    }
    return std::max(best, 1);
}

// peer halo possible for this strip (mapping exists, knob not off)
bool peer_possible(const DynState* d) { return !d->wrap && d->peer_ok && d->tune.halo_peer >= 0; }

// rows per band of the marching kernel; 0 = tiled pair kernel (nsdg_dyn_tuning.march_rows)
int march_rows(nsdg_ctx* ctx)
{
    DynState* d = ctx->dyn;
    if (d->tune.march_rows != 0)
        return d->tune.march_rows > 0 ? d->tune.march_rows : 0;
    // cached for the last geometry asked for (pure function of these four numbers)
    static thread_local int64_t c_ny = -1, c_nx = -1;
    static thread_local int c_sm = -1, c_strip = -1, c_ry = 64;
    // the "at least four waves" rule of strips exists for the NCCL exchange beside the interior
    // launch; with peer stores a strip takes the same band height as a single rank
    const int strip = (d->wrap || peer_possible(d)) ? 0 : 1;
    if (c_ny != d->nyl || c_nx != d->nx || c_sm != ctx->sm_count || c_strip != strip) {
        c_ny = d->nyl;
        c_nx = d->nx;
        c_sm = ctx->sm_count;
        c_strip = strip;
        c_ry = march_auto_rows(d->nyl, d->nx, ctx->sm_count, strip != 0);
    }
    return c_ry;
}
} // namespace

extern "C" int nsdg_mevp_band_rows(int64_t nrows, int64_t nx, int32_t sm_count, int32_t strip)
{
    if (nrows <= 0 || nx <= 0 || sm_count <= 0)
        return NSDG_EINVAL;
    return march_auto_rows(nrows, nx, sm_count, strip != 0);
}

int nsdg_mevp_pair_tile_rows(nsdg_ctx* ctx)
{
    static const int ty[6] = { 16, 8, 16, 16, 8, 12 };
    const int ry = march_rows(ctx);
    return ry ? ry : ty[ctx->dyn->tune.pair_tile];
}

namespace {
// one warp: returns when every flag of this rank (both directions) has reached `need`, i.e. all
// rows the ring neighbours store into this rank's ghost rows for pair `need` have landed
__global__ void halo_drain_kernel(const unsigned* flags, int n, unsigned need)
{
    const int lane = threadIdx.x;
    for (int base = 0; base < n; base += 32) {
        const int q = min(base + lane, n - 1);
        while (!__all_sync(0xffffffffu, (int)(ld_acquire_sys(flags + q) - need) >= 0))
            __nanosleep(128);
    }
}
} // namespace

// After the last peer pair of a sub-cycle: whatever follows on the stream (a fused sub-iteration of
// an odd count, a download, the release of the slab) must not overtake the neighbours' stores.
int nsdg_mevp_peer_drain(nsdg_ctx* ctx)
{
    DynState* d = ctx->dyn;
    NSDG_KERNEL(ctx, NSDG_K_OTHER, (halo_drain_kernel<<<1, 32, 0, ctx->stream>>>(d->flags, 2 * d->nflagw, d->halo_seq)));
    return NSDG_OK;
}

// (bands of at least two rows: the bottom band publishes the two bottom rows)
bool nsdg_mevp_peer_halo(nsdg_ctx* ctx) { return ctx->dyn && peer_possible(ctx->dyn) && march_rows(ctx) >= 2; }

int nsdg_mevp_pair(nsdg_ctx* ctx, double dt, const double* u, const double* v, double* unew,
    double* vnew, double* const* Sold, double* const* Snew, int64_t r0, int64_t r1, cudaStream_t st,
    unsigned peer_seq)
{
    DynState* d = ctx->dyn;
    if (!st)
        st = ctx->stream;
    MevpArgs A = base_args(d, dt);
    A.u = u;
    A.v = v;
    A.out0 = unew;
    A.out1 = vnew;
    set_stress(A, Sold, Snew);
    A.r0 = (int)r0;
    A.r1 = (int)r1;
    const int tile = d->tune.pair_tile;
    const int nrows = (int)(r1 - r0);
    if (nrows <= 0)
        return NSDG_OK;
    if (const int ry = march_rows(ctx)) {
        const int nwarps = (int)((d->nx + MARCH_OWN - 1) / MARCH_OWN);
        A.band = ry;
        dim3 block(MARCH_WPB * 32), grid((unsigned)((nwarps + MARCH_WPB - 1) / MARCH_WPB), (unsigned)((nrows + ry - 1) / ry));
        // shared-memory ring variant (nsdg_dyn_tuning.march_ring; needs an even nx): CTAs of eight
        // warps, one per SM, so that a bulk copy moves 1.9 KB
        const bool ring = d->tune.march_ring == 1 && d->nx % 16 == 0;
        const bool lring = d->tune.march_ring == 2 && !peer_seq; // per-lane cp.async variant (single launch path)
        if (lring) {
            if (!ctx->lring_attr_set) {
                NSDG_CUDA(ctx, cudaFuncSetAttribute(mevp_march_kernel<false, false, 2, MARCH_WPB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LaneRing)));
                ctx->lring_attr_set = true;
            }
            if (st == ctx->stream) {
                NSDG_KERNEL(ctx, NSDG_K_MEVP_MARCH, (mevp_march_kernel<false, false, 2, MARCH_WPB><<<grid, block, sizeof(LaneRing), st>>>(A)));
            } else {
                mevp_march_kernel<false, false, 2, MARCH_WPB><<<grid, block, sizeof(LaneRing), st>>>(A);
                NSDG_LAUNCHED(ctx);
            }
            return NSDG_OK;
        }
        if (d->tune.march_ring == 3 && !peer_seq && d->nx % 2 == 0) { // L2 prefetch arm
            PfMaps M;
            NSDG_REQUIRE(ctx, pf_maps(d, A, M), NSDG_ECUDA, "cuTensorMapEncodeTiled unavailable or failed (march_ring = 3)");
            A.pf_dist = 2; // rows ahead (1 ... 4 measured: 0.985, 1.000, 1.016, 1.052 ms per launch)
            A.pf_align = d->nx % 16 == 0; // boxes on 256-byte boundaries of the row: no line fetched twice
            if (st == ctx->stream) {
                NSDG_KERNEL(ctx, NSDG_K_MEVP_MARCH, (mevp_march_pf_kernel<false, false, MARCH_WPB><<<grid, block, 0, st>>>(A, M)));
            } else {
                mevp_march_pf_kernel<false, false, MARCH_WPB><<<grid, block, 0, st>>>(A, M);
                NSDG_LAUNCHED(ctx);
            }
            return NSDG_OK;
        }
        constexpr int RWPB = 4;
        if (ring) {
            bool& attr_set = ctx->ring_attr_set; // per device: the attribute belongs to the context's GPU
            if (!attr_set) {
                NSDG_CUDA(ctx, cudaFuncSetAttribute(mevp_march_kernel<false, false, 1, RWPB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RingSmem<RWPB>)));
                NSDG_CUDA(ctx, cudaFuncSetAttribute(mevp_march_kernel<true, true, 1, RWPB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RingSmem<RWPB>)));
                attr_set = true;
            }
            block = dim3(RWPB * 32);
            grid.x = (unsigned)((nwarps + RWPB - 1) / RWPB);
        }
        if (peer_seq) {
            // the two top rows go to the rank above and are published by the LAST band: a
            // remainder band of one row is merged into the band below it
            if (grid.y > 1 && nrows - (int)(grid.y - 1) * ry < 2)
                grid.y -= 1;
            NSDG_REQUIRE(ctx, peer_possible(d) && r0 == NSDG_GHOST && r1 == NSDG_GHOST + d->nyl && nwarps == d->nflagw,
                NSDG_ESTATE, "peer halo launch needs the mapped neighbours and the whole strip");
            // the neighbours' copies of the five output arrays sit at the same plane offsets of
            // their slabs (every rank performs the same buffer swaps)
            double* const outs[5] = { A.S11n, A.S12n, A.S22n, A.out0, A.out1 };
            for (int q = 0; q < 5; ++q) {
                const int64_t off = (outs[q] - d->slab) / d->plane;
                NSDG_REQUIRE(ctx, outs[q] >= d->slab && off < NSDG_SLAB_PLANES, NSDG_ESTATE, "output array outside the slab");
                A.up[q] = d->up.slab + off * d->up.plane;
                A.dn[q] = d->dn.slab + off * d->dn.plane;
            }
            A.plane_up = (int)d->up.plane;
            A.plane_dn = (int)d->dn.plane;
            A.nyl_dn = (int)d->dn.nyl;
            A.nflagw = d->nflagw;
            A.flag_mine = d->flags;
            A.flag_up = d->up.flags;
            A.flag_dn = d->dn.flags;
            A.seq = peer_seq;
            if (ring)
                NSDG_KERNEL(ctx, NSDG_K_MEVP_MARCH, (mevp_march_kernel<true, true, 1, RWPB><<<grid, block, sizeof(RingSmem<RWPB>), st>>>(A)));
            else
                NSDG_KERNEL(ctx, NSDG_K_MEVP_MARCH, (mevp_march_kernel<true, true, 0, MARCH_WPB><<<grid, block, 0, st>>>(A)));
        } else if (ring) {
            if (st == ctx->stream) {
                NSDG_KERNEL(ctx, NSDG_K_MEVP_MARCH, (mevp_march_kernel<false, false, 1, RWPB><<<grid, block, sizeof(RingSmem<RWPB>), st>>>(A)));
            } else {
                mevp_march_kernel<false, false, 1, RWPB><<<grid, block, sizeof(RingSmem<RWPB>), st>>>(A);
                NSDG_LAUNCHED(ctx);
            }
        } else if (st == ctx->stream) {
            NSDG_KERNEL(ctx, NSDG_K_MEVP_MARCH, (mevp_march_kernel<false, false, 0, MARCH_WPB><<<grid, block, 0, st>>>(A)));
        } else { /* the per-kernel profiler brackets the compute stream only */
            mevp_march_kernel<false, false, 0, MARCH_WPB><<<grid, block, 0, st>>>(A);
            NSDG_LAUNCHED(ctx);
        }
        return NSDG_OK;
    }
#define NSDG_PAIR_LAUNCH(TX, TY, MINB)                                                            \
    do {                                                                                          \
        dim3 block(PairCfg<TX, TY>::NT),                                                          \
            grid((unsigned)((d->nx + TX - 1) / TX), (unsigned)((nrows + TY - 1) / TY));           \
        if (st == ctx->stream) {                                                                  \
            NSDG_KERNEL(ctx, NSDG_K_MEVP_PAIR, (mevp_pair_kernel<TX, TY, MINB><<<grid, block, 0, st>>>(A))); \
        } else { /* the per-kernel profiler brackets the compute stream only */                   \
            mevp_pair_kernel<TX, TY, MINB><<<grid, block, 0, st>>>(A);                            \
            NSDG_LAUNCHED(ctx);                                                                   \
        }                                                                                         \
    } while (0)
    switch (tile) {
    case 1:
        NSDG_PAIR_LAUNCH(32, 8, 2);
        break;
    case 2:
        NSDG_PAIR_LAUNCH(32, 16, 1);
        break;
    case 3:
        NSDG_PAIR_LAUNCH(16, 16, 3);
        break;
    case 4:
        NSDG_PAIR_LAUNCH(16, 8, 4);
        break;
    case 5:
        NSDG_PAIR_LAUNCH(16, 12, 3);
        break;
    default:
        NSDG_PAIR_LAUNCH(16, 16, 2);
    }
#undef NSDG_PAIR_LAUNCH
    return NSDG_OK;
}
