// mEVP stress / velocity sub-cycle kernels.
//
// NOT IN THE REFERENCE SNAPSHOT (CMakeLists.txt:43-46 comments "dynamics" out; the only trace
// is the comment "Ice momentum fluxes are handled by the dynamics" at
// physics/src/modules/NextsimPhysics.cpp:124,129).  Scheme: oracle/mevp.c and DESIGN.md.
//
// Per sub-iteration two launches:
//   stress   one thread per cell: strain rates of the bilinear CG1 velocity (exact dG1
//            coefficients), viscous-plastic law at the 2x2 Gauss points, L2 projection to dG1,
//            relaxation S += (proj - S)/alpha.     reads S(9) u,v(2) P_g(4), writes S(9)
//   velocity one thread per node: -(S, grad phi) gathered from the 4 adjacent cells with the
//            analytic dG1 x Q1 integrals, implicit ocean drag, explicit Coriolis.
//            reads S(7 of 9 used) u,v u^n,v^n tau_a(2) u_o(2) H_n,A_n, writes u,v
// Both are HBM-bound streams over component-major planes (coalesced, neighbours via L1/L2).
#include "dyn.h"
#include <cuda.h> // CUtensorMap (types only; the encoder is resolved through the runtime)
#include <cstdlib>
#include <string>

namespace {

constexpr double GQ = 0.28867513459481287; // 1/(2 sqrt 3)

struct MevpArgs {
    nsdg_mevp_params P;
    int64_t nx, nyl, plane, r0, r1, j0, ny_global;
    double hx, hy, dt;
    int wrap;
    int nc;
    int selfhalo; // bulk-copy kernel, single rank: also write u,v rows nyl -> 0 and 1 -> nyl+1
    const double* H;
    const double* A;
    const double* u;
    const double* v;
    const double* un;
    const double* vn;
    const double* hn;
    const double* an;
    const double* taux;
    const double* tauy;
    const double* uo;
    const double* vo;
    const double* ua;
    const double* va;
    const double* pg;
    double* S11;
    double* S12;
    double* S22;
    double* S11n; // fused kernel: new stress (ping-pong partner)
    double* S12n;
    double* S22n;
    double* out0;
    double* out1;
    double* out2;
    double* out3;
    double* out4;
    double* out5;
};

__device__ __forceinline__ int64_t row_below(const MevpArgs& A, int64_t r)
{
    int64_t rm = r - 1;
    if (A.wrap && rm < 1)
        rm = A.nyl;
    return rm;
}
__device__ __forceinline__ int64_t row_above(const MevpArgs& A, int64_t r)
{
    int64_t rp = r + 1;
    if (A.wrap && rp > A.nyl)
        rp = 1;
    return rp;
}

__device__ __forceinline__ double eval_dg(const double* f, int64_t plane, int64_t k, int nc, double a, double b)
{
    double s = __ldg(f + k);
    if (nc >= 3)
        s += __ldg(f + plane + k) * a + __ldg(f + 2 * plane + k) * b;
    if (nc >= 6)
        s += __ldg(f + 3 * plane + k) * (a * a - 1. / 12.) + __ldg(f + 4 * plane + k) * (b * b - 1. / 12.)
            + __ldg(f + 5 * plane + k) * a * b;
    return s;
}

// ice strength at the Gauss points: out0 = pg (4 planes), g = 2*b + a
__global__ void __launch_bounds__(256) pressure_kernel(const __grid_constant__ MevpArgs A)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t r = A.r0 + blockIdx.y;
    if (i >= A.nx || r >= A.r1)
        return;
    const int64_t k = r * A.nx + i;
#pragma unroll
    for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int a = 0; a < 2; ++a) {
            const double xa = a ? GQ : -GQ, yb = b ? GQ : -GQ;
            double h = eval_dg(A.H, A.plane, k, A.nc, xa, yb);
            double c = eval_dg(A.A, A.plane, k, A.nc, xa, yb);
            A.out0[(2 * b + a) * A.plane + k] = A.P.p_star * h * exp(-A.P.c_star * (1. - c));
        }
}

// nodal H, A (mean of the 4 adjacent cell means), wind stress, copy of the start velocity
// out0=hn out1=an out2=taux out3=tauy out4=un out5=vn
__global__ void __launch_bounds__(256) nodal_kernel(const __grid_constant__ MevpArgs A)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t r = A.r0 + blockIdx.y;
    if (i >= A.nx || r >= A.r1)
        return;
    const int64_t rm = row_below(A, r);
    const int64_t im = (i == 0) ? A.nx - 1 : i - 1;
    const int64_t n = r * A.nx + i;
    A.out0[n] = 0.25 * (__ldg(A.H + rm * A.nx + im) + __ldg(A.H + rm * A.nx + i) + __ldg(A.H + r * A.nx + im) + __ldg(A.H + n));
    A.out1[n] = 0.25 * (__ldg(A.A + rm * A.nx + im) + __ldg(A.A + rm * A.nx + i) + __ldg(A.A + r * A.nx + im) + __ldg(A.A + n));
    const double ua = __ldg(A.ua + n), va = __ldg(A.va + n);
    const double wa = sqrt(ua * ua + va * va);
    A.out2[n] = A.P.c_atm * A.P.rho_atm * wa * ua;
    A.out3[n] = A.P.c_atm * A.P.rho_atm * wa * va;
    A.out4[n] = __ldg(A.u + n);
    A.out5[n] = __ldg(A.v + n);
}

__global__ void __launch_bounds__(256) stress_kernel(const __grid_constant__ MevpArgs A)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t r = A.r0 + (int64_t)blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= A.nx || r >= A.r1)
        return;
    const int64_t rp = row_above(A, r);
    const int64_t ip = (i == A.nx - 1) ? 0 : i + 1;
    const int64_t k = r * A.nx + i;
    const double u00 = __ldg(A.u + k), u10 = __ldg(A.u + r * A.nx + ip);
    const double u01 = __ldg(A.u + rp * A.nx + i), u11 = __ldg(A.u + rp * A.nx + ip);
    const double v00 = __ldg(A.v + k), v10 = __ldg(A.v + r * A.nx + ip);
    const double v01 = __ldg(A.v + rp * A.nx + i), v11 = __ldg(A.v + rp * A.nx + ip);
    const double ihx = 1. / A.hx, ihy = 1. / A.hy;
    // dG1 coefficients (mean, xi, eta) of the strain rates, exact for bilinear velocity
    const double ax = u10 - u00, bx = u11 - u01, cy = u01 - u00, dy = u11 - u10;
    const double avx = v10 - v00, bvx = v11 - v01, cvy = v01 - v00, dvy = v11 - v10;
    const double e11_0 = 0.5 * (ax + bx) * ihx, e11_2 = (bx - ax) * ihx;
    const double e22_0 = 0.5 * (cvy + dvy) * ihy, e22_1 = (dvy - cvy) * ihy;
    const double e12_0 = 0.5 * (0.5 * (cy + dy) * ihy + 0.5 * (avx + bvx) * ihx);
    const double e12_1 = 0.5 * (dy - cy) * ihy, e12_2 = 0.5 * (bvx - avx) * ihx;
    const double ie2 = 1. / (A.P.ellipse_e * A.P.ellipse_e);
    const double dmin2 = A.P.delta_min * A.P.delta_min;
    double p11[3] = { 0, 0, 0 }, p12[3] = { 0, 0, 0 }, p22[3] = { 0, 0, 0 };
#pragma unroll
    for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int a = 0; a < 2; ++a) {
            const double xa = a ? GQ : -GQ, yb = b ? GQ : -GQ;
            const double g11 = e11_0 + e11_2 * yb;
            const double g22 = e22_0 + e22_1 * xa;
            const double g12 = e12_0 + e12_1 * xa + e12_2 * yb;
            const double pr = __ldg(A.pg + (2 * b + a) * A.plane + k);
            const double tr = g11 + g22, df = g11 - g22;
            const double delta = sqrt(dmin2 + tr * tr + ie2 * (df * df + 4. * g12 * g12));
            const double zeta = pr / (2. * delta);
            const double eta = zeta * ie2;
            const double s11 = 2. * eta * g11 + (zeta - eta) * tr - 0.5 * pr;
            const double s22 = 2. * eta * g22 + (zeta - eta) * tr - 0.5 * pr;
            const double s12 = 2. * eta * g12;
            p11[0] += 0.25 * s11;
            p11[1] += 3. * xa * s11;
            p11[2] += 3. * yb * s11;
            p12[0] += 0.25 * s12;
            p12[1] += 3. * xa * s12;
            p12[2] += 3. * yb * s12;
            p22[0] += 0.25 * s22;
            p22[1] += 3. * xa * s22;
            p22[2] += 3. * yb * s22;
        }
    const double ia = 1. / A.P.alpha;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const int64_t q = c * A.plane + k;
        double s = A.S11[q];
        A.S11[q] = s + ia * (p11[c] - s);
        s = A.S12[q];
        A.S12[q] = s + ia * (p12[c] - s);
        s = A.S22[q];
        A.S22[q] = s + ia * (p22[c] - s);
    }
}

// out0 = unew, out1 = vnew
__global__ void __launch_bounds__(256) velocity_kernel(const __grid_constant__ MevpArgs A)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t r = A.r0 + (int64_t)blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= A.nx || r >= A.r1)
        return;
    const int64_t n = r * A.nx + i;
    if (A.P.bc == NSDG_BC_BOX) {
        const int64_t jg = A.j0 + (r - 1);
        if (i == 0 || jg == 0) {
            A.out0[n] = 0.;
            A.out1[n] = 0.;
            return;
        }
    }
    const int64_t rm = row_below(A, r);
    const int64_t im = (i == 0) ? A.nx - 1 : i - 1;
    const int64_t cells[4] = { rm * A.nx + im, rm * A.nx + i, r * A.nx + im, n };
    double dx = 0., dy = 0.;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const double sa = (q & 1) ? -1. : 1.; // node is the right corner of the left cells
        const double sb = (q & 2) ? -1. : 1.; // node is the top corner of the lower cells
        const int64_t k = cells[q];
        const double s11_0 = __ldg(A.S11 + k), s11_2 = __ldg(A.S11 + 2 * A.plane + k);
        const double s12_0 = __ldg(A.S12 + k), s12_1 = __ldg(A.S12 + A.plane + k), s12_2 = __ldg(A.S12 + 2 * A.plane + k);
        const double s22_0 = __ldg(A.S22 + k), s22_1 = __ldg(A.S22 + A.plane + k);
        const double ix11 = sa * A.hy * (0.5 * s11_0 + sb * s11_2 / 12.);
        const double iy12 = sb * A.hx * (0.5 * s12_0 + sa * s12_1 / 12.);
        const double ix12 = sa * A.hy * (0.5 * s12_0 + sb * s12_2 / 12.);
        const double iy22 = sb * A.hx * (0.5 * s22_0 + sa * s22_1 / 12.);
        dx -= ix11 + iy12;
        dy -= ix12 + iy22;
    }
    const double iarea = 1. / (A.hx * A.hy);
    const double u = __ldg(A.u + n), v = __ldg(A.v + n);
    const double uo = __ldg(A.uo + n), vo = __ldg(A.vo + n);
    const double m = A.P.rho_ice * fmax(__ldg(A.hn + n), A.P.h_min);
    const double a = __ldg(A.an + n);
    const double du = uo - u, dv = vo - v;
    const double cd = a * A.P.c_ocean * A.P.rho_ocean * sqrt(du * du + dv * dv);
    const double mdt = m / A.dt;
    const double denom = mdt * (1. + A.P.beta) + cd;
    const double rx = mdt * (A.P.beta * u + __ldg(A.un + n)) + a * __ldg(A.taux + n) + cd * uo
        + m * A.P.f_coriolis * (v - vo) + dx * iarea;
    const double ry = mdt * (A.P.beta * v + __ldg(A.vn + n)) + a * __ldg(A.tauy + n) + cd * vo
        - m * A.P.f_coriolis * (u - uo) + dy * iarea;
    A.out0[n] = rx / denom;
    A.out1[n] = ry / denom;
}

// ---------------------------------------------------------------------------------------------
// Fused sub-iteration: stress update + velocity update in ONE kernel.
//
// A CTA owns a TX x TY tile of nodes (and the cells of the same indices).  One thread per cell
// of the (TX+1) x (TY+1) cells that touch its nodes computes the new stress (the left column and
// bottom row are recomputed redundantly: ~16 % extra flops, their old-stress reads hit L2; all
// 21 loads of a cell are issued before the first use), keeps the 7 stress coefficients the
// divergence needs in shared memory, stores all 9 to HBM for the owned cells only; after one
// barrier the first TX*TY threads update their nodes from shared memory (the node streams were
// prefetched to L1 at kernel entry).  The velocity kernel's re-read of S from HBM disappears:
// 34 instead of 45 doubles per cell and sub-iteration.  Stress is ping-ponged (old -> new) so the
// redundant halo cells always see the OLD stress.
__device__ __forceinline__ void prefetch_l1(const double* p)
{
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
}

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
    constexpr int CXS = C::CXS, CYS = C::CYS;
    __shared__ double ss[7][CYS * CXS];
    const int tid = threadIdx.x;
    const int64_t i0 = (int64_t)blockIdx.x * TX;
    const int64_t r0 = A.r0 + (int64_t)blockIdx.y * TY;

    // node owned by this thread in phase 3 (threads >= TX*TY own none)
    const int tx = tid % TX, ty = tid / TX;
    const int64_t ni = i0 + tx, nr = r0 + ty;
    const bool has_node = tid < TX * TY && ni < A.nx && nr < A.r1;
    const int64_t n = nr * A.nx + ni;
    if (has_node) { // pull the node streams towards L1 while the stress phase runs
        prefetch_l1(A.un + n);
        prefetch_l1(A.vn + n);
        prefetch_l1(A.taux + n);
        prefetch_l1(A.tauy + n);
        prefetch_l1(A.uo + n);
        prefetch_l1(A.vo + n);
        prefetch_l1(A.hn + n);
        prefetch_l1(A.an + n);
    }

    // ---- phase 2: new stress of the (TX+1) x (TY+1) cells touching this tile's nodes
    const double ihx = 1. / A.hx, ihy = 1. / A.hy;
    const double ie2 = 1. / (A.P.ellipse_e * A.P.ellipse_e);
    const double dmin2 = A.P.delta_min * A.P.delta_min;
    const double ia = 1. / A.P.alpha;
    for (int t = tid; t < CYS * CXS; t += C::NT) { // executes at most once per thread
        const int cy = t / CXS - 1, cx = t % CXS - 1;
        int64_t gi = i0 + cx, gr = r0 + cy;
        if (gi >= A.nx || gr >= A.r1)
            continue; // beyond the grid (partial tile); never read by a valid node
        if (gi < 0)
            gi += A.nx;
        const int64_t gip = (gi == A.nx - 1) ? 0 : gi + 1;
        bool ghost_row = false;
        if (gr < 1) {
            if (A.wrap)
                gr = A.nyl;
            else
                ghost_row = true; // strip decomposition: row 0 is kept current redundantly
        }
        const int64_t k = gr * A.nx + gi;
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
        // old velocity at the 4 corner nodes (row above the cell: ghost or wrapped)
        int64_t grp = gr + 1;
        if (A.wrap && grp > A.nyl)
            grp = 1;
        const double u00 = __ldg(A.u + k), u10 = __ldg(A.u + gr * A.nx + gip);
        const double u01 = __ldg(A.u + grp * A.nx + gi), u11 = __ldg(A.u + grp * A.nx + gip);
        const double v00 = __ldg(A.v + k), v10 = __ldg(A.v + gr * A.nx + gip);
        const double v01 = __ldg(A.v + grp * A.nx + gi), v11 = __ldg(A.v + grp * A.nx + gip);
        const double ax = u10 - u00, bx = u11 - u01, cyy = u01 - u00, dyy = u11 - u10;
        const double avx = v10 - v00, bvx = v11 - v01, cvy = v01 - v00, dvy = v11 - v10;
        const double e11_0 = 0.5 * (ax + bx) * ihx, e11_2 = (bx - ax) * ihx;
        const double e22_0 = 0.5 * (cvy + dvy) * ihy, e22_1 = (dvy - cvy) * ihy;
        const double e12_0 = 0.5 * (0.5 * (cyy + dyy) * ihy + 0.5 * (avx + bvx) * ihx);
        const double e12_1 = 0.5 * (dyy - cyy) * ihy, e12_2 = 0.5 * (bvx - avx) * ihx;
        double p11[3] = { 0, 0, 0 }, p12[3] = { 0, 0, 0 }, p22[3] = { 0, 0, 0 };
#pragma unroll
        for (int b = 0; b < 2; ++b)
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                const double xa = a ? GQ : -GQ, yb = b ? GQ : -GQ;
                const double g11 = e11_0 + e11_2 * yb;
                const double g22 = e22_0 + e22_1 * xa;
                const double g12 = e12_0 + e12_1 * xa + e12_2 * yb;
                const double pp = pr[2 * b + a];
                const double tr = g11 + g22, df = g11 - g22;
                // zeta = P/(2 Delta) with one reciprocal square root instead of sqrt + divide
                const double zeta = 0.5 * pp * rsqrt(dmin2 + tr * tr + ie2 * (df * df + 4. * g12 * g12));
                const double eta = zeta * ie2;
                const double s11 = 2. * eta * g11 + (zeta - eta) * tr - 0.5 * pp;
                const double s22 = 2. * eta * g22 + (zeta - eta) * tr - 0.5 * pp;
                const double s12 = 2. * eta * g12;
                p11[0] += 0.25 * s11;
                p11[1] += 3. * xa * s11;
                p11[2] += 3. * yb * s11;
                p12[0] += 0.25 * s12;
                p12[1] += 3. * xa * s12;
                p12[2] += 3. * yb * s12;
                p22[0] += 0.25 * s22;
                p22[1] += 3. * xa * s22;
                p22[2] += 3. * yb * s22;
            }
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            so11[c] += ia * (p11[c] - so11[c]);
            so12[c] += ia * (p12[c] - so12[c]);
            so22[c] += ia * (p22[c] - so22[c]);
        }
        ss[0][t] = so11[0];
        ss[1][t] = so11[2];
        ss[2][t] = so12[0];
        ss[3][t] = so12[1];
        ss[4][t] = so12[2];
        ss[5][t] = so22[0];
        ss[6][t] = so22[1];
        if ((cx >= 0 && cy >= 0) || (ghost_row && cx >= 0)) {
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                A.S11n[c * A.plane + k] = so11[c];
                A.S12n[c * A.plane + k] = so12[c];
                A.S22n[c * A.plane + k] = so22[c];
            }
        }
    }
    __syncthreads();

    // ---- phase 3: velocity of the owned nodes
    if (!has_node)
        return;
    if (A.P.bc == NSDG_BC_BOX) {
        const int64_t jg = A.j0 + (nr - 1);
        if (ni == 0 || jg == 0) {
            A.out0[n] = 0.;
            A.out1[n] = 0.;
            return;
        }
    }
    // -(S, grad phi_n)/(hx hy) from the 4 adjacent cells; with sa = +,-,+,- and sb = +,+,-,- for
    // the cells (tx-1,ty-1), (tx,ty-1), (tx-1,ty), (tx,ty) the analytic integrals collapse to
    // signed sums of shared-memory values times four constants
    const int t00 = ty * CXS + tx, t10 = t00 + 1, t01 = t00 + CXS, t11 = t01 + 1;
    const double c1 = 0.5 * ihx, c2 = ihx * (1. / 12.), c3 = 0.5 * ihy, c4 = ihy * (1. / 12.);
#define SUM_A(c) (ss[c][t00] - ss[c][t10] + ss[c][t01] - ss[c][t11]) /* sum sa X */
#define SUM_B(c) (ss[c][t00] + ss[c][t10] - ss[c][t01] - ss[c][t11]) /* sum sb X */
#define SUM_AB(c) (ss[c][t00] - ss[c][t10] - ss[c][t01] + ss[c][t11]) /* sum sa sb X */
    const double dxa = -(c1 * SUM_A(0) + c2 * SUM_AB(1) + c3 * SUM_B(2) + c4 * SUM_AB(3));
    const double dya = -(c1 * SUM_A(2) + c2 * SUM_AB(4) + c3 * SUM_B(5) + c4 * SUM_AB(6));
#undef SUM_A
#undef SUM_B
#undef SUM_AB
    const double u = __ldg(A.u + n), v = __ldg(A.v + n);
    const double uo = __ldg(A.uo + n), vo = __ldg(A.vo + n);
    const double m = A.P.rho_ice * fmax(__ldg(A.hn + n), A.P.h_min);
    const double a = __ldg(A.an + n);
    const double du = uo - u, dv = vo - v;
    const double cd = a * A.P.c_ocean * A.P.rho_ocean * sqrt(du * du + dv * dv);
    const double mdt = m / A.dt;
    const double denom = mdt * (1. + A.P.beta) + cd;
    const double rx = mdt * (A.P.beta * u + __ldg(A.un + n)) + a * __ldg(A.taux + n) + cd * uo
        + m * A.P.f_coriolis * (v - vo) + dxa;
    const double ry = mdt * (A.P.beta * v + __ldg(A.vn + n)) + a * __ldg(A.tauy + n) + cd * vo
        - m * A.P.f_coriolis * (u - uo) + dya;
    const double inv = 1. / denom;
    A.out0[n] = rx * inv;
    A.out1[n] = ry * inv;
}

// ---------------------------------------------------------------------------------------------
// Fused sub-iteration, bulk-copy (TMA) pipelined variant: the fast path when nx % TX == 0.
//
// Persistent CTAs (2 per SM) walk the tiles in row-major order.  The (TX+1) x (TY+1) cell
// footprint of a tile -- 9 old-stress planes, 4 strength planes, and the (TX+2) x (TY+2) node
// footprint of u, v -- is brought into shared memory by cp.async.bulk row copies (one 16-byte
// aligned row segment per copy, completion counted on an mbarrier), two stages deep: while the
// threads compute tile t from stage t&1, the copies of tile t+1 are in flight and those of
// tile t+2 are issued as soon as the stage is drained.  No registers are held by loads in flight,
// and DRAM reads continue through the FP64 phases.  Periodic wrap in x is a second 16-byte copy
// on the first / last tile column; wrap (or ghost rows) in y is a row index choice per copy.
// The arithmetic is the same, in the same order, as mevp_fused_kernel.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "WAIT_%=:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                 "@p bra DONE_%=;\n"
                 "bra WAIT_%=;\n"
                 "DONE_%=:\n"
                 "}" ::"r"(smem_u32(bar)),
                 "r"(parity)
                 : "memory");
}
__device__ __forceinline__ void bulk_g2s(double* dst, const double* src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

template <int TX, int TY> struct TmaCfg {
    static constexpr int CXS = TX + 1, CYS = TY + 1;
    static constexpr int NCELL = CXS * CYS;
    static constexpr int NT = (NCELL + 31) / 32 * 32;
    static constexpr int WS = TX + 2; // cell row in smem: columns i0-2 .. i0+TX-1
    static constexpr int WU = TX + 4; // node row in smem: columns i0-2 .. i0+TX+1
    static constexpr int NROW_S = 13 * CYS; // 9 stress + 4 strength planes
    static constexpr int NROW_U = 2 * (TY + 2);
    static constexpr int NCOPY = NROW_S + NROW_U; // row copies per tile == arriving threads
    // every box destination is 128-byte aligned (16 doubles)
    static constexpr int r16(int n) { return (n + 15) / 16 * 16; }
    static constexpr int SZ_S = r16(3 * CYS * WS); // one stress component: [3][CYS][WS]
    static constexpr int SZ_P = r16(4 * CYS * WS); // strength: [4][CYS][WS]
    static constexpr int SZ_U = r16((TY + 2) * WU); // u or v: [TY+2][WU]
    static constexpr int OFF_P = 3 * SZ_S, OFF_U = OFF_P + SZ_P, OFF_V = OFF_U + SZ_U;
    static constexpr int STAGE = OFF_V + SZ_U; // doubles
    static constexpr uint32_t TILE_BYTES = 8u * (NROW_S * WS + NROW_U * WU);
    static constexpr int STAGES = 2;
    static constexpr size_t SMEM = sizeof(double) * (STAGES * STAGE + 2 * 7 * NCELL) + 64;
    static_assert(NCOPY <= NT, "one copy per thread");
};

// tensor maps of the six arrays a sub-iteration reads through boxes: S11, S12, S22 (3 planes
// each), pg (4 planes) as 3-d (x, row, plane); u, v as 2-d (x, row)
struct TmaMaps {
    CUtensorMap m[6];
};

__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_box3(double* dst, const CUtensorMap* map, int x, int y, int z, uint64_t* bar)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(smem_u32(dst)),
                 "l"(map), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tma_box2(double* dst, const CUtensorMap* map, int x, int y, uint64_t* bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_u32(dst)),
                 "l"(map), "r"(x), "r"(y), "r"(smem_u32(bar))
                 : "memory");
}

// Called by the threads q < NCOPY; together they arrive NCOPY times on `bar` and post TILE_BYTES.
// Interior tile columns: thread 0 issues six tensor boxes (out-of-range rows of a partial last
// tile row are zero-filled and never used).  First / last tile column: one row copy per thread,
// with the periodic column brought by a second 16-byte copy.
template <int TX, int TY>
__device__ __forceinline__ void tma_issue_tile(const MevpArgs& A, const TmaMaps& M, int64_t tile, int64_t ntx, double* stage, uint64_t* bar, int q)
{
    using C = TmaCfg<TX, TY>;
    const int64_t tcol = tile % ntx;
    const int64_t i0 = tcol * TX;
    const int64_t r0 = A.r0 + (tile / ntx) * TY;
    if (tcol != 0 && tcol != ntx - 1) {
        if (q == 0) {
            mbar_expect_tx(bar, C::TILE_BYTES);
            const int x = (int)(i0 - 2), y = (int)(r0 - 1);
            tma_box3(stage, &M.m[0], x, y, 0, bar);
            tma_box3(stage + C::SZ_S, &M.m[1], x, y, 0, bar);
            tma_box3(stage + 2 * C::SZ_S, &M.m[2], x, y, 0, bar);
            tma_box3(stage + C::OFF_P, &M.m[3], x, y, 0, bar);
            tma_box2(stage + C::OFF_U, &M.m[4], x, y, bar);
            tma_box2(stage + C::OFF_V, &M.m[5], x, y, bar);
        } else
            mbar_arrive(bar);
        return;
    }
    const double* src;
    double* dst;
    int64_t g;
    int W;
    if (q < C::NROW_S) {
        const int pl = q / C::CYS, cy = q % C::CYS;
        switch (pl) {
        case 0: case 1: case 2: src = A.S11 + pl * A.plane; break;
        case 3: case 4: case 5: src = A.S12 + (pl - 3) * A.plane; break;
        case 6: case 7: case 8: src = A.S22 + (pl - 6) * A.plane; break;
        default: src = A.pg + (pl - 9) * A.plane; break;
        }
        g = r0 + cy - 1;
        dst = stage + (pl < 9 ? (pl / 3) * C::SZ_S + ((pl % 3) * C::CYS + cy) * C::WS
                              : C::OFF_P + ((pl - 9) * C::CYS + cy) * C::WS);
        W = C::WS;
    } else {
        const int q2 = q - C::NROW_S;
        const int pl = q2 / (TY + 2), ry = q2 % (TY + 2);
        src = pl ? A.v : A.u;
        g = r0 + ry - 1;
        dst = stage + (pl ? C::OFF_V : C::OFF_U) + ry * C::WU;
        W = C::WU;
    }
    if (g > A.nyl + 1)
        g = A.nyl + 1; // rows past a partial last tile row: valid memory, never used
    src += g * A.nx;
    mbar_expect_tx(bar, (uint32_t)W * 8u);
    const int64_t cs = i0 - 2;
    if (cs < 0) { // first tile column: cell/node column -1 is column nx-1
        bulk_g2s(dst, src + A.nx - 2, 16u, bar);
        bulk_g2s(dst + 2, src, (uint32_t)(W - 2) * 8u, bar);
    } else if (cs + W > A.nx) { // last tile column, node rows only: node column nx is column 0
        bulk_g2s(dst, src + cs, (uint32_t)(A.nx - cs) * 8u, bar);
        bulk_g2s(dst + (A.nx - cs), src, (uint32_t)(cs + W - A.nx) * 8u, bar);
    } else
        bulk_g2s(dst, src + cs, (uint32_t)W * 8u, bar);
}

template <int TX, int TY>
__global__ void __launch_bounds__(TmaCfg<TX, TY>::NT, 2) mevp_fused_tma_kernel(const __grid_constant__ MevpArgs A, const __grid_constant__ TmaMaps M)
{
    using C = TmaCfg<TX, TY>;
    constexpr int CXS = C::CXS, CYS = C::CYS, WS = C::WS, WU = C::WU;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double* stage0 = reinterpret_cast<double*>(smem_raw);
    double* ssb = stage0 + C::STAGES * C::STAGE; // [2][7][NCELL]
    uint64_t* bars = reinterpret_cast<uint64_t*>(ssb + 2 * 7 * C::NCELL);
    const int tid = threadIdx.x;
    const int64_t ntx = A.nx / TX;
    const int64_t ntiles = ntx * ((A.r1 - A.r0 + TY - 1) / TY);
    if (tid == 0) {
        mbar_init(&bars[0], C::NCOPY);
        mbar_init(&bars[1], C::NCOPY);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    // prologue: two tiles in flight
    if (tid < C::NCOPY) {
        if ((int64_t)blockIdx.x < ntiles)
            tma_issue_tile<TX, TY>(A, M, blockIdx.x, ntx, stage0, &bars[0], tid);
        if ((int64_t)blockIdx.x + gridDim.x < ntiles)
            tma_issue_tile<TX, TY>(A, M, (int64_t)blockIdx.x + gridDim.x, ntx, stage0 + C::STAGE, &bars[1], tid);
    }
    const int tx = tid % TX, ty = tid / TX; // node owned in the velocity phase (tid < TX*TY)
    const int cx = tid % CXS, cy = tid / CXS; // cell owned in the stress phase (0 = halo column/row)
    const double ihx = 1. / A.hx, ihy = 1. / A.hy;
    const double ie2 = 1. / (A.P.ellipse_e * A.P.ellipse_e);
    const double dmin2 = A.P.delta_min * A.P.delta_min;
    const double ia = 1. / A.P.alpha;
    int it = 0;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int st = it & 1;
        const uint32_t parity = (it >> 1) & 1;
        const double* sg = stage0 + st * C::STAGE;
        double* ss = ssb + st * 7 * C::NCELL;
        const int64_t i0 = (tile % ntx) * TX;
        const int64_t r0 = A.r0 + (tile / ntx) * TY;
        // node streams of the velocity phase: issued now, consumed after the barrier
        const int64_t ni = i0 + tx, nr = r0 + ty;
        const bool has_node = tid < TX * TY && nr < A.r1;
        const int64_t n = nr * A.nx + ni;
        double n_un = 0, n_vn = 0, n_tx = 0, n_ty = 0, n_uo = 0, n_vo = 0, n_hn = 0, n_an = 0;
        if (has_node) {
            n_un = __ldg(A.un + n);
            n_vn = __ldg(A.vn + n);
            n_tx = __ldg(A.taux + n);
            n_ty = __ldg(A.tauy + n);
            n_uo = __ldg(A.uo + n);
            n_vo = __ldg(A.vo + n);
            n_hn = __ldg(A.hn + n);
            n_an = __ldg(A.an + n);
        }
        mbar_wait(&bars[st], parity);
        const double* su = sg + C::OFF_U;
        const double* sv = sg + C::OFF_V;
        double u_own = 0, v_own = 0;
        if (has_node) { // node (tx,ty) sits at smem node row ty+1, column tx+2
            u_own = su[(ty + 1) * WU + tx + 2];
            v_own = sv[(ty + 1) * WU + tx + 2];
        }
        // ---- stress of the (TX+1) x (TY+1) cells touching this tile's nodes
        if (tid < C::NCELL) {
            const int64_t gr = r0 + cy - 1; // row 0 only in strip mode (ghost row kept current)
            if (gr < A.r1) {
                const int co = cy * WS + cx + 1; // cell column i0+cx-1 is smem column cx+1
                double so11[3], so12[3], so22[3], pr[4];
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    so11[c] = sg[(c * CYS) * WS + co];
                    so12[c] = sg[C::SZ_S + (c * CYS) * WS + co];
                    so22[c] = sg[2 * C::SZ_S + (c * CYS) * WS + co];
                }
#pragma unroll
                for (int g = 0; g < 4; ++g)
                    pr[g] = sg[C::OFF_P + (g * CYS) * WS + co];
                const int no = cy * WU + cx + 1;
                const double u00 = su[no], u10 = su[no + 1], u01 = su[no + WU], u11 = su[no + WU + 1];
                const double v00 = sv[no], v10 = sv[no + 1], v01 = sv[no + WU], v11 = sv[no + WU + 1];
                const double ax = u10 - u00, bx = u11 - u01, cyy = u01 - u00, dyy = u11 - u10;
                const double avx = v10 - v00, bvx = v11 - v01, cvy = v01 - v00, dvy = v11 - v10;
                const double e11_0 = 0.5 * (ax + bx) * ihx, e11_2 = (bx - ax) * ihx;
                const double e22_0 = 0.5 * (cvy + dvy) * ihy, e22_1 = (dvy - cvy) * ihy;
                const double e12_0 = 0.5 * (0.5 * (cyy + dyy) * ihy + 0.5 * (avx + bvx) * ihx);
                const double e12_1 = 0.5 * (dyy - cyy) * ihy, e12_2 = 0.5 * (bvx - avx) * ihx;
                double p11[3] = { 0, 0, 0 }, p12[3] = { 0, 0, 0 }, p22[3] = { 0, 0, 0 };
#pragma unroll
                for (int b = 0; b < 2; ++b)
#pragma unroll
                    for (int a = 0; a < 2; ++a) {
                        const double xa = a ? GQ : -GQ, yb = b ? GQ : -GQ;
                        const double g11 = e11_0 + e11_2 * yb;
                        const double g22 = e22_0 + e22_1 * xa;
                        const double g12 = e12_0 + e12_1 * xa + e12_2 * yb;
                        const double pp = pr[2 * b + a];
                        const double tr = g11 + g22, df = g11 - g22;
                        const double zeta = 0.5 * pp * rsqrt(dmin2 + tr * tr + ie2 * (df * df + 4. * g12 * g12));
                        const double eta = zeta * ie2;
                        const double s11 = 2. * eta * g11 + (zeta - eta) * tr - 0.5 * pp;
                        const double s22 = 2. * eta * g22 + (zeta - eta) * tr - 0.5 * pp;
                        const double s12 = 2. * eta * g12;
                        p11[0] += 0.25 * s11;
                        p11[1] += 3. * xa * s11;
                        p11[2] += 3. * yb * s11;
                        p12[0] += 0.25 * s12;
                        p12[1] += 3. * xa * s12;
                        p12[2] += 3. * yb * s12;
                        p22[0] += 0.25 * s22;
                        p22[1] += 3. * xa * s22;
                        p22[2] += 3. * yb * s22;
                    }
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    so11[c] += ia * (p11[c] - so11[c]);
                    so12[c] += ia * (p12[c] - so12[c]);
                    so22[c] += ia * (p22[c] - so22[c]);
                }
                ss[0 * C::NCELL + tid] = so11[0];
                ss[1 * C::NCELL + tid] = so11[2];
                ss[2 * C::NCELL + tid] = so12[0];
                ss[3 * C::NCELL + tid] = so12[1];
                ss[4 * C::NCELL + tid] = so12[2];
                ss[5 * C::NCELL + tid] = so22[0];
                ss[6 * C::NCELL + tid] = so22[1];
                // owned cells, plus (strip mode) the redundantly evolved ghost row 0
                if (cx >= 1 && (cy >= 1 || (!A.wrap && gr == 0))) {
                    const int64_t k = gr * A.nx + i0 + cx - 1;
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        A.S11n[c * A.plane + k] = so11[c];
                        A.S12n[c * A.plane + k] = so12[c];
                        A.S22n[c * A.plane + k] = so22[c];
                    }
                }
            }
        }
        // all reads of stage st are done and ss[st] is complete
        __syncthreads();
        // refill the drained stage with tile t+2
        if (tid < C::NCOPY && tile + 2 * (int64_t)gridDim.x < ntiles) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            tma_issue_tile<TX, TY>(A, M, tile + 2 * (int64_t)gridDim.x, ntx, stage0 + st * C::STAGE, &bars[st], tid);
        }
        // ---- velocity of the owned nodes
        if (has_node) {
            double un_ = 0., vn_ = 0.;
            if (!(A.P.bc == NSDG_BC_BOX && (ni == 0 || A.j0 + (nr - 1) == 0))) {
                const int t00 = ty * CXS + tx, t10 = t00 + 1, t01 = t00 + CXS, t11 = t01 + 1;
                const double c1 = 0.5 * ihx, c2 = ihx * (1. / 12.), c3 = 0.5 * ihy, c4 = ihy * (1. / 12.);
#define SSV(c, t) ss[(c)*C::NCELL + (t)]
#define SUM_A(c) (SSV(c, t00) - SSV(c, t10) + SSV(c, t01) - SSV(c, t11)) /* sum sa X */
#define SUM_B(c) (SSV(c, t00) + SSV(c, t10) - SSV(c, t01) - SSV(c, t11)) /* sum sb X */
#define SUM_AB(c) (SSV(c, t00) - SSV(c, t10) - SSV(c, t01) + SSV(c, t11)) /* sum sa sb X */
                const double dxa = -(c1 * SUM_A(0) + c2 * SUM_AB(1) + c3 * SUM_B(2) + c4 * SUM_AB(3));
                const double dya = -(c1 * SUM_A(2) + c2 * SUM_AB(4) + c3 * SUM_B(5) + c4 * SUM_AB(6));
#undef SUM_A
#undef SUM_B
#undef SUM_AB
#undef SSV
                const double u = u_own, v = v_own;
                const double uo = n_uo, vo = n_vo;
                const double m = A.P.rho_ice * fmax(n_hn, A.P.h_min);
                const double a = n_an;
                const double du = uo - u, dv = vo - v;
                const double cd = a * A.P.c_ocean * A.P.rho_ocean * sqrt(du * du + dv * dv);
                const double mdt = m / A.dt;
                const double denom = mdt * (1. + A.P.beta) + cd;
                const double rx = mdt * (A.P.beta * u + n_un) + a * n_tx + cd * uo + m * A.P.f_coriolis * (v - vo) + dxa;
                const double ry = mdt * (A.P.beta * v + n_vn) + a * n_ty + cd * vo - m * A.P.f_coriolis * (u - uo) + dya;
                const double inv = 1. / denom;
                un_ = rx * inv;
                vn_ = ry * inv;
            }
            A.out0[n] = un_;
            A.out1[n] = vn_;
            if (A.selfhalo) { // single rank in ghost-row mode: the ring neighbour is this rank
                if (nr == A.nyl) {
                    A.out0[ni] = un_;
                    A.out1[ni] = vn_;
                }
                if (nr == 1) {
                    A.out0[(A.nyl + 1) * A.nx + ni] = un_;
                    A.out1[(A.nyl + 1) * A.nx + ni] = vn_;
                }
            }
        }
    }
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link against libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn()
{
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess
            && qr == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

// planes of (rows x nx) doubles, nplanes of them `plane` elements apart; box = bx x by x nplanes
bool make_map(CUtensorMap* m, const double* base, int64_t nx, int64_t rows, int64_t plane, int nplanes, int bx, int by)
{
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn)
        return false;
    const cuuint32_t rank = nplanes > 1 ? 3 : 2;
    cuuint64_t dims[3] = { (cuuint64_t)nx, (cuuint64_t)rows, (cuuint64_t)nplanes };
    cuuint64_t strides[2] = { (cuuint64_t)nx * 8, (cuuint64_t)plane * 8 };
    cuuint32_t box[3] = { (cuuint32_t)bx, (cuuint32_t)by, (cuuint32_t)nplanes };
    cuuint32_t estr[3] = { 1, 1, 1 };
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, rank, (void*)base, dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)
        == CUDA_SUCCESS;
}

MevpArgs base_args(DynState* d, bool ghost)
{
    MevpArgs A = {};
    A.P = d->mp;
    A.nx = d->nx;
    A.nyl = d->nyl;
    A.plane = d->plane;
    A.j0 = d->cfg.j0;
    A.ny_global = d->cfg.ny_global;
    A.hx = d->cfg.hx;
    A.hy = d->cfg.hy;
    A.wrap = ghost ? 0 : 1;
    A.nc = d->nc;
    A.H = d->f[NSDG_DF_H];
    A.A = d->f[NSDG_DF_A];
    A.u = d->f[NSDG_DF_U];
    A.v = d->f[NSDG_DF_V];
    A.un = d->un;
    A.vn = d->vn;
    A.hn = d->hn;
    A.an = d->an;
    A.taux = d->taux;
    A.tauy = d->tauy;
    A.uo = d->f[NSDG_DF_UOCN];
    A.vo = d->f[NSDG_DF_VOCN];
    A.ua = d->f[NSDG_DF_UATM];
    A.va = d->f[NSDG_DF_VATM];
    A.pg = d->pg;
    A.S11 = d->f[NSDG_DF_S11];
    A.S12 = d->f[NSDG_DF_S12];
    A.S22 = d->f[NSDG_DF_S22];
    return A;
}

} // namespace

// Requires (multi-rank): ghost rows of H, A current.  Single rank: kernels wrap.
bool nsdg_mevp_tma_eligible(int64_t nx)
{
    const char* e = getenv("NSDG_FUSED_TMA");
    // opt-in: measured 1.06 ms/sub-iteration at 4096^2 against 0.95-0.98 ms of the register-path
    // kernel (profiles/r1i_tma_fused_ncu.txt), so the register path stays the default
    return e && atoi(e) != 0 && nx % 32 == 0 && nx >= 96 && encode_tiled_fn() != nullptr;
}

int nsdg_mevp_prepare(nsdg_ctx* ctx, bool ghost)
{
    DynState* d = ctx->dyn;
    MevpArgs A = base_args(d, ghost);
    // strength on rows 0..nyl when the lower ghost row's stress is computed redundantly
    A.r0 = ghost ? 0 : 1;
    A.r1 = d->nyl + 1;
    A.out0 = d->pg;
    dim3 block(256), grid((unsigned)((d->nx + 255) / 256), (unsigned)(A.r1 - A.r0));
    NSDG_KERNEL(ctx, NSDG_K_PRESSURE, pressure_kernel<<<grid, block, 0, ctx->stream>>>(A));
    A.r0 = 1;
    A.r1 = d->nyl + 1;
    A.out0 = d->hn;
    A.out1 = d->an;
    A.out2 = d->taux;
    A.out3 = d->tauy;
    A.out4 = d->un;
    A.out5 = d->vn;
    grid = dim3((unsigned)((d->nx + 255) / 256), (unsigned)d->nyl);
    NSDG_KERNEL(ctx, NSDG_K_NODAL, nodal_kernel<<<grid, block, 0, ctx->stream>>>(A));
    return NSDG_OK;
}

int nsdg_mevp_stress(nsdg_ctx* ctx, const double* u, const double* v, int64_t r0, int64_t r1)
{
    DynState* d = ctx->dyn;
    if (r1 <= r0)
        return NSDG_OK;
    MevpArgs A = base_args(d, !d->wrap);
    A.u = u;
    A.v = v;
    A.r0 = r0;
    A.r1 = r1;
    dim3 block(64, 4), grid((unsigned)((d->nx + 63) / 64), (unsigned)((r1 - r0 + 3) / 4));
    NSDG_KERNEL(ctx, NSDG_K_STRESS, stress_kernel<<<grid, block, 0, ctx->stream>>>(A));
    return NSDG_OK;
}

int nsdg_mevp_velocity(nsdg_ctx* ctx, double dt, const double* u, const double* v, double* unew,
    double* vnew, int64_t r0, int64_t r1)
{
    DynState* d = ctx->dyn;
    if (r1 <= r0)
        return NSDG_OK;
    MevpArgs A = base_args(d, !d->wrap);
    A.u = u;
    A.v = v;
    A.dt = dt;
    A.out0 = unew;
    A.out1 = vnew;
    A.r0 = r0;
    A.r1 = r1;
    dim3 block(64, 4), grid((unsigned)((d->nx + 63) / 64), (unsigned)((r1 - r0 + 3) / 4));
    NSDG_KERNEL(ctx, NSDG_K_VELOCITY, velocity_kernel<<<grid, block, 0, ctx->stream>>>(A));
    return NSDG_OK;
}

// one fused sub-iteration: (u,v,S) -> (unew,vnew,Snew) on node rows [r0,r1)
int nsdg_mevp_fused(nsdg_ctx* ctx, double dt, const double* u, const double* v, double* unew,
    double* vnew, double* const* Sold, double* const* Snew, int64_t r0, int64_t r1, bool ghost)
{
    DynState* d = ctx->dyn;
    if (r1 <= r0)
        return NSDG_OK;
    MevpArgs A = base_args(d, ghost);
    A.u = u;
    A.v = v;
    A.dt = dt;
    A.out0 = unew;
    A.out1 = vnew;
    A.S11 = Sold[0];
    A.S12 = Sold[1];
    A.S22 = Sold[2];
    A.S11n = Snew[0];
    A.S12n = Snew[1];
    A.S22n = Snew[2];
    A.r0 = r0;
    A.r1 = r1;
    // fast path: bulk-copy (TMA) pipelined persistent kernel; needs ghost-row mode, whole tiles
    // in x and 16-byte aligned rows.  NSDG_FUSED_TMA=0 forces the register-path kernel.
    if (ghost && nsdg_mevp_tma_eligible(d->nx)) {
        using C = TmaCfg<32, 8>;
        static int tma_ctas = 0;
        if (!tma_ctas) {
            const char* c = getenv("NSDG_FUSED_CTAS");
            tma_ctas = (c && atoi(c) > 0) ? atoi(c) : 2;
            NSDG_CUDA(ctx, cudaFuncSetAttribute(mevp_fused_tma_kernel<32, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM));
        }
        TmaMaps M;
        const double* bases[6] = { A.S11, A.S12, A.S22, A.pg, A.u, A.v };
        const int npl[6] = { 3, 3, 3, 4, 1, 1 };
        for (int q = 0; q < 6; ++q)
            NSDG_REQUIRE(ctx, make_map(&M.m[q], bases[q], d->nx, d->rows, d->plane, npl[q], q < 4 ? C::WS : C::WU, q < 4 ? C::CYS : 8 + 2),
                NSDG_ECUDA, "cuTensorMapEncodeTiled failed for mEVP array %d", q);
        A.selfhalo = d->ghost_single ? 1 : 0;
        const int64_t ntiles = (d->nx / 32) * ((r1 - r0 + 7) / 8);
        const int64_t resident = (int64_t)ctx->sm_count * tma_ctas;
        dim3 block(C::NT), grid((unsigned)(ntiles < resident ? ntiles : resident));
        NSDG_KERNEL(ctx, NSDG_K_MEVP_FUSED, (mevp_fused_tma_kernel<32, 8><<<grid, block, C::SMEM, ctx->stream>>>(A, M)));
        return NSDG_OK;
    }
    // tile shape: 32x8 by default; NSDG_FUSED_TILE=WxH selects another compiled variant (tuning)
    static int tile = -1;
    if (tile < 0) {
        const char* e = getenv("NSDG_FUSED_TILE");
        std::string t = e ? e : "32x8";
        tile = t == "64x4" ? 1 : t == "32x16" ? 2 : t == "64x8" ? 3 : t == "16x16" ? 4 : 0;
    }
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
