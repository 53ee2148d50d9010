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

MevpArgs base_args(DynState* d)
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
    A.wrap = d->wrap ? 1 : 0;
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
int nsdg_mevp_prepare(nsdg_ctx* ctx)
{
    DynState* d = ctx->dyn;
    MevpArgs A = base_args(d);
    // strength on rows 0..nyl when the lower ghost row's stress is computed redundantly
    A.r0 = d->wrap ? 1 : 0;
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
    MevpArgs A = base_args(d);
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
    MevpArgs A = base_args(d);
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
    double* vnew, double* const* Sold, double* const* Snew, int64_t r0, int64_t r1)
{
    DynState* d = ctx->dyn;
    if (r1 <= r0)
        return NSDG_OK;
    MevpArgs A = base_args(d);
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
