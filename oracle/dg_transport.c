/* TEST INFRASTRUCTURE (oracle). Plain-C CPU restatement of the DG tracer transport.
 *
 * Parity status: PARITY UNPINNED against draenog/nextsimdg.  The mounted reference snapshot has
 * no dynamics code at all ("dynamics" is commented out of its build, CMakeLists.txt:43-46), so
 * there is no reference file:line to cite and no golden vector.  This file states the scheme
 * (after Richter et al., GMD 2023, "A dynamical core based on a discontinuous Galerkin method
 * for higher-order finite-element sea ice modeling", sect. on DG advection) and is pinned only
 * by analytic known-answer tests (tests/test_oracle_dynamics.py): constant preservation, mass
 * conservation, exact translation of polynomials at order k+1, convergence order.
 *
 * Scheme
 *   grid       nx*ny uniform cells (hx,hy), doubly periodic; cell (i,j) -> index j*nx+i
 *   tracer     dG_k, k=0,1,2: ncomp = 1,3,6 modal coefficients per cell, component-major
 *              phi[c][j][i]; basis on the reference square (xi,eta) in [0,1]^2:
 *                psi0=1, psi1=xi-1/2, psi2=eta-1/2, psi3=(xi-1/2)^2-1/12,
 *                psi4=(eta-1/2)^2-1/12, psi5=(xi-1/2)(eta-1/2)
 *              (orthogonal; mass = hx*hy*{1,1/12,1/12,1/180,1/180,1/144})
 *   velocity   CG1: nodal values u[j][i], v[j][i] at the lower-left corner of cell (i,j),
 *              bilinear inside each cell, periodic
 *   operator   M dphi_i/dt = int_K phi (vel.grad psi_i) - sum_edges int_e (vel.n) phi^ psi_i
 *              phi^ = upwind trace evaluated pointwise at the edge Gauss points:
 *              flux = max(vn,0)*phi_minus + min(vn,0)*phi_plus
 *   quadrature Gauss-Legendre with nq = 1,2,3 points per direction for ncomp = 1,3,6
 *              (exact for the polynomial integrands of that order)
 *   time       RK1 (forward Euler), Heun RK2, Shu-Osher SSP-RK3
 *   limiter    optional (off by default, as in the published dynamical core): rescale the
 *              higher moments about the cell mean so that the values on the 3x3 point lattice
 *              {0,1/2,1}^2 stay inside [lo,hi]; the mean is untouched (conservative)
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load this code.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

static const double INV_MASS[6] = { 1., 12., 12., 180., 180., 144. };

static int gauss_rule(int nq, double* x, double* w)
{
    if (nq == 1) {
        x[0] = 0.5;
        w[0] = 1.;
    } else if (nq == 2) {
        double d = 0.5 / sqrt(3.);
        x[0] = 0.5 - d;
        x[1] = 0.5 + d;
        w[0] = w[1] = 0.5;
    } else {
        double d = 0.5 * sqrt(3. / 5.);
        x[0] = 0.5 - d;
        x[1] = 0.5;
        x[2] = 0.5 + d;
        w[0] = w[2] = 5. / 18.;
        w[1] = 8. / 18.;
    }
    return nq;
}

static void basis(double xi, double eta, double* psi)
{
    double a = xi - 0.5, b = eta - 0.5;
    psi[0] = 1.;
    psi[1] = a;
    psi[2] = b;
    psi[3] = a * a - 1. / 12.;
    psi[4] = b * b - 1. / 12.;
    psi[5] = a * b;
}
static void basis_dxi(double xi, double eta, double* d)
{
    d[0] = 0;
    d[1] = 1;
    d[2] = 0;
    d[3] = 2 * (xi - 0.5);
    d[4] = 0;
    d[5] = eta - 0.5;
}
static void basis_deta(double xi, double eta, double* d)
{
    d[0] = 0;
    d[1] = 0;
    d[2] = 1;
    d[3] = 0;
    d[4] = 2 * (eta - 0.5);
    d[5] = xi - 0.5;
}

static double eval_cell(const double* phi, long ncell, long k, int ncomp, double xi, double eta)
{
    double psi[6], s = 0;
    basis(xi, eta, psi);
    for (int c = 0; c < ncomp; ++c)
        s += phi[(size_t)c * ncell + k] * psi[c];
    return s;
}

/* out = L(phi): the spatial operator (already multiplied by the inverse mass) */
void nsdg_oracle_transport_rhs(long nx, long ny, double hx, double hy, int ncomp,
    const double* phi, const double* u, const double* v, double* out)
{
    const long ncell = nx * ny;
    const int nq = (ncomp == 1) ? 1 : (ncomp == 3 ? 2 : 3);
    double gx[3], gw[3];
    gauss_rule(nq, gx, gw);
    memset(out, 0, sizeof(double) * (size_t)ncomp * ncell);

    /* volume term */
#pragma omp parallel for schedule(static)
    for (long j = 0; j < ny; ++j) {
        long jp = (j + 1) % ny;
        for (long i = 0; i < nx; ++i) {
            long ip = (i + 1) % nx, k = j * nx + i;
            double u00 = u[j * nx + i], u10 = u[j * nx + ip], u01 = u[jp * nx + i],
                   u11 = u[jp * nx + ip];
            double v00 = v[j * nx + i], v10 = v[j * nx + ip], v01 = v[jp * nx + i],
                   v11 = v[jp * nx + ip];
            for (int a = 0; a < nq; ++a)
                for (int b = 0; b < nq; ++b) {
                    double xi = gx[a], eta = gx[b], w = gw[a] * gw[b];
                    double uu = u00 * (1 - xi) * (1 - eta) + u10 * xi * (1 - eta)
                        + u01 * (1 - xi) * eta + u11 * xi * eta;
                    double vv = v00 * (1 - xi) * (1 - eta) + v10 * xi * (1 - eta)
                        + v01 * (1 - xi) * eta + v11 * xi * eta;
                    double p = eval_cell(phi, ncell, k, ncomp, xi, eta);
                    double dx[6], dy[6];
                    basis_dxi(xi, eta, dx);
                    basis_deta(xi, eta, dy);
                    for (int c = 0; c < ncomp; ++c)
                        out[(size_t)c * ncell + k] += w * p * (uu / hx * dx[c] + vv / hy * dy[c]);
                }
        }
    }
    /* vertical edges: edge (i,j) is the left edge of cell (i,j), between cells (i-1,j) and (i,j) */
#pragma omp parallel for schedule(static)
    for (long j = 0; j < ny; ++j) {
        long jp = (j + 1) % ny;
        for (long i = 0; i < nx; ++i) {
            long im = (i + nx - 1) % nx;
            long kL = j * nx + im, kR = j * nx + i;
            double ua = u[j * nx + i], ub = u[jp * nx + i];
            for (int b = 0; b < nq; ++b) {
                double eta = gx[b], w = gw[b];
                double un = ua * (1 - eta) + ub * eta;
                double pL = eval_cell(phi, ncell, kL, ncomp, 1., eta);
                double pR = eval_cell(phi, ncell, kR, ncomp, 0., eta);
                double flux = fmax(un, 0.) * pL + fmin(un, 0.) * pR;
                double psiL[6], psiR[6];
                basis(1., eta, psiL);
                basis(0., eta, psiR);
                for (int c = 0; c < ncomp; ++c) {
                    /* the two cells of one edge lie in the same row: no race between rows */
                    out[(size_t)c * ncell + kL] -= w * flux * psiL[c] / hx;
                    out[(size_t)c * ncell + kR] += w * flux * psiR[c] / hx;
                }
            }
        }
    }
    /* horizontal edges: edge (i,j) is the bottom edge of cell (i,j); serial over rows so that the
     * two cells of an edge (rows j-1 and j) are never updated concurrently */
    for (long j = 0; j < ny; ++j) {
        long jm = (j + ny - 1) % ny;
#pragma omp parallel for schedule(static)
        for (long i = 0; i < nx; ++i) {
            long ip = (i + 1) % nx;
            long kB = jm * nx + i, kT = j * nx + i;
            double va = v[j * nx + i], vb = v[j * nx + ip];
            for (int a = 0; a < nq; ++a) {
                double xi = gx[a], w = gw[a];
                double vn = va * (1 - xi) + vb * xi;
                double pB = eval_cell(phi, ncell, kB, ncomp, xi, 1.);
                double pT = eval_cell(phi, ncell, kT, ncomp, xi, 0.);
                double flux = fmax(vn, 0.) * pB + fmin(vn, 0.) * pT;
                double psiB[6], psiT[6];
                basis(xi, 1., psiB);
                basis(xi, 0., psiT);
                for (int c = 0; c < ncomp; ++c) {
                    out[(size_t)c * ncell + kB] -= w * flux * psiB[c] / hy;
                    out[(size_t)c * ncell + kT] += w * flux * psiT[c] / hy;
                }
            }
        }
    }
#pragma omp parallel for schedule(static)
    for (long k = 0; k < ncell; ++k)
        for (int c = 0; c < ncomp; ++c)
            out[(size_t)c * ncell + k] *= INV_MASS[c];
}

/* bounds-preserving rescaling of the higher moments (see header) */
void nsdg_oracle_limit(long ncell, int ncomp, double* phi, double lo, double hi)
{
    if (ncomp == 1)
        return;
    static const double P3[3] = { 0., 0.5, 1. };
#pragma omp parallel for schedule(static)
    for (long k = 0; k < ncell; ++k) {
        double mean = phi[k];
        double mn = mean, mx = mean;
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) {
                double p = eval_cell(phi, ncell, k, ncomp, P3[a], P3[b]);
                mn = fmin(mn, p);
                mx = fmax(mx, p);
            }
        double theta = 1.;
        if (mn < lo)
            theta = fmin(theta, fmax(0., (mean - lo) / (mean - mn)));
        if (mx > hi)
            theta = fmin(theta, fmax(0., (hi - mean) / (mx - mean)));
        if (theta < 1.)
            for (int c = 1; c < ncomp; ++c)
                phi[(size_t)c * ncell + k] *= theta;
    }
}

/* one time step of length dt; rk in {1,2,3}; limiter: 0 off, 1 on with bounds [lo,hi] */
void nsdg_oracle_transport_step(long nx, long ny, double hx, double hy, int ncomp, int rk,
    double dt, double* phi, const double* u, const double* v, int limiter, double lo, double hi)
{
    const size_t n = (size_t)ncomp * nx * ny;
    double* L = (double*)malloc(n * sizeof(double));
    double* p1 = (double*)malloc(n * sizeof(double));
    if (rk == 0)
        rk = (ncomp == 1) ? 1 : (ncomp == 3 ? 2 : 3);
    /* stage 1: p1 = phi + dt L(phi) */
    nsdg_oracle_transport_rhs(nx, ny, hx, hy, ncomp, phi, u, v, L);
    for (size_t q = 0; q < n; ++q)
        p1[q] = phi[q] + dt * L[q];
    if (limiter)
        nsdg_oracle_limit(nx * ny, ncomp, p1, lo, hi);
    if (rk == 1) {
        memcpy(phi, p1, n * sizeof(double));
    } else if (rk == 2) {
        /* Heun: phi = 1/2 phi + 1/2 (p1 + dt L(p1)) */
        nsdg_oracle_transport_rhs(nx, ny, hx, hy, ncomp, p1, u, v, L);
        for (size_t q = 0; q < n; ++q)
            phi[q] = 0.5 * phi[q] + 0.5 * (p1[q] + dt * L[q]);
        if (limiter)
            nsdg_oracle_limit(nx * ny, ncomp, phi, lo, hi);
    } else {
        /* SSP-RK3: p2 = 3/4 phi + 1/4 (p1 + dt L(p1)); phi = 1/3 phi + 2/3 (p2 + dt L(p2)) */
        nsdg_oracle_transport_rhs(nx, ny, hx, hy, ncomp, p1, u, v, L);
        for (size_t q = 0; q < n; ++q)
            p1[q] = 0.75 * phi[q] + 0.25 * (p1[q] + dt * L[q]);
        if (limiter)
            nsdg_oracle_limit(nx * ny, ncomp, p1, lo, hi);
        nsdg_oracle_transport_rhs(nx, ny, hx, hy, ncomp, p1, u, v, L);
        for (size_t q = 0; q < n; ++q)
            phi[q] = (1. / 3.) * phi[q] + (2. / 3.) * (p1[q] + dt * L[q]);
        if (limiter)
            nsdg_oracle_limit(nx * ny, ncomp, phi, lo, hi);
    }
    free(L);
    free(p1);
}
