/* TEST INFRASTRUCTURE (oracle). Plain-C CPU restatement of the mEVP stress/velocity sub-cycle.
 *
 * Parity status: PARITY UNPINNED against draenog/nextsimdg (no dynamics code in the mounted
 * reference snapshot, CMakeLists.txt:43-46).  Scheme after Kimmritz et al. (2016) / Bouillon
 * et al. (2013) for mEVP, Richter et al. (GMD 2023) for the CG1-velocity / dG1-stress pairing,
 * and Mehlmann et al. (2021) for the benchmark parameters and momentum equation.  Pinned only
 * by analytic KATs (tests/test_oracle_dynamics.py): uniform velocity => zero strain; zero strain =>
 * stress relaxes to -P/2 I; hand-computed linear-shear strain rates; residual decrease.
 *
 * Discretisation
 *   velocity  CG1, node (i,j) = lower-left corner of cell (i,j), periodic topology; with
 *             bc == NSDG_BC_BOX the node lines i==0 and j==0 are Dirichlet (u=v=0), which makes
 *             the periodic grid equal to a closed box of nx*ny cells
 *   strain    e = sym grad u of the bilinear velocity, exactly representable per cell in
 *             span{1, xi-1/2, eta-1/2} (3 coefficients per component)
 *   stress    S11,S12,S22 in dG1 (3 coefficients each, 9 doubles per cell)
 *   VP law    at the 2x2 Gauss points g of each cell:
 *               Delta = sqrt(dmin^2 + (e11+e22)^2 + ((e11-e22)^2 + 4 e12^2)/e^2)
 *               zeta = P_g/(2 Delta), eta = zeta/e^2, P_g = P* H_g exp(-C (1 - A_g))
 *               s11 = 2 eta e11 + (zeta-eta)(e11+e22) - P_g/2, s22 likewise, s12 = 2 eta e12
 *             L2-projected back to dG1, then S <- S + (proj(s) - S)/alpha
 *   momentum  node n, lumped mass hx*hy, m = rho_i max(H_n,h_min), H_n,A_n = mean of the four
 *             adjacent cell means, D = -(S, grad phi_n) assembled from the 4 adjacent cells:
 *               cd   = A_n C_o rho_o |u_o - u^p|
 *               u^{p+1} = [ m/dt (beta u^p + u^n) + A_n tau_ax + cd u_o + m f (v^p - v_o)
 *                           + D_x/(hx hy) ] / ( m (1+beta)/dt + cd )
 *               v^{p+1} = [ m/dt (beta v^p + v^n) + A_n tau_ay + cd v_o - m f (u^p - u_o)
 *                           + D_y/(hx hy) ] / ( m (1+beta)/dt + cd )
 *             tau_a = C_a rho_a |u_a| u_a
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load this code.
 */
#include "../include/nsdg_gpu.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

void nsdg_oracle_mevp_default_params(nsdg_mevp_params* p)
{
    p->rho_ice = 900.;
    p->rho_atm = 1.3;
    p->rho_ocean = 1026.;
    p->c_atm = 1.2e-3;
    p->c_ocean = 5.5e-3;
    p->f_coriolis = 1.46e-4;
    p->p_star = 27.5e3;
    p->c_star = 20.;
    p->ellipse_e = 2.;
    p->delta_min = 2e-9;
    p->alpha = 1500.;
    p->beta = 1500.;
    p->h_min = 1e-3;
    p->bc = NSDG_BC_PERIODIC;
    p->_pad = 0;
}

static double eval_dg(const double* f, long ncell, long k, int ncomp, double xi, double eta)
{
    double a = xi - 0.5, b = eta - 0.5;
    double s = f[k];
    if (ncomp >= 3)
        s += f[ncell + k] * a + f[2 * ncell + k] * b;
    if (ncomp >= 6)
        s += f[3 * ncell + k] * (a * a - 1. / 12.) + f[4 * ncell + k] * (b * b - 1. / 12.)
            + f[5 * ncell + k] * a * b;
    return s;
}

/* ice strength at the 2x2 Gauss points, pg[g][j][i], g = 2*b + a (a: xi index, b: eta index) */
void nsdg_oracle_mevp_pressure(const nsdg_mevp_params* P, long nx, long ny, int ncomp,
    const double* H, const double* A, double* pg)
{
    const long ncell = nx * ny;
    const double d = 0.5 / sqrt(3.);
    const double gx[2] = { 0.5 - d, 0.5 + d };
#pragma omp parallel for schedule(static)
    for (long k = 0; k < ncell; ++k)
        for (int b = 0; b < 2; ++b)
            for (int a = 0; a < 2; ++a) {
                double h = eval_dg(H, ncell, k, ncomp, gx[a], gx[b]);
                double c = eval_dg(A, ncell, k, ncomp, gx[a], gx[b]);
                pg[(size_t)(2 * b + a) * ncell + k] = P->p_star * h * exp(-P->c_star * (1. - c));
            }
}

/* nodal thickness and concentration: mean of the 4 adjacent cell means */
void nsdg_oracle_mevp_nodal(long nx, long ny, const double* H, const double* A, double* Hn,
    double* An)
{
#pragma omp parallel for schedule(static)
    for (long j = 0; j < ny; ++j) {
        long jm = (j + ny - 1) % ny;
        for (long i = 0; i < nx; ++i) {
            long im = (i + nx - 1) % nx;
            Hn[j * nx + i]
                = 0.25 * (H[jm * nx + im] + H[jm * nx + i] + H[j * nx + im] + H[j * nx + i]);
            An[j * nx + i]
                = 0.25 * (A[jm * nx + im] + A[jm * nx + i] + A[j * nx + im] + A[j * nx + i]);
        }
    }
}

/* dG1 coefficients (mean, xi, eta) of the strain-rate components of cell (i,j) */
void nsdg_oracle_strain_cell(long nx, long ny, double hx, double hy, const double* u,
    const double* v, long i, long j, double* e11, double* e12, double* e22)
{
    long ip = (i + 1) % nx, jp = (j + 1) % ny;
    double u00 = u[j * nx + i], u10 = u[j * nx + ip], u01 = u[jp * nx + i], u11 = u[jp * nx + ip];
    double v00 = v[j * nx + i], v10 = v[j * nx + ip], v01 = v[jp * nx + i], v11 = v[jp * nx + ip];
    /* du/dx = [(u10-u00)(1-eta) + (u11-u01) eta]/hx ; du/dy = [(u01-u00)(1-xi) + (u11-u10) xi]/hy */
    double ax = u10 - u00, bx = u11 - u01, cy = u01 - u00, dy = u11 - u10;
    double avx = v10 - v00, bvx = v11 - v01, cvy = v01 - v00, dvy = v11 - v10;
    e11[0] = 0.5 * (ax + bx) / hx;
    e11[1] = 0.;
    e11[2] = (bx - ax) / hx;
    e22[0] = 0.5 * (cvy + dvy) / hy;
    e22[1] = (dvy - cvy) / hy;
    e22[2] = 0.;
    e12[0] = 0.5 * (0.5 * (cy + dy) / hy + 0.5 * (avx + bvx) / hx);
    e12[1] = 0.5 * (dy - cy) / hy;
    e12[2] = 0.5 * (bvx - avx) / hx;
}

/* S <- S + (proj(sigma_VP(u)) - S)/alpha for all cells; S = [3 comps][ny][nx] each */
void nsdg_oracle_mevp_stress(const nsdg_mevp_params* P, long nx, long ny, double hx, double hy,
    const double* u, const double* v, const double* pg, double* S11, double* S12, double* S22)
{
    const long ncell = nx * ny;
    const double d = 0.5 / sqrt(3.);
    const double gx[2] = { -d, d }; /* xi - 1/2 at the Gauss points */
    const double ie2 = 1. / (P->ellipse_e * P->ellipse_e);
    const double ialpha = 1. / P->alpha;
#pragma omp parallel for schedule(static)
    for (long j = 0; j < ny; ++j)
        for (long i = 0; i < nx; ++i) {
            long k = j * nx + i;
            double e11[3], e12[3], e22[3];
            nsdg_oracle_strain_cell(nx, ny, hx, hy, u, v, i, j, e11, e12, e22);
            double p11[3] = { 0, 0, 0 }, p12[3] = { 0, 0, 0 }, p22[3] = { 0, 0, 0 };
            for (int b = 0; b < 2; ++b)
                for (int a = 0; a < 2; ++a) {
                    double xa = gx[a], yb = gx[b];
                    double g11 = e11[0] + e11[1] * xa + e11[2] * yb;
                    double g12 = e12[0] + e12[1] * xa + e12[2] * yb;
                    double g22 = e22[0] + e22[1] * xa + e22[2] * yb;
                    double pr = pg[(size_t)(2 * b + a) * ncell + k];
                    double tr = g11 + g22, df = g11 - g22;
                    double delta = sqrt(P->delta_min * P->delta_min + tr * tr
                        + ie2 * (df * df + 4. * g12 * g12));
                    double zeta = pr / (2. * delta);
                    double eta = zeta * ie2;
                    double s11 = 2. * eta * g11 + (zeta - eta) * tr - 0.5 * pr;
                    double s22 = 2. * eta * g22 + (zeta - eta) * tr - 0.5 * pr;
                    double s12 = 2. * eta * g12;
                    /* L2 projection onto {1, xi-1/2, eta-1/2}: weights 1/4, inverse mass 1,12,12 */
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
            for (int c = 0; c < 3; ++c) {
                size_t q = (size_t)c * ncell + k;
                S11[q] += ialpha * (p11[c] - S11[q]);
                S12[q] += ialpha * (p12[c] - S12[q]);
                S22[q] += ialpha * (p22[c] - S22[q]);
            }
        }
}

/* -(S, grad phi_n) for node (i,j): contributions of the 4 adjacent cells, analytic integrals.
 * For the node at local corner (a,b) of a cell, sa = +1 if a==1 else -1 (sb likewise):
 *   int_K S dphi/dx = sa*hy*(s0/2 + sb*s_eta/12),  int_K S dphi/dy = sb*hx*(s0/2 + sa*s_xi/12) */
void nsdg_oracle_stress_div_node(long nx, long ny, double hx, double hy, const double* S11,
    const double* S12, const double* S22, long i, long j, double* Dx, double* Dy)
{
    const long ncell = nx * ny;
    long im = (i + nx - 1) % nx, jm = (j + ny - 1) % ny;
    const long cells[4] = { jm * nx + im, jm * nx + i, j * nx + im, j * nx + i };
    const double sa[4] = { 1., -1., 1., -1. }; /* node is the right corner of the left cells */
    const double sb[4] = { 1., 1., -1., -1. }; /* node is the top corner of the lower cells */
    double dx = 0, dy = 0;
    for (int q = 0; q < 4; ++q) {
        long k = cells[q];
        double ix11 = sa[q] * hy * (0.5 * S11[k] + sb[q] * S11[2 * ncell + k] / 12.);
        double iy12 = sb[q] * hx * (0.5 * S12[k] + sa[q] * S12[ncell + k] / 12.);
        double ix12 = sa[q] * hy * (0.5 * S12[k] + sb[q] * S12[2 * ncell + k] / 12.);
        double iy22 = sb[q] * hx * (0.5 * S22[k] + sa[q] * S22[ncell + k] / 12.);
        dx -= ix11 + iy12;
        dy -= ix12 + iy22;
    }
    *Dx = dx;
    *Dy = dy;
}

/* velocity update of one sub-iteration; (u,v) -> (unew,vnew) (distinct arrays) */
void nsdg_oracle_mevp_velocity(const nsdg_mevp_params* P, long nx, long ny, double hx, double hy,
    double dt, const double* u, const double* v, const double* un, const double* vn,
    const double* Hn, const double* An, const double* uatm, const double* vatm,
    const double* uocn, const double* vocn, const double* S11, const double* S12,
    const double* S22, double* unew, double* vnew)
{
    const double iarea = 1. / (hx * hy);
#pragma omp parallel for schedule(static)
    for (long j = 0; j < ny; ++j)
        for (long i = 0; i < nx; ++i) {
            long n = j * nx + i;
            if (P->bc == NSDG_BC_BOX && (i == 0 || j == 0)) {
                unew[n] = 0.;
                vnew[n] = 0.;
                continue;
            }
            double Dx, Dy;
            nsdg_oracle_stress_div_node(nx, ny, hx, hy, S11, S12, S22, i, j, &Dx, &Dy);
            double m = P->rho_ice * fmax(Hn[n], P->h_min);
            double a = An[n];
            double wa = sqrt(uatm[n] * uatm[n] + vatm[n] * vatm[n]);
            double tax = P->c_atm * P->rho_atm * wa * uatm[n];
            double tay = P->c_atm * P->rho_atm * wa * vatm[n];
            double du = uocn[n] - u[n], dv = vocn[n] - v[n];
            double cd = a * P->c_ocean * P->rho_ocean * sqrt(du * du + dv * dv);
            double mdt = m / dt;
            double denom = mdt * (1. + P->beta) + cd;
            double rx = mdt * (P->beta * u[n] + un[n]) + a * tax + cd * uocn[n]
                + m * P->f_coriolis * (v[n] - vocn[n]) + Dx * iarea;
            double ry = mdt * (P->beta * v[n] + vn[n]) + a * tay + cd * vocn[n]
                - m * P->f_coriolis * (u[n] - uocn[n]) + Dy * iarea;
            unew[n] = rx / denom;
            vnew[n] = ry / denom;
        }
}

/* nsub sub-iterations of one momentum step of length dt. H,A: DG tracers (ncomp comps).
 * resid (may be NULL, length nsub): max-norm of u^{p+1}-u^p per sub-iteration. */
void nsdg_oracle_mevp_step(const nsdg_mevp_params* P, long nx, long ny, double hx, double hy,
    int ncomp, double dt, int nsub, const double* H, const double* A, double* u, double* v,
    double* S11, double* S12, double* S22, const double* uatm, const double* vatm,
    const double* uocn, const double* vocn, double* resid)
{
    const size_t nn = (size_t)nx * ny;
    double* pg = (double*)malloc(4 * nn * sizeof(double));
    double* Hn = (double*)malloc(nn * sizeof(double));
    double* An = (double*)malloc(nn * sizeof(double));
    double* un = (double*)malloc(nn * sizeof(double));
    double* vn = (double*)malloc(nn * sizeof(double));
    double* u2 = (double*)malloc(nn * sizeof(double));
    double* v2 = (double*)malloc(nn * sizeof(double));
    nsdg_oracle_mevp_pressure(P, nx, ny, ncomp, H, A, pg);
    nsdg_oracle_mevp_nodal(nx, ny, H, A, Hn, An);
    memcpy(un, u, nn * sizeof(double));
    memcpy(vn, v, nn * sizeof(double));
    for (int p = 0; p < nsub; ++p) {
        nsdg_oracle_mevp_stress(P, nx, ny, hx, hy, u, v, pg, S11, S12, S22);
        nsdg_oracle_mevp_velocity(P, nx, ny, hx, hy, dt, u, v, un, vn, Hn, An, uatm, vatm, uocn,
            vocn, S11, S12, S22, u2, v2);
        if (resid) {
            double r = 0;
            for (size_t q = 0; q < nn; ++q)
                r = fmax(r, fmax(fabs(u2[q] - u[q]), fabs(v2[q] - v[q])));
            resid[p] = r;
        }
        memcpy(u, u2, nn * sizeof(double));
        memcpy(v, v2, nn * sizeof(double));
    }
    free(pg);
    free(Hn);
    free(An);
    free(un);
    free(vn);
    free(u2);
    free(v2);
}
