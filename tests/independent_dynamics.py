"""A second, independently structured statement of the dynamics scheme of DESIGN.md section 2,
used to pin oracle/dg_transport.c and oracle/mevp.c (tests/test_oracle_independent.py).

It is written from the equations, not from the C loops:
  * the modal basis is built from numpy's Legendre polynomials (values and derivatives), the mass
    matrix comes out of a quadrature instead of the constants 1, 1/12, 1/180, 1/144;
  * everything is in PHYSICAL coordinates (gradients 1/hx, 1/hy, area hx hy, edge lengths), where
    the C code works on the reference square with pre-scaled velocities;
  * integrals are global array expressions over all cells / all edges at once, assembled by
    scatter-add (np.add.at) from the side of the integral, where the C code gathers per cell;
  * smooth integrands are over-integrated (6-point Gauss) to check the claimed exactness of the
    scheme's own 1/2/3-point rules; only the upwind edge fluxes use the scheme's rule, because a
    sign change of the normal velocity inside an edge makes the rule part of the scheme;
  * the mEVP stress divergence is the weak form -(S, grad phi_n) integrated numerically against
    the bilinear hat functions, where the C code uses closed-form dG1 x Q1 integrals.

Equation references: Richter et al., GMD 2023 (DG advection, eq. for the upwind flux; CG1 / dG1
velocity-stress pairing); Kimmritz et al., Ocean Modelling 2016, eqs. (7)-(10) (mEVP stress and
momentum sub-iteration with alpha, beta); Hibler 1979 (VP law with the elliptic yield curve);
Mehlmann et al. 2021 (momentum equation and forcing of the box benchmark).
"""
import numpy as np
from numpy.polynomial import legendre as L


# ------------------------------------------------------------------ basis from Legendre polynomials
def _leg(n, x, deriv=0):
    c = np.zeros(n + 1)
    c[n] = 1.0
    if deriv:
        c = L.legder(c, deriv)
    return L.legval(x, c)


# psi_k(xi, eta) on [0,1]^2 as products of shifted Legendre polynomials P_n(2 xi - 1), scaled so
# that they equal 1, xi-1/2, eta-1/2, (xi-1/2)^2-1/12, (eta-1/2)^2-1/12, (xi-1/2)(eta-1/2)
_MODES = ((0, 0, 1.0), (1, 0, 0.5), (0, 1, 0.5), (2, 0, 1.0 / 6.0), (0, 2, 1.0 / 6.0), (1, 1, 0.25))


def psi(k, xi, eta):
    a, b, s = _MODES[k]
    return s * _leg(a, 2 * xi - 1) * _leg(b, 2 * eta - 1)


def dpsi(k, xi, eta):
    """(d/dxi, d/deta) on the reference square"""
    a, b, s = _MODES[k]
    return (s * 2 * _leg(a, 2 * xi - 1, 1) * _leg(b, 2 * eta - 1),
            s * 2 * _leg(a, 2 * xi - 1) * _leg(b, 2 * eta - 1, 1))


def gauss01(n):
    x, w = L.leggauss(n)
    return 0.5 * (x + 1), 0.5 * w


def mass_diag(ncomp, hx, hy):
    x, w = gauss01(5)
    X, Y = np.meshgrid(x, x, indexing="ij")
    W = np.outer(w, w)
    m = np.array([[np.sum(W * psi(i, X, Y) * psi(j, X, Y)) for j in range(ncomp)] for i in range(ncomp)])
    assert np.abs(m - np.diag(np.diag(m))).max() < 1e-15  # orthogonal basis
    return np.diag(m) * hx * hy


def eval_dg(phi, xi, eta):
    return sum(phi[k] * psi(k, xi, eta) for k in range(phi.shape[0]))


def _q1(u, xi, eta):
    """bilinear field with nodal values u[j][i] (node = lower-left corner), evaluated at the
    reference point (xi, eta) of every cell, periodic"""
    u10 = np.roll(u, -1, axis=1)
    u01 = np.roll(u, -1, axis=0)
    u11 = np.roll(u10, -1, axis=0)
    return u * (1 - xi) * (1 - eta) + u10 * xi * (1 - eta) + u01 * (1 - xi) * eta + u11 * xi * eta


# ------------------------------------------------------------------ DG transport
def transport_rhs(phi, u, v, hx, hy):
    """d(phi)/dt of  M dphi_i/dt = int_K phi v.grad(psi_i) dx - sum_e int_e (v.n) phi^ psi_i ds
    with the pointwise upwind trace phi^ (Richter et al. 2023)."""
    nc, ny, nx = phi.shape
    nq = {1: 1, 3: 2, 6: 3}[nc]
    r = np.zeros_like(phi)
    # volume term, over-integrated
    x, w = gauss01(6)
    for a, wa in zip(x, w):
        for b, wb in zip(x, w):
            f = eval_dg(phi, a, b)
            uu, vv = _q1(u, a, b), _q1(v, a, b)
            for k in range(nc):
                dxi, deta = dpsi(k, a, b)
                r[k] += wa * wb * f * (uu * dxi / hx + vv * deta / hy) * (hx * hy)
    # edges, the scheme's own rule.  Vertical edge i: between cell (i-1, j) [minus side, its xi = 1]
    # and cell (i, j) [plus side, xi = 0], normal +x, length hy; normal velocity = u on the edge,
    # linear between node (i, j) and node (i, j+1).
    s, ws = gauss01(nq)
    for t, wt in zip(s, ws):
        un = u * (1 - t) + np.roll(u, -1, axis=0) * t
        minus = np.roll(eval_dg(phi, 1.0, t), 1, axis=1)  # value of the left cell on this edge
        plus = eval_dg(phi, 0.0, t)
        flux = np.where(un > 0, un * minus, un * plus) * wt * hy
        for k in range(nc):
            r[k] += flux * psi(k, 0.0, t)  # inflow into the right cell through its left edge
            r[k] -= np.roll(flux, -1, axis=1) * psi(k, 1.0, t)  # outflow of the cell through its right edge
        vn = v * (1 - t) + np.roll(v, -1, axis=1) * t
        minus = np.roll(eval_dg(phi, t, 1.0), 1, axis=0)
        plus = eval_dg(phi, t, 0.0)
        flux = np.where(vn > 0, vn * minus, vn * plus) * wt * hx
        for k in range(nc):
            r[k] += flux * psi(k, t, 0.0)
            r[k] -= np.roll(flux, -1, axis=0) * psi(k, t, 1.0)
    return r / mass_diag(nc, hx, hy)[:, None, None]


def transport_step(phi, u, v, hx, hy, dt):
    """RK1 / Heun / Shu-Osher SSP-RK3 for dG0 / dG1 / dG2"""
    Lh = lambda p: transport_rhs(p, u, v, hx, hy)
    nc = phi.shape[0]
    if nc == 1:
        return phi + dt * Lh(phi)
    p1 = phi + dt * Lh(phi)
    if nc == 3:
        return 0.5 * phi + 0.5 * (p1 + dt * Lh(p1))
    p2 = 0.75 * phi + 0.25 * (p1 + dt * Lh(p1))
    return phi / 3.0 + (2.0 / 3.0) * (p2 + dt * Lh(p2))


# ------------------------------------------------------------------ mEVP
def _hat_grad(c, xi, eta, hx, hy):
    """physical gradient of the bilinear hat function of corner c = (cx, cy) of a cell"""
    cx, cy = c
    nx_ = (1.0 if cx else -1.0) * (eta if cy else 1 - eta) / hx
    ny_ = (xi if cx else 1 - xi) * (1.0 if cy else -1.0) / hy
    return nx_, ny_


CORNERS = ((0, 0), (1, 0), (0, 1), (1, 1))


def _corner(u, c):
    return np.roll(np.roll(u, -c[0], axis=1), -c[1], axis=0)


def strain_at(u, v, xi, eta, hx, hy):
    e11 = e22 = e12 = 0.0
    for c in CORNERS:
        gx, gy = _hat_grad(c, xi, eta, hx, hy)
        uc, vc = _corner(u, c), _corner(v, c)
        e11 = e11 + uc * gx
        e22 = e22 + vc * gy
        e12 = e12 + 0.5 * (uc * gy + vc * gx)
    return e11, e12, e22


def vp_stress(P, e11, e12, e22, strength):
    """Hibler's viscous-plastic law with the elliptic yield curve and the replacement pressure
    P/2 subtracted as in Kimmritz et al. 2016 eq. (3)-(5)"""
    e2 = P.ellipse_e ** 2
    delta = np.sqrt(P.delta_min ** 2 + (e11 + e22) ** 2 + ((e11 - e22) ** 2 + 4 * e12 ** 2) / e2)
    zeta = strength / (2 * delta)
    eta = zeta / e2
    tr = e11 + e22
    return (2 * eta * e11 + (zeta - eta) * tr - strength / 2, 2 * eta * e12,
            2 * eta * e22 + (zeta - eta) * tr - strength / 2)


def stress_target(P, H, A, u, v, hx, hy):
    """L2 projection to dG1 of the VP stress evaluated at the 2x2 Gauss points (the points are
    part of the scheme: the strength P_g and the nonlinear law are sampled there)"""
    x, w = gauss01(2)
    num = [np.zeros((3,) + u.shape) for _ in range(3)]
    den = np.zeros(3)
    for a, wa in zip(x, w):
        for b, wb in zip(x, w):
            strength = P.p_star * eval_dg(H, a, b) * np.exp(-P.c_star * (1 - eval_dg(A, a, b)))
            s = vp_stress(P, *strain_at(u, v, a, b, hx, hy), strength)
            for k in range(3):
                pk = psi(k, a, b)
                den[k] += wa * wb * pk * pk
                for q in range(3):
                    num[q][k] += wa * wb * s[q] * pk
    return [n / den[:, None, None] for n in num]


def stress_divergence(S11, S12, S22, hx, hy):
    """D_n = -(S, grad phi_n): integral over the four cells around node n of -S : grad(phi_n),
    assembled from the cells' side by scatter-add, 3x3 Gauss (over-integrated)"""
    ny, nx = S11.shape[1:]
    dx = np.zeros((ny, nx))
    dy = np.zeros((ny, nx))
    jj, ii = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
    x, w = gauss01(3)
    for a, wa in zip(x, w):
        for b, wb in zip(x, w):
            s11, s12, s22 = eval_dg(S11, a, b), eval_dg(S12, a, b), eval_dg(S22, a, b)
            for c in CORNERS:
                gx, gy = _hat_grad(c, a, b, hx, hy)
                tgt = ((jj + c[1]) % ny, (ii + c[0]) % nx)
                np.add.at(dx, tgt, -(s11 * gx + s12 * gy) * wa * wb * hx * hy)
                np.add.at(dy, tgt, -(s12 * gx + s22 * gy) * wa * wb * hx * hy)
    return dx, dy


def node_mean(Hmean):
    """mean of the four cells around node (i, j): cells (i-1, j-1), (i, j-1), (i-1, j), (i, j)"""
    return 0.25 * (Hmean + np.roll(Hmean, 1, axis=1) + np.roll(Hmean, 1, axis=0)
                   + np.roll(np.roll(Hmean, 1, axis=0), 1, axis=1))


def dirichlet_mask(P, ny, nx):
    m = np.zeros((ny, nx), bool)
    if P.bc == 1:
        m[0, :] = True
        m[:, 0] = True
    return m


def momentum_terms(P, H, A, u, v, un, vn, S, forcing, hx, hy, dt):
    """the pieces of Kimmritz et al. 2016 eq. (10) at every node"""
    uatm, vatm, uocn, vocn = forcing
    Hn, An = node_mean(H[0]), node_mean(A[0])
    m = P.rho_ice * np.maximum(Hn, P.h_min)
    wind = np.hypot(uatm, vatm)
    tax, tay = P.c_atm * P.rho_atm * wind * uatm, P.c_atm * P.rho_atm * wind * vatm
    cd = An * P.c_ocean * P.rho_ocean * np.hypot(uocn - u, vocn - v)
    dx, dy = stress_divergence(*S, hx, hy)
    dx, dy = dx / (hx * hy), dy / (hx * hy)  # lumped mass: one cell area per node
    return m, An, tax, tay, cd, dx, dy


def mevp_subiteration(P, H, A, u, v, un, vn, S, forcing, hx, hy, dt):
    """one sub-iteration p -> p+1 (Kimmritz et al. 2016 eqs. (9), (10)): stress first, then the
    velocity with the NEW stress; drag and Coriolis explicit in u^p except the implicit part of
    the ocean drag"""
    target = stress_target(P, H, A, u, v, hx, hy)
    S = [s + (t - s) / P.alpha for s, t in zip(S, target)]
    uatm, vatm, uocn, vocn = forcing
    m, An, tax, tay, cd, dx, dy = momentum_terms(P, H, A, u, v, un, vn, S, forcing, hx, hy, dt)
    f = P.f_coriolis
    den = m * (1 + P.beta) / dt + cd
    u1 = (m / dt * (P.beta * u + un) + An * tax + cd * uocn + m * f * (v - vocn) + dx) / den
    v1 = (m / dt * (P.beta * v + vn) + An * tay + cd * vocn - m * f * (u - uocn) + dy) / den
    mask = dirichlet_mask(P, *u.shape)
    u1[mask] = 0.0
    v1[mask] = 0.0
    return u1, v1, S


def mevp_step(P, H, A, u, v, S11, S12, S22, forcing, hx, hy, dt, nsub):
    un, vn = u.copy(), v.copy()
    S = [S11, S12, S22]
    for _ in range(nsub):
        u, v, S = mevp_subiteration(P, H, A, u, v, un, vn, S, forcing, hx, hy, dt)
    return u, v, S[0], S[1], S[2]


def vp_residual(P, H, A, u, v, un, vn, forcing, hx, hy, dt):
    """residual of the implicit VP momentum equation the sub-cycle converges to (Kimmritz et al.
    2016 eq. (6)): stress = projected VP stress of (u, v), everything at the new time level"""
    S = stress_target(P, H, A, u, v, hx, hy)
    uatm, vatm, uocn, vocn = forcing
    m, An, tax, tay, cd, dx, dy = momentum_terms(P, H, A, u, v, un, vn, S, forcing, hx, hy, dt)
    f = P.f_coriolis
    rx = m / dt * (u - un) - (An * tax + cd * (uocn - u) + m * f * (v - vocn) + dx)
    ry = m / dt * (v - vn) - (An * tay + cd * (vocn - v) - m * f * (u - uocn) + dy)
    mask = dirichlet_mask(P, *u.shape)
    rx[mask] = 0.0
    ry[mask] = 0.0
    return rx, ry
