"""Analytic known-answer tests that pin the DG transport and mEVP restatements
(oracle/dg_transport.c, oracle/mevp.c).  These parts are NOT in the reference snapshot:
parity is unpinned against draenog/nextsimdg, the KATs below are what anchors the oracle."""
import numpy as np
import pytest

from nextsimdg_b200 import synthetic as S
from nextsimdg_b200._abi import BC_BOX, MevpParams
from oracle import cpu_oracle as O


def smooth(x, y):
    return np.sin(2 * np.pi * x) * np.cos(2 * np.pi * y) + 2.0


@pytest.mark.parametrize("ncomp", [1, 3, 6])
def test_constant_field_is_steady(ncomp):
    nx, ny = 24, 16
    c = np.zeros((ncomp, ny, nx))
    c[0] = 3.0
    # uniform flow: every component of the RHS vanishes
    r = O.transport_rhs(c, np.full((ny, nx), 0.7), np.full((ny, nx), -0.3), 1 / nx, 1 / ny)
    assert np.abs(r).max() < 1e-11


@pytest.mark.parametrize("ncomp", [1, 3, 6])
def test_mass_conservation(ncomp):
    nx, ny = 20, 28
    hx, hy = 1 / nx, 1 / ny
    rng = np.random.default_rng(1)
    phi = S.project(smooth, nx, ny, hx, hy, ncomp) + 0.1 * rng.standard_normal((ncomp, ny, nx))
    u, v = rng.standard_normal((ny, nx)), rng.standard_normal((ny, nx))
    m0 = phi[0].sum()
    for _ in range(5):
        phi = O.transport_step(phi, u, v, hx, hy, 0.002)
    assert abs(phi[0].sum() - m0) / abs(m0) < 1e-13


@pytest.mark.parametrize("ncomp,order", [(1, 1), (3, 2), (6, 3)])
def test_convergence_order(ncomp, order):
    errs = []
    T = 0.1
    for n in (16, 32):
        hx = hy = 1 / n
        phi = S.project(smooth, n, n, hx, hy, ncomp)
        u, v = np.full((n, n), 0.7), np.full((n, n), -0.3)
        nst = int(round(T / (0.05 * hx)))
        for _ in range(nst):
            phi = O.transport_step(phi, u, v, hx, hy, T / nst)
        errs.append(S.l2_error(phi, lambda x, y: smooth(x - 0.7 * T, y + 0.3 * T), hx, hy))
    assert np.log2(errs[0] / errs[1]) > order - 0.15


def test_polynomial_translation_exact_in_space():
    """dG2 represents quadratics exactly: L(phi) of a quadratic under uniform flow equals the
    projection of -v.grad(phi) away from the periodic seam."""
    nx = ny = 16
    hx = hy = 1.0 / nx
    f = lambda x, y: 1 + x * x - 0.5 * x * y + 2 * y
    phi = S.project(f, nx, ny, hx, hy, 6)
    u, v = np.full((ny, nx), 0.3), np.full((ny, nx), 0.2)
    r = O.transport_rhs(phi, u, v, hx, hy)
    exact = S.project(lambda x, y: -(0.3 * (2 * x - 0.5 * y) + 0.2 * (-0.5 * x + 2)), nx, ny,
                      hx, hy, 6)
    assert np.abs(r[:, 2:-2, 2:-2] - exact[:, 2:-2, 2:-2]).max() < 1e-10


def test_limiter_keeps_mean_and_bounds():
    nx = ny = 12
    rng = np.random.default_rng(2)
    phi = rng.standard_normal((6, ny, nx))
    phi[0] = rng.uniform(0.05, 0.95, (ny, nx))
    out = np.ascontiguousarray(phi.copy())
    import ctypes
    O.lib().nsdg_oracle_limit(nx * ny, 6, out.ctypes.data_as(ctypes.POINTER(ctypes.c_double)),
                              0.0, 1.0)
    assert np.array_equal(out[0], phi[0])
    for a in (0, 0.5, 1):
        for b in (0, 0.5, 1):
            val = S.evaluate(out, a, b)
            assert val.min() > -1e-12 and val.max() < 1 + 1e-12


# ------------------------------------------------------------------------- mEVP
def test_uniform_velocity_relaxes_to_minus_half_pressure():
    nx, ny = 16, 12
    F, hx, hy = S.box_case(nx, ny, 6)
    p = MevpParams.defaults()
    p.alpha = 1.0  # one relaxation step lands on sigma_VP
    pg = O.mevp_pressure(p, F["H"], F["A"])
    u, v = np.full((ny, nx), 0.1), np.full((ny, nx), -0.2)
    s11, s12, s22 = O.mevp_stress(p, u, v, pg, F["S11"], F["S12"], F["S22"], hx, hy)
    # zero strain: sigma = -P/2 I; its dG1 projection has mean = mean of the 4 Gauss values
    assert np.allclose(s11[0], -0.5 * pg.mean(0), rtol=1e-14)
    assert np.allclose(s22[0], -0.5 * pg.mean(0), rtol=1e-14)
    assert np.abs(s12).max() == 0.0


def test_linear_shear_strain_and_stress():
    """u = g*y, v = 0 (away from the periodic seam): e12 = g/2, e11 = e22 = 0, hence
    Delta = sqrt(dmin^2 + g^2/e^2), s12 = 2 eta e12 = P g /(2 e^2 Delta), s11 = -P/2."""
    nx, ny = 8, 8
    hx = hy = 1000.0
    g = 1e-6
    yn = (np.arange(ny) * hy)[:, None] + 0 * np.arange(nx)[None, :]
    u, v = g * yn, np.zeros((ny, nx))
    p = MevpParams.defaults()
    p.alpha = 1.0
    H = np.zeros((1, ny, nx))
    H[0] = 2.0
    A = np.ones((1, ny, nx))
    pg = O.mevp_pressure(p, H, A)
    P = p.p_star * 2.0
    assert np.allclose(pg, P)
    z = np.zeros((3, ny, nx))
    s11, s12, s22 = O.mevp_stress(p, u, v, pg, z, z, z, hx, hy)
    delta = np.sqrt(p.delta_min ** 2 + g * g / p.ellipse_e ** 2)
    inner = (slice(1, ny - 1), slice(None))
    assert np.allclose(s12[0][inner], P * g / (2 * p.ellipse_e ** 2 * delta), rtol=1e-13)
    assert np.allclose(s11[0][inner], -0.5 * P, rtol=1e-13)
    assert np.abs(s12[1:][:, 1:ny - 1]).max() < 1e-9 * P


def test_box_subcycle_residual_decreases_and_walls_hold():
    nx = ny = 32
    F, hx, hy = S.box_case(nx, ny, 6)
    p = MevpParams.defaults()
    p.bc = BC_BOX
    u, v, s11, s12, s22, res = O.mevp_step(p, F["H"], F["A"], F["U"], F["V"], F["S11"], F["S12"],
                                           F["S22"], F["UATM"], F["VATM"], F["UOCN"], F["VOCN"],
                                           hx, hy, 120.0, 100, want_resid=True)
    assert np.all(np.diff(res) < 0), "sub-cycle residual must decrease monotonically"
    assert np.all(u[0] == 0) and np.all(u[:, 0] == 0) and np.all(v[0] == 0) and np.all(v[:, 0] == 0)
    assert 0 < np.abs(u).max() < 0.1


def test_stress_divergence_of_constant_stress_vanishes():
    nx, ny = 10, 10
    F, hx, hy = S.box_case(nx, ny, 1)
    p = MevpParams.defaults()
    p.c_atm = p.c_ocean = p.f_coriolis = 0.0
    z = np.zeros((ny, nx))
    S0 = np.zeros((3, ny, nx))
    S0[0] = 123.0
    # huge alpha: stress stays (almost) the constant S0 during the one sub-iteration
    p.alpha = 1e30
    u, v, *_ = O.mevp_step(p, F["H"], F["A"], z, z, S0, S0, S0, z, z, z, z, hx, hy, 10.0, 1)
    assert np.abs(u).max() < 1e-15 and np.abs(v).max() < 1e-15
