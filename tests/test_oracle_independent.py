"""Independent pins of the dynamics oracle (oracle/dg_transport.c, oracle/mevp.c).

The reference snapshot has no dynamics code, so the C oracle cannot be checked against it.
Instead it is checked here against
  (a) a second implementation written from the equations with a different structure
      (tests/independent_dynamics.py: Legendre-built basis, physical coordinates, global
      quadrature with scatter-add assembly, over-integration) on random inputs with EVERY term
      switched on (Coriolis, wind, ocean current, ocean drag, non-uniform H and A, dG1 stress with
      all nine coefficients, box and periodic topology);
  (b) manufactured smooth solutions of the CONTINUOUS equations: the discrete stress update must
      converge to Hibler's VP stress of the exact strain rates and the discrete momentum update to
      the one with the analytic stress divergence, both at second order under refinement;
  (c) the fixed point of the sub-cycle: a converged mEVP iteration satisfies the implicit VP
      momentum equation (Kimmritz et al. 2016, eq. (6)) as evaluated by (a).
CPU only; the CUDA kernels are tied to the oracle by tests/test_gpu_dynamics.py.
"""
import numpy as np
import pytest

import independent_dynamics as I
from nextsimdg_b200._abi import BC_BOX, BC_PERIODIC, MevpParams
from oracle import cpu_oracle as O


def rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def random_state(nx, ny, bc, seed, nc=6):
    r = np.random.default_rng(seed)
    H = np.zeros((nc, ny, nx))
    A = np.zeros((nc, ny, nx))
    H[0] = r.uniform(0.2, 2.0, (ny, nx))
    A[0] = r.uniform(0.6, 0.98, (ny, nx))
    for k in range(1, nc):
        H[k] = 0.05 * r.standard_normal((ny, nx))
        A[k] = 0.01 * r.standard_normal((ny, nx))
    u = 0.1 * r.standard_normal((ny, nx))
    v = 0.1 * r.standard_normal((ny, nx))
    if bc == BC_BOX:
        u[0, :] = u[:, 0] = v[0, :] = v[:, 0] = 0.0
    S = [2e3 * r.standard_normal((3, ny, nx)) for _ in range(3)]
    forcing = [8.0 * r.standard_normal((ny, nx)), 8.0 * r.standard_normal((ny, nx)),
               0.1 * r.standard_normal((ny, nx)), 0.1 * r.standard_normal((ny, nx))]
    return H, A, u, v, S, forcing


# ---------------------------------------------------------------- (a) second implementation
@pytest.mark.parametrize("nc", [1, 3, 6])
def test_transport_rhs_and_step_against_independent_implementation(nc):
    """24 x 16 cells, random coefficients in every mode, random sign-changing velocity (the
    upwind side switches inside edges), anisotropic cells."""
    nx, ny, hx, hy = 24, 16, 1.3, 0.7
    r = np.random.default_rng(100 + nc)
    phi = r.standard_normal((nc, ny, nx))
    u, v = r.standard_normal((ny, nx)), r.standard_normal((ny, nx))
    got = O.transport_rhs(phi, u, v, hx, hy)
    want = I.transport_rhs(phi, u, v, hx, hy)
    assert rel(got, want) < 1e-12
    dt = 0.02
    assert rel(O.transport_step(phi, u, v, hx, hy, dt), I.transport_step(phi, u, v, hx, hy, dt)) < 1e-12


def test_mass_matrix_constants():
    """the diagonal mass the C code hard-wires (1, 1/12, 1/12, 1/180, 1/180, 1/144) comes out of
    the quadrature of the Legendre-built basis"""
    m = I.mass_diag(6, 2.0, 3.0) / 6.0
    assert np.allclose(m, [1, 1 / 12, 1 / 12, 1 / 180, 1 / 180, 1 / 144], rtol=1e-14)


@pytest.mark.parametrize("bc", [BC_PERIODIC, BC_BOX])
@pytest.mark.parametrize("nc", [3, 6])
def test_mevp_subiterations_against_independent_implementation(bc, nc):
    """three sub-iterations (so that u^p differs from u^n and the beta terms matter) with every
    term of the momentum equation on"""
    nx, ny, hx, hy, dt = 24, 16, 4000.0, 3000.0, 120.0
    H, A, u, v, S, forcing = random_state(nx, ny, bc, 7 + nc, nc)
    P = MevpParams.defaults()
    P.bc = bc
    P.alpha, P.beta = 40.0, 60.0  # distinct, so that a swapped alpha / beta cannot pass
    P.f_coriolis = 1.46e-4
    got = O.mevp_step(P, H, A, u, v, *S, *forcing, hx, hy, dt, 3)
    want = I.mevp_step(P, H, A, u, v, *S, forcing, hx, hy, dt, 3)
    for name, g, w in zip(("u", "v", "S11", "S12", "S22"), got, want):
        assert rel(g, w) < 1e-12, name
    # each term matters at this tolerance: dropping any of them changes the answer by much more
    for knob, val in (("f_coriolis", 0.0), ("c_atm", 0.0), ("c_ocean", 0.0)):
        Q = MevpParams.defaults()
        for fld, _ in MevpParams._fields_:
            setattr(Q, fld, getattr(P, fld))
        setattr(Q, knob, val)
        other = O.mevp_step(Q, H, A, u, v, *S, *forcing, hx, hy, dt, 3)
        assert rel(other[0], want[0]) > 1e-6, knob


def test_full_step_against_independent_implementation():
    """the north-star unit (transport of H and A, then the sub-cycle) end to end"""
    nx, ny, hx, hy, dt = 20, 12, 4000.0, 3000.0, 120.0
    H, A, u, v, S, forcing = random_state(nx, ny, BC_BOX, 3)
    P = MevpParams.defaults()
    P.bc = BC_BOX
    P.alpha = P.beta = 50.0
    got = O.dyn_step(P, H, A, u, v, *S, *forcing, hx, hy, dt, 4)
    H2, A2 = I.transport_step(H, u, v, hx, hy, dt), I.transport_step(A, u, v, hx, hy, dt)
    want = (H2, A2) + tuple(I.mevp_step(P, H2, A2, u, v, *S, forcing, hx, hy, dt, 4))
    for name, g, w in zip(("H", "A", "u", "v", "S11", "S12", "S22"), got, want):
        assert rel(g, w) < 1e-12, name


# ---------------------------------------------------------------- (b) manufactured solutions
LX = LY = 1.0e5


def smooth_fields(n):
    """smooth periodic u, v, stress, H, A, forcing on an n x n grid of a 100 km square, sampled
    at the nodes / projected to DG; returns also the analytic stress divergence at the nodes"""
    hx, hy = LX / n, LY / n
    kx, ky = 2 * np.pi / LX, 2 * np.pi / LY
    xn = (np.arange(n) * hx)[None, :] + 0 * np.arange(n)[:, None]
    yn = (np.arange(n) * hy)[:, None] + 0 * np.arange(n)[None, :]
    f = dict(
        u=lambda x, y: 0.3 * np.sin(kx * x) * np.cos(ky * y) + 0.1,
        v=lambda x, y: 0.2 * np.cos(kx * x + 0.3) * np.sin(2 * ky * y),
        s11=lambda x, y: 4e3 * np.sin(kx * x + 0.2) * np.cos(ky * y),
        s12=lambda x, y: 3e3 * np.cos(2 * kx * x) * np.sin(ky * y + 0.4),
        s22=lambda x, y: 5e3 * np.sin(kx * x) * np.sin(ky * y + 0.1),
        # analytic divergence: (d s11/dx + d s12/dy, d s12/dx + d s22/dy)
        divx=lambda x, y: 4e3 * kx * np.cos(kx * x + 0.2) * np.cos(ky * y)
        + 3e3 * ky * np.cos(2 * kx * x) * np.cos(ky * y + 0.4),
        divy=lambda x, y: -3e3 * 2 * kx * np.sin(2 * kx * x) * np.sin(ky * y + 0.4)
        + 5e3 * ky * np.sin(kx * x) * np.cos(ky * y + 0.1),
        H=lambda x, y: 1.0 + 0.3 * np.sin(kx * x) * np.sin(ky * y),
        A=lambda x, y: 0.8 + 0.1 * np.cos(kx * x) * np.cos(ky * y),
        ua=lambda x, y: 10.0 * np.cos(ky * y), va=lambda x, y: 6.0 * np.sin(kx * x),
        uo=lambda x, y: 0.1 * np.sin(ky * y), vo=lambda x, y: -0.1 * np.sin(kx * x))
    from nextsimdg_b200 import synthetic as S
    proj = lambda fn, nc: S.project(fn, n, n, hx, hy, nc, nq=4)
    return f, xn, yn, hx, hy, proj


def convergence_order(errs):
    return [np.log2(errs[i] / errs[i + 1]) for i in range(len(errs) - 1)]


def test_momentum_update_converges_to_the_continuous_equation():
    """Prescribed smooth stress, velocity, H, A and forcing: one velocity update of the oracle
    (alpha = 1e300 freezes the stress) against the same formula with the ANALYTIC stress
    divergence and point values of H, A at the nodes: second-order convergence, all terms on."""
    errs = []
    for n in (16, 32, 64, 128):
        f, xn, yn, hx, hy, proj = smooth_fields(n)
        P = MevpParams.defaults()
        P.alpha, P.beta, P.bc = 1e300, 300.0, BC_PERIODIC
        dt = 120.0
        H, A = proj(f["H"], 3), proj(f["A"], 3)
        u, v = f["u"](xn, yn), f["v"](xn, yn)
        S = [proj(f[k], 3) for k in ("s11", "s12", "s22")]
        forcing = [f[k](xn, yn) for k in ("ua", "va", "uo", "vo")]
        got_u, got_v = O.mevp_step(P, H, A, u, v, *S, *forcing, hx, hy, dt, 1)[:2]
        m = P.rho_ice * f["H"](xn, yn)
        An = f["A"](xn, yn)
        wind = np.hypot(forcing[0], forcing[1])
        cd = An * P.c_ocean * P.rho_ocean * np.hypot(forcing[2] - u, forcing[3] - v)
        den = m * (1 + P.beta) / dt + cd
        fc = P.f_coriolis
        ex_u = (m / dt * (P.beta * u + u) + An * P.c_atm * P.rho_atm * wind * forcing[0] + cd * forcing[2]
                + m * fc * (v - forcing[3]) + f["divx"](xn, yn)) / den
        ex_v = (m / dt * (P.beta * v + v) + An * P.c_atm * P.rho_atm * wind * forcing[1] + cd * forcing[3]
                - m * fc * (u - forcing[2]) + f["divy"](xn, yn)) / den
        # compare the INCREMENTS: u^{p+1} - u^p is what the terms produce
        errs.append(max(np.abs((got_u - u) - (ex_u - u)).max(), np.abs((got_v - v) - (ex_v - v)).max())
                    / max(np.abs(ex_u - u).max(), np.abs(ex_v - v).max()))
    orders = convergence_order(errs)
    assert errs[-1] < 2e-3 and all(o > 1.8 for o in orders), (errs, orders)


def test_stress_update_converges_to_the_continuous_vp_law():
    """alpha = 1: the new stress is the projected VP stress of the current velocity.  Its cell
    mean must converge at second order to Hibler's law evaluated with the exact strain rates and
    strength at the cell centre."""
    errs = []
    kx, ky = 2 * np.pi / LX, 2 * np.pi / LY
    for n in (16, 32, 64, 128):
        f, xn, yn, hx, hy, proj = smooth_fields(n)
        P = MevpParams.defaults()
        P.alpha, P.bc = 1.0, BC_PERIODIC
        # the strain rates of the manufactured velocity are ~2e-5 1/s and vanish together at
        # isolated points, where the law with the default Delta_min is not smooth (zeta jumps
        # by orders of magnitude within a cell); a Delta_min of the size of the strain rates
        # keeps the full nonlinearity of the ellipse and a smooth exact solution
        P.delta_min = 1e-5
        H, A = proj(f["H"], 6), proj(f["A"], 6)
        u, v = f["u"](xn, yn), f["v"](xn, yn)
        zero = [np.zeros((3, n, n)) for _ in range(3)]
        pg = O.mevp_pressure(P, H, A)
        S11, S12, S22 = O.mevp_stress(P, u, v, pg, *zero, hx, hy)
        xc, yc = xn + 0.5 * hx, yn + 0.5 * hy
        e11 = 0.3 * kx * np.cos(kx * xc) * np.cos(ky * yc)
        e22 = 0.2 * 2 * ky * np.cos(kx * xc + 0.3) * np.cos(2 * ky * yc)
        e12 = 0.5 * (-0.3 * ky * np.sin(kx * xc) * np.sin(ky * yc) - 0.2 * kx * np.sin(kx * xc + 0.3) * np.sin(2 * ky * yc))
        strength = P.p_star * f["H"](xc, yc) * np.exp(-P.c_star * (1 - f["A"](xc, yc)))
        w11, w12, w22 = I.vp_stress(P, e11, e12, e22, strength)
        scale = np.abs(w11).max()
        errs.append(max(np.abs(S11[0] - w11).max(), np.abs(S12[0] - w12).max(), np.abs(S22[0] - w22).max()) / scale)
    orders = convergence_order(errs)
    assert errs[-1] < 2e-3 and all(o > 1.8 for o in orders), (errs, orders)


# ---------------------------------------------------------------- (c) fixed point of the sub-cycle
def test_subcycle_converges_to_the_implicit_vp_equation():
    """A converged sub-cycle solves the implicit VP momentum equation.  Small closed box, wind and
    ocean forcing, Coriolis on; the residual of that equation, evaluated by the independent
    implementation, falls below 1e-8 of its initial value."""
    from nextsimdg_b200 import synthetic as S
    nx = ny = 12
    F, hx, hy = S.box_case(nx, ny, 3, L=96e3)
    P = MevpParams.defaults()
    P.bc = BC_BOX
    P.alpha = P.beta = 120.0
    P.delta_min = 2e-7  # keeps the viscosities (and the iteration count) of a 12 x 12 grid bounded
    dt = 120.0
    forcing = [F[k] for k in ("UATM", "VATM", "UOCN", "VOCN")]
    u0, v0 = F["U"], F["V"]
    S0 = [F[k] for k in ("S11", "S12", "S22")]
    r0 = I.vp_residual(P, F["H"], F["A"], u0, v0, u0, v0, forcing, hx, hy, dt)
    n0 = max(np.abs(r0[0]).max(), np.abs(r0[1]).max())
    u, v, *_ = O.mevp_step(P, F["H"], F["A"], u0, v0, *S0, *forcing, hx, hy, dt, 5000)
    r = I.vp_residual(P, F["H"], F["A"], u, v, u0, v0, forcing, hx, hy, dt)
    n1 = max(np.abs(r[0]).max(), np.abs(r[1]).max())
    assert n0 > 0 and n1 / n0 < 1e-8, (n0, n1)
    assert np.abs(u).max() > 1e-4  # it did go somewhere
