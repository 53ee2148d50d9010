"""Synthetic inputs for the BASELINE.json configs. numpy only."""
import numpy as np

THERMO_SEED = 20240611
INV_MASS = np.array([1., 12., 12., 180., 180., 144.])


def dev1_inputs(n=100):
    """run/dev1.cfg + dev1.res.nc content, fully determined by the reference's
    run/dev_res.py:9-20 and core/src/include/DummyExternalData.hpp:25-32 (wind stays 0)."""
    f = lambda v: np.full(n, float(v))
    state = dict(hice=f(0.1), cice=f(0.5), hsnow=f(0.0), sst=f(-1.0), sss=f(32.0), tice0=f(-1.0))
    forcing = dict(tair=f(-1), tdew=f(-4), slp=f(1e5), qsw=f(0), qlw=f(311), mld=f(10),
                   snowfall=f(0), wind=f(0))
    return state, forcing


def thermo_inputs(n, seed=THERMO_SEED):
    """Randomised column-physics inputs that exercise every branch."""
    r = np.random.default_rng(seed)
    cice = np.where(r.random(n) < 0.10, 0.0, r.uniform(0.05, 1.0, n))
    htrue = r.uniform(0.01, 3.0, n)
    thin = r.random(n) < 0.05
    htrue = np.where(thin, r.uniform(0.005, 0.02, n), htrue)
    hice = cice * htrue
    hsnow = np.where(r.random(n) < 0.30, 0.0, cice * r.uniform(0.0, 0.3, n))
    state = dict(hice=hice, cice=cice, hsnow=hsnow, sst=r.uniform(-1.8, 2.0, n),
                 sss=r.uniform(30.0, 35.0, n), tice0=r.uniform(-20.0, -0.3, n))
    tair = r.uniform(-30.0, 5.0, n)
    forcing = dict(tair=tair, tdew=tair - r.uniform(0.0, 5.0, n),
                   slp=r.uniform(98000.0, 103000.0, n), qsw=r.uniform(0.0, 300.0, n),
                   qlw=r.uniform(150.0, 350.0, n), mld=r.uniform(10.0, 50.0, n),
                   snowfall=np.where(r.random(n) < 0.5, 0.0, r.uniform(0.0, 1e-4, n)),
                   wind=r.uniform(0.0, 20.0, n))
    return state, forcing


# ------------------------------------------------------------------ DG helpers
def basis(xi, eta):
    a, b = xi - 0.5, eta - 0.5
    return [np.ones_like(a), a, b, a * a - 1 / 12, b * b - 1 / 12, a * b]


def project(fn, nx, ny, hx, hy, ncomp, nq=4, j0=0):
    """L2 projection of fn(x,y) onto the modal dG basis: [ncomp][ny][nx] (rows j0..j0+ny)."""
    gx, gw = np.polynomial.legendre.leggauss(nq)
    gx, gw = 0.5 * (gx + 1), 0.5 * gw
    i = np.arange(nx)[None, :]
    j = (j0 + np.arange(ny))[:, None]
    out = np.zeros((ncomp, ny, nx))
    for a in range(nq):
        for b in range(nq):
            x = (i + gx[a]) * hx + 0 * j
            y = (j + gx[b]) * hy + 0 * i
            f = fn(x, y)
            psi = basis(np.full_like(x, gx[a]), np.full_like(x, gx[b]))
            for c in range(ncomp):
                out[c] += gw[a] * gw[b] * f * psi[c]
    return out * INV_MASS[:ncomp, None, None]


def evaluate(phi, xi, eta):
    """Value of the DG field at reference point (xi,eta) of every cell."""
    ncomp = phi.shape[0]
    psi = basis(np.asarray(xi, float), np.asarray(eta, float))
    return sum(phi[c] * psi[c] for c in range(ncomp))


def l2_error(phi, fn, hx, hy, nq=4):
    ncomp, ny, nx = phi.shape
    gx, gw = np.polynomial.legendre.leggauss(nq)
    gx, gw = 0.5 * (gx + 1), 0.5 * gw
    i = np.arange(nx)[None, :]
    j = np.arange(ny)[:, None]
    e = 0.0
    for a in range(nq):
        for b in range(nq):
            x = (i + gx[a]) * hx + 0 * j
            y = (j + gx[b]) * hy + 0 * i
            d = evaluate(phi, gx[a], gx[b]) - fn(x, y)
            e += gw[a] * gw[b] * np.sum(d * d) * hx * hy
    return np.sqrt(e)


def cone_case(nx, ny, ncomp):
    """Config 2: unit square, smooth bump at (0.5,0.75) carried by a solid-body rotation about
    the centre whose angular velocity is tapered to zero before the (periodic) boundary, so that
    periodicity and rotation are consistent.  Returns phi0, u, v, hx, hy, f0."""
    hx, hy = 1.0 / nx, 1.0 / ny
    xn = (np.arange(nx) * hx)[None, :] + 0 * np.arange(ny)[:, None]
    yn = (np.arange(ny) * hy)[:, None] + 0 * np.arange(nx)[None, :]
    r = np.sqrt((xn - 0.5) ** 2 + (yn - 0.5) ** 2)
    taper = np.where(r < 0.4, 1.0, np.where(r < 0.48, 0.5 * (1 + np.cos(np.pi * (r - 0.4) / 0.08)),
                                            0.0))
    omega = 2 * np.pi
    u = -omega * (yn - 0.5) * taper
    v = omega * (xn - 0.5) * taper

    def f0(x, y):
        d = np.sqrt((x - 0.5) ** 2 + (y - 0.75) ** 2) / 0.15
        return np.where(d < 1.0, 0.25 * (1 + np.cos(np.pi * np.minimum(d, 1.0))) ** 2, 0.0)

    return project(f0, nx, ny, hx, hy, ncomp), u, v, hx, hy, f0


def box_case(nx, ny, ncomp, L=512e3, t=0.0, rows=None, Ly=None, separable=False, state=True):
    """Configs 3/5: Mehlmann et al. (2021)-style box test scaled to the grid: 512 km square,
    H = 0.3 + 0.005 (sin(6e-5 x) + sin(3e-5 y)), A = 1, anticyclonic ocean gyre, moving
    cyclone wind; closed box (NSDG_BC_BOX).  `rows=(j0, nl)` returns only that strip of the
    global nx x ny grid.  Returns dict of fields + hx, hy.
    separable=True projects the x- and the y-part of H in 1-D and adds them (the projection is
    linear; equal to the 2-D projection to rounding, ~30x faster: the 16384^2 bench grid);
    state=False leaves out the all-zero U, V, S arrays (device buffers start zeroed)."""
    Ly = L if Ly is None else Ly
    hx, hy = L / nx, Ly / ny
    j0, nl = (0, ny) if rows is None else rows
    if separable:
        hx_part = project(lambda x, y: 0.3 + 0.005 * np.sin(6e-5 * x), nx, 1, hx, hy, ncomp, nq=3)
        hy_part = project(lambda x, y: 0.005 * np.sin(3e-5 * y), 1, nl, hx, hy, ncomp, nq=3, j0=j0)
        H = hx_part + hy_part
    else:
        H = project(lambda x, y: 0.3 + 0.005 * (np.sin(6e-5 * x) + np.sin(3e-5 * y)),
                    nx, nl, hx, hy, ncomp, nq=3, j0=j0)
    A = np.zeros_like(H)
    A[0] = 1.0
    xn = (np.arange(nx) * hx)[None, :] + 0.0 * np.arange(nl)[:, None]
    yn = ((j0 + np.arange(nl)) * hy)[:, None] + 0.0 * np.arange(nx)[None, :]
    ny = nl
    uocn = 0.1 * (2 * yn - Ly) / Ly
    vocn = -0.1 * (2 * xn - L) / L
    # cyclone centred at (256 km + 51.2 km/day * t) moving along the diagonal
    tday = t / 86400.0
    cx = cy = 256e3 + 51.2e3 * tday
    alpha = np.deg2rad(72.0)
    rr = np.sqrt((xn - cx) ** 2 + (yn - cy) ** 2)
    # speed(r) = 15 m/s * e * (r/100km) * exp(-r/100km): maximum 15 m/s at r = 100 km
    s = 15.0 * np.e / 100.0 * np.exp(-rr / 100e3) / 1e3  # [1/s] per metre of distance
    uatm = -s * (np.cos(alpha) * (xn - cx) + np.sin(alpha) * (yn - cy))
    vatm = -s * (-np.sin(alpha) * (xn - cx) + np.cos(alpha) * (yn - cy))
    if not state:
        return dict(H=H, A=A, UATM=uatm, VATM=vatm, UOCN=uocn, VOCN=vocn), hx, hy
    z = np.zeros((ny, nx))
    return dict(H=H, A=A, U=z.copy(), V=z.copy(), S11=np.zeros((3, ny, nx)),
                S12=np.zeros((3, ny, nx)), S22=np.zeros((3, ny, nx)), UATM=uatm, VATM=vatm,
                UOCN=uocn, VOCN=vocn), hx, hy
