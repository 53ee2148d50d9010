"""GPU parity tests of DG transport and the mEVP sub-cycle, through the C ABI, against the CPU
restatement (oracle/dg_transport.c, oracle/mevp.c).  Tolerance (BASELINE.json north_star):
1e-11 relative (to the field's max-norm) on the DG coefficients after N steps.
PARITY UNPINNED against draenog/nextsimdg: its snapshot has no dynamics code."""
import numpy as np
import pytest

from nextsimdg_b200 import synthetic as S
from nextsimdg_b200._abi import (BC_BOX, BC_PERIODIC, DYN_SINGLE_MEVP, DYN_SPLIT_MEVP, MevpParams,
                                   NsdgError)
from nextsimdg_b200.gpu import GpuDynamics
from oracle import cpu_oracle as O

pytestmark = pytest.mark.gpu
RTOL = 1e-11


def maxrel(got, ref):
    return np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-300)


def random_velocity(nx, ny, seed, scale=1.0):
    r = np.random.default_rng(seed)
    return scale * r.standard_normal((ny, nx)), scale * r.standard_normal((ny, nx))


@pytest.mark.parametrize("ncomp", [1, 3, 6])
@pytest.mark.parametrize("shape", [(64, 48), (37, 29), (130, 5)])
def test_transport_matches_oracle(ctx, ncomp, shape):
    """dG0/dG1/dG2, ragged grid sizes (not multiples of the CTA tile), random velocity with
    sign changes along every edge (pointwise upwinding), N = 5 steps."""
    nx, ny = shape
    hx, hy = 1.0 / nx, 1.0 / ny
    rng = np.random.default_rng(7)
    f = lambda x, y: np.sin(2 * np.pi * x) * np.cos(2 * np.pi * y) + 2
    phi = S.project(f, nx, ny, hx, hy, ncomp) + 0.05 * rng.standard_normal((ncomp, ny, nx))
    u, v = random_velocity(nx, ny, 9)
    dt = 0.02 * min(hx, hy)
    dyn = GpuDynamics(ctx, nx, ny, hx, hy, ncomp)
    dyn.upload("H", phi)
    dyn.upload("U", u)
    dyn.upload("V", v)
    ref = phi
    m0 = dyn.mass("H")
    for _ in range(5):
        dyn.transport_step(dt, "H")
        ref = O.transport_step(ref, u, v, hx, hy, dt)
    got = dyn.download("H")
    assert maxrel(got, ref) < RTOL
    assert abs(dyn.mass("H") - m0) / abs(m0) < 1e-13  # conservation on the GPU itself
    assert dyn.mass("H") == pytest.approx(ref[0].sum() * hx * hy, rel=1e-12)


@pytest.mark.parametrize("rk", [1, 2, 3])
def test_rk_orders_and_two_field_kernel(ctx, rk):
    nx, ny, nc = 48, 40, 6
    hx, hy = 1.0 / nx, 1.0 / ny
    rng = np.random.default_rng(rk)
    H = rng.standard_normal((nc, ny, nx))
    A = rng.standard_normal((nc, ny, nx))
    u, v = random_velocity(nx, ny, 4)
    dyn = GpuDynamics(ctx, nx, ny, hx, hy, nc, rk_order=rk)
    dyn.upload_all(dict(H=H, A=A, U=u, V=v))
    dt = 0.01 * hx
    for _ in range(3):
        dyn.transport_step(dt)  # H and A in one launch per stage
        H = O.transport_step(H, u, v, hx, hy, dt, rk)
        A = O.transport_step(A, u, v, hx, hy, dt, rk)
    assert maxrel(dyn.download("H"), H) < RTOL
    assert maxrel(dyn.download("A"), A) < RTOL


def test_limiter_matches_oracle(ctx):
    nx, ny, nc = 40, 32, 6
    hx, hy = 1.0 / nx, 1.0 / ny
    f = lambda x, y: ((np.abs(x - 0.5) < 0.2) & (np.abs(y - 0.5) < 0.2)).astype(float)
    A = S.project(f, nx, ny, hx, hy, nc)
    u, v = np.full((ny, nx), 1.0), np.full((ny, nx), 0.5)
    dyn = GpuDynamics(ctx, nx, ny, hx, hy, nc, limiter=1)
    dyn.upload_all(dict(A=A, U=u, V=v))
    dt = 0.05 * hx
    ref = A
    for _ in range(10):
        dyn.transport_step(dt, "A")
        ref = O.transport_step(ref, u, v, hx, hy, dt, 0, 1, 0.0, 1.0)
    got = dyn.download("A")
    assert maxrel(got, ref) < 1e-10
    for a in (0, 0.5, 1):
        for b in (0, 0.5, 1):
            val = S.evaluate(got, a, b)
            assert val.min() > -1e-12 and val.max() < 1 + 1e-12


def test_cone_rotation_dg1_quarter_turn(ctx):
    """BASELINE config 2 at reduced size (128^2 dG1): compact bump in a tapered rotation;
    mass conserved and GPU == oracle after the run."""
    n, nc = 128, 3
    phi, u, v, hx, hy, f0 = S.cone_case(n, n, nc)
    dyn = GpuDynamics(ctx, n, n, hx, hy, nc)
    dyn.upload_all(dict(H=phi, U=u, V=v))
    nsteps = 60
    dt = 0.25 / nsteps / 4
    ref = phi
    m0 = dyn.mass("H")
    for _ in range(nsteps):
        dyn.transport_step(dt, "H")
        ref = O.transport_step(ref, u, v, hx, hy, dt)
    got = dyn.download("H")
    assert maxrel(got, ref) < RTOL
    assert abs(dyn.mass("H") - m0) / m0 < 1e-13


@pytest.mark.parametrize("flags", [0, DYN_SINGLE_MEVP, DYN_SPLIT_MEVP], ids=["pair", "fused", "split"])
@pytest.mark.parametrize("bc", [BC_PERIODIC, BC_BOX])
@pytest.mark.parametrize("shape", [(64, 48), (33, 21), (130, 19), (64, 8), (65, 9), (128, 21), (96, 16), (2400 // 25 * 3 + 32, 11)])
def test_mevp_matches_oracle(ctx, bc, shape, flags):
    nx, ny = shape
    F, hx, hy = S.box_case(nx, ny, 6)
    rng = np.random.default_rng(5)
    F["U"] = 0.05 * rng.standard_normal((ny, nx))
    F["V"] = 0.05 * rng.standard_normal((ny, nx))
    F["S11"] = 100 * rng.standard_normal((3, ny, nx))
    F["S12"] = 100 * rng.standard_normal((3, ny, nx))
    F["S22"] = 100 * rng.standard_normal((3, ny, nx))
    F["A"][0] = rng.uniform(0.7, 1.0, (ny, nx))
    mp = MevpParams.defaults()
    mp.bc = bc
    mp.alpha = mp.beta = 300.0
    dyn = GpuDynamics(ctx, nx, ny, hx, hy, 6, mevp=mp, flags=flags)
    dyn.upload_all(F)
    for nsub in (1, 7):  # odd counts exercise the ping-pong pointer swap
        dyn.mevp_step(120.0, nsub)
        ref = O.mevp_step(mp, F["H"], F["A"], F["U"], F["V"], F["S11"], F["S12"], F["S22"],
                          F["UATM"], F["VATM"], F["UOCN"], F["VOCN"], hx, hy, 120.0, nsub)
        for name, r in zip(("U", "V", "S11", "S12", "S22"), ref):
            assert maxrel(dyn.download(name), r) < RTOL, (name, nsub)
            F[name] = r
    if bc == BC_BOX:
        u = dyn.download("U")
        assert np.all(u[0] == 0) and np.all(u[:, 0] == 0)


def test_north_star_step_box_100_substeps(ctx):
    """BASELINE config 3 at reduced size (96^2): 2 x (dG2 transport of H,A + 100 mEVP
    sub-iterations), every state field against the oracle."""
    n = 96
    F, hx, hy = S.box_case(n, n, 6)
    mp = MevpParams.defaults()
    mp.bc = BC_BOX
    dyn = GpuDynamics(ctx, n, n, hx, hy, 6, mevp=mp)
    dyn.upload_all(F)
    st = [F[k] for k in ("H", "A", "U", "V", "S11", "S12", "S22")]
    for _ in range(2):
        dyn.step(120.0, 100)
        st = list(O.dyn_step(mp, *st, F["UATM"], F["VATM"], F["UOCN"], F["VOCN"], hx, hy, 120.0,
                             100))
    for name, r in zip(("H", "A", "U", "V", "S11", "S12", "S22"), st):
        assert maxrel(dyn.download(name), r) < RTOL, name
    assert 1e-4 < np.abs(st[2]).max() < 0.1  # the ice actually moves


def test_step_host_equals_resident_path(ctx):
    """nsdg_dyn_step_host (host buffers, copies overlapped with compute) == upload, step, download."""
    nx, ny = 96, 40
    F, hx, hy = S.box_case(nx, ny, 6)
    F["U"], F["V"] = random_velocity(nx, ny, 3, 0.05)
    mp = MevpParams.defaults()
    mp.bc = BC_PERIODIC
    a = GpuDynamics(ctx, nx, ny, hx, hy, 6, mevp=mp)
    a.upload_all(F)
    a.step(120.0, 9)
    a.step(120.0, 9)
    names = ("H", "A", "U", "V", "S11", "S12", "S22")
    ref = {k: a.download(k) for k in names}
    b = GpuDynamics(ctx, nx, ny, hx, hy, 6, mevp=mp)
    out = {k: np.empty(b.field_shape(k)) for k in names}
    b.step_host(120.0, 9, {k: np.ascontiguousarray(v, dtype=np.float64) for k, v in F.items()}, out)
    # second step: state stays resident, only the forcing goes up again
    b.step_host(120.0, 9, {k: np.ascontiguousarray(F[k]) for k in ("UATM", "VATM", "UOCN", "VOCN")}, out)
    for k in names:
        assert np.array_equal(out[k], ref[k]), k
    with pytest.raises(ValueError):
        b.step_host(120.0, 1, {"H": np.zeros((3, 3))}, None)


def test_deferred_host_steps_equal_resident_path(ctx):
    """nsdg_dyn_step_host_async: downloads of step k overlap step k+1, forcing of step k+1 goes
    up into alternate buffers while step k computes; every step's output must equal the resident
    path's, with a forcing that changes from step to step.  Resident steps are enqueued right
    before the first deferred call, so the compute stream has a backlog when the alternate
    forcing buffers are created: their clearing must be ordered before the forcing upload on the
    copy stream (it would zero the wind otherwise)."""
    nx, ny, nsteps = 320, 256, 5
    F, hx, hy = S.box_case(nx, ny, 6)
    F["U"], F["V"] = random_velocity(nx, ny, 5, 0.05)
    mp = MevpParams.defaults()
    mp.bc = BC_PERIODIC
    names = ("H", "A", "U", "V", "S11", "S12", "S22")
    forcing = ("UATM", "VATM", "UOCN", "VOCN")
    fk = [{k: np.ascontiguousarray(F[k] * (1.0 + 0.1 * s)) for k in forcing} for s in range(nsteps)]

    def preamble(d):
        d.upload_all(F)
        for _ in range(2):
            d.step(120.0, 100)

    a = GpuDynamics(ctx, nx, ny, hx, hy, 6, mevp=mp)
    preamble(a)
    ref = []
    for s in range(nsteps):
        for k in forcing:
            a.upload(k, fk[s][k])
        a.step(120.0, 7)
        ref.append({k: a.download(k) for k in names})
    a.step(120.0, 7)  # one more with the last forcing
    ref_last = {k: a.download(k) for k in names}
    b = GpuDynamics(ctx, nx, ny, hx, hy, 6, mevp=mp)  # takes the context's dynamics slot over from a
    preamble(b)  # still running when the first deferred step is enqueued
    outs = [{k: np.full(b.field_shape(k), np.nan) for k in names} for _ in range(nsteps)]
    for s in range(nsteps):
        b.step_host(120.0, 7, fk[s], outs[s], deferred=True)
    b.host_wait()
    for s in range(nsteps):
        for k in names:
            assert np.array_equal(outs[s][k], ref[s][k]), (s, k)
    # a blocking step after deferred ones continues from the same state
    last = {k: np.empty(b.field_shape(k)) for k in names}
    b.step_host(120.0, 7, None, last)
    for k in names:
        assert np.array_equal(last[k], ref_last[k]), k
        assert np.array_equal(last[k], b.download(k)), k


def test_dynamics_forcing_levels_blend(ctx):
    nx, ny = 48, 20
    F, hx, hy = S.box_case(nx, ny, 3)
    dyn = GpuDynamics(ctx, nx, ny, hx, hy, 3)
    dyn.upload_all(F)
    r = np.random.default_rng(4)
    f0 = {k: r.standard_normal((ny, nx)) for k in ("UATM", "VATM", "UOCN", "VOCN")}
    f1 = {k: r.standard_normal((ny, nx)) for k in f0}
    with pytest.raises(NsdgError):
        dyn.blend_forcing(0.5)
    dyn.set_forcing_levels(f0, f1)
    dyn.blend_forcing(0.3)
    for k in f0:
        want = f0[k] + 0.3 * (f1[k] - f0[k])
        assert np.abs(dyn.download(k) - want).max() < 1e-15


def test_size_independent_properties_large_grid(ctx):
    """2048^2 dG2 (too large for the oracle in seconds): constants are preserved, mass is
    conserved to round-off, zero forcing keeps a state of rest."""
    n = 2048
    F, hx, hy = S.box_case(n, n, 6)
    mp = MevpParams.defaults()
    mp.bc = BC_BOX
    dyn = GpuDynamics(ctx, n, n, hx, hy, 6, mevp=mp)
    dyn.upload_all(F)
    dyn.step(120.0, 10)
    mH, mA = dyn.mass("H"), dyn.mass("A")
    dyn.step(120.0, 10)
    assert abs(dyn.mass("H") - mH) / mH < 1e-13
    assert abs(dyn.mass("A") - mA) / mA < 1e-13
    A = dyn.download("A")
    assert np.abs(A[0] - 1).max() < 1e-3
    # state of rest
    z = np.zeros((n, n))
    dyn2 = GpuDynamics(ctx, n, n, hx, hy, 6, mevp=mp)
    H = np.zeros((6, n, n))
    H[0] = 1.0
    dyn2.upload_all(dict(H=H, A=F["A"], U=z, V=z, UATM=z, VATM=z, UOCN=z, VOCN=z))
    dyn2.step(120.0, 5)
    assert np.abs(dyn2.download("U")).max() == 0.0
    assert np.array_equal(dyn2.download("H")[0], H[0])


@pytest.mark.parametrize("flags", [0, DYN_SINGLE_MEVP, DYN_SPLIT_MEVP], ids=["pair", "fused", "split"])
def test_gpu_mevp_analytic_uniform_flow(ctx, flags):
    """Independent of the oracle: ice drifting with a uniform ocean current, no wind, no Coriolis,
    uniform H = 2, A = 1 on a periodic grid.  Zero strain => sigma_VP = -P/2 I, so after n
    sub-iterations S11_0 = S22_0 = -(P/2)(1 - (1 - 1/alpha)^n), every other coefficient 0, and
    the velocity does not change (uniform stress has no divergence, no relative drag)."""
    nx, ny, n = 80, 48, 10
    mp = MevpParams.defaults()
    mp.bc = BC_PERIODIC
    mp.alpha = mp.beta = 300.0
    mp.f_coriolis = 0.0
    H = np.zeros((6, ny, nx))
    H[0] = 2.0
    A = np.zeros((6, ny, nx))
    A[0] = 1.0
    one = np.ones((ny, nx))
    dyn = GpuDynamics(ctx, nx, ny, 1000.0, 1000.0, 6, mevp=mp, flags=flags)
    dyn.upload_all(dict(H=H, A=A, U=0.1 * one, V=-0.2 * one, UOCN=0.1 * one, VOCN=-0.2 * one,
                        UATM=0 * one, VATM=0 * one))
    dyn.mevp_step(120.0, n)
    P = mp.p_star * 2.0
    want = -0.5 * P * (1.0 - (1.0 - 1.0 / mp.alpha) ** n)
    for k in ("S11", "S22"):
        s = dyn.download(k)
        assert np.abs(s[0] / want - 1).max() < 1e-12, k
        assert np.abs(s[1:]).max() < 1e-9 * P, k
    assert np.abs(dyn.download("S12")).max() < 1e-9 * P
    assert np.abs(dyn.download("U") - 0.1).max() < 1e-13
    assert np.abs(dyn.download("V") + 0.2).max() < 1e-13


@pytest.mark.parametrize("flags", [DYN_SINGLE_MEVP, DYN_SPLIT_MEVP], ids=["fused", "split"])
def test_gpu_mevp_analytic_linear_shear(ctx, flags):
    """Independent of the oracle: u = g y, v = 0 (away from the periodic seam) has e12 = g/2,
    e11 = e22 = 0, hence Delta = sqrt(dmin^2 + g^2/e^2), s12 = P g/(2 e^2 Delta), s11 = -P/2;
    one sub-iteration from S = 0 leaves exactly 1/alpha of it in the dG1 means."""
    nx, ny, g = 64, 40, 1e-6
    hx = hy = 1000.0
    mp = MevpParams.defaults()
    mp.bc = BC_PERIODIC
    mp.alpha = mp.beta = 300.0
    H = np.zeros((6, ny, nx))
    H[0] = 2.0
    A = np.zeros((6, ny, nx))
    A[0] = 1.0
    yn = (np.arange(ny) * hy)[:, None] + 0.0 * np.arange(nx)[None, :]
    z = np.zeros((ny, nx))
    dyn = GpuDynamics(ctx, nx, ny, hx, hy, 6, mevp=mp, flags=flags)
    dyn.upload_all(dict(H=H, A=A, U=g * yn, V=z, UOCN=z, VOCN=z, UATM=z, VATM=z))
    dyn.mevp_step(120.0, 1)
    P = mp.p_star * 2.0
    delta = np.sqrt(mp.delta_min ** 2 + g * g / mp.ellipse_e ** 2)
    inner = (slice(1, ny - 1), slice(None))
    s12, s11, s22 = dyn.download("S12"), dyn.download("S11"), dyn.download("S22")
    assert np.allclose(s12[0][inner], P * g / (2 * mp.ellipse_e ** 2 * delta) / mp.alpha, rtol=1e-12)
    assert np.allclose(s11[0][inner], -0.5 * P / mp.alpha, rtol=1e-12)
    assert np.allclose(s22[0][inner], -0.5 * P / mp.alpha, rtol=1e-12)
    assert np.abs(s12[1:][:, 1:ny - 1]).max() < 1e-9 * P


@pytest.mark.parametrize("ncomp,order", [(1, 1), (3, 2), (6, 3)])
def test_gpu_transport_convergence_against_the_analytic_solution(ctx, ncomp, order):
    """Independent of the oracle: uniform flow on the periodic unit square translates a smooth
    field exactly; the CUDA transport must converge to the shifted field at order k + 1, and
    reach a small absolute error at 256^2 (dG2: < 1e-7)."""
    def smooth(x, y):
        return 1.0 + 0.5 * np.sin(2 * np.pi * x) * np.cos(2 * np.pi * y)

    errs, T = [], 0.1
    for n in (64, 128, 256):
        hx = hy = 1.0 / n
        dyn = GpuDynamics(ctx, n, n, hx, hy, ncomp)
        u, v = np.full((n, n), 0.7), np.full((n, n), -0.3)
        dyn.upload_all(dict(H=S.project(smooth, n, n, hx, hy, ncomp), U=u, V=v))
        nst = int(round(T / (0.05 * hx)))
        for _ in range(nst):
            dyn.transport_step(T / nst, "H")
        errs.append(S.l2_error(dyn.download("H"), lambda x, y: smooth(x - 0.7 * T, y + 0.3 * T), hx, hy))
    assert np.log2(errs[0] / errs[1]) > order - 0.15, errs
    assert np.log2(errs[1] / errs[2]) > order - 0.15, errs
    if ncomp == 6:
        assert errs[2] < 1e-7, errs


def test_full_size_4096sq_is_the_tiled_small_problem(ctx):
    """BASELINE configs[2] size.  A periodic 4096^2 problem whose inputs are a 256^2 periodic
    problem tiled 16 x 16 must reproduce the tiled 256^2 solution BIT FOR BIT (every cell sees the
    same neighbourhood; tile, wave and wrap position must not matter), and the 256^2 solution is
    within tolerance of the oracle: the full-size run is pinned to the oracle through both."""
    m, reps, nsub = 256, 16, 8
    n = m * reps
    F, hx, hy = S.box_case(m, m, 6)
    F["U"], F["V"] = random_velocity(m, m, 17, 0.05)
    r = np.random.default_rng(23)
    for k in ("S11", "S12", "S22"):
        F[k] = 30.0 * r.standard_normal((3, m, m))
    mp = MevpParams.defaults()
    mp.bc = BC_PERIODIC
    names = ("H", "A", "U", "V", "S11", "S12", "S22")
    # (the default transport stage kernel differs between the two sizes: one thread per cell below
    # 2^20 cells, marching above; they are bit-identical as well)
    small = GpuDynamics(ctx, m, m, hx, hy, 6, mevp=mp)
    small.upload_all(F)
    small.step(120.0, nsub)
    got_small = {k: small.download(k) for k in names}
    ref = O.dyn_step(mp, *[F[k] for k in names], F["UATM"], F["VATM"], F["UOCN"], F["VOCN"], hx, hy,
                     120.0, nsub)
    for k, a in zip(names, ref):
        assert maxrel(got_small[k], a) < RTOL, k
    big = GpuDynamics(ctx, n, n, hx, hy, 6, mevp=mp)
    big.upload_all({k: np.tile(v, (1, reps, reps) if v.ndim == 3 else (reps, reps)) for k, v in F.items()})
    big.step(120.0, nsub)
    for k in names:
        a = big.download(k)
        t = np.tile(got_small[k], (1, reps, reps) if a.ndim == 3 else (reps, reps))
        assert np.array_equal(a, t), k
    mH = big.mass("H")
    big.step(120.0, nsub)
    assert abs(big.mass("H") - mH) / mH < 1e-13


def test_error_behaviour(ctx):
    with pytest.raises(NsdgError):
        GpuDynamics(ctx, 16, 16, 1.0, 1.0, ncomp=4)
    with pytest.raises(NsdgError):  # a strip needs a communicator first
        GpuDynamics(ctx, 16, 32, 1.0, 1.0, 6, ny_local=16)
    dyn = GpuDynamics(ctx, 16, 16, 1.0, 1.0, 3)
    with pytest.raises(NsdgError):
        dyn.transport_step(0.1, "U")
    with pytest.raises(ValueError):
        dyn.upload("H", np.zeros((3, 16, 15)))


def test_pair_kernel_variants_bit_identical(ctx):
    """The sub-iteration kernels share their arithmetic: the marching pair kernel (default, band
    heights chosen per grid / 64 / 16 / 5 rows), the tiled pair kernel in two tile shapes,
    single fused sub-iterations, CUDA-graph replay, the shared-memory-ring variant of the marching
    kernel (rows ahead brought in by cp.async.bulk), and the marching / one-thread-per-cell dG2
    transport stage kernels give the same bits on periodic and box grids, ragged sizes and a
    grid narrower than the halo.  The variant is a per-context knob of the C ABI
    (nsdg_dyn_set_tuning), so all of them run in this one process."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import dyn_digest
    variants = {"auto": {}, "march64": {"march_rows": 64}, "march16": {"march_rows": 16},
                "march5": {"march_rows": 5}, "tiled": {"march_rows": -1},
                "tiled32x8": {"march_rows": -1, "pair_tile": "32x8"}, "single": {"pair": -1},
                "graph": {"graph": 1}, "nograph": {"graph": -1},
                "transport-march": {"transport_march": 1}, "transport-tiled": {"transport_march": -1},
                "transport-march-rows16": {"transport_march": 1, "march_rows": 16},
                "transport-fused": {"transport_fused": 1}, "transport-unfused": {"transport_fused": -1},
                "ring": {"march_ring": 1}, "ring-rows5": {"march_ring": 1, "march_rows": 5},
                "lane-ring": {"march_ring": 2}, "lane-ring-rows5": {"march_ring": 2, "march_rows": 5},
                "l2-prefetch": {"march_ring": 3}, "l2-prefetch-rows7": {"march_ring": 3, "march_rows": 7}}
    digests = {k: dyn_digest.digest(ctx, v) for k, v in variants.items()}
    assert len(set(digests.values())) == 1, digests


@pytest.mark.parametrize("nc", [6, 3])
@pytest.mark.parametrize("shape,bc", [((300, 211), BC_PERIODIC), ((64, 48), BC_BOX), ((17, 5), BC_PERIODIC),
                                      ((1024, 1024), BC_BOX), ((29, 70), BC_BOX)])
def test_transport_marching_kernel_against_tiled_kernel_and_oracle(ctx, shape, bc, nc):
    """The marching dG2 / dG1 stage kernel (every edge flux computed once, neighbours by shuffle and
    from registers) against the one-thread-per-cell kernel: same integrals with every fused
    multiply-add written out (transport.cu is compiled without automatic contraction), so the
    two agree BIT FOR BIT after three SSP-RK3 steps, and both are within 1e-11 of the oracle.
    Third arm (the default from 2^20 cells): TWO stages in one launch (transport_pair_march_kernel,
    stage 1 kept in registers on a one-cell halo): the whole Heun step of dG1 (tracer and stage
    buffer trade places) / stages 1 + 2 of the dG2 step, also bit for bit, also through graph
    replay and as single-tracer steps (nsdg_transport_step).
    Ragged sizes, a grid narrower than a warp's 28 columns, fewer rows than the pipeline depth."""
    nx, ny = shape
    F, hx, hy = S.box_case(nx, ny, nc)
    r = np.random.default_rng(9)
    for k in range(1, nc):
        F["H"][k] = 0.02 * r.standard_normal((ny, nx))
        F["A"][k] = 0.01 * r.standard_normal((ny, nx))
    F["U"], F["V"] = random_velocity(nx, ny, 3, 0.3)
    if bc == BC_BOX:
        F["U"][0, :] = F["U"][:, 0] = F["V"][0, :] = F["V"][:, 0] = 0.0
    mp = MevpParams.defaults()
    mp.bc = bc
    dt = 0.1 * min(hx, hy) / 1.5
    out = {}
    arms = [("march", {"transport_march": 1, "transport_fused": -1}),
            ("tiled", {"transport_march": -1, "transport_fused": -1})]
    # two stages per launch: the whole Heun step (dG1) / stages 1 + 2 of SSP-RK3 (dG2)
    arms += [("fused", {"transport_fused": 1}), ("fused-graph", {"transport_fused": 1, "graph": 1}),
             ("fused-single", {"transport_fused": 1}), ("fused-march3", {"transport_fused": 1, "transport_march": 1})]
    for name, tune in arms:
        dyn = GpuDynamics(ctx, nx, ny, hx, hy, nc, mevp=mp, tuning=tune)
        dyn.upload_all({k: v for k, v in F.items() if k in ("H", "A", "U", "V")})
        for _ in range(3):
            if name == "fused-single":  # one tracer at a time: nsdg_transport_step
                dyn.transport_step(dt, "A")
                dyn.transport_step(dt, "H")
            else:
                dyn.transport_step(dt)
        out[name] = {k: dyn.download(k) for k in ("H", "A")}
        dyn.close()
    for k in ("H", "A"):
        for name, _ in arms[1:]:
            assert np.array_equal(out["march"][k], out[name][k]), (k, name)
        if nx * ny <= 300 * 211:
            ref = F[k]
            for _ in range(3):
                ref = O.transport_step(ref, F["U"], F["V"], hx, hy, dt)
            assert maxrel(out["march"][k], ref) < RTOL, k


def test_tuning_roundtrip_and_state_ownership(ctx):
    """nsdg_dyn_set_tuning / get_tuning; a second nsdg_dyn_init on a live context is NSDG_ESTATE
    until nsdg_dyn_release; the Python classes hand the context's slot over explicitly and a
    released object raises instead of touching its successor's state."""
    import ctypes
    from nextsimdg_b200._abi import NSDG_ESTATE, DynConfig
    a = GpuDynamics(ctx, 32, 16, 1.0, 1.0, 3, tuning={"march_rows": 16, "halo_overlap": -1})
    t = a.get_tuning()
    assert (t.march_rows, t.halo_overlap, t.pair) == (16, -1, 0)
    with pytest.raises(NsdgError):
        a.set_tuning(march_rows=5000)
    with pytest.raises(NsdgError):
        a.set_tuning(pair=2)
    cfg = DynConfig(32, 16, 16, 0, 1.0, 1.0, 3, 0, 0, 0)
    assert ctx.lib.nsdg_dyn_init(ctx.h, ctypes.byref(cfg)) == NSDG_ESTATE
    assert ctx.lib.nsdg_dyn_release(ctx.h) == 0
    assert ctx.lib.nsdg_dyn_init(ctx.h, ctypes.byref(cfg)) == 0
    assert ctx.lib.nsdg_dyn_release(ctx.h) == 0
    a._released = True  # its state is gone (released behind its back above)
    b = GpuDynamics(ctx, 32, 16, 1.0, 1.0, 3)
    c = GpuDynamics(ctx, 32, 16, 1.0, 1.0, 3)  # takes the slot over: b is released explicitly
    with pytest.raises(RuntimeError):
        b.download("H")
    assert c.download("H").shape == (3, 16, 32)
    assert ctx.lib.nsdg_thermo_release(ctx.h) == 0
    assert ctx.lib.nsdg_thermo_init(ctx.h, 10, 1) == 0
    assert ctx.lib.nsdg_thermo_init(ctx.h, 10, 1) == NSDG_ESTATE
    assert ctx.lib.nsdg_thermo_release(ctx.h) == 0


def test_row_range_upload_download(ctx):
    """nsdg_dyn_upload_rows / download_rows: chunked I/O of a strip equals the whole-field calls
    (DG and node fields), and out-of-range rows are rejected."""
    nx, ny = 40, 23
    r = np.random.default_rng(2)
    dyn = GpuDynamics(ctx, nx, ny, 1.0, 1.0, 6)
    H = r.standard_normal((6, ny, nx))
    U = r.standard_normal((ny, nx))
    for c0 in range(0, ny, 7):
        n = min(7, ny - c0)
        dyn.upload_rows("H", c0, np.ascontiguousarray(H[:, c0:c0 + n]))
        dyn.upload_rows("U", c0, np.ascontiguousarray(U[c0:c0 + n]))
    assert np.array_equal(dyn.download("H"), H) and np.array_equal(dyn.download("U"), U)
    assert np.array_equal(dyn.download_rows("H", 5, 9), H[:, 5:14])
    assert np.array_equal(dyn.download_rows("U", 20, 3), U[20:23])
    with pytest.raises(NsdgError):
        dyn.download_rows("U", 20, 4)
    with pytest.raises(ValueError):
        dyn.upload_rows("H", 0, np.zeros((6, 2, nx - 1)))


def test_graph_replay_follows_parameter_and_buffer_changes(ctx):
    """CUDA-graph replay (forced on): captured launches bake in kernel arguments and buffer roles.
    Changing the mEVP parameters, uploading new fields, an odd sub-iteration count (buffer roles
    end up swapped) and deferred host steps (forcing buffers swap) must all give what plain
    launches give."""
    nx, ny = 72, 40
    F, hx, hy = S.box_case(nx, ny, 6)
    F["U"], F["V"] = random_velocity(nx, ny, 8, 0.05)
    out = {}
    for name, g in (("graph", 1), ("plain", -1)):
        mp = MevpParams.defaults()
        mp.bc = BC_BOX
        dyn = GpuDynamics(ctx, nx, ny, hx, hy, 6, mevp=mp, tuning={"graph": g})
        dyn.upload_all(F)
        for nsub in (6, 7, 7, 6):  # odd counts leave the ping-pong buffers swapped
            dyn.step(120.0, nsub)
        mp.alpha = mp.beta = 200.0
        dyn.set_mevp_params(mp)  # cached graphs must not survive this
        dyn.step(120.0, 6)
        dyn.upload("UATM", 0.5 * F["UATM"])
        dyn.step(120.0, 6)
        outs = [{k: np.empty(dyn.field_shape(k)) for k in ("H", "U", "S11")} for _ in range(3)]
        for q in range(3):
            dyn.step_host(120.0, 6, {"UATM": np.ascontiguousarray((1 + 0.1 * q) * F["UATM"])}, outs[q], deferred=True)
        dyn.host_wait()
        out[name] = [dyn.download(k) for k in ("H", "A", "U", "V", "S11", "S12", "S22")] + [o[k] for o in outs for k in o]
        dyn.close()
    for a, b in zip(out["graph"], out["plain"]):
        assert np.array_equal(a, b)
