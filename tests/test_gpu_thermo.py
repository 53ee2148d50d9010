"""GPU parity tests of the column thermodynamics, through the C ABI, against the oracle and the
reference-generated golden fixtures.  Tolerance (BASELINE.json north_star): 1e-10 relative on
the thermodynamic state; cells whose data-dependent branch resolved differently from the CPU
(differences of O(1), not O(ulp)) are counted separately and must be (almost) absent."""
import os
import sys

import numpy as np
import pytest

from nextsimdg_b200 import synthetic as S
from nextsimdg_b200._abi import NsdgError, ThermoParams
from nextsimdg_b200.config import thermo_params_from_ini
from nextsimdg_b200.gpu import GpuStep
from oracle import cpu_oracle as O

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
RTOL = 1e-10
ATOL = 1e-14


def rel_err(got, ref):
    return np.abs(got - ref) / np.maximum(np.abs(ref), 1e-3)


def compare_state(got, ref, max_flips, newice=None):
    """returns number of branch-divergent cells; asserts all others within RTOL.  newice =
    (got, ref): the hidden new-ice state is checked on every cell that did not flip."""
    flips = np.zeros(ref["hice"].size, bool)
    for k, kk in (("hice", "hice"), ("cice", "cice"), ("hsnow", "hsnow"), ("tice", "tice0")):
        g = got[k] if k in got else got[kk]
        e = np.abs(g - ref[kk])
        bad = e > (ATOL + RTOL * np.maximum(np.abs(ref[kk]), 1e-3))
        flips |= bad
    n = int(flips.sum())
    assert n <= max_flips, f"{n} cells outside tolerance (allowed branch flips: {max_flips})"
    if newice is not None:
        g, r = newice
        assert rel_err(g[~flips], r[~flips]).max(initial=0.0) < RTOL
    return n


def params(name):
    return {"linear_smu": thermo_params_from_ini(""),
            "unesco_ccsm": thermo_params_from_ini(
                "[Modules]\nNextsim::IFreezingPoint = Nextsim::UnescoFreezing\n"
                "Nextsim::IIceAlbedo = Nextsim::CCSMIceAlbedo\n[CCSMIceAlbedo]\n"
                "iceAlbedo = 0.63\nsnowAlbedo = 0.88\n"),
            "linear_smu2": thermo_params_from_ini(
                "[Modules]\nNextsim::IIceAlbedo = Nextsim::SMU2IceAlbedo\n"),
            "custom_keys": custom_keys_params()}[name]


def custom_keys_params():
    sys.path.insert(0, GOLD)
    import make_thermo_golden as G
    return thermo_params_from_ini("[Modules]\nNextsim::IFreezingPoint = Nextsim::UnescoFreezing\n"
                                  "Nextsim::IIceAlbedo = Nextsim::SMUIceAlbedo\n" + G.CUSTOM_KEYS_INI)


def run_gpu(ctx, p, st, fo, dt, nsteps, nlayers=1, diag=False):
    n = st["hice"].size
    step = GpuStep(ctx, n, nlayers, p)
    prog = dict(st)
    prog["tice"] = np.repeat(st["tice0"], nlayers)
    step.setInitialData(prog, fo)
    step.iterate(dt, nsteps, diagnostics=diag)
    step.stop()
    return step


def test_reference_golden_cases(ctx):
    """The melting and freezing cases of physics/test/NextsimPhysics_test.cpp:176-312 and the
    dev1 config, on the GPU, against the reference's own test constants."""
    p = params("unesco_ccsm")
    a = lambda *v: np.array(v, float)
    st = dict(hice=a(0.1, 0.1), cice=a(0.5, 0.5), hsnow=a(0.01, 0.01), sst=a(-1, -1.75),
              sss=a(32, 32), tice0=a(-1, -9))
    fo = dict(tair=a(3, -12), tdew=a(2, -12), slp=a(1e5, 1e5), qsw=a(50, 0), qlw=a(330, 265),
              mld=a(10, 10), snowfall=a(0, 1e-3), wind=a(5, 5))
    step = run_gpu(ctx, p, st, fo, 600.0, 1, diag=True)
    g, d, ni = step.get_prognostic(), step.get_diagnostics(), step.get_newice()
    c = g["cice"]
    assert g["hice"][0] / c[0] == pytest.approx(0.12846, rel=1e-4)
    assert g["hsnow"][0] / c[0] == pytest.approx(0.01957732, rel=1e-4)
    assert c[0] == pytest.approx(0.368269, rel=1e-4) and g["tice"][0] == 0.0 and ni[0] == 0.0
    assert d["qia"][0] == pytest.approx(-84.6156, rel=1e-2)
    assert d["qio"][0] == pytest.approx(53717.8, rel=1e-2)
    assert d["subl"][0] == pytest.approx(-7.3858e-06, rel=1e-4)
    assert d["dqdt"][0] == pytest.approx(19.7013, rel=1e-2)
    assert g["hice"][1] / c[1] == pytest.approx(0.199998, rel=1e-4)
    assert g["hsnow"][1] / c[1] == pytest.approx(0.02179357, rel=1e-4)
    assert c[1] == pytest.approx(0.5002, rel=1e-4)
    assert g["tice"][1] == pytest.approx(-8.90443, rel=1e-4)
    assert ni[1] == pytest.approx(6.79707e-5, rel=1e-2)
    # full-precision pins of the reference build, at the stated tolerance
    assert c[0] == pytest.approx(0.36826938031129286, rel=RTOL)
    assert g["tice"][1] == pytest.approx(-8.9044278062473516, rel=RTOL)
    assert ni[1] == pytest.approx(6.799058277473154e-05, rel=RTOL)
    assert d["qio"][0] == pytest.approx(53717.834007845158, rel=RTOL)
    # dev1 (BASELINE config 1): 10x10 cells, dt = 1 s, one step
    st, fo = S.dev1_inputs(100)
    g = run_gpu(ctx, params("linear_smu"), st, fo, 1.0, 1).get_prognostic()
    assert np.all(g["hice"] == g["hice"][0])
    assert g["hice"][0] == pytest.approx(0.04668325240678619, rel=RTOL)
    assert g["cice"][0] == pytest.approx(0.36670813101696548, rel=RTOL)
    assert g["tice"][0] == pytest.approx(-1.444501803353837, rel=RTOL)


@pytest.mark.parametrize("name", ["linear_smu", "unesco_ccsm", "linear_smu2", "custom_keys"])
def test_against_reference_fixtures(ctx, name):
    """Multi-step runs against fixtures generated by the reference's own sources."""
    sys.path.insert(0, GOLD)
    import make_thermo_golden as G
    gold = np.load(os.path.join(GOLD, f"thermo_ref_{name}.npz"))
    p = params(name)
    for tag, (st, fo), dt in (("rand", S.thermo_inputs(G.N, G.SEED), G.DT),
                              ("cold", G.cold_inputs(G.N, G.SEED + 1), G.DT),
                              ("dev1", S.dev1_inputs(100), 1.0)):
        step = run_gpu(ctx, p, st, fo, dt, 1, diag=True)
        ref = {k: gold[f"{tag}_s0_{kk}"] for k, kk in (("hice", "hice"), ("cice", "cice"),
                                                       ("hsnow", "hsnow"), ("tice0", "tice"))}
        flips = compare_state(step.get_prognostic(), ref, max_flips=0)
        d = step.get_diagnostics()
        for k in ("qia", "qio", "subl", "dqdt", "hifroms"):
            assert rel_err(d[k], gold[f"{tag}_s0_{k}"]).max() < RTOL, (tag, k)
        assert rel_err(step.get_newice(), gold[f"{tag}_s0_newice"]).max() < RTOL
        step.iterate(dt, G.NSTEPS - 1)
        last = G.NSTEPS - 1
        ref = {k: gold[f"{tag}_s{last}_{kk}"] for k, kk in (("hice", "hice"), ("cice", "cice"),
                                                            ("hsnow", "hsnow"), ("tice0", "tice"))}
        compare_state(step.get_prognostic(), ref, max_flips=0)
        assert rel_err(step.get_newice(), gold[f"{tag}_s{last}_newice"]).max() < RTOL


@pytest.mark.parametrize("n", [0, 1, 2, 3, 255, 256, 257, 100003])
def test_ragged_sizes_against_oracle(ctx, n):
    """Empty, odd and non-multiple-of-block sizes (vector path + scalar tail)."""
    p = params("unesco_ccsm")
    st, fo = S.thermo_inputs(max(n, 1), seed=n + 11)
    st = {k: v[:n] for k, v in st.items()}
    fo = {k: v[:n] for k, v in fo.items()}
    step = run_gpu(ctx, p, st, fo, 600.0, 2)
    got = step.get_prognostic()
    if n == 0:
        assert got["hice"].size == 0
        return
    ref, ni, _ = O.thermo_step(p, 600.0, 2, st, fo)
    compare_state(got, ref, max_flips=0)
    assert rel_err(step.get_newice(), ni).max() < RTOL


def test_million_cells_three_configs(ctx):
    """1 M-cell subsample of config 4, 3 steps, oracle on the host cores."""
    n = 1 << 20
    st, fo = S.thermo_inputs(n)
    for name in ("linear_smu", "unesco_ccsm"):
        p = params(name)
        step = run_gpu(ctx, p, st, fo, 600.0, 3)
        ref, ni, _ = O.thermo_step(p, 600.0, 3, st, fo, nthreads=os.cpu_count() or 1)
        compare_state(step.get_prognostic(), ref, max_flips=2, newice=(step.get_newice(), ni))


def test_full_size_4096sq_against_oracle(ctx):
    """BASELINE configs[3] at its full size: 16.8 M cells of the synthetic distribution, 2 sweeps,
    against the C port of the reference's column physics on all host cores; the first 2^18 cells
    also against a run of just those cells (cell independence: position must not matter)."""
    n = 4096 * 4096
    st, fo = S.thermo_inputs(n)
    p = params("linear_smu")
    step = run_gpu(ctx, p, st, fo, 600.0, 2)
    got = step.get_prognostic()
    ref, ni, _ = O.thermo_step(p, 600.0, 2, st, fo, nthreads=os.cpu_count() or 1)
    compare_state(got, ref, max_flips=8, newice=(step.get_newice(), ni))
    m = 1 << 18
    sub = run_gpu(ctx, p, {k: v[:m] for k, v in st.items()}, {k: v[:m] for k, v in fo.items()}, 600.0, 2)
    g2 = sub.get_prognostic()
    for k in ("hice", "cice", "hsnow", "tice"):
        assert np.array_equal(g2[k], got[k][:m]), k


def test_layers_collapse_and_sst_untouched(ctx):
    """Q5/Q9 of SURVEY.md 8a: SST/SSS never evolve; every ice layer takes the surface value."""
    n, nl = 1000, 3
    st, fo = S.thermo_inputs(n, 3)
    step = run_gpu(ctx, params("linear_smu"), st, fo, 600.0, 1, nlayers=nl)
    g = step.get_prognostic()
    assert np.array_equal(g["sst"], st["sst"]) and np.array_equal(g["sss"], st["sss"])
    t = g["tice"].reshape(n, nl)
    assert np.array_equal(t[:, 0], t[:, 1]) and np.array_equal(t[:, 0], t[:, 2])
    ref, _, _ = O.thermo_step(params("linear_smu"), 600.0, 1, st, fo)
    assert rel_err(t[:, 0], ref["tice0"]).max() < RTOL


def test_hidden_newice_state_persists(ctx):
    """Q1: newice written in a freezing step still feeds the next, non-freezing step."""
    p = params("unesco_ccsm")
    a = lambda v: np.array([float(v)] * 4)
    st = dict(hice=a(0.1), cice=a(0.5), hsnow=a(0.01), sst=a(-1.75), sss=a(32), tice0=a(-9))
    cold = dict(tair=a(-12), tdew=a(-12), slp=a(1e5), qsw=a(0), qlw=a(265), mld=a(10),
                snowfall=a(1e-3), wind=a(5))
    warm = dict(cold, tair=a(5), tdew=a(4), qsw=a(300), qlw=a(400))
    step = run_gpu(ctx, p, st, cold, 600.0, 1)
    ni1 = step.get_newice()
    assert ni1[0] > 0
    step.setForcing(warm)
    step.iterate(600.0, 1)
    assert np.array_equal(step.get_newice(), ni1), "newice must be carried, not reset"
    s1, n1, _ = O.thermo_step(p, 600.0, 1, st, cold)
    s2, n2, _ = O.thermo_step(p, 600.0, 1, s1, warm, newice=n1)
    compare_state(step.get_prognostic(), s2, 0)
    # checkpoint / resume of the hidden state
    # (a second context: one thermodynamics state per context)
    from nextsimdg_b200.gpu import Context
    ctx2 = Context(0)
    step2 = GpuStep(ctx2, 4, 1, p)
    step2.setInitialData(step.get_prognostic(), warm)
    step2.set_newice(step.get_newice())
    step.iterate(600.0, 1)
    step2.iterate(600.0, 1)
    for k in ("hice", "cice", "hsnow", "tice"):
        assert np.array_equal(step.get_prognostic()[k], step2.get_prognostic()[k]), k
    assert np.array_equal(step.get_newice(), step2.get_newice())
    ctx2.close()


def test_forcing_time_interpolation_on_device(ctx):
    """two resident forcing levels blended on the GPU == host-side blend."""
    n = 4099
    st, f0 = S.thermo_inputs(n, 21)
    _, f1 = S.thermo_inputs(n, 22)
    p = params("unesco_ccsm")
    # one context holds one thermodynamics state: run the device-blended sequence first
    a = run_gpu(ctx, p, st, f0, 600.0, 0)
    a.setForcingLevels(f0, f1)
    seq = []
    for w in (0.0, 0.25, 1.0):
        a.blendForcing(w)
        a.iterate(600.0, 1)
        seq.append(a.get_prognostic())
    with pytest.raises(NsdgError):
        a.blendForcing(1.5)
    b = run_gpu(ctx, p, st, f0, 600.0, 0)
    for w, ga in zip((0.0, 0.25, 1.0), seq):
        b.setForcing({k: f0[k] + w * (f1[k] - f0[k]) for k in f0})
        b.iterate(600.0, 1)
        gb = b.get_prognostic()
        compare_state(ga, {"hice": gb["hice"], "cice": gb["cice"], "hsnow": gb["hsnow"],
                           "tice0": gb["tice"]}, max_flips=0)
    c = GpuStep(ctx, 8, 1, p)
    with pytest.raises(NsdgError):  # levels not uploaded
        c.blendForcing(0.5)


def test_slab_ocean_closure(ctx):
    """SURVEY.md 8f rank 3 (opt-in; the reference drops Qow/Qio, quirk Q5): with the closure on,
    SST moves by exactly the two fluxes of the diagnostics, the ice state is unchanged."""
    n = 20000
    st, fo = S.thermo_inputs(n, 31)
    p = params("linear_smu")
    off = run_gpu(ctx, p, st, fo, 600.0, 1, diag=True)
    d, g_off = off.get_diagnostics(), off.get_prognostic()
    on = GpuStep(ctx, n, 1, p)  # (replaces `off`: one thermodynamics state per context)
    prog = dict(st)
    prog["tice"] = st["tice0"]
    on.setInitialData(prog, fo)
    on.setSlabOcean(True)
    on.iterate(600.0, 1)
    g_on = on.get_prognostic()
    for k in ("hice", "cice", "hsnow", "tice"):  # (another template instance: FMA contraction
        assert rel_err(g_on[k], g_off[k]).max() < 1e-12, k  # may differ in the last bit)
    assert np.array_equal(g_off["sst"], st["sst"])
    bulk = fo["mld"] * 1025.0 * 4186.84
    expect = st["sst"] - 600.0 * ((1 - st["cice"]) * d["qow"] + st["cice"] * d["qio"]) / bulk
    assert np.abs(g_on["sst"] - expect).max() < 1e-10 * np.abs(expect).max()
    assert np.abs(g_on["sst"] - st["sst"]).max() > 1e-3
    on.setSlabOcean(False)
    before = on.get_prognostic()["sst"]
    on.iterate(600.0, 1)
    assert np.array_equal(on.get_prognostic()["sst"], before)


@pytest.mark.parametrize("name", ["mid.nc", "mid.nsdgr"])
def test_restart_write_resume_is_bit_identical(ctx, tmp_path, name):
    """Restart written from the SoA device mirror and read back (incl. hidden newice): the
    resumed run continues exactly like the uninterrupted one.  Both formats: the reference's
    netCDF-4 layout (hdf5write / hdf5lite) and the NSDGR1 container."""
    st, fo = S.thermo_inputs(64 * 64, 5)
    p = params("unesco_ccsm")
    step = run_gpu(ctx, p, st, fo, 600.0, 2)
    path = str(tmp_path / name)
    step.writeRestartFile(path)
    step.iterate(600.0, 2)
    want, want_ni = step.get_prognostic(), step.get_newice()
    resumed = GpuStep.from_restart(ctx, path, fo, p)
    assert resumed.n == 64 * 64
    resumed.iterate(600.0, 2)
    got = resumed.get_prognostic()
    for k in ("hice", "cice", "hsnow", "tice", "sst", "sss"):
        assert np.array_equal(got[k], want[k]), k
    assert np.array_equal(resumed.get_newice(), want_ni)


def test_dev1_config_end_to_end(ctx, tmp_path):
    """BASELINE configs[0]: run/dev1.cfg with the content of the reference's dev1.res.nc
    (tests/golden/dev1_restart.npz, extracted from the shipped file) through run.run_config:
    restart -> SoA -> GPU step -> restart.  Pins: SURVEY.md 8c (compiled reference)."""
    from nextsimdg_b200 import restart as R, run as RUN
    g = np.load(os.path.join(GOLD, "dev1_restart.npz"))
    nx, ny, nl = (int(x) for x in g["dims"])
    R.write_restart(str(tmp_path / "dev1.nsdgr"), {k: g[k] for k in R.PROGNOSTIC}, nx, ny, nl)
    cfg = tmp_path / "dev1.cfg"  # run/dev1.cfg with the container's file name
    cfg.write_text("[model]\ninit_file = dev1.nsdgr\nstart = 0\nstop = 1\ntime_step = 1\n")
    final, plan = RUN.run_config(str(cfg), ctx, base_dir=str(tmp_path))
    assert plan["nsteps"] == 1
    pins = {"hice": 0.04668325240678619, "cice": 0.36670813101696548, "hsnow": 0.0,
            "tice": -1.444501803353837, "sst": -1.0, "sss": 32.0}
    for k, v in pins.items():
        assert final[k].shape == (nx * ny * (nl if k == "tice" else 1),)
        assert np.all(np.abs(final[k] - v) <= ATOL + RTOL * abs(v)), (k, final[k][0], v)
    assert RUN.FINAL_FILE_NAME == "restart.nc"  # core/src/Model.cpp:33
    back, meta = R.read_netcdf(str(tmp_path / RUN.FINAL_FILE_NAME))
    assert meta["nx"] == nx and meta["structure"] == "devgrid"
    for k in R.PROGNOSTIC:
        assert np.array_equal(back[k], final[k]), k
    # a second run resumes from the written restart (t = 1 -> 3, two more steps) and equals
    # three steps of the oracle from the initial file
    cfg.write_text("[model]\ninit_file = %s\nstart = 1\nrun_length = 2\ntime_step = 1\n"
                   % RUN.FINAL_FILE_NAME)
    final3, plan = RUN.run_config(str(cfg), ctx, out_path=str(tmp_path / "r3.nsdgr"),
                                  base_dir=str(tmp_path))
    assert plan["nsteps"] == 2
    st, fo = S.dev1_inputs(nx * ny)
    ref, _, _ = O.thermo_step(ThermoParams.defaults(), 1.0, 3, st, fo)
    compare_state(final3, ref, 0)


@pytest.mark.parametrize("n", [5000, 300_001])  # the larger one takes the chunked 3-stream pipeline
def test_run_host_equals_resident_path(ctx, n):
    st, fo = S.thermo_inputs(n, 8)
    p = params("linear_smu")
    step = run_gpu(ctx, p, st, fo, 600.0, 2)
    a = step.get_prognostic()
    step2 = GpuStep(ctx, n, 1, p)
    prog = dict(st)
    prog["tice"] = st["tice0"].copy()
    b = step2.run_host(600.0, 2, {k: np.array(v) for k, v in prog.items()}, fo)
    for k in ("hice", "cice", "hsnow", "tice"):
        assert np.array_equal(a[k], b[k])
    # the device copy is the same state as well (the call leaves it resident)
    c = step2.get_prognostic()
    for k in ("hice", "cice", "hsnow", "tice", "sst", "sss"):
        assert np.array_equal(a[k], c[k]), k


def test_error_behaviour(ctx):
    step = GpuStep(ctx, 10, 1)
    with pytest.raises(NsdgError):  # step before data
        step.iterate(1.0)
    p = ThermoParams.defaults()
    p.albedo_kind = 7
    with pytest.raises(NsdgError):  # unknown module id: error, never a fallback
        step.set_params(p)
    with pytest.raises(ValueError):
        step.setInitialData({k: np.zeros(9) for k in ("hice", "cice", "hsnow", "sst", "sss", "tice")},
                            {})


def test_zero_mixed_layer_depth_follows_the_reference(ctx):
    """mld == 0: the reference divides by a zero mixed-layer heat capacity (deltaTml = -/+inf,
    NextsimPhysics.cpp:236) and still produces finite ice state through the new-ice branch; the
    kernel uses an IEEE division there and must agree cell by cell."""
    n = 4096
    st, fo = S.thermo_inputs(n, seed=77)
    fo = dict(fo)
    fo["mld"] = fo["mld"].copy()
    fo["mld"][::3] = 0.0
    p = params("linear_smu")
    step = run_gpu(ctx, p, st, fo, 600.0, 2)
    ref, ni, _ = O.thermo_step(p, 600.0, 2, st, fo)
    got = step.get_prognostic()
    for k in ("hice", "cice", "hsnow", "tice0"):
        assert np.array_equal(np.isfinite(ref[k]), np.isfinite(got[k if k in got else "tice"])), k
    fin = np.isfinite(ref["hice"]) & np.isfinite(ni)
    assert fin.sum() > n // 2
    sub = lambda d, m: {k: v[m] for k, v in d.items()}
    compare_state({k: (got[k] if k in got else got["tice"])[fin] for k in ("hice", "cice", "hsnow", "tice")},
                  sub(ref, fin), max_flips=0, newice=(step.get_newice()[fin], ni[fin]))


def test_zero_cooling_flux_below_freezing(ctx):
    """KNOWN DIVERGENCE (documented in DESIGN.md 4.1): with Qow == 0 exactly and sst < tf the
    reference evaluates (tf - sst)/deltaTml * Qow = x/(-0) * 0 = NaN (NextsimPhysics.cpp:244-247)
    and poisons the cell; the kernel's closed form of that quotient, (sst - tf) C/dt, stays
    finite.  Qow is made exactly zero: no wind (no turbulent fluxes), no short wave, and a
    long-wave forcing equal to the surface emission at a temperature whose fourth power is exact
    in binary64 (271 K), so pow(T, 4) of the reference and the two squarings of the kernel agree
    to the bit."""
    n = 64
    st, fo = S.thermo_inputs(n, seed=5)
    st, fo = dict(st), dict(fo)
    st["sst"] = np.full(n, 271.0 - 273.15)
    assert (st["sst"] + 273.15 == 271.0).all()
    st["sss"] = np.full(n, 32.0)  # tf = -1.76 > sst
    fo["wind"] = np.zeros(n)
    fo["qsw"] = np.zeros(n)
    tk = st["sst"] + 273.15
    fo["qlw"] = 0.996 * 5.670374419e-8 * ((tk * tk) * (tk * tk))
    p = params("linear_smu")
    ref, ni, _ = O.thermo_step(p, 600.0, 1, st, fo)
    assert np.isnan(ni).all()  # the reference poisons the cell's new ice (and with it hice where cice > 0)
    assert np.isnan(ref["hice"][st["cice"] > 0.2]).all()
    step = run_gpu(ctx, p, st, fo, 600.0, 1, diag=True)
    gni, got = step.get_newice(), step.get_prognostic()
    assert np.isfinite(gni).all() and all(np.isfinite(v).all() for v in got.values())
    # closed form: Qow = 0, so the latent flux is -(sst - tf) C/dt > 0 and new ice forms
    tf = -0.055 * 32.0
    want = -(st["sst"] - tf) * fo["mld"] * 1025.0 * 4186.84 * (1 - st["cice"]) / (333.55e3 * 917.0)
    assert (gni > 0).all()
    assert np.abs(gni - want).max() / np.abs(want).max() < 1e-12
