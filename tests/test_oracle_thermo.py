"""Pins the thermodynamics oracle (oracle/thermo_port.c): against the golden values of the
reference's own tests, against fixtures produced by the reference's own sources
(tests/golden/thermo_ref_*.npz, generator committed beside them) and, where oracle/_ref was
built in this container, against that library live."""
import os
import subprocess
import sys

import numpy as np
import pytest

from nextsimdg_b200 import synthetic as S
from nextsimdg_b200._abi import ThermoParams
from oracle import cpu_oracle as O

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def params(freezing=0, albedo=0, ice=0.538, snow=0.8256):
    p = ThermoParams.defaults()
    p.freezing_kind, p.albedo_kind = freezing, albedo
    p.ccsm_ice_albedo, p.ccsm_snow_albedo = ice, snow
    return p


def custom_keys_params():
    """Every configuration key of the path away from its default (the INI text the golden
    generator fed to the reference), parsed by the product's own config reader."""
    sys.path.insert(0, GOLD)
    import make_thermo_golden as G
    from nextsimdg_b200.config import thermo_params_from_ini
    return thermo_params_from_ini("[Modules]\nNextsim::IFreezingPoint = Nextsim::UnescoFreezing\n"
                                  "Nextsim::IIceAlbedo = Nextsim::SMUIceAlbedo\n" + G.CUSTOM_KEYS_INI)


CONFIGS = {"linear_smu": params(0, 0), "unesco_ccsm": params(1, 2, 0.63, 0.88),
           "linear_smu2": params(0, 1), "custom_keys": custom_keys_params()}


def one_cell(p, dt, hice, cice, hsnow, sst, sss, tice, tair, tdew, slp, qsw, qlw, mld, snowfall,
             wind):
    a = lambda v: np.array([float(v)])
    st = dict(hice=a(hice), cice=a(cice), hsnow=a(hsnow), sst=a(sst), sss=a(sss), tice0=a(tice))
    fo = dict(tair=a(tair), tdew=a(tdew), slp=a(slp), qsw=a(qsw), qlw=a(qlw), mld=a(mld),
              snowfall=a(snowfall), wind=a(wind))
    s, ni, d = O.thermo_step(p, dt, 1, st, fo, diag=True)
    return s, ni[0], {k: v[0] for k, v in d.items()}


def test_reference_derived_data():
    # physics/test/NextsimPhysics_test.cpp:47-78 (tolerance of the reference test: 1e-4 rel)
    _, _, d = one_cell(params(), 1.0, 0.1, 0.5, 0.0, -1, 32, -2, tair=-3, tdew=0.1, slp=100000,
                       qsw=0, qlw=0, mld=10, snowfall=0, wind=0)
    assert d["rho"] == pytest.approx(1.29253, rel=1e-4)
    assert d["sphuma"] == pytest.approx(0.00385326, rel=1e-4)
    assert d["sphumw"] == pytest.approx(0.00349446, rel=1e-4)
    assert d["sphumi"] == pytest.approx(0.00323958, rel=1e-4)
    assert d["cspec"] == pytest.approx(1011.81, rel=1e-4)
    # full-precision pins from the reference build
    assert d["rho"] == pytest.approx(1.2925314542542079, rel=1e-13)
    assert d["sphuma"] == pytest.approx(0.003853256222259654, rel=1e-13)
    assert d["sphumi"] == pytest.approx(0.0032395781371107875, rel=1e-13)


def test_reference_drag_pressure():
    # NextsimPhysics_test.cpp:126-174
    for w, tau, full in ((1.5, 0.00126936, 0.0012693596387566633),
                         (8, 0.00141407, 0.0014140666375749228),
                         (23, 0.00253872, 0.0025387192775133266)):
        _, _, d = one_cell(params(), 1.0, 0.1, 0.5, 0.0, -1.5, 32, -1, tair=2, tdew=1.5, slp=100000,
                           qsw=0, qlw=0, mld=10, snowfall=0, wind=w)
        assert d["tau"] == pytest.approx(tau, rel=1e-4)
        assert d["tau"] == pytest.approx(full, rel=1e-13)


def test_reference_melting():
    # NextsimPhysics_test.cpp:176-243 and core/test/ElementData_test.cpp:21-86
    p = CONFIGS["unesco_ccsm"]
    s, ni, d = one_cell(p, 600.0, 0.1, 0.5, 0.01, -1, 32, -1, tair=3, tdew=2, slp=100000, qsw=50,
                        qlw=330, mld=10, snowfall=0, wind=5)
    c = s["cice"][0]
    assert s["hice"][0] / c == pytest.approx(0.12846, rel=1e-4)
    assert s["hsnow"][0] / c == pytest.approx(0.01957732, rel=1e-4)
    assert c == pytest.approx(0.368269, rel=1e-4)
    assert s["tice0"][0] == 0.0
    assert ni == 0.0
    assert d["qia"] == pytest.approx(-84.6156, rel=1e-2)
    assert d["qio"] == pytest.approx(53717.8, rel=1e-2)
    assert d["subl"] == pytest.approx(-7.3858e-06, rel=1e-4)
    assert d["dqdt"] == pytest.approx(19.7013, rel=1e-2)
    assert d["hifroms"] == 0.0
    assert d["rho"] == pytest.approx(1.26488, rel=1e-4)
    assert d["cspec"] == pytest.approx(1012.86, rel=1e-4)
    # full precision
    assert c == pytest.approx(0.36826938031129286, rel=1e-13)
    assert d["qia"] == pytest.approx(-84.595181806888306, rel=1e-13)
    assert d["qio"] == pytest.approx(53717.834007845158, rel=1e-13)


def test_reference_freezing():
    # NextsimPhysics_test.cpp:245-312
    p = CONFIGS["unesco_ccsm"]
    s, ni, d = one_cell(p, 600.0, 0.1, 0.5, 0.01, -1.75, 32, -9, tair=-12, tdew=-12, slp=100000,
                        qsw=0, qlw=265, mld=10, snowfall=1e-3, wind=5)
    c = s["cice"][0]
    assert s["hice"][0] / c == pytest.approx(0.199998, rel=1e-4)
    assert s["hsnow"][0] / c == pytest.approx(0.02179357, rel=1e-4)
    assert c == pytest.approx(0.5002, rel=1e-4)
    assert s["tice0"][0] == pytest.approx(-8.90443, rel=1e-4)
    assert ni == pytest.approx(6.79707e-5, rel=1e-2)
    assert d["qia"] == pytest.approx(42.2955, rel=1e-2)
    assert d["qio"] == pytest.approx(73.9465, rel=1e-2)
    assert d["subl"] == pytest.approx(2.15132e-06, rel=1e-4)
    assert d["dqdt"] == pytest.approx(16.7615, rel=1e-2)
    assert ni == pytest.approx(6.799058277473154e-05, rel=1e-13)
    assert s["tice0"][0] == pytest.approx(-8.9044278062473516, rel=1e-13)


def test_reference_new_ice():
    # NextsimPhysics_test.cpp:80-124: one-day step, UNESCO freezing
    p = params(freezing=1)
    s, ni, d = one_cell(p, 86400.0, 0.1, 0.5, 0.0, -1.5, 32, -2.0, tair=-3, tdew=0.1, slp=1e5,
                        qsw=0, qlw=0, mld=10, snowfall=0, wind=0)
    # the reference test's constant times its Stefan-Boltzmann correction factor (:120-123)
    assert ni == pytest.approx(0.0258236 * 5.670374419e-8 / 5.67e-8, rel=1e-4)
    assert ni == pytest.approx(0.025826427498618871, rel=1e-13)


def test_dev1_config():
    # BASELINE config 1 (run/dev1.cfg): oracle-derived pins from the reference build
    st, fo = S.dev1_inputs(100)
    s, ni, _ = O.thermo_step(params(), 1.0, 1, st, fo)
    assert np.all(s["hice"] == s["hice"][0])
    assert s["hice"][0] == pytest.approx(0.04668325240678619, rel=1e-13)
    assert s["cice"][0] == pytest.approx(0.36670813101696548, rel=1e-13)
    assert s["hsnow"][0] == 0.0
    assert s["tice0"][0] == pytest.approx(-1.444501803353837, rel=1e-13)


@pytest.mark.parametrize("name", sorted(CONFIGS))
def test_port_matches_reference_fixtures(name):
    """Bitwise agreement with the reference's own sources on randomized multi-step inputs."""
    sys.path.insert(0, GOLD)
    import make_thermo_golden as G
    gold = np.load(os.path.join(GOLD, f"thermo_ref_{name}.npz"))
    p = CONFIGS[name]
    for tag, (st, fo), dt in (("rand", S.thermo_inputs(G.N, G.SEED), G.DT),
                              ("cold", G.cold_inputs(G.N, G.SEED + 1), G.DT),
                              ("dev1", S.dev1_inputs(100), 1.0)):
        ni = None
        for s in range(G.NSTEPS):
            st2, ni, d = O.thermo_step(p, dt, 1, st, fo, newice=ni, diag=True)
            st = dict(st2)
            if s in (0, G.NSTEPS - 1):
                for k, kk in (("hice", "hice"), ("cice", "cice"), ("hsnow", "hsnow"),
                              ("tice", "tice0")):
                    np.testing.assert_array_equal(st[kk], gold[f"{tag}_s{s}_{k}"], err_msg=f"{tag} {s} {k}")
                np.testing.assert_array_equal(ni, gold[f"{tag}_s{s}_newice"])
            if s == 0:
                for k in ("qia", "qio", "subl", "dqdt", "hifroms"):
                    np.testing.assert_array_equal(d[k], gold[f"{tag}_s0_{k}"], err_msg=f"{tag} {k}")
    # the fixtures exercise every branch
    if name == "custom_keys":
        assert not np.any(gold["cold_s0_hifroms"] > 0), "flooding = false must switch snow-ice off"
    else:
        assert np.any(gold["cold_s0_hifroms"] > 0), "flooding branch not covered"
    assert np.any(gold["cold_s3_newice"] > 0), "new-ice branch not covered"
    assert np.any(gold["rand_s0_cice"] == 0) and np.any(gold["rand_s3_cice"] > 0)


def test_port_matches_live_reference_library():
    """Where the reference compiled here, run it live beside the port (subprocess: the
    reference's configuration is process-global)."""
    from oracle import refthermo
    if not refthermo.available():
        pytest.skip("oracle/_ref/libnsdg_ref.so not built (no /root/reference on this box)")
    code = r"""
import sys, numpy as np
sys.path.insert(0, %r)
from oracle import refthermo as R, cpu_oracle as O
from nextsimdg_b200 import synthetic as S
from nextsimdg_b200._abi import ThermoParams
R.configure(R.ini("UnescoFreezing", "SMU2IceAlbedo"))
p = ThermoParams.defaults(); p.freezing_kind = 1; p.albedo_kind = 1
st, fo = S.thermo_inputs(30011, 5)
g = R.RefGrid(30011, 2)
g.set_prognostic(st["hice"], st["cice"], st["hsnow"], st["sst"], st["sss"], np.repeat(st["tice0"], 2))
g.set_forcing(**fo)
g.step(450.0, 3, 2, 0)
ref = g.get_prognostic()
s, ni, _ = O.thermo_step(p, 450.0, 3, st, fo)
for k, kk in (("hice", "hice"), ("cice", "cice"), ("hsnow", "hsnow")):
    assert np.array_equal(ref[k], s[kk]), k
assert np.array_equal(ref["tice"].reshape(-1, 2)[:, 0], s["tice0"])
assert np.array_equal(ref["tice"].reshape(-1, 2)[:, 1], s["tice0"])  # layers collapse (Q9)
print("ok")
""" % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))),)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout + r.stderr
