"""Generates tests/golden/thermo_ref.npz from the reference's OWN sources compiled in place
(oracle/_ref/libnsdg_ref.so, see oracle/Makefile).  Run in the build container only:
    python tests/golden/make_thermo_golden.py
The .npz travels to the GPU box; /root/reference does not.  One subprocess per module
configuration because the reference's configuration is process-global."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

CUSTOM_KEYS_INI = (
    "[nextsim_thermo]\ndrag_ocean_q = 1.7e-3\ndrag_ocean_t = 0.9e-3\ndrag_ice_t = 1.1e-3\n"
    "albedoW = 0.09\nI_0 = 0.2\nmin_conc = 1e-6\nmin_thick = 0.02\n"
    "[thermoice0]\nks = 0.28\nflooding = false\n[Hibler]\nh0 = 0.4\nphiM = 0.6\n")

CONFIGS = {
    # name: (freezing, albedo, extra ini)
    "linear_smu": ("LinearFreezing", "SMUIceAlbedo", ""),
    "unesco_ccsm": ("UnescoFreezing", "CCSMIceAlbedo",
                    "[CCSMIceAlbedo]\niceAlbedo = 0.63\nsnowAlbedo = 0.88\n"),
    "linear_smu2": ("LinearFreezing", "SMU2IceAlbedo", ""),
    # every configuration key of the path away from its default, flooding off
    "custom_keys": ("UnescoFreezing", "SMUIceAlbedo", CUSTOM_KEYS_INI),
}
N, NSTEPS, DT, SEED = 2000, 4, 600.0, 777


def worker(name):
    from nextsimdg_b200 import synthetic as S
    from oracle import refthermo as R
    fr, al, extra = CONFIGS[name]
    R.configure(R.ini(fr, al, extra))
    out = {}
    for tag, (st, fo), dt in (("rand", S.thermo_inputs(N, SEED), DT),
                              ("cold", cold_inputs(N, SEED + 1), DT),
                              ("dev1", S.dev1_inputs(100), 1.0)):
        n = st["hice"].size
        g = R.RefGrid(n, 1)
        g.set_prognostic(st["hice"], st["cice"], st["hsnow"], st["sst"], st["sss"], st["tice0"])
        g.set_forcing(**fo)
        for s in range(NSTEPS):
            g.step(dt, 1, 1, 1)
            if s not in (0, NSTEPS - 1):
                continue
            pr = g.get_prognostic()
            pv = g.get_private()
            for k in ("hice", "cice", "hsnow", "tice"):
                out[f"{tag}_s{s}_{k}"] = pr[k].copy()
            for k in (("newice", "qia", "qio", "subl", "dqdt", "hifroms") if s == 0
                      else ("newice",)):
                out[f"{tag}_s{s}_{k}"] = pv[k].copy()
        g.close()
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", f"thermo_ref_{name}.npz"), **out)


def cold_inputs(n, seed):
    """Near-freezing ocean so that ice survives and the growth / flooding / new-ice branches
    are exercised for several steps."""
    from nextsimdg_b200 import synthetic as S
    st, fo = S.thermo_inputs(n, seed)
    r = np.random.default_rng(seed + 99)
    st["sst"] = -0.055 * st["sss"] + r.uniform(-0.02, 0.05, n)
    fo["mld"] = r.uniform(10.0, 20.0, n)
    st["hsnow"] = np.where(r.random(n) < 0.3, st["cice"] * r.uniform(0.2, 1.0, n), st["hsnow"])
    return st, fo


if __name__ == "__main__":
    if len(sys.argv) > 1:
        worker(sys.argv[1])
    else:
        for name in CONFIGS:
            subprocess.check_call([sys.executable, os.path.abspath(__file__), name])
            print("wrote", name)
