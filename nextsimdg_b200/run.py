"""The caller of the path: what `nextsim --config-file run/dev1.cfg` does around `DevStep`
(BASELINE.json configs[0]), with the GPU step in its place.

Mirrors `Model::configure` / `Model::run` / `~Model` (core/src/Model.cpp:57-91) and
`Iterator::parseAndSet` / `Iterator::run` (core/src/Iterator.cpp:29-62):

* `[model] init_file, start, stop, run_length, time_step` — integer seconds
  (core/src/include/Iterator.hpp:22-23); `run_length`, when present, wins over `stop`
  (Iterator.cpp:42-50); `init_file` is resolved relative to the working directory, as the
  reference does, or to `base_dir` when given;
* the restart is read into the SoA mirror (netCDF-4 through `restart.read_netcdf`, or the NSDGR1
  container), never into per-cell objects;
* every cell gets the constant forcing of `DummyExternalData::setAll`
  (core/src/include/DummyExternalData.hpp:22-34; wind stays 0);
* the step is `iterate(time_step)` for `t = start; t < stop; t += time_step` (Iterator.cpp:57-59);
* on exit the state is written to `restart.nc`'s stand-in (Model.cpp:33 `finalFileName`).

This is host plumbing around the C ABI, not a rebuild of the reference's `Model`; there is no CPU
path: `run_config` needs a `gpu.Context`.
"""
import configparser
import os

import numpy as np

from . import restart
from .config import thermo_params_from_ini

# core/src/include/DummyExternalData.hpp:25-32
DUMMY_FORCING = dict(tair=-1.0, tdew=-4.0, slp=1e5, qsw=0.0, qlw=311.0, mld=10.0, snowfall=0.0,
                     wind=0.0)
FINAL_FILE_NAME = "restart.nc"  # Model.cpp:33; netCDF-4 layout of DevGridIO::dump via hdf5write


def _int(text, key):
    # Iterator::parseAndSet reads with operator>> into an int; an absent key leaves 0.  Text that
    # is not an integer is an error here instead of the reference's silent 0
    t = (text or "").strip()
    if not t:
        return 0
    try:
        return int(t)
    except ValueError:
        raise ValueError(f"model.{key} = {text!r}: integer seconds expected "
                         "(core/src/include/Iterator.hpp:22-23)") from None


def plan_from_config(text):
    """INI text -> dict(init_file, start, stop, time_step, nsteps, params)."""
    cp = configparser.ConfigParser(delimiters=("=",), comment_prefixes=("#",), strict=False,
                                   interpolation=None)
    cp.optionxform = str
    cp.read_string(text or "")
    m = dict(cp.items("model")) if cp.has_section("model") else {}
    start = _int(m.get("start"), "start")
    step = _int(m.get("time_step"), "time_step")
    if m.get("run_length", "").strip():
        stop = start + _int(m["run_length"], "run_length")
    else:
        stop = _int(m.get("stop"), "stop")
    if stop > start and step <= 0:
        raise ValueError("model.time_step must be positive (the reference's loop "
                         "`for (t = start; t < stop; t += timestep)` would not end)")
    nsteps = 0 if stop <= start else -(-(stop - start) // step)
    return {"init_file": m.get("init_file", "").strip(), "start": start, "stop": stop,
            "time_step": step, "nsteps": nsteps, "params": thermo_params_from_ini(text)}


def read_initial_state(path):
    """Restart in either format -> (state, meta); the format is told from the file's magic."""
    with open(path, "rb") as f:
        magic = f.read(8)
    if magic.startswith(restart.MAGIC):
        return restart.read_restart(path)
    return restart.read_netcdf(path)


def dummy_forcing(n):
    return {k: np.full(n, v) for k, v in DUMMY_FORCING.items()}


def run_config(cfg_path, ctx, out_path=None, base_dir=None):
    """Run a reference configuration file on the GPU step.  Returns (final state, plan)."""
    from .gpu import GpuStep
    with open(cfg_path) as f:
        plan = plan_from_config(f.read())
    if not plan["init_file"]:
        raise ValueError(f"{cfg_path}: model.init_file is not set")
    init = plan["init_file"]
    if base_dir is not None and not os.path.isabs(init):
        init = os.path.join(base_dir, init)
    state, meta = read_initial_state(init)
    n = meta["nx"] * meta["ny"]
    step = GpuStep(ctx, n, meta["nlayers"], plan["params"])
    step.setInitialData(state, dummy_forcing(n))
    if "newice" in state:
        step.set_newice(state["newice"])
    step.init()
    step.start(plan["start"])
    if plan["nsteps"]:
        step.iterate(float(plan["time_step"]), plan["nsteps"])
    step.stop(plan["stop"])
    final = step.get_prognostic()
    final["newice"] = step.get_newice()
    if out_path is None:
        out_path = FINAL_FILE_NAME if base_dir is None else os.path.join(base_dir, FINAL_FILE_NAME)
    writer = restart.write_restart if out_path.endswith(".nsdgr") else restart.write_netcdf
    writer(out_path, final, meta["nx"], meta["ny"], meta["nlayers"], structure=meta["structure"])
    return final, plan


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser(description="run a reference config file on the GPU step")
    ap.add_argument("--config-file", required=True)  # core/src/main.cpp's own option name
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    from .gpu import Context
    c = Context(0)
    fin, pl = run_config(a.config_file, c, a.out, base_dir=os.path.dirname(a.config_file) or ".")
    print({k: float(v[0]) for k, v in fin.items()}, "steps:", pl["nsteps"])
    c.close()
