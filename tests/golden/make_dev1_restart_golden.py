"""Extracts tests/golden/dev1_restart.npz from the reference's shipped restart file
run/dev1.res.nc (NETCDF4/HDF5, written by run/dev_res.py:1-21, read by
core/src/DevGridIO.cpp:101-138) with this package's own HDF5 subset reader.  Run in the build
container only (the GPU box has no /root/reference):
    python tests/golden/make_dev1_restart_golden.py
The arrays are stored in DevGrid cell order (i*nx + j, x outer) exactly as found in the file."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
SRC = "/root/reference/run/dev1.res.nc"

if __name__ == "__main__":
    from nextsimdg_b200 import hdf5lite, restart
    state, meta = restart.read_netcdf(SRC)
    f = hdf5lite.File(SRC)
    out = {k: v for k, v in state.items()}
    out["dims"] = np.array([meta["nx"], meta["ny"], meta["nlayers"]], dtype=np.int64)
    out["structure_type"] = np.array(meta["structure"])
    out["data_names"] = np.array(f.names("/data"))
    out["nc_properties"] = np.array(f.attrs("/")["_NCProperties"])
    np.savez(os.path.join(ROOT, "tests", "golden", "dev1_restart.npz"), **out)
    print({k: (v.shape, v.ravel()[:1]) for k, v in out.items()})
