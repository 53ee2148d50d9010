"""Restart files as structure-of-arrays, read and written straight from/to the SoA host mirror
(SURVEY.md 8f rank 1).

The reference reads/writes netCDF-4 through netCDF-cxx4 (core/src/DevGridIO.cpp:101-138 read,
:160-201 write): group `structure` (attribute `type` = "devgrid"), group `data` with variables
hice, cice, hsnow, sst, sss [x][y] and tice [x][y][nLayers]; DevGrid stores cell (i, j) at
`i*nx + j`, x outer (DevGridIO.cpp:107-109), tice layer-fastest (:193-199).  Neither netCDF nor
HDF5 exists in this image or on the GPU box, so the same content travels in a minimal
self-describing container instead (same names, same dimension order, same element order):

    b"NSDGR1\\n" | uint64 little-endian header length | JSON header | raw little-endian float64

The JSON header holds {"structure": {"type": ...}, "dims": {"x", "y", "nLayers"}, "data":
[{"name", "dims", "offset", "count"}]}.  `from_netcdf` converts a reference restart when the
netCDF4 python package is importable (it is not here: the call then raises).  The hidden
per-cell state `newice`, which the reference does not write, is an
optional extra variable so that a resumed GPU run continues bit-identically.
"""
import json
import struct

import numpy as np

MAGIC = b"NSDGR1\n"
PROGNOSTIC = ("hice", "cice", "hsnow", "sst", "sss", "tice")
STRUCTURE_TYPE = "devgrid"  # core/src/modules/include/DevGrid.hpp:51 (ourStructureName)


def write_restart(path, state, nx, ny=None, nlayers=1, structure=STRUCTURE_TYPE):
    """state: name -> flat array in DevGrid cell order (length nx*ny; tice nx*ny*nlayers)."""
    ny = nx if ny is None else ny
    n = nx * ny
    entries, blobs, off = [], [], 0
    for name, arr in state.items():
        a = np.ascontiguousarray(arr, dtype="<f8").ravel()
        want = n * nlayers if name == "tice" else n
        if a.size != want:
            raise ValueError(f"{name}: expected {want} values, got {a.size}")
        dims = ["x", "y", "nLayers"] if name == "tice" else ["x", "y"]
        entries.append({"name": name, "dims": dims, "offset": off, "count": int(a.size)})
        blobs.append(a.tobytes())
        off += a.size * 8
    missing = [k for k in PROGNOSTIC if k not in state]
    if missing:
        raise ValueError(f"restart state lacks {missing}")
    header = json.dumps({"structure": {"type": structure},
                         "dims": {"x": nx, "y": ny, "nLayers": nlayers},
                         "data": entries}).encode()
    with open(path, "wb") as f:
        f.write(MAGIC)
        f.write(struct.pack("<Q", len(header)))
        f.write(header)
        for b in blobs:
            f.write(b)


def read_restart(path):
    """Returns (state dict of flat float64 arrays, meta dict with nx, ny, nlayers, structure)."""
    with open(path, "rb") as f:
        if f.read(len(MAGIC)) != MAGIC:
            raise ValueError(f"{path}: not an NSDGR1 restart file")
        (hlen,) = struct.unpack("<Q", f.read(8))
        header = json.loads(f.read(hlen).decode())
        base = f.tell()
        state = {}
        for e in header["data"]:
            f.seek(base + e["offset"])
            a = np.frombuffer(f.read(e["count"] * 8), dtype="<f8")
            if a.size != e["count"]:
                raise ValueError(f"{path}: truncated variable {e['name']}")
            state[e["name"]] = a.astype(np.float64)
    d = header["dims"]
    meta = {"nx": d["x"], "ny": d["y"], "nlayers": d["nLayers"],
            "structure": header["structure"]["type"]}
    return state, meta


def dev_restart(nx=10, nlayers=1):
    """The content of run/dev1.res.nc as run/dev_res.py:9-20 generates it."""
    n = nx * nx
    return {"hice": np.full(n, 0.1), "cice": np.full(n, 0.5), "hsnow": np.zeros(n),
            "sst": np.full(n, -1.0), "sss": np.full(n, 32.0), "tice": np.full(n * nlayers, -1.0)}


def from_netcdf(nc_path, out_path):
    """Convert a reference netCDF-4 restart (DevGridIO layout) to NSDGR1."""
    try:
        import netCDF4  # noqa: F401
    except ImportError as e:
        raise RuntimeError("from_netcdf needs the netCDF4 python package, which this image does "
                           "not have; generate the state with dev_restart() or write_restart()") from e
    ds = netCDF4.Dataset(nc_path)
    data = ds.groups["data"]
    nx, ny = data.dimensions["x"].size, data.dimensions["y"].size
    nl = data.dimensions["nLayers"].size
    state = {k: np.asarray(data.variables[k][:], dtype=np.float64).ravel() for k in PROGNOSTIC}
    write_restart(out_path, state, nx, ny, nl, structure=ds.groups["structure"].type)


if __name__ == "__main__":
    import sys
    if len(sys.argv) == 3:
        from_netcdf(sys.argv[1], sys.argv[2])
    else:
        print("usage: python -m nextsimdg_b200.restart in.nc out.nsdgr")
