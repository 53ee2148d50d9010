"""Restart files as structure-of-arrays, read and written straight from/to the SoA host mirror
(SURVEY.md 8f rank 1).

The reference reads/writes netCDF-4 through netCDF-cxx4 (core/src/DevGridIO.cpp:101-138 read,
:160-201 write): group `structure` (attribute `type` = "devgrid"), group `data` with variables
hice, cice, hsnow, sst, sss [x][y] and tice [x][y][nLayers]; DevGrid stores cell (i, j) at
`i*nx + j`, x outer (DevGridIO.cpp:107-109), tice layer-fastest (:193-199).  Neither netCDF nor
HDF5 exists in this image or on the GPU box.  `write_netcdf` writes that layout as an HDF5 file
through this package's own minimal writer (`hdf5write`), `read_netcdf` reads it (and the
reference's own files) through `hdf5lite`.  A second, trivially simple self-describing container
carries the same content (same names, same dimension order, same element order):

    b"NSDGR1\\n" | uint64 little-endian header length | JSON header | raw little-endian float64

The JSON header holds {"structure": {"type": ...}, "dims": {"x", "y", "nLayers"}, "data":
[{"name", "dims", "offset", "count"}]}.  `read_netcdf` / `from_netcdf` read a reference restart
directly: through the netCDF4 python package when importable, else through `hdf5lite`, this
package's own reader for the HDF5 subset such files use (it reads the reference's shipped
`run/dev1.res.nc`; tests/golden/dev1_restart.npz holds what it found there).  The hidden
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


def write_netcdf(path, state, nx, ny=None, nlayers=1, structure=STRUCTURE_TYPE):
    """The reference's own restart layout (core/src/DevGridIO.cpp:149-201; run/dev_res.py:1-21)
    as a netCDF-4 / HDF5 file written by `hdf5write` (no netCDF or HDF5 library exists here):
    group `structure` with attribute `type`, group `data` with dimensions x, y, nLayers and
    float64 variables hice, cice, hsnow, sst, sss [x][y], tice [x][y][nLayers], in the order
    dumpData creates them.  A `newice` entry (hidden state of the physics, not a reference
    variable) is written as one more [x][y] variable; the reference's reader ignores it.
    Readable by `read_netcdf` / `hdf5lite`; NOT checked against netCDF-C (none available)."""
    from . import hdf5write
    ny = nx if ny is None else ny
    n = nx * ny
    missing = [k for k in PROGNOSTIC if k not in state]
    if missing:
        raise ValueError(f"restart state lacks {missing}")
    # DevGridIO.cpp:161-173: the five 2-D fields in the order of its name -> accessor map, then tice
    order = [k for k in ("cice", "hice", "hsnow", "sss", "sst") ] + ["tice"] + \
            [k for k in state if k not in PROGNOSTIC]
    variables = []
    for name in order:
        a = np.ascontiguousarray(state[name], dtype=np.float64).ravel()
        want = n * nlayers if name == "tice" else n
        if a.size != want:
            raise ValueError(f"{name}: expected {want} values, got {a.size}")
        if name == "tice":
            variables.append((name, ("x", "y", "nLayers"), a.reshape(nx, ny, nlayers)))
        else:
            variables.append((name, ("x", "y"), a.reshape(nx, ny)))
    hdf5write.write_netcdf4(path, {
        "structure": {"attrs": {"type": structure}},
        "data": {"dims": [("x", nx), ("y", ny), ("nLayers", nlayers)], "vars": variables}})


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


def read_netcdf(nc_path):
    """Read a reference netCDF-4 restart (the layout core/src/DevGridIO.cpp:101-138 reads:
    group `structure` with attribute `type`, group `data` with hice, cice, hsnow, sst, sss [x][y]
    and tice [x][y][nLayers]) -> (state, meta) as `read_restart` returns them.  Uses the netCDF4
    python package when importable, else the built-in HDF5 subset reader (`hdf5lite`), which
    handles the uncompressed files `run/dev_res.py` style scripts write."""
    try:
        import netCDF4
    except ImportError:
        netCDF4 = None
    if netCDF4 is not None:
        ds = netCDF4.Dataset(nc_path)
        data = ds.groups["data"]
        arrays = {k: np.asarray(data.variables[k][:], dtype=np.float64) for k in PROGNOSTIC}
        structure = ds.groups["structure"].type
    else:
        from .hdf5lite import File
        f = File(nc_path)
        missing = [k for k in PROGNOSTIC if k not in f.names("/data")]
        if missing:
            raise ValueError(f"{nc_path}: group 'data' lacks {missing}")
        arrays = {k: f.read("/data/" + k).astype(np.float64) for k in PROGNOSTIC}
        if "newice" in f.names("/data"):  # this package's extra variable (write_netcdf)
            arrays["newice"] = f.read("/data/newice").astype(np.float64)
        structure = f.attrs("/structure").get("type")
    if structure != STRUCTURE_TYPE:
        # core/src/StructureFactory.cpp:41-43 throws std::invalid_argument for an unknown structure
        raise ValueError(f"{nc_path}: structure type {structure!r}, expected {STRUCTURE_TYPE!r}")
    if arrays["tice"].ndim != 3:
        raise ValueError(f"{nc_path}: tice must be [x][y][nLayers]")
    nx, ny, nl = arrays["tice"].shape
    for k in [q for q in arrays if q != "tice"]:
        if arrays[k].shape != (nx, ny):
            raise ValueError(f"{nc_path}: {k} has shape {arrays[k].shape}, expected {(nx, ny)}")
    state = {k: np.ascontiguousarray(a).ravel() for k, a in arrays.items()}
    return state, {"nx": nx, "ny": ny, "nlayers": nl, "structure": structure}


def from_netcdf(nc_path, out_path):
    """Convert a reference netCDF-4 restart (DevGridIO layout) to NSDGR1."""
    state, m = read_netcdf(nc_path)
    write_restart(out_path, state, m["nx"], m["ny"], m["nlayers"], structure=m["structure"])


if __name__ == "__main__":
    import sys
    if len(sys.argv) == 3:
        from_netcdf(sys.argv[1], sys.argv[2])
    else:
        print("usage: python -m nextsimdg_b200.restart in.nc out.nsdgr")
