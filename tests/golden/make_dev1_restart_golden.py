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
    # object tree of the file + every checksummed metadata block (superblock, object headers) with
    # the checksum stored in the file: pins hdf5write's lookup3 and its output's structure
    import json
    import struct
    b = f.b

    def block(addr):
        flags = b[addr + 5]
        p = addr + 6 + (16 if flags & 0x20 else 0) + (4 if flags & 0x10 else 0)
        szlen = 1 << (flags & 3)
        end = p + szlen + int.from_bytes(b[p:p + szlen], "little")
        return {"bytes": b[addr:end].hex(), "checksum": int.from_bytes(b[end:end + 4], "little")}

    def describe(path):
        o = f.group(path)
        d = {"attrs": {k: (v.tolist() if hasattr(v, "tolist") else v) for k, v in o.attrs.items() if v is not None}}
        if o.shape is not None and o.dtype is not None:
            d.update(shape=list(o.shape), dtype=o.dtype.str, layout=o.layout[0])
        else:
            d["members"] = {n: describe(path.rstrip("/") + "/" + n) for n in sorted(o.links)}
        return d

    blocks = [{"bytes": b[:44].hex(), "checksum": int.from_bytes(b[44:48], "little")},
              block(struct.unpack_from("<Q", b, 36)[0])]
    blocks += [block(a) for a in f.root.links.values()] + [block(a) for a in f.group("/data").links.values()]
    with open(os.path.join(ROOT, "tests", "golden", "dev1_restart_tree.json"), "w") as fo:
        json.dump({"tree": describe("/"), "checksummed_blocks": blocks}, fo, indent=1)
    print(len(blocks), "checksummed blocks")
