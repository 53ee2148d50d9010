"""Minimal HDF5 writer for netCDF-4 style restart files (the counterpart of `hdf5lite`).

The image has neither netCDF nor HDF5, so a GPU run could not hand a restart back to the
reference (`core/src/DevGridIO.cpp:65-78,101-138` opens it with netCDF-cxx4).  This module writes
the subset of HDF5 that the reference's own restart (`run/dev1.res.nc`, produced by
`run/dev_res.py:1-21`) consists of, from the published file-format specification (version 3.0)
and from the byte layout of that file:

* superblock version 2; version-2 object headers (one chunk each, Jenkins lookup3 checksums),
  attribute creation order tracked, as netCDF-4 writes them;
* groups with COMPACT link storage (link messages with creation order; the group-info message
  raises the compact limit to 32 links so that no fractal heap / B-tree is needed);
* contiguous little-endian float64 datasets with the netCDF default fill value, dataspaces with
  maximum dimensions;
* netCDF-4 dimensions as HDF5 dimension-scale datasets (`CLASS`, `NAME`, `_Netcdf4Dimid`) and
  the hidden `_Netcdf4Coordinates` attribute on every variable, from which netCDF-C assigns the
  dimension ids without walking `DIMENSION_LIST` / `REFERENCE_LIST` (object-reference types that
  are NOT written here);
* fixed-length string and int32 attributes.

What is checked here: `hdf5lite` reads the files back, the checksum routine reproduces every
checksum stored in `run/dev1.res.nc`, and the object tree (groups, names, shapes, types,
attributes) equals that file's (tests/test_host.py).  What CANNOT be checked here: that
netCDF-C / netCDF-cxx4 opens the file, because no netCDF library exists in this image.
"""
import struct

import numpy as np

SIGNATURE = b"\x89HDF\r\n\x1a\n"
UNDEF = 0xFFFFFFFFFFFFFFFF
NC_FILL_DOUBLE = 9.9692099683868690e+36


def lookup3(data, initval=0):
    """Bob Jenkins' lookup3 hashlittle(), the checksum of HDF5 metadata (H5_checksum_lookup3)."""
    M = 0xFFFFFFFF

    def rot(x, k):
        return ((x << k) & M) | (x >> (32 - k))

    n = len(data)
    a = b = c = (0xDEADBEEF + n + initval) & M
    p = 0
    while n > 12:
        a = (a + int.from_bytes(data[p:p + 4], "little")) & M
        b = (b + int.from_bytes(data[p + 4:p + 8], "little")) & M
        c = (c + int.from_bytes(data[p + 8:p + 12], "little")) & M
        a = (a - c) & M; a ^= rot(c, 4); c = (c + b) & M
        b = (b - a) & M; b ^= rot(a, 6); a = (a + c) & M
        c = (c - b) & M; c ^= rot(b, 8); b = (b + a) & M
        a = (a - c) & M; a ^= rot(c, 16); c = (c + b) & M
        b = (b - a) & M; b ^= rot(a, 19); a = (a + c) & M
        c = (c - b) & M; c ^= rot(b, 4); b = (b + a) & M
        p += 12
        n -= 12
    if n == 0:
        return c
    tail = bytes(data[p:p + n]) + b"\0" * (12 - n)
    a = (a + int.from_bytes(tail[0:4], "little")) & M
    b = (b + int.from_bytes(tail[4:8], "little")) & M
    c = (c + int.from_bytes(tail[8:12], "little")) & M
    c ^= b; c = (c - rot(b, 14)) & M
    a ^= c; a = (a - rot(c, 11)) & M
    b ^= a; b = (b - rot(a, 25)) & M
    c ^= b; c = (c - rot(b, 16)) & M
    a ^= c; a = (a - rot(c, 4)) & M
    b ^= a; b = (b - rot(a, 14)) & M
    c ^= b; c = (c - rot(b, 24)) & M
    return c


# ---- message bodies -----------------------------------------------------------------------
F64_LE = bytes.fromhex("11203f000800000000004000340b0034ff030000")  # IEEE binary64, little-endian
F32_BE = bytes.fromhex("11211f000400000000002000170800177f000000")  # netCDF's dimension-only scales
I32_LE = bytes.fromhex("1008000004000000" "00002000")  # fixed point, signed, 32 bits


def _string_type(n):
    return bytes([0x13, 0x00, 0x00, 0x00]) + struct.pack("<I", n)  # null-terminated ASCII, n bytes


def _dataspace(shape, with_max=True):
    if shape == ():
        return bytes([2, 0, 0, 0])  # scalar
    body = bytes([2, len(shape), 1 if with_max else 0, 1]) + struct.pack(f"<{len(shape)}Q", *shape)
    if with_max:
        body += struct.pack(f"<{len(shape)}Q", *shape)
    return body


def _attribute(name, dtype, space, raw):
    nm = name.encode() + b"\0"
    return (bytes([3, 0]) + struct.pack("<3H", len(nm), len(dtype), len(space)) + b"\0" + nm + dtype
            + space + raw)


def attr_string(name, text):
    raw = text.encode() + b"\0"
    return _attribute(name, _string_type(len(raw)), _dataspace(()), raw)


def attr_int32(name, values):
    a = np.atleast_1d(np.asarray(values, dtype="<i4"))
    space = _dataspace(()) if np.ndim(values) == 0 else _dataspace(a.shape, with_max=False)
    return _attribute(name, I32_LE, space, a.tobytes())


class _Header:
    """one version-2 object header: messages in creation order, one chunk, checksum"""

    def __init__(self):
        self.msgs = []  # (type, flags, body)
        self.nattr = 0

    def add(self, mtype, body, flags=0):
        self.msgs.append((mtype, flags, body))

    def add_attr(self, body):
        self.add(0x0C, body)
        self.nattr += 1

    def size(self):
        return 4 + 1 + 1 + 2 + sum(6 + len(b) for _, _, b in self.msgs) + 4

    def encode(self):
        out = b"OHDR" + bytes([2, 0x0D])  # attr creation order tracked + indexed, 2-byte chunk size
        out += struct.pack("<H", sum(6 + len(b) for _, _, b in self.msgs))
        for order, (t, fl, b) in enumerate(self.msgs):
            out += bytes([t]) + struct.pack("<H", len(b)) + bytes([fl]) + struct.pack("<H", order) + b
        return out + struct.pack("<I", lookup3(out))


def _attr_info(n):
    return bytes([0, 3]) + struct.pack("<H", n) + struct.pack("<3Q", UNDEF, UNDEF, UNDEF)


def _group_header(nlinks):
    h = _Header()
    # link info: creation order tracked + indexed, compact storage (no heap, no B-trees)
    h.add(0x02, bytes([0, 3]) + struct.pack("<Q", nlinks) + struct.pack("<3Q", UNDEF, UNDEF, UNDEF))
    # group info: link phase change stored: up to 32 links stay compact
    h.add(0x0A, bytes([0, 1]) + struct.pack("<HH", 32, 16), flags=1)
    return h


def _link(name, order, addr):
    nm = name.encode()
    return bytes([1, 0x04]) + struct.pack("<Q", order) + bytes([len(nm)]) + nm + struct.pack("<Q", addr)


def _dataset_header(shape, dtype_body, fill, data_addr, nbytes):
    h = _Header()
    h.add(0x01, _dataspace(shape))
    h.add(0x03, dtype_body, flags=1)
    if fill is None:
        h.add(0x05, bytes([3, 0x0A]), flags=1)  # late allocation, fill if set, no value
    else:
        h.add(0x05, bytes([3, 0x2A]) + struct.pack("<I", len(fill)) + fill, flags=1)
    h.add(0x08, bytes([3, 1]) + struct.pack("<QQ", data_addr, nbytes))
    return h


def write_netcdf4(path, groups):
    """groups: {group name: {"attrs": {name: str}, "dims": [(name, size)], "vars":
    [(name, dims tuple of names, float64 array)]}}.  Dimension ids are numbered in the order
    the dimensions are declared, over all groups (netCDF-4 ids are file-wide)."""
    # ---- plan: collect objects, then assign addresses
    objects = []  # (kind, ...), each gets an address
    dimid = {}
    plan = {}
    for gname, g in groups.items():
        entries = []  # (link name, header builder taking the data address -> _Header, data bytes)
        for dname, size in g.get("dims", []):
            did = dimid.setdefault((gname, dname), len(dimid))
            entries.append((dname, ("dim", size, did), np.zeros(size, ">f4").tobytes()))
        for vname, vdims, arr in g.get("vars", []):
            a = np.ascontiguousarray(arr, dtype="<f8")
            want = tuple(dict(g["dims"])[d] for d in vdims)
            if a.shape != want:
                raise ValueError(f"{gname}/{vname}: shape {a.shape}, dimensions say {want}")
            entries.append((vname, ("var", a.shape, [dimid[(gname, d)] for d in vdims]), a.tobytes()))
        # netCDF-4 lists a group's members alphabetically in dev1.res.nc; creation order is kept
        plan[gname] = entries

    def dataset_header(desc, addr, nbytes):
        if desc[0] == "dim":
            _, size, did = desc
            h = _dataset_header((size,), F32_BE, None, addr, nbytes)
            h.add(0x15, _attr_info(3), flags=4)
            h.add_attr(attr_string("CLASS", "DIMENSION_SCALE"))
            h.add_attr(attr_string("NAME", "This is a netCDF dimension but not a netCDF variable.%10d" % size))
            h.add_attr(attr_int32("_Netcdf4Dimid", np.int32(did)))
        else:
            _, shape, ids = desc
            h = _dataset_header(shape, F64_LE, struct.pack("<d", NC_FILL_DOUBLE), addr, nbytes)
            h.add(0x15, _attr_info(1), flags=4)
            h.add_attr(attr_int32("_Netcdf4Coordinates", ids))
        return h

    def group_header(gname, link_addrs):
        g = groups[gname]
        h = _group_header(len(link_addrs))
        for order, (nm, addr) in enumerate(link_addrs):
            h.add(0x06, _link(nm, order, addr))
        h.add(0x15, _attr_info(len(g.get("attrs", {}))), flags=4)
        for k, v in g.get("attrs", {}).items():
            h.add_attr(attr_string(k, v))
        return h

    # sizes do not depend on the addresses, so one dry pass fixes the layout
    pos = 48
    root_addr = pos
    root = _group_header(len(groups))
    for order, gname in enumerate(groups):
        root.add(0x06, _link(gname, order, 0))
    root.add(0x15, _attr_info(0), flags=4)
    pos += root.size()
    gaddr, daddr = {}, {}
    for gname, entries in plan.items():
        gaddr[gname] = pos
        pos += group_header(gname, [(nm, 0) for nm, _, _ in entries]).size()
        for nm, desc, raw in entries:
            daddr[(gname, nm)] = pos
            pos += dataset_header(desc, 0, len(raw)).size()
    pos = (pos + 7) & ~7
    raw_addr = {}
    for gname, entries in plan.items():
        for nm, _, raw in entries:
            raw_addr[(gname, nm)] = pos
            pos = (pos + len(raw) + 7) & ~7
    eof = pos

    out = bytearray(eof)
    sb = SIGNATURE + bytes([2, 8, 8, 0]) + struct.pack("<4Q", 0, UNDEF, eof, root_addr)
    out[0:48] = sb + struct.pack("<I", lookup3(sb))
    root = _group_header(len(groups))
    for order, gname in enumerate(groups):
        root.add(0x06, _link(gname, order, gaddr[gname]))
    root.add(0x15, _attr_info(0), flags=4)
    enc = root.encode()
    out[root_addr:root_addr + len(enc)] = enc
    for gname, entries in plan.items():
        enc = group_header(gname, [(nm, daddr[(gname, nm)]) for nm, _, _ in entries]).encode()
        out[gaddr[gname]:gaddr[gname] + len(enc)] = enc
        for nm, desc, raw in entries:
            enc = dataset_header(desc, raw_addr[(gname, nm)], len(raw)).encode()
            a = daddr[(gname, nm)]
            out[a:a + len(enc)] = enc
            out[raw_addr[(gname, nm)]:raw_addr[(gname, nm)] + len(raw)] = raw
    with open(path, "wb") as f:
        f.write(out)
