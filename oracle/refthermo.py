"""TEST INFRASTRUCTURE: ctypes binding of oracle/_ref/libnsdg_ref.so.

That library is the reference's own column-physics sources compiled in place (see
oracle/Makefile, oracle/ref_harness.cpp).  Only tests/, __graft_entry__.smoke() and the
CPU-baseline legs of bench.py may import this module; the product package never does.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libnsdg_ref.so")

_dp = ctypes.POINTER(ctypes.c_double)
_lib = None

PROG = ("hice", "cice", "hsnow", "sst", "sss", "tice")
FORC = ("tair", "tdew", "slp", "qsw", "qlw", "mld", "snowfall", "wind")


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(LIB_PATH)
        L.nsref_last_error.restype = ctypes.c_char_p
        L.nsref_configure.argtypes = [ctypes.c_char_p]
        L.nsref_grid_create.restype = ctypes.c_void_p
        L.nsref_grid_create.argtypes = [ctypes.c_int, ctypes.c_int]
        L.nsref_grid_destroy.argtypes = [ctypes.c_void_p]
        L.nsref_grid_set_prognostic.argtypes = [ctypes.c_void_p] + [_dp] * 6
        L.nsref_grid_set_forcing.argtypes = [ctypes.c_void_p] + [_dp] * 8
        L.nsref_grid_step.argtypes = [ctypes.c_void_p, ctypes.c_double, ctypes.c_int,
                                      ctypes.c_int, ctypes.c_int]
        L.nsref_grid_get_prognostic.argtypes = [ctypes.c_void_p] + [_dp] * 6
        L.nsref_grid_get_physics.argtypes = [ctypes.c_void_p] + [_dp] * 9
        L.nsref_grid_get_private.argtypes = [ctypes.c_void_p] + [_dp] * 7
        L.nsref_sizeof_elementdata.restype = ctypes.c_size_t
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(_dp)


def _c(a, n):
    a = np.ascontiguousarray(a, dtype=np.float64)
    assert a.size == n, (a.size, n)
    return a


def configure(ini_text=""):
    """Process-global, like the reference."""
    if lib().nsref_configure(ini_text.encode()) != 0:
        raise RuntimeError(lib().nsref_last_error().decode())


def ini(freezing="LinearFreezing", albedo="SMUIceAlbedo", extra=""):
    return ("[Modules]\n"
            f"Nextsim::IFreezingPoint = Nextsim::{freezing}\n"
            f"Nextsim::IIceAlbedo = Nextsim::{albedo}\n\n" + extra)


class RefGrid:
    """std::vector<ElementData> driven by the DevStep::iterate body."""

    def __init__(self, n, nlayers=1):
        self.n, self.nlayers = int(n), int(nlayers)
        self.h = lib().nsref_grid_create(self.n, self.nlayers)
        if not self.h:
            raise RuntimeError(lib().nsref_last_error().decode())

    def close(self):
        if self.h:
            lib().nsref_grid_destroy(self.h)
            self.h = None

    __del__ = close

    def set_prognostic(self, hice, cice, hsnow, sst, sss, tice):
        n = self.n
        arrs = [_c(hice, n), _c(cice, n), _c(hsnow, n), _c(sst, n), _c(sss, n),
                _c(tice, n * self.nlayers)]
        lib().nsref_grid_set_prognostic(self.h, *[_p(a) for a in arrs])

    def set_forcing(self, tair, tdew, slp, qsw, qlw, mld, snowfall, wind):
        n = self.n
        arrs = [_c(a, n) for a in (tair, tdew, slp, qsw, qlw, mld, snowfall, wind)]
        lib().nsref_grid_set_forcing(self.h, *[_p(a) for a in arrs])

    def step(self, dt, nsteps=1, nthreads=1, mode=0):
        if lib().nsref_grid_step(self.h, float(dt), int(nsteps), int(nthreads), int(mode)) != 0:
            raise RuntimeError(lib().nsref_last_error().decode())

    def get_prognostic(self):
        n = self.n
        out = {k: np.empty(n if k != "tice" else n * self.nlayers) for k in PROG}
        lib().nsref_grid_get_prognostic(self.h, *[_p(out[k]) for k in PROG])
        return out

    def get_physics(self):
        names = ("rho", "sphuma", "sphumw", "sphumi", "cspec", "tau", "hi_new", "hs_new", "c_new")
        out = {k: np.empty(self.n) for k in names}
        lib().nsref_grid_get_physics(self.h, *[_p(out[k]) for k in names])
        return out

    def get_private(self):
        names = ("newice", "qia", "qio", "subl", "dqdt", "hifroms", "t_new")
        out = {k: np.empty(self.n) for k in names}
        lib().nsref_grid_get_private(self.h, *[_p(out[k]) for k in names])
        return out
