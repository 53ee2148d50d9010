"""TEST INFRASTRUCTURE: ctypes binding of oracle/_ref/libnsdg_oracle.so (plain-C restatement:
thermo_port.c, dg_transport.c, mevp.c).  Only tests/, __graft_entry__.smoke() and the
CPU-baseline legs of bench.py may import this module; the product package never does.
"""
import ctypes
import os

import numpy as np

from nextsimdg_b200._abi import MevpParams, ThermoParams, THERMO_DIAG

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libnsdg_oracle.so")
_dp = ctypes.POINTER(ctypes.c_double)
_lib = None


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(LIB_PATH)
        L.nsdg_oracle_thermo_step.argtypes = (
            [ctypes.POINTER(ThermoParams), ctypes.c_double, ctypes.c_int, ctypes.c_long]
            + [_dp] * 15 + [_dp, ctypes.c_int])
        L.nsdg_oracle_thermo_step.restype = None
        L.nsdg_oracle_transport_rhs.argtypes = [ctypes.c_long, ctypes.c_long, ctypes.c_double,
                                                ctypes.c_double, ctypes.c_int] + [_dp] * 4
        L.nsdg_oracle_transport_rhs.restype = None
        L.nsdg_oracle_transport_step.argtypes = [
            ctypes.c_long, ctypes.c_long, ctypes.c_double, ctypes.c_double, ctypes.c_int,
            ctypes.c_int, ctypes.c_double, _dp, _dp, _dp, ctypes.c_int, ctypes.c_double,
            ctypes.c_double]
        L.nsdg_oracle_transport_step.restype = None
        L.nsdg_oracle_limit.argtypes = [ctypes.c_long, ctypes.c_int, _dp, ctypes.c_double,
                                        ctypes.c_double]
        L.nsdg_oracle_limit.restype = None
        L.nsdg_oracle_mevp_step.argtypes = [
            ctypes.POINTER(MevpParams), ctypes.c_long, ctypes.c_long, ctypes.c_double,
            ctypes.c_double, ctypes.c_int, ctypes.c_double, ctypes.c_int] + [_dp] * 12
        L.nsdg_oracle_mevp_step.restype = None
        L.nsdg_oracle_mevp_stress.argtypes = [
            ctypes.POINTER(MevpParams), ctypes.c_long, ctypes.c_long, ctypes.c_double,
            ctypes.c_double] + [_dp] * 6
        L.nsdg_oracle_mevp_stress.restype = None
        L.nsdg_oracle_mevp_pressure.argtypes = [
            ctypes.POINTER(MevpParams), ctypes.c_long, ctypes.c_long, ctypes.c_int] + [_dp] * 3
        L.nsdg_oracle_mevp_pressure.restype = None
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(_dp) if a is not None else None


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


# ---------------------------------------------------------------- thermodynamics
def thermo_step(params, dt, nsteps, state, forcing, newice=None, diag=False, nthreads=1):
    """state: dict hice,cice,hsnow,sst,sss,tice0; forcing: dict tair..wind (arrays of n).
    Returns (new_state, newice, diag_dict|None). Inputs are not modified."""
    s = {k: _f64(state[k]).copy() for k in ("hice", "cice", "hsnow", "sst", "sss", "tice0")}
    f = {k: _f64(forcing[k]) for k in ("tair", "tdew", "slp", "qsw", "qlw", "mld", "snowfall",
                                       "wind")}
    n = s["hice"].size
    ni = np.zeros(n) if newice is None else _f64(newice).copy()
    d = np.empty((len(THERMO_DIAG), n)) if diag else None
    lib().nsdg_oracle_thermo_step(
        ctypes.byref(params), float(dt), int(nsteps), n, _p(s["hice"]), _p(s["cice"]),
        _p(s["hsnow"]), _p(s["sst"]), _p(s["sss"]), _p(s["tice0"]), _p(f["tair"]), _p(f["tdew"]),
        _p(f["slp"]), _p(f["qsw"]), _p(f["qlw"]), _p(f["mld"]), _p(f["snowfall"]), _p(f["wind"]),
        _p(ni), _p(d), int(nthreads))
    dd = {k: d[q] for q, k in enumerate(THERMO_DIAG)} if diag else None
    return s, ni, dd


# ---------------------------------------------------------------- transport
def transport_rhs(phi, u, v, hx, hy):
    phi = _f64(phi)
    ncomp, ny, nx = phi.shape
    out = np.empty_like(phi)
    lib().nsdg_oracle_transport_rhs(nx, ny, hx, hy, ncomp, _p(phi), _p(_f64(u)), _p(_f64(v)),
                                    _p(out))
    return out


def transport_step(phi, u, v, hx, hy, dt, rk=0, limiter=0, lo=0.0, hi=1e300):
    """phi [ncomp][ny][nx] -> advanced copy."""
    phi = _f64(phi).copy()
    ncomp, ny, nx = phi.shape
    lib().nsdg_oracle_transport_step(nx, ny, hx, hy, ncomp, rk, dt, _p(phi), _p(_f64(u)),
                                     _p(_f64(v)), limiter, lo, hi)
    return phi


# ---------------------------------------------------------------- mEVP
def mevp_step(params, H, A, u, v, S11, S12, S22, uatm, vatm, uocn, vocn, hx, hy, dt, nsub,
              want_resid=False):
    """Returns advanced copies (u, v, S11, S12, S22[, resid])."""
    H, A = _f64(H), _f64(A)
    ncomp, ny, nx = H.shape
    u, v = _f64(u).copy(), _f64(v).copy()
    S11, S12, S22 = _f64(S11).copy(), _f64(S12).copy(), _f64(S22).copy()
    resid = np.zeros(nsub) if want_resid else None
    lib().nsdg_oracle_mevp_step(ctypes.byref(params), nx, ny, hx, hy, ncomp, dt, nsub, _p(H),
                                _p(A), _p(u), _p(v), _p(S11), _p(S12), _p(S22), _p(_f64(uatm)),
                                _p(_f64(vatm)), _p(_f64(uocn)), _p(_f64(vocn)), _p(resid))
    if want_resid:
        return u, v, S11, S12, S22, resid
    return u, v, S11, S12, S22


def mevp_stress(params, u, v, pg, S11, S12, S22, hx, hy):
    ny, nx = u.shape
    S11, S12, S22 = _f64(S11).copy(), _f64(S12).copy(), _f64(S22).copy()
    lib().nsdg_oracle_mevp_stress(ctypes.byref(params), nx, ny, hx, hy, _p(_f64(u)), _p(_f64(v)),
                                  _p(_f64(pg)), _p(S11), _p(S12), _p(S22))
    return S11, S12, S22


def mevp_pressure(params, H, A):
    H, A = _f64(H), _f64(A)
    ncomp, ny, nx = H.shape
    pg = np.empty((4, ny, nx))
    lib().nsdg_oracle_mevp_pressure(ctypes.byref(params), nx, ny, ncomp, _p(H), _p(A), _p(pg))
    return pg


def dyn_step(params, H, A, u, v, S11, S12, S22, uatm, vatm, uocn, vocn, hx, hy, dt, nsub, rk=0):
    """North-star unit: transport of H and A with the current velocity, then the sub-cycle."""
    H2 = transport_step(H, u, v, hx, hy, dt, rk)
    A2 = transport_step(A, u, v, hx, hy, dt, rk)
    u2, v2, s11, s12, s22 = mevp_step(params, H2, A2, u, v, S11, S12, S22, uatm, vatm, uocn, vocn,
                                      hx, hy, dt, nsub)
    return H2, A2, u2, v2, s11, s12, s22
