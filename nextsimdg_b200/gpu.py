"""Python host side above the C ABI.  Mirrors the reference's step interface for the path:
`GpuStep` has the IModelStep verbs (core/src/include/IModelStep.hpp:16-34: setInitialData,
init, start, iterate, stop, writeRestartFile) over structure-of-arrays state; `GpuDynamics`
drives DG transport + mEVP.  All compute goes through csrc/libnsdg_gpu.so; nothing here
computes on the CPU and there is no fallback.
"""
import ctypes

import numpy as np

from . import _abi
from ._abi import (DYN_FIELDS, THERMO_DIAG, THERMO_FIELDS, DynConfig, DynTuning, MevpParams,
                   ThermoParams, c_double_p, check)

PROGNOSTIC = ("hice", "cice", "hsnow", "sst", "sss", "tice")
FORCING = ("tair", "tdew", "slp", "qsw", "qlw", "mld", "snowfall", "wind")


def _f64(a, n=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if n is not None and a.size != n:
        raise ValueError(f"expected {n} values, got {a.size}")
    return a


def _p(a):
    return a.ctypes.data_as(c_double_p)


class Context:
    """One nsdg_ctx: one GPU, one compute stream."""

    def __init__(self, device=0):
        self.lib = _abi.load()
        h = ctypes.c_void_p()
        rc = self.lib.nsdg_create(ctypes.byref(h), int(device))
        if rc != 0:
            raise _abi.NsdgError(rc, f"nsdg_create(device={device}) failed: no CUDA device; "
                                     "this package has no CPU path")
        self.h = h
        self.device = device
        # the one thermodynamics / dynamics state this context holds (C ABI: a second *_init on a
        # live state is NSDG_ESTATE); the Python classes hand the slot over explicitly
        self._owner = {"thermo": None, "dyn": None}

    def _take(self, kind, obj):
        prev = self._owner[kind]
        if prev is not None and prev is not obj:
            prev.close()
        self._owner[kind] = obj

    def close(self):
        if getattr(self, "h", None):
            for kind in ("thermo", "dyn"):
                if self._owner[kind] is not None:
                    self._owner[kind]._released = True
                    self._owner[kind] = None
            self.lib.nsdg_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def synchronize(self):
        check(self.h, self.lib.nsdg_synchronize(self.h))

    @property
    def launches(self):
        return int(self.lib.nsdg_launch_count(self.h))

    def timer_start(self):
        check(self.h, self.lib.nsdg_timer_start(self.h))

    def timer_stop(self):
        ms = ctypes.c_double()
        check(self.h, self.lib.nsdg_timer_stop(self.h, ctypes.byref(ms)))
        return ms.value

    def profile_begin(self):
        check(self.h, self.lib.nsdg_profile_begin(self.h))

    def profile_end(self):
        """-> {kernel name: (total_ms, launches)} for the launches since profile_begin"""
        check(self.h, self.lib.nsdg_profile_end(self.h))
        out, k = {}, 0
        while True:
            name = self.lib.nsdg_profile_kernel_name(k)
            if not name:
                return out
            ms, cnt = ctypes.c_double(), ctypes.c_int64()
            self.lib.nsdg_profile_query(self.h, k, ctypes.byref(ms), ctypes.byref(cnt))
            if cnt.value:
                out[name.decode()] = (ms.value, cnt.value)
            k += 1

    # strip decomposition -----------------------------------------------------------------
    def comm_init(self, rank, nranks, unique_id=None):
        buf = (ctypes.c_char * _abi.COMM_ID_BYTES).from_buffer_copy(
            unique_id if unique_id is not None else bytes(_abi.COMM_ID_BYTES))
        check(self.h, self.lib.nsdg_comm_init(self.h, ctypes.cast(buf, ctypes.c_void_p), rank,
                                              nranks))

    @staticmethod
    def comm_unique_id():
        buf = (ctypes.c_char * _abi.COMM_ID_BYTES)()
        rc = _abi.load().nsdg_comm_get_unique_id(ctypes.cast(buf, ctypes.c_void_p))
        if rc != 0:
            raise _abi.NsdgError(rc, "ncclGetUniqueId failed")
        return bytes(buf)

    @property
    def comm_bytes_sent(self):
        return int(self.lib.nsdg_comm_bytes_sent(self.h))

    @property
    def halo_peer(self):
        """True when the sub-cycle's halo goes through NVLink peer stores (no NCCL in the loop)."""
        return bool(self.lib.nsdg_comm_halo_peer(self.h))


class _State:
    """A state object of a context.  Creating another one of the same kind on the context
    releases this one (explicitly, through nsdg_*_release); a released object raises on use."""
    _kind = None
    _released = False

    @property
    def h(self):
        if self._released:
            raise RuntimeError(f"{type(self).__name__}: this state was released (a newer "
                               f"{type(self).__name__} took over its context, or close() was called)")
        return self._h

    def close(self):
        if self._released:
            return
        self._released = True
        ctx = self.ctx
        if ctx._owner.get(self._kind) is self:
            ctx._owner[self._kind] = None
        if getattr(ctx, "h", None):
            fn = ctx.lib.nsdg_thermo_release if self._kind == "thermo" else ctx.lib.nsdg_dyn_release
            check(ctx.h, fn(ctx.h))


class GpuStep(_State):
    """Drop-in for the reference's DevStep (core/src/DevStep.cpp:14-23) over SoA arrays.

    setInitialData gathers/uploads, iterate(dt) launches the fused kernel on device-resident
    state, stop()/get_prognostic() scatter back.  Cell order = the reference's cursor order
    (DevGrid: index i*nx + j, core/src/DevGridIO.cpp:107-109); tice is [cell][layer].
    """

    _kind = "thermo"

    def __init__(self, ctx, ncells, nlayers=1, params=None):
        self.ctx, self.lib, self._h = ctx, ctx.lib, ctx.h
        self.n, self.nlayers = int(ncells), int(nlayers)
        ctx._take("thermo", None)
        check(self.h, self.lib.nsdg_thermo_init(self.h, self.n, self.nlayers))
        ctx._take("thermo", self)
        self.set_params(params if params is not None else ThermoParams.defaults())

    def set_params(self, params):
        self.params = params
        check(self.h, self.lib.nsdg_thermo_set_params(self.h, ctypes.byref(params)))

    # IModelStep verbs ----------------------------------------------------------------------
    def setInitialData(self, prognostic, forcing):
        n = self.n
        pr = [_f64(prognostic[k], n * (self.nlayers if k == "tice" else 1)) for k in PROGNOSTIC]
        check(self.h, self.lib.nsdg_thermo_upload_prognostic(self.h, *[_p(a) for a in pr]))
        self.setForcing(forcing)

    def setForcing(self, forcing):
        fo = [_f64(forcing[k], self.n) for k in FORCING]
        check(self.h, self.lib.nsdg_thermo_upload_forcing(self.h, *[_p(a) for a in fo]))

    def setForcingLevels(self, forcing0, forcing1):
        """Two time levels of the forcing stay on the device (nsdg_thermo_upload_forcing_level)."""
        for lvl, fo in enumerate((forcing0, forcing1)):
            a = [_f64(fo[k], self.n) for k in FORCING]
            check(self.h, self.lib.nsdg_thermo_upload_forcing_level(self.h, lvl, *[_p(x) for x in a]))

    def blendForcing(self, w):
        """active forcing = f0 + w (f1 - f0), interpolated on the device"""
        check(self.h, self.lib.nsdg_thermo_blend_forcing(self.h, float(w)))

    def setSlabOcean(self, enable=True):
        check(self.h, self.lib.nsdg_thermo_set_slab_ocean(self.h, int(bool(enable))))

    def init(self):
        pass

    def start(self, start_time=0):
        pass

    def iterate(self, dt, nsteps=1, diagnostics=False):
        fn = self.lib.nsdg_thermo_step_diag if diagnostics else self.lib.nsdg_thermo_step
        check(self.h, fn(self.h, float(dt), int(nsteps)))

    def stop(self, stop_time=0):
        self.ctx.synchronize()

    def writeRestartFile(self, path, nx=None):
        """SoA restart (nextsimdg_b200/restart.py; the reference's structure/data layout of
        core/src/DevGridIO.cpp:160-201) straight from the device mirror, plus the hidden newice
        state: netCDF-4 / HDF5 (restart.write_netcdf) unless the name ends in .nsdgr (NSDGR1
        container).  nx: edge of the square DevGrid (default: sqrt(ncells), else a 1 x n grid)."""
        from . import restart
        d = self.get_prognostic()
        d["newice"] = self.get_newice()
        if nx is None:
            r = int(round(self.n ** 0.5))
            nx = r if r * r == self.n else 1
        writer = restart.write_restart if str(path).endswith(".nsdgr") else restart.write_netcdf
        writer(path, d, nx, self.n // nx, self.nlayers)

    @classmethod
    def from_restart(cls, ctx, path, forcing, params=None):
        """Model::configure's restart read (core/src/Model.cpp:62-73) without ElementData."""
        from . import restart
        with open(path, "rb") as f:
            is_container = f.read(len(restart.MAGIC)) == restart.MAGIC
        st, meta = restart.read_restart(path) if is_container else restart.read_netcdf(path)
        step = cls(ctx, meta["nx"] * meta["ny"], meta["nlayers"], params)
        step.setInitialData(st, forcing)
        if "newice" in st:
            step.set_newice(st["newice"])
        return step

    # accessors -----------------------------------------------------------------------------
    def get_prognostic(self):
        n = self.n
        out = {k: np.empty(n * (self.nlayers if k == "tice" else 1)) for k in PROGNOSTIC}
        check(self.h, self.lib.nsdg_thermo_download_prognostic(
            self.h, *[_p(out[k]) for k in PROGNOSTIC]))
        return out

    def get_newice(self):
        a = np.empty(self.n)
        check(self.h, self.lib.nsdg_thermo_download_newice(self.h, _p(a)))
        return a

    def set_newice(self, a):
        a = _f64(a, self.n)
        check(self.h, self.lib.nsdg_thermo_upload_newice(self.h, _p(a)))

    def get_diagnostics(self):
        d = np.empty((len(THERMO_DIAG), self.n))
        check(self.h, self.lib.nsdg_thermo_download_diagnostics(self.h, _p(d)))
        return {k: d[q] for q, k in enumerate(THERMO_DIAG)}

    def field_ptr(self, name):
        p = ctypes.c_void_p()
        check(self.h, self.lib.nsdg_thermo_field_ptr(self.h, THERMO_FIELDS.index(name),
                                                     ctypes.byref(p)))
        return p.value

    def run_host(self, dt, nsteps, prognostic, forcing):
        """One C-ABI call with host buffers: upload, nsteps sweeps, download (in place)."""
        n = self.n
        pr = {k: _f64(prognostic[k], n * (self.nlayers if k == "tice" else 1)) for k in PROGNOSTIC}
        fo = [_f64(forcing[k], n) for k in FORCING]
        check(self.h, self.lib.nsdg_thermo_run_host(
            self.h, float(dt), int(nsteps), _p(pr["hice"]), _p(pr["cice"]), _p(pr["hsnow"]),
            _p(pr["sst"]), _p(pr["sss"]), _p(pr["tice"]), *[_p(a) for a in fo]))
        return pr


class GpuDynamics(_State):
    """DG transport (dG0/1/2) + mEVP on one strip of a structured periodic grid."""

    _kind = "dyn"

    def __init__(self, ctx, nx, ny_global, hx, hy, ncomp=6, ny_local=None, j0=0, rk_order=0,
                 limiter=0, mevp=None, flags=0, tuning=None):
        self.ctx, self.lib, self._h = ctx, ctx.lib, ctx.h
        self.nx, self.ny_global = int(nx), int(ny_global)
        self.ny = int(ny_local if ny_local is not None else ny_global)
        self.ncomp = int(ncomp)
        self.cfg = DynConfig(self.nx, self.ny_global, self.ny, int(j0), float(hx), float(hy),
                             self.ncomp, int(rk_order), int(limiter), int(flags))
        ctx._take("dyn", None)
        check(self.h, self.lib.nsdg_dyn_init(self.h, ctypes.byref(self.cfg)))
        ctx._take("dyn", self)
        self.set_mevp_params(mevp if mevp is not None else MevpParams.defaults())
        if tuning:
            self.set_tuning(**tuning)

    def set_tuning(self, **knobs):
        """nsdg_dyn_set_tuning on top of what is in effect: march_rows (0 auto, -1 tiled kernel),
        pair, halo_overlap, halo_peer, transport_march, graph (-1 off / 0 default / 1 on),
        pair_tile / fused_tile (id or "WxH")."""
        t = self.get_tuning()
        for k, v in knobs.items():
            if k == "pair_tile" and isinstance(v, str):
                v = DynTuning.PAIR_TILES.index(v)
            if k == "fused_tile" and isinstance(v, str):
                v = DynTuning.FUSED_TILES.index(v)
            if k not in dict(DynTuning._fields_) or k == "reserved":
                raise KeyError(k)
            setattr(t, k, int(v))
        check(self.h, self.lib.nsdg_dyn_set_tuning(self.h, ctypes.byref(t)))

    def get_tuning(self):
        t = DynTuning()
        check(self.h, self.lib.nsdg_dyn_get_tuning(self.h, ctypes.byref(t)))
        return t

    def set_mevp_params(self, p):
        self.mevp = p
        check(self.h, self.lib.nsdg_dyn_set_mevp_params(self.h, ctypes.byref(p)))

    def field_shape(self, name):
        nc = self.lib.nsdg_dyn_field_ncomp(self.h, DYN_FIELDS.index(name))
        return (nc, self.ny, self.nx) if nc > 1 or name in ("H", "A") else (self.ny, self.nx)

    def upload(self, name, host):
        shape = self.field_shape(name)
        a = _f64(host, int(np.prod(shape)))
        check(self.h, self.lib.nsdg_dyn_upload(self.h, DYN_FIELDS.index(name), _p(a)))

    def download(self, name):
        out = np.empty(self.field_shape(name))
        check(self.h, self.lib.nsdg_dyn_download(self.h, DYN_FIELDS.index(name), _p(out)))
        return out

    def upload_rows(self, name, row0, host):
        """rows [row0, row0 + host.shape[-2]) of the strip; host is [ncomp][nrows][nx]"""
        nc = self.field_shape(name)[0] if len(self.field_shape(name)) == 3 else 1
        a = _f64(host)
        nrows = a.size // (nc * self.nx)
        if a.size != nc * nrows * self.nx:
            raise ValueError(f"{name}: {a.size} values are not whole rows of {nc} x {self.nx}")
        check(self.h, self.lib.nsdg_dyn_upload_rows(self.h, DYN_FIELDS.index(name), int(row0), nrows, _p(a)))

    def download_rows(self, name, row0, nrows):
        shp = self.field_shape(name)
        out = np.empty(shp[:-2] + (int(nrows), self.nx))
        check(self.h, self.lib.nsdg_dyn_download_rows(self.h, DYN_FIELDS.index(name), int(row0), int(nrows), _p(out)))
        return out

    def upload_all(self, fields):
        for k, v in fields.items():
            self.upload(k, v)

    def transport_step(self, dt, field=None):
        if field is None:
            check(self.h, self.lib.nsdg_transport_step_HA(self.h, float(dt)))
        else:
            check(self.h, self.lib.nsdg_transport_step(self.h, DYN_FIELDS.index(field), float(dt)))

    def mevp_step(self, dt, nsub=100):
        check(self.h, self.lib.nsdg_mevp_step(self.h, float(dt), int(nsub)))

    def step(self, dt, nsub=100):
        check(self.h, self.lib.nsdg_dyn_step(self.h, float(dt), int(nsub)))

    def set_forcing_levels(self, f0, f1):
        for lvl, fo in enumerate((f0, f1)):
            a = [_f64(fo[k], self.nx * self.ny) for k in ("UATM", "VATM", "UOCN", "VOCN")]
            check(self.h, self.lib.nsdg_dyn_upload_forcing_level(self.h, lvl, *[_p(x) for x in a]))

    def blend_forcing(self, w):
        check(self.h, self.lib.nsdg_dyn_blend_forcing(self.h, float(w)))

    def step_host(self, dt, nsub, inputs=None, outputs=None, deferred=False):
        """One step with host buffers (nsdg_dyn_step_host): `inputs` maps field names to arrays
        (or raw addresses of pinned buffers) uploaded for this step, `outputs` maps field names
        to writable buffers; copies overlap the compute.  Missing names stay resident / are
        not downloaded.  deferred=True (nsdg_dyn_step_host_async) returns once the step is
        enqueued: its downloads overlap the next step; the buffers are valid after host_wait()."""
        def table(m, writable):
            t = (c_double_p * len(DYN_FIELDS))()
            keep = []
            for k, a in (m or {}).items():
                if isinstance(a, int):
                    t[DYN_FIELDS.index(k)] = ctypes.cast(a, c_double_p)
                    continue
                if a.dtype != np.float64 or not a.flags.c_contiguous or a.size != int(np.prod(self.field_shape(k))):
                    raise ValueError(f"{k}: need a contiguous float64 array of shape {self.field_shape(k)}")
                if writable and not a.flags.writeable:
                    raise ValueError(f"{k}: output buffer is read-only")
                keep.append(a)
                t[DYN_FIELDS.index(k)] = _p(a)
            return t, keep
        tin, k1 = table(inputs, False)
        tout, k2 = table(outputs, True)
        fn = self.lib.nsdg_dyn_step_host_async if deferred else self.lib.nsdg_dyn_step_host
        if deferred:  # the library owns the buffers until host_wait
            self._pending = getattr(self, "_pending", []) + [k1, k2]
        check(self.h, fn(self.h, float(dt), int(nsub), tin if inputs else None,
                         tout if outputs else None))

    def host_wait(self):
        check(self.h, self.lib.nsdg_dyn_host_wait(self.h))
        self._pending = []

    def mass(self, name):
        m = ctypes.c_double()
        check(self.h, self.lib.nsdg_dyn_mass(self.h, DYN_FIELDS.index(name), ctypes.byref(m)))
        return m.value
