"""ctypes declaration of the C ABI in include/nsdg_gpu.h and the loader of the CUDA library.

There is no CPU fallback: if csrc/libnsdg_gpu.so is missing the loader raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libnsdg_gpu.so")

c_double_p = ctypes.POINTER(ctypes.c_double)

NSDG_OK, NSDG_EINVAL, NSDG_ECUDA, NSDG_ENOMEM, NSDG_ESTATE, NSDG_ECOMM = 0, -1, -2, -3, -4, -5
STATUS_NAMES = {0: "NSDG_OK", -1: "NSDG_EINVAL", -2: "NSDG_ECUDA", -3: "NSDG_ENOMEM",
                -4: "NSDG_ESTATE", -5: "NSDG_ECOMM"}

ALBEDO_KINDS = {"Nextsim::SMUIceAlbedo": 0, "Nextsim::SMU2IceAlbedo": 1, "Nextsim::CCSMIceAlbedo": 2}
FREEZING_KINDS = {"Nextsim::LinearFreezing": 0, "Nextsim::UnescoFreezing": 1}
OCEAN_FLUX_KINDS = {"Nextsim::BasicIceOceanHeatFlux": 0}

THERMO_FIELDS = ("hice", "cice", "hsnow", "sst", "sss", "tice0", "tair", "tdew", "slp", "qsw",
                 "qlw", "mld", "snowfall", "wind", "newice")
THERMO_DIAG = ("rho", "sphuma", "sphumw", "sphumi", "cspec", "tau", "qia", "qio", "qow", "subl",
               "dqdt", "hifroms", "t_new")
DYN_FIELDS = ("H", "A", "U", "V", "S11", "S12", "S22", "UATM", "VATM", "UOCN", "VOCN")
BC_PERIODIC, BC_BOX = 0, 1
COMM_ID_BYTES = 128


class ThermoParams(ctypes.Structure):
    """nsdg_thermo_params; defaults are the reference's (SURVEY.md section 5 key list)."""
    _fields_ = [(n, ctypes.c_double) for n in (
        "drag_ocean_q", "drag_ocean_t", "drag_ice_t", "albedo_w", "i0", "min_conc", "min_thick",
        "ks", "h0", "phi_m", "ccsm_ice_albedo", "ccsm_snow_albedo")] + [
        (n, ctypes.c_int32) for n in ("flooding", "albedo_kind", "freezing_kind",
                                      "ocean_flux_kind")]

    @classmethod
    def defaults(cls):
        return cls(1.5e-3, 0.83e-3, 1.3e-3, 0.07, 0.17, 1e-12, 0.01, 0.3096, 0.25, 0.5, 0.538,
                   0.8256, 1, 0, 0, 0)


class MevpParams(ctypes.Structure):
    _fields_ = [(n, ctypes.c_double) for n in (
        "rho_ice", "rho_atm", "rho_ocean", "c_atm", "c_ocean", "f_coriolis", "p_star", "c_star",
        "ellipse_e", "delta_min", "alpha", "beta", "h_min")] + [
        ("bc", ctypes.c_int32), ("_pad", ctypes.c_int32)]

    @classmethod
    def defaults(cls):
        return cls(900., 1.3, 1026., 1.2e-3, 5.5e-3, 1.46e-4, 27.5e3, 20., 2., 2e-9, 1500., 1500.,
                   1e-3, BC_PERIODIC, 0)


class DynConfig(ctypes.Structure):
    _fields_ = [("nx", ctypes.c_int64), ("ny_global", ctypes.c_int64), ("ny_local", ctypes.c_int64),
                ("j0", ctypes.c_int64), ("hx", ctypes.c_double), ("hy", ctypes.c_double),
                ("ncomp", ctypes.c_int32), ("rk_order", ctypes.c_int32),
                ("limiter", ctypes.c_int32), ("flags", ctypes.c_int32)]


DYN_SPLIT_MEVP = 1
DYN_SINGLE_MEVP = 2


class DynTuning(ctypes.Structure):
    """nsdg_dyn_tuning: 0 = default in every field (include/nsdg_gpu.h)."""
    _fields_ = [(n, ctypes.c_int32) for n in (
        "march_rows", "pair", "halo_overlap", "halo_peer", "pair_tile", "fused_tile",
        "transport_march", "graph", "march_ring", "transport_fused")] + [("reserved", ctypes.c_int32 * 6)]

    PAIR_TILES = ("16x16", "32x8", "32x16", "16x16m3", "16x8", "16x12")
    FUSED_TILES = ("32x8", "64x4", "32x16", "64x8", "16x16")


# every symbol include/nsdg_gpu.h declares: name -> (restype, argtypes)
_ctx = ctypes.c_void_p
_i32, _i64, _dbl = ctypes.c_int32, ctypes.c_int64, ctypes.c_double
SYMBOLS = {
    "nsdg_abi_version": (ctypes.c_int, []),
    "nsdg_create": (ctypes.c_int, [ctypes.POINTER(_ctx), ctypes.c_int]),
    "nsdg_destroy": (ctypes.c_int, [_ctx]),
    "nsdg_last_error": (ctypes.c_char_p, [_ctx]),
    "nsdg_synchronize": (ctypes.c_int, [_ctx]),
    "nsdg_launch_count": (_i64, [_ctx]),
    "nsdg_timer_start": (ctypes.c_int, [_ctx]),
    "nsdg_timer_stop": (ctypes.c_int, [_ctx, c_double_p]),
    "nsdg_profile_begin": (ctypes.c_int, [_ctx]),
    "nsdg_profile_end": (ctypes.c_int, [_ctx]),
    "nsdg_profile_query": (ctypes.c_int, [_ctx, _i32, c_double_p, ctypes.POINTER(_i64)]),
    "nsdg_profile_kernel_name": (ctypes.c_char_p, [_i32]),
    "nsdg_thermo_default_params": (None, [ctypes.POINTER(ThermoParams)]),
    "nsdg_thermo_init": (ctypes.c_int, [_ctx, _i64, _i32]),
    "nsdg_thermo_release": (ctypes.c_int, [_ctx]),
    "nsdg_thermo_set_params": (ctypes.c_int, [_ctx, ctypes.POINTER(ThermoParams)]),
    "nsdg_thermo_upload_prognostic": (ctypes.c_int, [_ctx] + [c_double_p] * 6),
    "nsdg_thermo_upload_forcing": (ctypes.c_int, [_ctx] + [c_double_p] * 8),
    "nsdg_thermo_upload_forcing_level": (ctypes.c_int, [_ctx, _i32] + [c_double_p] * 8),
    "nsdg_thermo_blend_forcing": (ctypes.c_int, [_ctx, _dbl]),
    "nsdg_thermo_set_slab_ocean": (ctypes.c_int, [_ctx, _i32]),
    "nsdg_thermo_upload_newice": (ctypes.c_int, [_ctx, c_double_p]),
    "nsdg_thermo_download_newice": (ctypes.c_int, [_ctx, c_double_p]),
    "nsdg_thermo_step": (ctypes.c_int, [_ctx, _dbl, _i32]),
    "nsdg_thermo_step_diag": (ctypes.c_int, [_ctx, _dbl, _i32]),
    "nsdg_thermo_download_prognostic": (ctypes.c_int, [_ctx] + [c_double_p] * 6),
    "nsdg_thermo_download_diagnostics": (ctypes.c_int, [_ctx, c_double_p]),
    "nsdg_thermo_field_ptr": (ctypes.c_int, [_ctx, _i32, ctypes.POINTER(ctypes.c_void_p)]),
    "nsdg_thermo_run_host": (ctypes.c_int, [_ctx, _dbl, _i32] + [c_double_p] * 14),
    "nsdg_mevp_default_params": (None, [ctypes.POINTER(MevpParams)]),
    "nsdg_dyn_init": (ctypes.c_int, [_ctx, ctypes.POINTER(DynConfig)]),
    "nsdg_dyn_release": (ctypes.c_int, [_ctx]),
    "nsdg_dyn_set_tuning": (ctypes.c_int, [_ctx, ctypes.POINTER(DynTuning)]),
    "nsdg_dyn_get_tuning": (ctypes.c_int, [_ctx, ctypes.POINTER(DynTuning)]),
    "nsdg_dyn_set_mevp_params": (ctypes.c_int, [_ctx, ctypes.POINTER(MevpParams)]),
    "nsdg_dyn_upload": (ctypes.c_int, [_ctx, _i32, c_double_p]),
    "nsdg_dyn_download": (ctypes.c_int, [_ctx, _i32, c_double_p]),
    "nsdg_dyn_upload_rows": (ctypes.c_int, [_ctx, _i32, _i64, _i64, c_double_p]),
    "nsdg_dyn_download_rows": (ctypes.c_int, [_ctx, _i32, _i64, _i64, c_double_p]),
    "nsdg_dyn_field_ncomp": (ctypes.c_int, [_ctx, _i32]),
    "nsdg_dyn_field_ptr": (ctypes.c_int, [_ctx, _i32, ctypes.POINTER(ctypes.c_void_p)]),
    "nsdg_transport_step": (ctypes.c_int, [_ctx, _i32, _dbl]),
    "nsdg_transport_step_HA": (ctypes.c_int, [_ctx, _dbl]),
    "nsdg_mevp_step": (ctypes.c_int, [_ctx, _dbl, _i32]),
    "nsdg_dyn_step": (ctypes.c_int, [_ctx, _dbl, _i32]),
    "nsdg_dyn_step_host": (ctypes.c_int, [_ctx, _dbl, _i32, ctypes.POINTER(c_double_p),
                                          ctypes.POINTER(c_double_p)]),
    "nsdg_dyn_step_host_async": (ctypes.c_int, [_ctx, _dbl, _i32, ctypes.POINTER(c_double_p),
                                                ctypes.POINTER(c_double_p)]),
    "nsdg_dyn_host_wait": (ctypes.c_int, [_ctx]),
    "nsdg_dyn_upload_forcing_level": (ctypes.c_int, [_ctx, _i32] + [c_double_p] * 4),
    "nsdg_dyn_blend_forcing": (ctypes.c_int, [_ctx, _dbl]),
    "nsdg_dyn_mass": (ctypes.c_int, [_ctx, _i32, c_double_p]),
    "nsdg_dyn_mass_noalloc": (ctypes.c_int, [_ctx, ctypes.c_void_p, _i64, _i64, _i64, _dbl, _dbl,
                                             ctypes.c_void_p]),
    "nsdg_mevp_band_rows": (ctypes.c_int, [_i64, _i64, _i32, _i32]),
    "nsdg_comm_get_unique_id": (ctypes.c_int, [ctypes.c_void_p]),
    "nsdg_comm_init": (ctypes.c_int, [_ctx, ctypes.c_void_p, _i32, _i32]),
    "nsdg_comm_rank": (ctypes.c_int, [_ctx, ctypes.POINTER(_i32), ctypes.POINTER(_i32)]),
    "nsdg_comm_bytes_sent": (_i64, [_ctx]),
    "nsdg_comm_halo_peer": (ctypes.c_int, [_ctx]),
}

_lib = None


class NsdgError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"{STATUS_NAMES.get(code, code)}: {text}")
        self.code = code


def load():
    """dlopen the CUDA library and type every entry point. Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; "
            "g.build()'` (nvcc, sm_100a). There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(ctx, code):
    if code != 0:
        msg = load().nsdg_last_error(ctx)
        raise NsdgError(code, msg.decode() if msg else "")
