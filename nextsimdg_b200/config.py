"""Reference-compatible configuration: the same INI keys and [Modules] strings the reference
reads (core/src/ConfiguredModule.cpp:19-56, physics/src/modules/
NextsimPhysics.cpp:50-58,76-82, ThermoIce0.cpp:23-31, HiblerConcentration.cpp:21-29,
CCSMIceAlbedo.cpp:22-23,40-41) mapped onto the POD parameter block of the C ABI."""
import configparser

from ._abi import ALBEDO_KINDS, FREEZING_KINDS, OCEAN_FLUX_KINDS, ThermoParams

# module interface -> (known implementations, params field)
_MODULES = {
    "Nextsim::IIceAlbedo": (ALBEDO_KINDS, "albedo_kind"),
    "Nextsim::IFreezingPoint": (FREEZING_KINDS, "freezing_kind"),
    "Nextsim::IIceOceanHeatFlux": (OCEAN_FLUX_KINDS, "ocean_flux_kind"),
    # single-implementation interfaces: accepted, nothing to select
    "Nextsim::IConcentrationModel": ({"Nextsim::HiblerConcentration": 0}, None),
    "Nextsim::IThermodynamics": ({"Nextsim::ThermoIce0": 0}, None),
    "Nextsim::IPhysics1d": ({"Nextsim::NextsimPhysics": 0}, None),
    "Nextsim::IStructure": ({"Nextsim::DevGrid": 0}, None),
}

_KEYS = {
    ("nextsim_thermo", "drag_ocean_q"): "drag_ocean_q",
    ("nextsim_thermo", "drag_ocean_t"): "drag_ocean_t",
    ("nextsim_thermo", "drag_ice_t"): "drag_ice_t",
    ("nextsim_thermo", "albedow"): "albedo_w",
    ("nextsim_thermo", "i_0"): "i0",
    ("nextsim_thermo", "min_conc"): "min_conc",
    ("nextsim_thermo", "min_thick"): "min_thick",
    ("thermoice0", "ks"): "ks",
    ("hibler", "h0"): "h0",
    ("hibler", "phim"): "phi_m",
    ("ccsmicealbedo", "icealbedo"): "ccsm_ice_albedo",
    ("ccsmicealbedo", "snowalbedo"): "ccsm_snow_albedo",
}


def _as_bool(text):
    t = text.strip().lower()
    if t in ("1", "true", "yes", "on"):
        return 1
    if t in ("0", "false", "no", "off"):
        return 0
    raise ValueError(f"not a boolean: {text!r}")


def thermo_params_from_ini(text=""):
    """Parse reference-style INI text.  Unknown implementation strings raise ValueError, as the
    reference raises std::domain_error (ConfiguredModule.cpp:47-53); unknown keys are ignored
    (allow_unregistered).  Missing keys keep the reference defaults."""
    p = ThermoParams.defaults()
    cp = configparser.ConfigParser(delimiters=("=",), comment_prefixes=("#",), strict=False,
                                   interpolation=None)
    cp.optionxform = str  # module names are case sensitive
    cp.read_string(text or "")
    for section in cp.sections():
        for key, val in cp.items(section):
            if section == "Modules":
                if key not in _MODULES:
                    continue  # not a module this path knows about: the reference ignores it too
                impls, field = _MODULES[key]
                if val not in impls:
                    raise ValueError(f'Invalid implementation "{val}" of module {key}.')
                if field:
                    setattr(p, field, impls[val])
                continue
            sk = (section.lower(), key.lower())
            if sk == ("thermoice0", "flooding"):
                p.flooding = _as_bool(val)
            elif sk in _KEYS:
                setattr(p, _KEYS[sk], float(val))
    return p
