/*!
 * @file GpuStep.cpp
 * See GpuStep.hpp.  Marshals only through the public cursor and accessors of the reference
 * (IStructure.hpp:96-122, PrognosticData.hpp:53-87, ExternalData.hpp:22-62,
 * PhysicsData.hpp:44-47, PrognosticGenerator.hpp) - no friend access, no layout assumptions.
 */
#include "GpuStep.hpp"

#include "include/CCSMIceAlbedo.hpp"
#include "include/Configured.hpp"
#include "include/ElementData.hpp"
#include "include/IFreezingPoint.hpp"
#include "include/IIceAlbedo.hpp"
#include "include/LinearFreezing.hpp"
#include "include/ModuleLoader.hpp"
#include "include/NextsimPhysics.hpp"
#include "include/PrognosticGenerator.hpp"
#include "include/SMU2IceAlbedo.hpp"
#include "include/SMUIceAlbedo.hpp"
#include "include/UnescoFreezing.hpp"

#include <stdexcept>

namespace Nextsim {

namespace {
    // tag type for Configured<>::getConfiguration: the keys are global, the class only names
    // the per-class option cache (core/src/include/Configured.hpp:73-79)
    struct GpuStepConfig : public Configured<GpuStepConfig> {
        void configure() override { }
    };
}
template <> const std::map<int, std::string> Configured<GpuStepConfig>::keyMap = {};

GpuStep::GpuStep(int device)
    : pStructure(nullptr)
    , device(device)
    , ctx(nullptr)
    , n(0)
    , nLayers(1)
    , dirty(false)
{
    nsdg_thermo_default_params(&params);
}

void GpuStep::setDevice(int dev)
{
    if (ctx)
        throw std::logic_error("GpuStep::setDevice after setInitialData");
    device = dev;
}

GpuStep::~GpuStep()
{
    if (ctx)
        nsdg_destroy(ctx);
}

void GpuStep::check(int rc, const char* what) const
{
    if (rc != NSDG_OK)
        throw std::runtime_error(std::string("GpuStep: ") + what + ": " + nsdg_last_error(ctx));
}

// Module selection from the same ModuleLoader state the CPU path uses
// (core/src/ModuleLoader.cpp:46-61, core/src/ConfiguredModule.cpp:19-56); scalar parameters from
// the same keys and defaults (NextsimPhysics.cpp:50-58,76-82; ThermoIce0.cpp:23-31;
// HiblerConcentration.cpp:21-29; CCSMIceAlbedo.cpp:22-23,40-41).
nsdg_thermo_params GpuStep::queryConfiguration()
{
    nsdg_thermo_params p;
    nsdg_thermo_default_params(&p);
    ModuleLoader& loader = ModuleLoader::getLoader();

    IIceAlbedo& albedo = loader.getImplementation<IIceAlbedo>();
    if (dynamic_cast<CCSMIceAlbedo*>(&albedo))
        p.albedo_kind = NSDG_ALBEDO_CCSM;
    else if (dynamic_cast<SMU2IceAlbedo*>(&albedo))
        p.albedo_kind = NSDG_ALBEDO_SMU2;
    else if (dynamic_cast<SMUIceAlbedo*>(&albedo))
        p.albedo_kind = NSDG_ALBEDO_SMU;
    else
        throw std::invalid_argument("GpuStep: no GPU implementation of the selected IIceAlbedo");

    IFreezingPoint& freezing = loader.getImplementation<IFreezingPoint>();
    if (dynamic_cast<UnescoFreezing*>(&freezing))
        p.freezing_kind = NSDG_FREEZING_UNESCO;
    else if (dynamic_cast<LinearFreezing*>(&freezing))
        p.freezing_kind = NSDG_FREEZING_LINEAR;
    else
        throw std::invalid_argument("GpuStep: no GPU implementation of the selected IFreezingPoint");

    typedef GpuStepConfig C;
    p.drag_ocean_q = C::getConfiguration(std::string("nextsim_thermo.drag_ocean_q"), 1.5e-3);
    p.drag_ocean_t = C::getConfiguration(std::string("nextsim_thermo.drag_ocean_t"), 0.83e-3);
    p.drag_ice_t = C::getConfiguration(std::string("nextsim_thermo.drag_ice_t"), 1.3e-3);
    p.albedo_w = C::getConfiguration(std::string("nextsim_thermo.albedoW"), 0.07);
    p.i0 = C::getConfiguration(std::string("nextsim_thermo.I_0"), 0.17);
    p.min_conc = C::getConfiguration(std::string("nextsim_thermo.min_conc"), 1e-12);
    p.min_thick = C::getConfiguration(std::string("nextsim_thermo.min_thick"), 0.01);
    p.ks = C::getConfiguration(std::string("thermoice0.ks"), 0.3096);
    p.flooding = C::getConfiguration(std::string("thermoice0.flooding"), true) ? 1 : 0;
    p.h0 = C::getConfiguration(std::string("Hibler.h0"), 0.25);
    p.phi_m = C::getConfiguration(std::string("Hibler.phiM"), 0.5);
    p.ccsm_ice_albedo = C::getConfiguration(std::string("CCSMIceAlbedo.iceAlbedo"), 0.538);
    p.ccsm_snow_albedo = C::getConfiguration(std::string("CCSMIceAlbedo.snowAlbedo"), 0.8256);
    return p;
}

void GpuStep::setInitialData(IStructure& dataStructure)
{
    if (!ctx && nsdg_create(&ctx, device) != NSDG_OK) {
        ctx = nullptr;
        throw std::runtime_error("GpuStep: no CUDA device (nsdg_create failed); there is no CPU "
                                 "fallback - select Nextsim::DevStep");
    }
    pStructure = &dataStructure;
    params = queryConfiguration();

    // gather AoS -> SoA in cursor order (== DevGrid's i*nx+j, core/src/DevGridIO.cpp:107-109)
    std::vector<double> hice, cice, hsnow, sst, sss, tice;
    nLayers = dataStructure.nIceLayers();
    for (pStructure->cursor = 0; pStructure->cursor; ++pStructure->cursor) {
        const ElementData& d = *pStructure->cursor;
        hice.push_back(d.iceThickness());
        cice.push_back(d.iceConcentration());
        hsnow.push_back(d.snowThickness());
        sst.push_back(d.seaSurfaceTemperature());
        sss.push_back(d.seaSurfaceSalinity());
        for (int l = 0; l < nLayers; ++l)
            tice.push_back(d.iceTemperature(l < d.nIceLayers() ? l : 0));
    }
    n = hice.size();
    // this adapter owns its context: a repeated setInitialData replaces its own state explicitly
    check(nsdg_thermo_release(ctx), "nsdg_thermo_release");
    check(nsdg_thermo_init(ctx, (int64_t)n, nLayers), "nsdg_thermo_init");
    check(nsdg_thermo_set_params(ctx, &params), "nsdg_thermo_set_params");
    check(nsdg_thermo_upload_prognostic(ctx, hice.data(), cice.data(), hsnow.data(), sst.data(),
              sss.data(), tice.data()),
        "nsdg_thermo_upload_prognostic");
    updateForcing();
    dirty = false;
}

void GpuStep::updateForcing()
{
    std::vector<double> f[8];
    for (auto& v : f)
        v.reserve(n);
    for (pStructure->cursor = 0; pStructure->cursor; ++pStructure->cursor) {
        ElementData& d = *pStructure->cursor;
        f[0].push_back(d.airTemperature());
        f[1].push_back(d.dewPoint2m());
        f[2].push_back(d.airPressure());
        f[3].push_back(d.incomingShortwave());
        f[4].push_back(d.incomingLongwave());
        f[5].push_back(d.mixedLayerDepth());
        f[6].push_back(d.snowfall());
        f[7].push_back(d.windSpeed());
    }
    check(nsdg_thermo_upload_forcing(ctx, f[0].data(), f[1].data(), f[2].data(), f[3].data(),
              f[4].data(), f[5].data(), f[6].data(), f[7].data()),
        "nsdg_thermo_upload_forcing");
}

void GpuStep::iterate(const Iterator::Duration& dt)
{
    if (!pStructure)
        throw std::logic_error("GpuStep::iterate before setInitialData");
    PrognosticData::setTimestep(dt); // keep the host-side static consistent (DevStep.cpp:16)
    check(nsdg_thermo_step(ctx, (double)dt, 1), "nsdg_thermo_step");
    dirty = true;
}

void GpuStep::sync()
{
    if (!pStructure || !dirty)
        return;
    std::vector<double> hice(n), cice(n), hsnow(n), sst(n), sss(n), tice(n * (size_t)nLayers);
    check(nsdg_thermo_download_prognostic(ctx, hice.data(), cice.data(), hsnow.data(), sst.data(),
              sss.data(), tice.data()),
        "nsdg_thermo_download_prognostic");
    size_t k = 0;
    std::vector<double> layers;
    for (pStructure->cursor = 0; pStructure->cursor; ++pStructure->cursor, ++k) {
        ElementData& d = *pStructure->cursor;
        layers.assign(tice.begin() + k * nLayers, tice.begin() + (k + 1) * nLayers);
        layers.resize(d.nIceLayers(), layers.empty() ? 0. : layers[0]);
        // PrognosticData::operator=(const PrognosticGenerator&), PrognosticData.cpp:42-54
        d = PrognosticGenerator()
                .hice(hice[k])
                .cice(cice[k])
                .hsnow(hsnow[k])
                .sst(sst[k])
                .sss(sss[k])
                .tice(layers);
    }
    dirty = false;
}

void GpuStep::stop(const Iterator::TimePoint& stopTime) { sync(); }

void GpuStep::writeRestartFile(const std::string& filePath)
{
    sync();
    if (pStructure)
        pStructure->dump(filePath);
}

} /* namespace Nextsim */
