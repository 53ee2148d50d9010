// TEST INFRASTRUCTURE (oracle). Not part of the product library; only tests/,
// __graft_entry__.smoke() and bench.py's CPU-baseline legs may load the library built
// from this file.
//
// C-callable harness around the reference's OWN column-physics sources, which are
// compiled in place from /root/reference by oracle/Makefile (nothing is copied).  The
// harness replays the body of DevStep::iterate (core/src/DevStep.cpp:14-23) over a
// std::vector<ElementData> (the storage DevGrid uses, core/src/modules/include/DevGrid.hpp:66)
// and gathers/scatters structure-of-arrays buffers the way DevGridIO does at the restart
// boundary (core/src/DevGridIO.cpp:101-138,160-201).
#include "include/ConfiguredModule.hpp"
#include "include/Configurator.hpp"
#include "include/ElementData.hpp"
#include "include/ModuleLoader.hpp"
#include "include/NextsimPhysics.hpp"
#include "include/PrognosticGenerator.hpp"

#include <algorithm>
#include <cstring>
#include <memory>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

using namespace Nextsim;

namespace {
struct RefGrid {
    int n = 0;
    int nLayers = 1;
    std::vector<ElementData> cells;
    // Side instances used in "diagnostic" mode: same call sequence as the reference's own
    // tests (physics/test/NextsimPhysics_test.cpp:225-229), which lets private per-cell
    // state (newIce, QIceAtmosphere, ...) be read through the public getters.
    std::vector<NextsimPhysics> side;
};
std::string g_error;
}

extern "C" {

const char* nsref_last_error() { return g_error.c_str(); }

// Configure the process-global module selection and parameters from INI text, exactly as
// main() does (core/src/main.cpp:22-27) followed by ElementData::configure
// (core/src/ElementData.cpp:69-73). NB Hibler.h0 is latched at the first freeze() call
// (physics/src/modules/HiblerConcentration.cpp:36): configure before stepping.
int nsref_configure(const char* iniText)
{
    try {
        Configurator::clear();
        std::unique_ptr<std::istream> ps(new std::stringstream(iniText ? iniText : ""));
        Configurator::addStream(std::move(ps));
        ModuleLoader::getLoader().setAllDefaults();
        ConfiguredModule::parseConfigurator();
        ElementData probe;
        probe.configure();
        return 0;
    } catch (std::exception& e) {
        g_error = e.what();
        return -1;
    }
}

void* nsref_grid_create(int n, int nLayers)
{
    try {
        auto* g = new RefGrid;
        g->n = n;
        g->nLayers = nLayers;
        g->cells.reserve(n);
        // DevGrid builds its elements with the default (1-layer) PhysicsData
        // (core/src/modules/include/DevGrid.hpp:49); nLayers only sizes the prognostic tice.
        for (int k = 0; k < n; ++k)
            g->cells.emplace_back(1);
        g->side.resize(n);
        return g;
    } catch (std::exception& e) {
        g_error = e.what();
        return nullptr;
    }
}

void nsref_grid_destroy(void* h) { delete static_cast<RefGrid*>(h); }

// tice is layer-fastest: tice[k*nLayers + l] (core/src/DevGridIO.cpp:193-199)
int nsref_grid_set_prognostic(void* h, const double* hice, const double* cice, const double* hsnow,
    const double* sst, const double* sss, const double* tice)
{
    auto* g = static_cast<RefGrid*>(h);
    std::vector<double> layers(g->nLayers);
    for (int k = 0; k < g->n; ++k) {
        for (int l = 0; l < g->nLayers; ++l)
            layers[l] = tice[(size_t)k * g->nLayers + l];
        g->cells[k] = PrognosticGenerator()
                          .hice(hice[k])
                          .cice(cice[k])
                          .hsnow(hsnow[k])
                          .sst(sst[k])
                          .sss(sss[k])
                          .tice(layers);
    }
    return 0;
}

int nsref_grid_set_forcing(void* h, const double* tair, const double* tdew, const double* slp,
    const double* qsw, const double* qlw, const double* mld, const double* snowfall,
    const double* wind)
{
    auto* g = static_cast<RefGrid*>(h);
    for (int k = 0; k < g->n; ++k) {
        ElementData& d = g->cells[k];
        d.airTemperature() = tair[k];
        d.dewPoint2m() = tdew[k];
        d.airPressure() = slp[k];
        d.ExternalData::mixingRatio() = -1.;
        d.incomingShortwave() = qsw[k];
        d.incomingLongwave() = qlw[k];
        d.mixedLayerDepth() = mld[k];
        d.snowfall() = snowfall[k];
        d.windSpeed() = wind[k];
    }
    return 0;
}

// mode 0: the DevStep::iterate body through the ElementData facade (faithful; timed baseline)
// mode 1: same arithmetic through side NextsimPhysics instances so diagnostics are readable
int nsref_grid_step(void* h, double dt, int nsteps, int nthreads, int mode)
{
    auto* g = static_cast<RefGrid*>(h);
    try {
        PrognosticData::setTimestep(dt);
        if (g->n == 0)
            return 0;
        // first freeze() call latches a function-local static: do it single-threaded once
        nthreads = std::max(1, std::min(nthreads, g->n));
        auto work = [g, mode, nsteps](int lo, int hi) {
            for (int s = 0; s < nsteps; ++s) {
                for (int k = lo; k < hi; ++k) {
                    ElementData& data = g->cells[k];
                    if (mode == 0) {
                        data.updateDerivedData(data, data, data);
                        data.calculate(data, data, data);
                    } else {
                        NextsimPhysics& p = g->side[k];
                        p.updateDerivedData(data, data, data);
                        p.calculate(data, data, data);
                    }
                    data.updateAndIntegrate(data);
                }
            }
        };
        if (nthreads == 1) {
            work(0, g->n);
        } else {
            std::vector<std::thread> pool;
            size_t chunk = ((size_t)g->n + nthreads - 1) / nthreads;
            for (int t = 0; t < nthreads; ++t) {
                int lo = (int)std::min<size_t>(g->n, t * chunk);
                int hi = (int)std::min<size_t>(g->n, (t + 1) * chunk);
                if (lo < hi)
                    pool.emplace_back(work, lo, hi);
            }
            for (auto& t : pool)
                t.join();
        }
        return 0;
    } catch (std::exception& e) {
        g_error = e.what();
        return -1;
    }
}

int nsref_grid_get_prognostic(void* h, double* hice, double* cice, double* hsnow, double* sst,
    double* sss, double* tice)
{
    auto* g = static_cast<RefGrid*>(h);
    for (int k = 0; k < g->n; ++k) {
        const ElementData& d = g->cells[k];
        hice[k] = d.iceThickness();
        cice[k] = d.iceConcentration();
        hsnow[k] = d.snowThickness();
        sst[k] = d.seaSurfaceTemperature();
        sss[k] = d.seaSurfaceSalinity();
        for (int l = 0; l < g->nLayers; ++l)
            tice[(size_t)k * g->nLayers + l] = d.iceTemperature(l);
    }
    return 0;
}

// PhysicsData derived/updated quantities after the last step: 9 arrays of length n
int nsref_grid_get_physics(void* h, double* rho, double* sphuma, double* sphumw, double* sphumi,
    double* cspec, double* tau, double* hiNew, double* hsNew, double* cNew)
{
    auto* g = static_cast<RefGrid*>(h);
    for (int k = 0; k < g->n; ++k) {
        ElementData& d = g->cells[k];
        rho[k] = d.airDensity();
        sphuma[k] = d.specificHumidityAir();
        sphumw[k] = d.specificHumidityWater();
        sphumi[k] = d.specificHumidityIce();
        cspec[k] = d.heatCapacityWetAir();
        tau[k] = d.dragPressure();
        hiNew[k] = d.updatedIceTrueThickness();
        hsNew[k] = d.updatedSnowTrueThickness();
        cNew[k] = static_cast<PhysicsData&>(d).updatedIceConcentration();
    }
    return 0;
}

// Per-cell private NextsimPhysics state (valid after mode-1 steps): 7 arrays of length n
int nsref_grid_get_private(void* h, double* newice, double* qia, double* qio, double* subl,
    double* dqdt, double* hifroms, double* tnew)
{
    auto* g = static_cast<RefGrid*>(h);
    for (int k = 0; k < g->n; ++k) {
        const NextsimPhysics& p = g->side[k];
        newice[k] = p.newIce();
        qia[k] = p.QIceAtmosphere();
        qio[k] = p.QIceOceanHeat();
        subl[k] = p.sublimationRate();
        dqdt[k] = p.QDerivativeWRTTemperature();
        hifroms[k] = p.totalIceFromSnow();
        tnew[k] = g->cells[k].updatedIceSurfaceTemperature();
    }
    return 0;
}

size_t nsref_sizeof_elementdata() { return sizeof(ElementData); }

} // extern "C"
