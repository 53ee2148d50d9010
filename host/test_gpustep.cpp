// Drop-in test: the reference's own DevStep (CPU) and GpuStep (B200) are driven through the
// reference's own Iterator / IModelStep interface on identical std::vector<ElementData>
// structures; the scattered-back ElementData must agree to 1e-10 relative.
// Built by host/Makefile against the reference's headers and sources (compiled in place,
// nothing copied); the binary travels to the GPU box (tests/test_gpu_host_cpp.py runs it).
#include "GpuStep.hpp"

#include "include/ConfiguredModule.hpp"
#include "include/Configurator.hpp"
#include "include/DevStep.hpp"
#include "include/ElementData.hpp"
#include "include/Iterator.hpp"
#include "include/ModuleLoader.hpp"
#include "include/PrognosticGenerator.hpp"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <sstream>
#include <vector>

using namespace Nextsim;

// test-side IStructure over a plain vector (DevGrid itself needs netCDF for init/dump)
class VectorGrid : public IStructure {
public:
    VectorGrid(size_t n, int layers)
        : nl(layers)
        , pos(0)
    {
        data.reserve(n);
        for (size_t k = 0; k < n; ++k)
            data.emplace_back(1); // as DevGrid: PhysicsData keeps one layer
    }
    void init(const std::string&) override { }
    void dump(const std::string& path) const override
    {
        FILE* f = fopen(path.c_str(), "w");
        if (!f)
            return;
        for (const ElementData& d : data)
            fprintf(f, "%.17g %.17g %.17g %.17g\n", d.iceThickness(), d.iceConcentration(),
                d.snowThickness(), d.iceTemperature(0));
        fclose(f);
    }
    int nIceLayers() const override { return nl; }
    int resetCursor() override
    {
        pos = 0;
        return 0;
    }
    bool validCursor() const override { return pos < data.size(); }
    ElementData& cursorData() override { return data[pos]; }
    const ElementData& cursorData() const override { return data[pos]; }
    void incrCursor() override { ++pos; }
    std::vector<ElementData> data;

private:
    int nl;
    size_t pos;
};

static double rnd(unsigned long long& s)
{ // splitmix64 -> [0,1)
    s += 0x9E3779B97F4A7C15ull;
    unsigned long long z = s;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return (z >> 11) * (1.0 / 9007199254740992.0);
}

static void fill(VectorGrid& g, int layers, bool cold)
{
    unsigned long long s = 20240611ull;
    for (ElementData& d : g.data) {
        double c = rnd(s) < 0.1 ? 0. : 0.05 + 0.95 * rnd(s);
        double h = (rnd(s) < 0.05 ? 0.005 + 0.015 * rnd(s) : 0.01 + 2.99 * rnd(s)) * c;
        double hs = rnd(s) < 0.3 ? 0. : c * 0.3 * rnd(s);
        double sss = 30 + 5 * rnd(s);
        double sst = cold ? -0.055 * sss + (-0.02 + 0.07 * rnd(s)) : -1.8 + 3.8 * rnd(s);
        double t0 = -20 + 19.7 * rnd(s);
        d = PrognosticGenerator().hice(h).cice(c).hsnow(hs).sst(sst).sss(sss).tice(
            std::vector<double>(layers, t0));
        double tair = -30 + 35 * rnd(s);
        d.airTemperature() = tair;
        d.dewPoint2m() = tair - 5 * rnd(s);
        d.airPressure() = 98000 + 5000 * rnd(s);
        d.ExternalData::mixingRatio() = -1.;
        d.incomingShortwave() = 300 * rnd(s);
        d.incomingLongwave() = 150 + 200 * rnd(s);
        d.mixedLayerDepth() = 10 + 40 * rnd(s);
        d.snowfall() = rnd(s) < 0.5 ? 0. : 1e-4 * rnd(s);
        d.windSpeed() = 20 * rnd(s);
    }
}

int main(int argc, char** argv)
{
    size_t n = argc > 1 ? (size_t)atol(argv[1]) : 100000;
    int nsteps = argc > 2 ? atoi(argv[2]) : 3;
    int dt = argc > 3 ? atoi(argv[3]) : 600;
    const int layers = 2;
    int failures = 0;
    const char* configs[2] = { "",
        "[Modules]\nNextsim::IFreezingPoint = Nextsim::UnescoFreezing\n"
        "Nextsim::IIceAlbedo = Nextsim::CCSMIceAlbedo\n\n[CCSMIceAlbedo]\niceAlbedo = 0.63\n"
        "snowAlbedo = 0.88\n" };
    for (int cfg = 0; cfg < 2; ++cfg) {
        Configurator::clear();
        Configurator::addStream(std::unique_ptr<std::istream>(new std::stringstream(configs[cfg])));
        ModuleLoader::getLoader().setAllDefaults();
        ConfiguredModule::parseConfigurator();
        {
            ElementData probe;
            probe.configure();
        }
        VectorGrid cpuGrid(n, layers), gpuGrid(n, layers);
        fill(cpuGrid, layers, cfg == 1);
        fill(gpuGrid, layers, cfg == 1);

        DevStep cpuStep;
        cpuStep.setInitialData(cpuGrid);
        Iterator cpuIt(&cpuStep);
        cpuIt.setStartStopStep(0, nsteps * dt, dt);
        cpuIt.run();

        GpuStep gpuStep(0);
        gpuStep.setInitialData(gpuGrid);
        Iterator gpuIt(&gpuStep);
        gpuIt.setStartStopStep(0, nsteps * dt, dt);
        gpuIt.run(); // start, iterate x nsteps, stop (scatter back)

        double worst = 0;
        size_t flips = 0;
        for (size_t k = 0; k < n; ++k) {
            const ElementData& a = cpuGrid.data[k];
            const ElementData& b = gpuGrid.data[k];
            double va[5] = { a.iceThickness(), a.iceConcentration(), a.snowThickness(),
                a.iceTemperature(0), a.iceTemperature(1) };
            double vb[5] = { b.iceThickness(), b.iceConcentration(), b.snowThickness(),
                b.iceTemperature(0), b.iceTemperature(1) };
            bool bad = false;
            for (int q = 0; q < 5; ++q) {
                double e = std::fabs(va[q] - vb[q]) / std::fmax(std::fabs(va[q]), 1e-3);
                if (e > 1e-10)
                    bad = true;
                else
                    worst = std::fmax(worst, e);
            }
            flips += bad;
            if (a.seaSurfaceTemperature() != b.seaSurfaceTemperature())
                ++failures;
        }
        printf("config %d: cells=%zu steps=%d albedo_kind=%d freezing_kind=%d max_rel_err=%.3e "
               "cells_out_of_tolerance=%zu\n",
            cfg, n, nsteps, gpuStep.parameters().albedo_kind, gpuStep.parameters().freezing_kind,
            worst, flips);
        if (flips > n / 100000 + 0)
            ++failures;
        gpuStep.writeRestartFile("/tmp/gpustep_restart.txt");
    }
    // ---- the step chosen like every other plug-in: [Modules] Nextsim::IModelStep = Nextsim::GpuStep
    {
        Configurator::clear();
        Configurator::addStream(std::unique_ptr<std::istream>(
            new std::stringstream("[Modules]\nNextsim::IModelStep = Nextsim::GpuStep\n")));
        ModuleLoader::getLoader().setAllDefaults();
        IModelStep& dflt = ModuleLoader::getLoader().getImplementation<IModelStep>();
        if (!dynamic_cast<DevStep*>(&dflt)) {
            printf("default IModelStep is not DevStep\n");
            ++failures;
        }
        ConfiguredModule::parseConfigurator();
        {
            ElementData probe;
            probe.configure();
        }
        IModelStep& chosen = ModuleLoader::getLoader().getImplementation<IModelStep>();
        if (!dynamic_cast<GpuStep*>(&chosen)) {
            printf("[Modules] did not select GpuStep\n");
            ++failures;
        }
        size_t m = n < 5000 ? n : 5000;
        VectorGrid cpuGrid(m, layers), gpuGrid(m, layers);
        fill(cpuGrid, layers, false);
        fill(gpuGrid, layers, false);
        DevStep cpuStep;
        cpuStep.setInitialData(cpuGrid);
        Iterator cpuIt(&cpuStep);
        cpuIt.setStartStopStep(0, dt, dt);
        cpuIt.run();
        chosen.setInitialData(gpuGrid);
        Iterator gpuIt(&chosen);
        gpuIt.setStartStopStep(0, dt, dt);
        gpuIt.run();
        size_t bad = 0;
        for (size_t k = 0; k < m; ++k) {
            double a = cpuGrid.data[k].iceThickness(), b = gpuGrid.data[k].iceThickness();
            if (std::fabs(a - b) > 1e-10 * std::fmax(std::fabs(a), 1e-3))
                ++bad;
        }
        printf("module-loader selected step: cells=%zu out_of_tolerance=%zu\n", m, bad);
        if (bad)
            ++failures;
    }
    printf(failures ? "FAIL\n" : "PASS\n");
    return failures ? 1 : 0;
}
