/*!
 * @file GpuComponents.hpp
 * ModelComponents that run on the B200 through the C ABI (include/nsdg_gpu.h):
 *   GpuDynamics       DG transport of hice/cice + mEVP sub-cycle (NOT IN THE REFERENCE SNAPSHOT,
 *                     discretisation in DESIGN.md section 2; parity unpinned)
 *   GpuColumnPhysics  the column thermodynamics of DevStep::iterate (core/src/DevStep.cpp:14-23)
 *                     over H-fields instead of ElementData
 * State lives in HBM between calls; setData uploads the fields present in the ModelState,
 * getState downloads.  Field names follow later-upstream neXtSIM_DG ("hice", "cice", "u", ...).
 * Errors of the C ABI become std::runtime_error; there is no CPU fallback.
 */
#ifndef HOST_GPUCOMPONENTS_HPP
#define HOST_GPUCOMPONENTS_HPP

#include "ModelArray.hpp"
#include "nsdg_gpu.h"

#include <string>
#include <utility>
#include <vector>

namespace Nextsim {

class GpuContext {
public:
    explicit GpuContext(int device = 0)
        : ctx(nullptr)
    {
        if (nsdg_create(&ctx, device) != NSDG_OK)
            throw std::runtime_error("GpuContext: no CUDA device (nsdg_create failed); there is no CPU fallback");
    }
    ~GpuContext()
    {
        if (ctx)
            nsdg_destroy(ctx);
    }
    GpuContext(const GpuContext&) = delete;
    GpuContext& operator=(const GpuContext&) = delete;
    void check(int rc, const char* what) const
    {
        if (rc != NSDG_OK)
            throw std::runtime_error(std::string(what) + ": " + nsdg_last_error(ctx));
    }
    nsdg_ctx* ctx;
};

class GpuDynamics : public ModelComponent {
public:
    //! Grid size from ModelArray::setDimensions, DG order from ModelArray::nComponents(DG).
    GpuDynamics(GpuContext& c, double hx, double hy, const nsdg_mevp_params* mevp = nullptr)
        : gpu(c)
    {
        nsdg_dyn_config cfg = {};
        cfg.nx = (int64_t)ModelArray::nx();
        cfg.ny_global = cfg.ny_local = (int64_t)ModelArray::ny();
        cfg.hx = hx;
        cfg.hy = hy;
        cfg.ncomp = (int32_t)ModelArray::nComponents(ModelArray::Type::DG);
        gpu.check(nsdg_dyn_init(gpu.ctx, &cfg), "nsdg_dyn_init");
        if (mevp)
            gpu.check(nsdg_dyn_set_mevp_params(gpu.ctx, mevp), "nsdg_dyn_set_mevp_params");
        registerModule();
    }
    //! One dynamics state per context (a second nsdg_dyn_init is NSDG_ESTATE): released here.
    ~GpuDynamics() override { nsdg_dyn_release(gpu.ctx); }
    GpuDynamics(const GpuDynamics&) = delete;
    GpuDynamics& operator=(const GpuDynamics&) = delete;
    std::string getName() const override { return "GpuDynamics"; }

    void setData(const ModelState& state) override
    {
        for (const auto& f : fields()) {
            auto it = state.find(f.first);
            if (it == state.end())
                continue;
            const ModelArray& a = it->second;
            if ((int)a.nComponents() != nsdg_dyn_field_ncomp(gpu.ctx, f.second))
                throw std::invalid_argument("GpuDynamics::setData: wrong component count for " + f.first);
            std::vector<double> planes = a.toComponentMajor();
            gpu.check(nsdg_dyn_upload(gpu.ctx, f.second, planes.data()), "nsdg_dyn_upload");
        }
    }
    //! Prognostic fields only (the forcing is the caller's).
    ModelState getState() const override
    {
        ModelState s;
        for (const auto& f : fields()) {
            if (f.second >= NSDG_DF_UATM)
                continue;
            ModelArray a(typeOf(f.second));
            std::vector<double> planes(a.size());
            gpu.check(nsdg_dyn_download(gpu.ctx, f.second, planes.data()), "nsdg_dyn_download");
            a.fromComponentMajor(planes);
            s.emplace(f.first, std::move(a));
        }
        return s;
    }
    //! Transport of hice, cice with the current velocity, then nSub mEVP sub-iterations.
    void update(double dt, int nSub = 100) { gpu.check(nsdg_dyn_step(gpu.ctx, dt, nSub), "nsdg_dyn_step"); }

    static const std::vector<std::pair<std::string, int>>& fields()
    {
        static const std::vector<std::pair<std::string, int>> f = { { "hice", NSDG_DF_H },
            { "cice", NSDG_DF_A }, { "u", NSDG_DF_U }, { "v", NSDG_DF_V }, { "s11", NSDG_DF_S11 },
            { "s12", NSDG_DF_S12 }, { "s22", NSDG_DF_S22 }, { "uwind", NSDG_DF_UATM },
            { "vwind", NSDG_DF_VATM }, { "uocean", NSDG_DF_UOCN }, { "vocean", NSDG_DF_VOCN } };
        return f;
    }
    static ModelArray::Type typeOf(int field)
    {
        if (field == NSDG_DF_H || field == NSDG_DF_A)
            return ModelArray::Type::DG;
        if (field == NSDG_DF_S11 || field == NSDG_DF_S12 || field == NSDG_DF_S22)
            return ModelArray::Type::DGSTRESS;
        return ModelArray::Type::VERTEX;
    }

private:
    GpuContext& gpu;
};

class GpuColumnPhysics : public ModelComponent {
public:
    explicit GpuColumnPhysics(GpuContext& c, const nsdg_thermo_params* p = nullptr)
        : gpu(c)
        , n(ModelArray::ny() * ModelArray::nx())
    {
        gpu.check(nsdg_thermo_init(gpu.ctx, (int64_t)n, 1), "nsdg_thermo_init");
        if (p)
            gpu.check(nsdg_thermo_set_params(gpu.ctx, p), "nsdg_thermo_set_params");
        for (auto& v : prog)
            v.assign(n, 0.);
        for (auto& v : forc)
            v.assign(n, 0.);
        registerModule();
    }
    ~GpuColumnPhysics() override { nsdg_thermo_release(gpu.ctx); }
    GpuColumnPhysics(const GpuColumnPhysics&) = delete;
    GpuColumnPhysics& operator=(const GpuColumnPhysics&) = delete;
    std::string getName() const override { return "GpuColumnPhysics"; }

    void setData(const ModelState& state) override
    {
        static const char* pn[6] = { "hice_mean", "cice_mean", "hsnow", "sst", "sss", "tice" };
        static const char* fn[8] = { "tair", "tdew", "slp", "qsw", "qlw", "mld", "snowfall", "wind" };
        bool pchg = false, fchg = false;
        for (int q = 0; q < 6; ++q)
            pchg |= take(state, pn[q], prog[q]);
        for (int q = 0; q < 8; ++q)
            fchg |= take(state, fn[q], forc[q]);
        if (pchg)
            gpu.check(nsdg_thermo_upload_prognostic(gpu.ctx, prog[0].data(), prog[1].data(), prog[2].data(),
                          prog[3].data(), prog[4].data(), prog[5].data()),
                "nsdg_thermo_upload_prognostic");
        if (fchg)
            gpu.check(nsdg_thermo_upload_forcing(gpu.ctx, forc[0].data(), forc[1].data(), forc[2].data(),
                          forc[3].data(), forc[4].data(), forc[5].data(), forc[6].data(), forc[7].data()),
                "nsdg_thermo_upload_forcing");
    }
    ModelState getState() const override
    {
        static const char* pn[6] = { "hice_mean", "cice_mean", "hsnow", "sst", "sss", "tice" };
        std::vector<double> out[6];
        for (auto& v : out)
            v.resize(n);
        gpu.check(nsdg_thermo_download_prognostic(gpu.ctx, out[0].data(), out[1].data(), out[2].data(),
                      out[3].data(), out[4].data(), out[5].data()),
            "nsdg_thermo_download_prognostic");
        ModelState s;
        for (int q = 0; q < 6; ++q) {
            ModelArray a = ModelArray::HField();
            for (size_t k = 0; k < n; ++k)
                a[k] = out[q][k];
            s.emplace(pn[q], std::move(a));
        }
        return s;
    }
    //! One DevStep::iterate sweep.
    void update(double dt) { gpu.check(nsdg_thermo_step(gpu.ctx, dt, 1), "nsdg_thermo_step"); }

private:
    bool take(const ModelState& state, const char* name, std::vector<double>& dst)
    {
        auto it = state.find(name);
        if (it == state.end())
            return false;
        if (it->second.size() != n)
            throw std::invalid_argument(std::string("GpuColumnPhysics::setData: ") + name + " is not an HField");
        dst.assign(it->second.data(), it->second.data() + n);
        return true;
    }
    GpuContext& gpu;
    size_t n;
    std::vector<double> prog[6], forc[8];
};

} /* namespace Nextsim */

#endif /* HOST_GPUCOMPONENTS_HPP */
