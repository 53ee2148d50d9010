// Fused column-thermodynamics step: one launch = one DevStep::iterate sweep over all cells.
//
// Replaces, per cell, the reference call chain (paths relative to the reference tree)
//   DevStep::iterate                      core/src/DevStep.cpp:14-23
//   IPhysics1d::updateDerivedData         physics/src/modules/include/IPhysics1d.hpp:33-45
//   NextsimPhysics::calculate             physics/src/modules/NextsimPhysics.cpp:116-131
//   ThermoIce0::calculate                 physics/src/modules/ThermoIce0.cpp:34-133
//   PrognosticData::updateAndIntegrate    core/src/PrognosticData.cpp:63-71
// The per-cell virtual plug-ins (IIceAlbedo, IFreezingPoint, IIceOceanHeatFlux,
// IConcentrationModel, IThermodynamics) become switches on the enums of nsdg_thermo_params.
//
// Layout: 15 structure-of-arrays FP64 streams of length ncells (14 inputs + the hidden
// persistent newice); 5 are written back.  160 algorithmic bytes per
// cell-update.  The reference's ~38 per-cell divisions are the cost centre on the FP64 pipe, so
// quotients by constants and by dt are multiplications by reciprocals taken at compile time / on
// the host, quotients with a common denominator share one reciprocal, and the remaining 14 are
// Newton reciprocals (fp64_fast.cuh); exact zeros stay exact zeros (0*r == 0/d), so the
// data-dependent branches (t1<tf, hi<minh, draught>hi, c<minc, hs>0) resolve as on the CPU and
// the state agrees with the compiled reference to <= 1e-10 (tests/test_gpu_thermo.py).
#include "ctx.h"
#include <algorithm>
#include "fp64_fast.cuh"

namespace {

// core/src/include/constants.hpp:18-123
constexpr double SIGMA_SB = 5.670374419e-8;
constexpr double ICE_EPSILON = 0.996;
constexpr double ICE_KAPPA = 2.0334;
constexpr double ICE_LF = 333.55e3;
constexpr double ICE_RHO = 917.;
constexpr double ICE_RHOSNOW = 330.;
constexpr double ICE_S = 5.;
constexpr double AIR_CP = 1004.64;
constexpr double AIR_RA = 287.058;
constexpr double VAPOUR_CP = 1860.;
constexpr double VAPOUR_LV0 = 2500.79e3;
constexpr double VAPOUR_RA = 461.5;
constexpr double WATER_CP = 4186.84;
constexpr double WATER_MU = 0.055;
constexpr double WATER_RHOOCEAN = 1025.;
constexpr double WATER_TF = 273.15;

struct SphCoef {
    double a, b, c, d, bigA, bigB, bigC;
};
// NextsimPhysics.cpp:309-312 (water), :345-348 (ice)
__device__ constexpr SphCoef SPH_WATER = { 6.1121e2, 18.729, 257.87, 227.3, 7.2e-4, 3.20e-6, 5.9e-10 };
__device__ constexpr SphCoef SPH_ICE = { 6.1115e2, 23.036, 279.82, 333.7, 2.2e-4, 3.83e-6, 6.4e-10 };
constexpr double SPH_ALPHA = 0.62197; // :324-325
constexpr double SPH_BETA = 1 - 0.62197;

__device__ __forceinline__ double kelvin(double c) { return c + WATER_TF; }
__device__ __forceinline__ double sq(double x) { return x * x; }

// NextsimPhysics.cpp:371-375
__device__ __forceinline__ double sph_f(const SphCoef& k, double t, double pPa)
{
    double pmb = pPa * 0.01;
    return 1 + k.bigA + pmb * (k.bigB + k.bigC * t * t);
}
// NextsimPhysics.cpp:377-381; rtc = 1/(t + c)
__device__ __forceinline__ double sph_est(const SphCoef& k, double t, double sal, double rtc)
{
    double salFactor = 1 - 5.37e-4 * sal;
    return k.a * exp((k.b - t * (1. / k.d)) * t * rtc) * salFactor;
}
// NextsimPhysics.cpp:383-386; pow(x,4) as two squarings
__device__ __forceinline__ double stefan_boltzmann(double tC)
{
    return ICE_EPSILON * SIGMA_SB * sq(sq(kelvin(tC)));
}
// NextsimPhysics.cpp:297-302
__device__ __forceinline__ double latent_heat_water(double t)
{
    return VAPOUR_LV0 + t * (-2.36418e3 + t * (1.58927 + t * (-6.14342e-2)));
}
// NextsimPhysics.cpp:304-307
__device__ __forceinline__ double latent_heat_ice(double t)
{
    return VAPOUR_LV0 + ICE_LF - 240. + t * (-290. + t * (-4.));
}

struct CellIn {
    double hice, cice, hsnow, sst, sss, tice0, tair, tdew, slp, qsw, qlw, mld, snowfall, wind,
        newice;
};
struct CellOut {
    double hice, cice, hsnow, tice0, newice;
    double sst; // only meaningful with the slab-ocean closure
};

// idt = 1/dt
template <bool DIAG, bool SLAB>
__device__ __forceinline__ CellOut thermo_cell(
    const nsdg_thermo_params& P, double dt, double idt, const CellIn& in, double* diag)
{
    const double hice = in.hice, cice = in.cice, hsnow = in.hsnow, tice0 = in.tice0;
    const double sst = in.sst, sss = in.sss, slp = in.slp, wind = in.wind;
    double newice = in.newice;

    // ---- updateDerivedData: IPhysics1d.hpp:33-45, NextsimPhysics.cpp:85-114 ----
    // specific humidities (NextsimPhysics.cpp:333-343): q = alpha f est / (p - beta f est); est, f
    // and the reciprocals of the ice surface are reused by dq_dT below (the reference recomputes
    // them, :356-369)
    const double estA = sph_est(SPH_WATER, in.tdew, 0., rcp_pos(in.tdew + SPH_WATER.c));
    const double fA = sph_f(SPH_WATER, in.tdew, slp);
    const double sphuma = SPH_ALPHA * fA * estA * rcp_pos(slp - SPH_BETA * fA * estA);
    const double estW = sph_est(SPH_WATER, sst, sss, rcp_pos(sst + SPH_WATER.c));
    const double fW = sph_f(SPH_WATER, sst, slp);
    const double sphumw = SPH_ALPHA * fW * estW * rcp_pos(slp - SPH_BETA * fW * estW);
    const double rtcI = rcp_pos(tice0 + SPH_ICE.c);
    const double estI = sph_est(SPH_ICE, tice0, 0., rtcI);
    const double fI = sph_f(SPH_ICE, tice0, slp);
    const double rqI = rcp_pos(slp - SPH_BETA * fI * estI);
    const double sphumi = SPH_ALPHA * fI * estI * rqI;
    // rho = p / (R_wet T), R_wet = R_a / (1 - q_a (1 - R_v/R_a))   (:104-109)
    const double rho = slp * (1 - sphuma * (1 - VAPOUR_RA / AIR_RA)) * rcp_pos(AIR_RA * kelvin(in.tair));
    const double cspec = AIR_CP + sphuma * VAPOUR_CP;
    const double icice = (cice != 0) ? rcp_pos(cice) : 0.;
    const double hiTrue = hice * icice; // PrognosticData.hpp:56
    const double hsTrue = hsnow * icice; // PrognosticData.hpp:75
    double hs_new = hsTrue, hi_new = hiTrue;

    // ---- NextsimPhysics::calculate :116-131 ----
    const double rw = rho * wind;
    double evap = P.drag_ocean_q * rw * (sphumw - sphuma); // :133-137
    double tau = rho * (1e-3 * fmax(1.0, fmin(2.0, 0.61 + 0.063 * wind))); // :139-142,291-295
    // heatFluxOpenWater :144-162
    double Qlhow = evap * latent_heat_water(sst);
    double Qshow = P.drag_ocean_t * rw * cspec * (sst - in.tair);
    double Qswow = -in.qsw * (1 - P.albedo_w);
    double Qlwow = stefan_boltzmann(sst) - in.qlw;
    double Qow = Qlhow + Qshow + Qlwow + Qswow;
    // massFluxIceAtmosphere :164-168
    double subl = P.drag_ice_t * rw * (sphumi - sphuma);
    // heatFluxIceAtmosphere :170-198
    double lhi = latent_heat_ice(tice0);
    double Qlhi = subl * lhi;
    double dq_dT;
    { // SpecificHumidityIce::dq_dT :356-369
        const SphCoef& k = SPH_ICE;
        double df_dT = 2 * k.bigC * k.bigB * tice0;
        double numerator = k.b * k.c * k.d - tice0 * (2 * k.c + tice0);
        double dest_dT = numerator * (sq(rtcI) * (1. / k.d)) * estI; // / (d (c + T)^2)
        dq_dT = SPH_ALPHA * slp * (fI * dest_dT + estI * df_dT) * sq(rqI); // / (p - beta est f)^2
    }
    const double dsh = P.drag_ice_t * rw * cspec;
    double dQlh_dT = lhi * (P.drag_ice_t * rw * dq_dT);
    double Qshi = dsh * (tice0 - in.tair);
    double dQsh_dT = dsh;
    double albedoValue;
    { // IIceAlbedo::albedo, argument as NextsimPhysics.cpp:185-186 (hsnow/cice if cice > 0 else 0)
        const double hsArg = (cice > 0) ? hsTrue : 0.;
        const double ICE_ALBEDO = 0.64, SNOW_ALBEDO = 0.85;
        if (P.albedo_kind == NSDG_ALBEDO_CCSM) { // CCSMIceAlbedo.cpp:28-36
            double iceAlbedoT = P.ccsm_ice_albedo - fmax(0., 0.075 * (tice0 - (-1.)));
            double snowAlbedoT = P.ccsm_snow_albedo - fmax(0., 0.124 * (tice0 - (-1.)));
            double snowCoverFraction = hsArg * rcp_pos(hsArg + 0.02);
            albedoValue = snowCoverFraction * snowAlbedoT + (1 - snowCoverFraction) * iceAlbedoT;
        } else if (hsArg > 0.) {
            albedoValue = (P.albedo_kind == NSDG_ALBEDO_SMU) // SMUIceAlbedo.cpp:19-26
                ? SNOW_ALBEDO
                : fmin(SNOW_ALBEDO, // SMU2IceAlbedo.cpp:21-29
                    ICE_ALBEDO + (SNOW_ALBEDO - ICE_ALBEDO) * hsArg * (1. / 0.2));
        } else {
            albedoValue = ICE_ALBEDO + 0.4 * (1 - ICE_ALBEDO) * P.i0;
        }
    }
    double Qswi = -in.qsw * (1. - P.i0) * (1 - albedoValue);
    const double kI = kelvin(tice0);
    const double sb3 = ICE_EPSILON * SIGMA_SB * kI * sq(kI);
    double sbI = sb3 * kI;
    double Qlwi = sbI - in.qlw;
    double dQlw_dT = 4 * sb3; // 4/T * eps sigma T^4
    double Qia = Qlhi + Qshi + Qlwi + Qswi;
    double dQ_dT = dQlh_dT + dQsh_dT + dQlw_dT;
    // heatFluxIceOcean :222-226 -> BasicIceOceanHeatFlux.cpp:16-25, ExternalData.hpp:60
    double tf;
    if (P.freezing_kind == NSDG_FREEZING_UNESCO) { // UnescoFreezing.hpp:28-38
        tf = sss * (-0.0575 + 1.710523e-3 * sqrt(sss) + -2.154996e-4 * sss) + -7.53e-4 * 0.;
    } else { // LinearFreezing.hpp:30-34
        tf = -WATER_MU * sss;
    }
    const double mlBulkCp = in.mld * WATER_RHOOCEAN * WATER_CP;
    const double Qio0 = (sst - tf) * mlBulkCp * idt;
    double Qio = Qio0;

    // ---- massFluxIceOcean :200-220 ----
    double hifroms = 0;
    double t_new;
    { // ThermoIce0::calculate, ThermoIce0.cpp:34-133
        const double freezingPointIce = -WATER_MU * ICE_S;
        const double bulkLHFusionSnow = ICE_LF * ICE_RHOSNOW;
        const double bulkLHFusionIce = ICE_LF * ICE_RHO;
        if (hice == 0 || cice == 0) { // :45-51
            hi_new = 0;
            hs_new = 0;
            t_new = freezingPointIce;
        } else {
            double k_lSlab = P.ks * ICE_KAPPA * rcp_pos(P.ks * hiTrue + ICE_KAPPA * hsTrue);
            double QIceConduction = k_lSlab * (tf - tice0);
            double remainingFlux = QIceConduction - Qia;
            t_new = tice0 + remainingFlux * rcp_pos(k_lSlab + dQ_dT);
            double meltingLimit = (hsTrue > 0.) ? 0 : freezingPointIce;
            t_new = fmin(meltingLimit, t_new);
            double snowMeltRate = fmin(-remainingFlux, 0.) * (1. / bulkLHFusionSnow);
            double snowSublRate = subl * (1. / ICE_RHOSNOW);
            hs_new += (snowMeltRate - snowSublRate) * dt;
            double excessIceMelt = fmin(hs_new, 0.) * (bulkLHFusionSnow / bulkLHFusionIce);
            hs_new = fmax(hs_new, 0.);
            hs_new += in.snowfall * dt * (1. / ICE_RHOSNOW);
            double iceBottomChange = (QIceConduction - Qio) * dt * (1. / bulkLHFusionIce);
            double iceThicknessChange = excessIceMelt + iceBottomChange;
            hi_new += iceThicknessChange;
            double iceDraught = (hi_new * ICE_RHO + hs_new * ICE_RHOSNOW) * (1. / WATER_RHOOCEAN);
            if (P.flooding && iceDraught > hi_new) { // :95-106
                double newIce = iceDraught - hi_new;
                hifroms += newIce;
                hi_new = iceDraught;
                hs_new -= newIce * (ICE_RHO / ICE_RHOSNOW);
            }
            if (hi_new < P.min_thick) { // :108-132
                hifroms = 0;
                double deltaQio = (hi_new * bulkLHFusionIce + hs_new * bulkLHFusionSnow) * idt;
                Qio += deltaQio;
                hi_new = 0;
                hs_new = 0;
                t_new = freezingPointIce;
            }
        }
    }
    { // newIceFormation :228-254
        double coolingFlux = Qow;
        // IEEE division: a cell with mld == 0 must give deltaTml = -/+inf (and take the new-ice
        // branch with finite results) exactly as the reference does, not the NaN of the
        // Newton reciprocal
        double deltaTml = -coolingFlux / mlBulkCp * dt;
        double t1 = sst + deltaTml;
        if (t1 < tf) {
            // (tf - sst)/deltaTml * coolingFlux == (sst - tf) mlBulkCp / dt: the flux that takes
            // the mixed layer exactly to the freezing point.  KNOWN DIVERGENCE from the reference:
            // with Qow == 0 exactly and sst < tf the reference evaluates (tf - sst)/(-0) * 0 = NaN
            // (NextsimPhysics.cpp:244-247) and poisons the cell; the closed form stays finite
            // (tests/test_gpu_thermo.py::test_zero_cooling_flux_below_freezing documents it).
            double sensibleFlux = Qio0;
            double latentFlux = coolingFlux - sensibleFlux;
            Qow = sensibleFlux;
            newice = latentFlux * dt * (1 - cice) * (1. / (ICE_LF * ICE_RHO));
        }
    }
    double c_new;
    { // lateralGrowth :262-289, HiblerConcentration.cpp:32-47, updateThickness :257-260
        double del_c = newice * (1. / P.h0);
        if (hi_new < hiTrue) {
            double m = 0;
            if (!(cice >= 1)) {
                double del_hi = hi_new - hiTrue;
                m = del_hi * cice * P.phi_m * rcp_pos(hiTrue);
            }
            del_c += m;
        }
        c_new = cice + del_c;
        if (c_new >= P.min_conc) {
            const double rc = rcp_pos(c_new);
            hi_new += (newice - hi_new * del_c) * rc;
            if (del_c < 0) {
                Qow -= del_c * hs_new * ICE_LF * ICE_RHOSNOW * idt;
            } else {
                hs_new += (0. - hs_new * del_c) * rc;
            }
        }
    }
    if (c_new < P.min_conc || hi_new < P.min_thick) { // NextsimPhysics.cpp:210-219
        Qow += c_new * ICE_LF * (hi_new * ICE_RHO + hs_new * ICE_RHOSNOW) * idt;
        c_new = 0;
        hi_new = 0;
        hs_new = 0;
    }
    if (DIAG) {
        diag[0] = rho;
        diag[1] = sphuma;
        diag[2] = sphumw;
        diag[3] = sphumi;
        diag[4] = cspec;
        diag[5] = tau;
        diag[6] = Qia;
        diag[7] = Qio;
        diag[8] = Qow;
        diag[9] = subl;
        diag[10] = dQ_dT;
        diag[11] = hifroms;
        diag[12] = t_new;
    }
    // updateAndIntegrate: PrognosticData.cpp:63-71, PhysicsData.hpp:61,66
    CellOut o;
    o.hice = hi_new * c_new;
    o.cice = c_new;
    o.hsnow = hs_new * c_new;
    o.tice0 = t_new;
    o.newice = newice;
    // slab-ocean closure (opt-in): the two fluxes the reference drops
    if (SLAB)
        o.sst = sst - dt * ((1 - cice) * Qow + cice * Qio) / mlBulkCp;
    else
        o.sst = sst;
    return o;
}

struct ThermoPtrs {
    double* f[NSDG_TF_COUNT];
    double* diag;
};

// one cell per thread: 15 coalesced 8-byte loads issued up front, 5 stores
template <bool DIAG, bool SLAB>
__global__ void __launch_bounds__(256) thermo_kernel(const __grid_constant__ nsdg_thermo_params P,
    const __grid_constant__ ThermoPtrs ptr, double dt, double idt, int64_t n)
{
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n)
        return;
    double v[NSDG_TF_COUNT];
#pragma unroll
    for (int q = 0; q < NSDG_TF_COUNT; ++q)
        v[q] = __ldg(ptr.f[q] + k);
    const CellIn a = { v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7], v[8], v[9], v[10], v[11], v[12], v[13], v[14] };
    double da[NSDG_THERMO_NDIAG];
    const CellOut o = thermo_cell<DIAG, SLAB>(P, dt, idt, a, da);
    ptr.f[NSDG_TF_HICE][k] = o.hice;
    ptr.f[NSDG_TF_CICE][k] = o.cice;
    ptr.f[NSDG_TF_HSNOW][k] = o.hsnow;
    ptr.f[NSDG_TF_TICE0][k] = o.tice0;
    ptr.f[NSDG_TF_NEWICE][k] = o.newice;
    if (SLAB)
        ptr.f[NSDG_TF_SST][k] = o.sst;
    if (DIAG) {
#pragma unroll
        for (int q = 0; q < NSDG_THERMO_NDIAG; ++q)
            ptr.diag[(size_t)q * n + k] = da[q];
    }
}

// tice[k*nlayers + l] <-> tice0[k]
__global__ void gather_layer0(const double* tice, double* tice0, int64_t n, int nlayers)
{
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n)
        tice0[k] = tice[k * nlayers];
}
// all layers take the surface value (PrognosticData.cpp:82-94)
__global__ void broadcast_layers(const double* tice0, double* tice, int64_t n, int nlayers)
{
    int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q < n * nlayers)
        tice[q] = tice0[q / nlayers];
}

// out[q][k] = f0[q][k] + w (f1[q][k] - f0[q][k]) for the 8 forcing streams (blockIdx.y = q)
struct BlendPtrs {
    double* out[8];
};
__global__ void blend_kernel(const __grid_constant__ BlendPtrs o, const double* f0, const double* f1, double w, int64_t n)
{
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n)
        return;
    const size_t q = blockIdx.y;
    const double a = f0[q * n + k];
    o.out[q][k] = a + w * (f1[q * n + k] - a);
}

int thermo_launch(nsdg_ctx* ctx, double dt, int nsteps, bool diag)
{
    ThermoState* t = ctx->thermo;
    NSDG_REQUIRE(ctx, t, NSDG_ESTATE, "nsdg_thermo_init has not been called");
    NSDG_REQUIRE(ctx, t->have_prog && t->have_forcing, NSDG_ESTATE,
        "prognostic and forcing data must be uploaded before stepping");
    NSDG_REQUIRE(ctx, dt > 0 && nsteps >= 0, NSDG_EINVAL, "dt must be > 0 and nsteps >= 0");
    if (t->n == 0)
        return NSDG_OK;
    if (diag && !t->diag)
        NSDG_CUDA(ctx, cudaMalloc(&t->diag, sizeof(double) * NSDG_THERMO_NDIAG * (size_t)t->n));
    ThermoPtrs p;
    for (int q = 0; q < NSDG_TF_COUNT; ++q)
        p.f[q] = t->fields[q];
    p.diag = t->diag;
    const int block = 256;
    NSDG_REQUIRE(ctx, (t->n + block - 1) / block < ((int64_t)1 << 31), NSDG_EINVAL, "too many cells for one launch");
    const unsigned grid = (unsigned)((t->n + block - 1) / block);
    const double idt = 1. / dt;
    for (int s = 0; s < nsteps; ++s) {
        const bool dg = diag && s == nsteps - 1;
        nsdg_prof_begin(ctx, NSDG_K_THERMO);
        if (dg && t->slab_ocean)
            thermo_kernel<true, true><<<grid, block, 0, ctx->stream>>>(t->params, p, dt, idt, t->n);
        else if (dg)
            thermo_kernel<true, false><<<grid, block, 0, ctx->stream>>>(t->params, p, dt, idt, t->n);
        else if (t->slab_ocean)
            thermo_kernel<false, true><<<grid, block, 0, ctx->stream>>>(t->params, p, dt, idt, t->n);
        else
            thermo_kernel<false, false><<<grid, block, 0, ctx->stream>>>(t->params, p, dt, idt, t->n);
        nsdg_prof_end(ctx);
        NSDG_LAUNCHED(ctx);
    }
    return NSDG_OK;
}

} // namespace

void nsdg_thermo_free(nsdg_ctx* ctx)
{
    ThermoState* t = ctx->thermo;
    if (!t)
        return;
    for (int q = 0; q < NSDG_TF_COUNT; ++q)
        cudaFree(t->fields[q]);
    cudaFree(t->diag);
    cudaFree(t->stage);
    cudaFree(t->level[0]);
    cudaFree(t->level[1]);
    delete t;
    ctx->thermo = nullptr;
}

extern "C" {

void nsdg_thermo_default_params(nsdg_thermo_params* p)
{
    if (!p)
        return;
    p->drag_ocean_q = 1.5e-3;
    p->drag_ocean_t = 0.83e-3;
    p->drag_ice_t = 1.3e-3;
    p->albedo_w = 0.07;
    p->i0 = 0.17;
    p->min_conc = 1e-12;
    p->min_thick = 0.01;
    p->ks = 0.3096;
    p->h0 = 0.25;
    p->phi_m = 0.5;
    p->ccsm_ice_albedo = 0.538;
    p->ccsm_snow_albedo = 0.8256;
    p->flooding = 1;
    p->albedo_kind = NSDG_ALBEDO_SMU;
    p->freezing_kind = NSDG_FREEZING_LINEAR;
    p->ocean_flux_kind = NSDG_OCEANFLUX_BASIC;
}

int nsdg_thermo_init(nsdg_ctx* ctx, int64_t ncells, int32_t nlayers)
{
    NSDG_ENTER(ctx);
    NSDG_REQUIRE(ctx, ncells >= 0 && nlayers >= 1, NSDG_EINVAL, "ncells >= 0 and nlayers >= 1");
    NSDG_REQUIRE(ctx, !ctx->thermo, NSDG_ESTATE,
        "this context already holds a thermodynamics state: call nsdg_thermo_release first");
    ThermoState* t = new ThermoState;
    ctx->thermo = t;
    t->n = ncells;
    t->nlayers = nlayers;
    nsdg_thermo_default_params(&t->params);
    // even-padded so the vector path may read the pair that holds the last cell
    size_t bytes = sizeof(double) * (size_t)(ncells + 2);
    for (int q = 0; q < NSDG_TF_COUNT; ++q) {
        NSDG_CUDA(ctx, cudaMalloc(&t->fields[q], bytes));
        NSDG_CUDA(ctx, cudaMemsetAsync(t->fields[q], 0, bytes, ctx->stream));
    }
    if (nlayers > 1)
        NSDG_CUDA(ctx, cudaMalloc(&t->stage, sizeof(double) * (size_t)ncells * nlayers + 16));
    return NSDG_OK;
}

int nsdg_thermo_release(nsdg_ctx* ctx)
{
    NSDG_ENTER(ctx);
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    nsdg_thermo_free(ctx);
    return NSDG_OK;
}

int nsdg_thermo_set_params(nsdg_ctx* ctx, const nsdg_thermo_params* p)
{
    NSDG_ENTER(ctx);
    NSDG_REQUIRE(ctx, ctx->thermo, NSDG_ESTATE, "nsdg_thermo_init has not been called");
    NSDG_REQUIRE(ctx, p, NSDG_EINVAL, "params is null");
    // unknown module combination => error, never a fallback
    NSDG_REQUIRE(ctx, p->albedo_kind >= 0 && p->albedo_kind <= NSDG_ALBEDO_CCSM, NSDG_EINVAL,
        "unknown IIceAlbedo implementation id %d", p->albedo_kind);
    NSDG_REQUIRE(ctx, p->freezing_kind >= 0 && p->freezing_kind <= NSDG_FREEZING_UNESCO,
        NSDG_EINVAL, "unknown IFreezingPoint implementation id %d", p->freezing_kind);
    NSDG_REQUIRE(ctx, p->ocean_flux_kind == NSDG_OCEANFLUX_BASIC, NSDG_EINVAL,
        "unknown IIceOceanHeatFlux implementation id %d", p->ocean_flux_kind);
    NSDG_REQUIRE(ctx, p->h0 != 0., NSDG_EINVAL, "Hibler.h0 must be non-zero");
    // the kernel takes 1/c_new by a Newton reciprocal without special-case paths: c_new >= min_conc
    // must be strictly positive there
    NSDG_REQUIRE(ctx, p->min_conc > 0. && p->min_thick >= 0., NSDG_EINVAL,
        "min_conc must be > 0 and min_thick >= 0 (got %g, %g)", p->min_conc, p->min_thick);
    NSDG_REQUIRE(ctx, p->ks > 0., NSDG_EINVAL, "thermoice0.ks must be > 0 (got %g)", p->ks);
    ctx->thermo->params = *p;
    return NSDG_OK;
}

static int upload(nsdg_ctx* ctx, int field, const double* host)
{
    ThermoState* t = ctx->thermo;
    NSDG_REQUIRE(ctx, host || t->n == 0, NSDG_EINVAL, "null host array for field %d", field);
    if (t->n)
        NSDG_CUDA(ctx, cudaMemcpyAsync(t->fields[field], host, sizeof(double) * (size_t)t->n,
                           cudaMemcpyHostToDevice, ctx->stream));
    return NSDG_OK;
}
static int download(nsdg_ctx* ctx, int field, double* host)
{
    ThermoState* t = ctx->thermo;
    NSDG_REQUIRE(ctx, host || t->n == 0, NSDG_EINVAL, "null host array for field %d", field);
    if (t->n)
        NSDG_CUDA(ctx, cudaMemcpyAsync(host, t->fields[field], sizeof(double) * (size_t)t->n,
                           cudaMemcpyDeviceToHost, ctx->stream));
    return NSDG_OK;
}

int nsdg_thermo_upload_prognostic(nsdg_ctx* ctx, const double* hice, const double* cice,
    const double* hsnow, const double* sst, const double* sss, const double* tice)
{
    NSDG_ENTER(ctx);
    ThermoState* t = ctx->thermo;
    NSDG_REQUIRE(ctx, t, NSDG_ESTATE, "nsdg_thermo_init has not been called");
    int rc;
    if ((rc = upload(ctx, NSDG_TF_HICE, hice)) || (rc = upload(ctx, NSDG_TF_CICE, cice))
        || (rc = upload(ctx, NSDG_TF_HSNOW, hsnow)) || (rc = upload(ctx, NSDG_TF_SST, sst))
        || (rc = upload(ctx, NSDG_TF_SSS, sss)))
        return rc;
    if (t->nlayers == 1) {
        if ((rc = upload(ctx, NSDG_TF_TICE0, tice)))
            return rc;
    } else if (t->n) {
        NSDG_REQUIRE(ctx, tice, NSDG_EINVAL, "null tice");
        NSDG_CUDA(ctx, cudaMemcpyAsync(t->stage, tice, sizeof(double) * (size_t)t->n * t->nlayers,
                           cudaMemcpyHostToDevice, ctx->stream));
        gather_layer0<<<(unsigned)((t->n + 255) / 256), 256, 0, ctx->stream>>>(
            t->stage, t->fields[NSDG_TF_TICE0], t->n, t->nlayers);
        NSDG_LAUNCHED(ctx);
    }
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); // host arrays are borrowed for the call only
    t->have_prog = true;
    return NSDG_OK;
}

int nsdg_thermo_upload_forcing(nsdg_ctx* ctx, const double* tair, const double* tdew,
    const double* slp, const double* qsw, const double* qlw, const double* mld,
    const double* snowfall, const double* wind)
{
    NSDG_ENTER(ctx);
    ThermoState* t = ctx->thermo;
    NSDG_REQUIRE(ctx, t, NSDG_ESTATE, "nsdg_thermo_init has not been called");
    const double* src[8] = { tair, tdew, slp, qsw, qlw, mld, snowfall, wind };
    for (int q = 0; q < 8; ++q) {
        int rc = upload(ctx, NSDG_TF_TAIR + q, src[q]);
        if (rc)
            return rc;
    }
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    t->have_forcing = true;
    return NSDG_OK;
}

int nsdg_thermo_upload_forcing_level(nsdg_ctx* ctx, int32_t level, const double* tair,
    const double* tdew, const double* slp, const double* qsw, const double* qlw,
    const double* mld, const double* snowfall, const double* wind)
{
    NSDG_ENTER(ctx);
    ThermoState* t = ctx->thermo;
    NSDG_REQUIRE(ctx, t, NSDG_ESTATE, "nsdg_thermo_init has not been called");
    NSDG_REQUIRE(ctx, level == 0 || level == 1, NSDG_EINVAL, "forcing level must be 0 or 1");
    const double* src[8] = { tair, tdew, slp, qsw, qlw, mld, snowfall, wind };
    for (const double* p : src)
        NSDG_REQUIRE(ctx, p || t->n == 0, NSDG_EINVAL, "null forcing array");
    if (!t->level[level])
        NSDG_CUDA(ctx, cudaMalloc(&t->level[level], sizeof(double) * 8 * (size_t)(t->n + 1)));
    for (int q = 0; q < 8 && t->n > 0; ++q)
        NSDG_CUDA(ctx, cudaMemcpyAsync(t->level[level] + (size_t)q * t->n, src[q], sizeof(double) * (size_t)t->n,
                           cudaMemcpyHostToDevice, ctx->stream));
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    t->have_level[level] = true;
    return NSDG_OK;
}

int nsdg_thermo_blend_forcing(nsdg_ctx* ctx, double w)
{
    NSDG_ENTER(ctx);
    ThermoState* t = ctx->thermo;
    NSDG_REQUIRE(ctx, t, NSDG_ESTATE, "nsdg_thermo_init has not been called");
    NSDG_REQUIRE(ctx, t->have_level[0] && t->have_level[1], NSDG_ESTATE, "both forcing levels must be uploaded first");
    NSDG_REQUIRE(ctx, w >= 0. && w <= 1., NSDG_EINVAL, "interpolation weight %g outside [0,1]", w);
    if (t->n > 0) {
        BlendPtrs o;
        for (int q = 0; q < 8; ++q)
            o.out[q] = t->fields[NSDG_TF_TAIR + q];
        dim3 block(256), grid((unsigned)((t->n + 255) / 256), 8);
        NSDG_KERNEL(ctx, NSDG_K_BLEND, blend_kernel<<<grid, block, 0, ctx->stream>>>(o, t->level[0], t->level[1], w, t->n));
    }
    t->have_forcing = true;
    return NSDG_OK;
}

int nsdg_thermo_set_slab_ocean(nsdg_ctx* ctx, int32_t enable)
{
    NSDG_ENTER(ctx);
    NSDG_REQUIRE(ctx, ctx->thermo, NSDG_ESTATE, "nsdg_thermo_init has not been called");
    ctx->thermo->slab_ocean = enable != 0;
    return NSDG_OK;
}

int nsdg_thermo_upload_newice(nsdg_ctx* ctx, const double* newice)
{
    NSDG_ENTER(ctx);
    NSDG_REQUIRE(ctx, ctx->thermo, NSDG_ESTATE, "nsdg_thermo_init has not been called");
    int rc = upload(ctx, NSDG_TF_NEWICE, newice);
    if (rc)
        return rc;
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return NSDG_OK;
}

int nsdg_thermo_download_newice(nsdg_ctx* ctx, double* newice)
{
    NSDG_ENTER(ctx);
    NSDG_REQUIRE(ctx, ctx->thermo, NSDG_ESTATE, "nsdg_thermo_init has not been called");
    int rc = download(ctx, NSDG_TF_NEWICE, newice);
    if (rc)
        return rc;
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return NSDG_OK;
}

int nsdg_thermo_step(nsdg_ctx* ctx, double dt, int32_t nsteps)
{
    NSDG_ENTER(ctx);
    return thermo_launch(ctx, dt, nsteps, false);
}

int nsdg_thermo_step_diag(nsdg_ctx* ctx, double dt, int32_t nsteps)
{
    NSDG_ENTER(ctx);
    return thermo_launch(ctx, dt, nsteps, true);
}

int nsdg_thermo_download_prognostic(nsdg_ctx* ctx, double* hice, double* cice, double* hsnow,
    double* sst, double* sss, double* tice)
{
    NSDG_ENTER(ctx);
    ThermoState* t = ctx->thermo;
    NSDG_REQUIRE(ctx, t, NSDG_ESTATE, "nsdg_thermo_init has not been called");
    int rc;
    if ((rc = download(ctx, NSDG_TF_HICE, hice)) || (rc = download(ctx, NSDG_TF_CICE, cice))
        || (rc = download(ctx, NSDG_TF_HSNOW, hsnow)))
        return rc;
    if (sst && (rc = download(ctx, NSDG_TF_SST, sst)))
        return rc;
    if (sss && (rc = download(ctx, NSDG_TF_SSS, sss)))
        return rc;
    if (t->nlayers == 1) {
        if ((rc = download(ctx, NSDG_TF_TICE0, tice)))
            return rc;
    } else if (t->n) {
        NSDG_REQUIRE(ctx, tice, NSDG_EINVAL, "null tice");
        int64_t tot = t->n * t->nlayers;
        broadcast_layers<<<(unsigned)((tot + 255) / 256), 256, 0, ctx->stream>>>(
            t->fields[NSDG_TF_TICE0], t->stage, t->n, t->nlayers);
        NSDG_LAUNCHED(ctx);
        NSDG_CUDA(ctx, cudaMemcpyAsync(tice, t->stage, sizeof(double) * (size_t)tot,
                           cudaMemcpyDeviceToHost, ctx->stream));
    }
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return NSDG_OK;
}

int nsdg_thermo_download_diagnostics(nsdg_ctx* ctx, double* diag)
{
    NSDG_ENTER(ctx);
    ThermoState* t = ctx->thermo;
    NSDG_REQUIRE(ctx, t && t->diag, NSDG_ESTATE, "no diagnostics: call nsdg_thermo_step_diag first");
    NSDG_REQUIRE(ctx, diag, NSDG_EINVAL, "diag is null");
    NSDG_CUDA(ctx, cudaMemcpyAsync(diag, t->diag, sizeof(double) * NSDG_THERMO_NDIAG * (size_t)t->n,
                       cudaMemcpyDeviceToHost, ctx->stream));
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return NSDG_OK;
}

int nsdg_thermo_field_ptr(nsdg_ctx* ctx, int32_t field, void** device_ptr)
{
    NSDG_ENTER(ctx);
    ThermoState* t = ctx->thermo;
    NSDG_REQUIRE(ctx, t, NSDG_ESTATE, "nsdg_thermo_init has not been called");
    NSDG_REQUIRE(ctx, field >= 0 && field < NSDG_TF_COUNT && device_ptr, NSDG_EINVAL,
        "bad field id %d", field);
    *device_ptr = t->fields[field];
    // a caller that writes through these pointers owns the "uploaded" state
    t->have_prog = t->have_forcing = true;
    return NSDG_OK;
}

int nsdg_thermo_run_host(nsdg_ctx* ctx, double dt, int32_t nsteps, double* hice, double* cice,
    double* hsnow, const double* sst, const double* sss, double* tice, const double* tair,
    const double* tdew, const double* slp, const double* qsw, const double* qlw,
    const double* mld, const double* snowfall, const double* wind)
{
    NSDG_ENTER(ctx);
    ThermoState* t = ctx->thermo;
    NSDG_REQUIRE(ctx, t, NSDG_ESTATE, "nsdg_thermo_init has not been called");
    int rc;
    if (t->nlayers != 1 || t->n < (1 << 16) || nsteps < 1) { // small or layered: plain sequence
        if ((rc = nsdg_thermo_upload_prognostic(ctx, hice, cice, hsnow, sst, sss, tice)))
            return rc;
        if ((rc = nsdg_thermo_upload_forcing(ctx, tair, tdew, slp, qsw, qlw, mld, snowfall, wind)))
            return rc;
        if ((rc = nsdg_thermo_step(ctx, dt, nsteps)))
            return rc;
        return nsdg_thermo_download_prognostic(ctx, hice, cice, hsnow, nullptr, nullptr, tice);
    }
    // Cells are independent: cut them into chunks and run upload -> sweeps -> download as a
    // three-stage pipeline over three streams, so that both PCIe directions and the SMs work at
    // the same time (the call is bound by the 14-array upload).
    NSDG_REQUIRE(ctx, dt > 0, NSDG_EINVAL, "dt must be > 0");
    const double* in[14] = { hice, cice, hsnow, sst, sss, tice, tair, tdew, slp, qsw, qlw, mld, snowfall, wind };
    const int inf[14] = { NSDG_TF_HICE, NSDG_TF_CICE, NSDG_TF_HSNOW, NSDG_TF_SST, NSDG_TF_SSS, NSDG_TF_TICE0,
        NSDG_TF_TAIR, NSDG_TF_TAIR + 1, NSDG_TF_TAIR + 2, NSDG_TF_TAIR + 3, NSDG_TF_TAIR + 4,
        NSDG_TF_TAIR + 5, NSDG_TF_TAIR + 6, NSDG_TF_TAIR + 7 };
    for (const double* a : in)
        NSDG_REQUIRE(ctx, a, NSDG_EINVAL, "null host array");
    double* out[4] = { hice, cice, hsnow, tice };
    const int outf[4] = { NSDG_TF_HICE, NSDG_TF_CICE, NSDG_TF_HSNOW, NSDG_TF_TICE0 };
    const int nchunks = 8;
    const int64_t per = ((t->n + nchunks - 1) / nchunks + 255) / 256 * 256;
    const double idt = 1. / dt;
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); // earlier work on the fields is finished
    for (int64_t c0 = 0; c0 < t->n; c0 += per) {
        const int64_t cn = std::min<int64_t>(per, t->n - c0);
        for (int q = 0; q < 14; ++q)
            NSDG_CUDA(ctx, cudaMemcpyAsync(t->fields[inf[q]] + c0, in[q] + c0, sizeof(double) * (size_t)cn,
                               cudaMemcpyHostToDevice, ctx->h2d_stream));
        NSDG_CUDA(ctx, cudaEventRecord(ctx->ev_pipe[0], ctx->h2d_stream));
        NSDG_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_pipe[0], 0));
        ThermoPtrs p;
        for (int q = 0; q < NSDG_TF_COUNT; ++q)
            p.f[q] = t->fields[q] + c0;
        p.diag = nullptr;
        const unsigned grid = (unsigned)((cn + 255) / 256);
        for (int s = 0; s < nsteps; ++s) {
            nsdg_prof_begin(ctx, NSDG_K_THERMO);
            if (t->slab_ocean)
                thermo_kernel<false, true><<<grid, 256, 0, ctx->stream>>>(t->params, p, dt, idt, cn);
            else
                thermo_kernel<false, false><<<grid, 256, 0, ctx->stream>>>(t->params, p, dt, idt, cn);
            nsdg_prof_end(ctx);
            NSDG_LAUNCHED(ctx);
        }
        NSDG_CUDA(ctx, cudaEventRecord(ctx->ev_pipe[1], ctx->stream));
        NSDG_CUDA(ctx, cudaStreamWaitEvent(ctx->d2h_stream, ctx->ev_pipe[1], 0));
        for (int q = 0; q < 4; ++q)
            NSDG_CUDA(ctx, cudaMemcpyAsync(out[q] + c0, t->fields[outf[q]] + c0, sizeof(double) * (size_t)cn,
                               cudaMemcpyDeviceToHost, ctx->d2h_stream));
    }
    t->have_prog = t->have_forcing = true;
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->d2h_stream));
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return NSDG_OK;
}

} // extern "C"
