/* TEST INFRASTRUCTURE (oracle). Plain-C restatement of the reference's column physics.
 *
 * Parity status: PINNED.  tests/test_oracle_thermo.py checks this port against (i) the golden
 * values of the reference's own tests (physics/test/NextsimPhysics_test.cpp:47-312,
 * core/test/ElementData_test.cpp:21-86) and (ii) oracle/_ref/libnsdg_ref.so, i.e. the
 * reference's own sources compiled in place, on randomized inputs (bitwise on x86-64/glibc).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load this code.
 * Every function cites the reference lines it follows (paths relative to the reference tree).
 * Operation order is kept identical to the reference so that results agree to the last bit
 * when compiled without FMA contraction.
 */
#include "../include/nsdg_gpu.h"
#include <math.h>
#include <stddef.h>

/* core/src/include/constants.hpp:18-123 */
static const double SIGMA_SB = 5.670374419e-8;
static const double ICE_EPSILON = 0.996;
static const double ICE_KAPPA = 2.0334;
static const double ICE_LF = 333.55e3;
static const double ICE_RHO = 917.;
static const double ICE_RHOSNOW = 330.;
static const double ICE_S = 5.;
static const double AIR_CP = 1004.64;
static const double AIR_RA = 287.058;
static const double VAPOUR_CP = 1860.;
static const double VAPOUR_LV0 = 2500.79e3;
static const double VAPOUR_RA = 461.5;
static const double WATER_CP = 4186.84;
static const double WATER_MU = 0.055;
static const double WATER_RHOOCEAN = 1025.;
static const double WATER_TF = 273.15;

static double kelvin(double c) { return c + WATER_TF; } /* constants.hpp:128 */

/* NextsimPhysics.cpp:309-312 (water), :345-348 (ice), alpha/beta :324-325 */
typedef struct {
    double a, b, c, d, bigA, bigB, bigC, alpha, beta;
} sphum_coef;
static const sphum_coef SPH_WATER = { 6.1121e2, 18.729, 257.87, 227.3, 7.2e-4, 3.20e-6, 5.9e-10,
    0.62197, 1 - 0.62197 };
static const sphum_coef SPH_ICE = { 6.1115e2, 23.036, 279.82, 333.7, 2.2e-4, 3.83e-6, 6.4e-10,
    0.62197, 1 - 0.62197 };

/* NextsimPhysics.cpp:371-375 */
static double sph_f(const sphum_coef* k, double t, double pPa)
{
    double pmb = pPa * 0.01;
    return 1 + k->bigA + pmb * (k->bigB + k->bigC * t * t);
}
/* NextsimPhysics.cpp:377-381 */
static double sph_est(const sphum_coef* k, double t, double sal)
{
    double salFactor = 1 - 5.37e-4 * sal;
    return k->a * exp((k->b - t / k->d) * t / (t + k->c)) * salFactor;
}
/* NextsimPhysics.cpp:333-343 */
static double sph(const sphum_coef* k, double t, double p, double sal)
{
    double est = sph_est(k, t, sal);
    double f = sph_f(k, t, p);
    return k->alpha * f * est / (p - k->beta * f * est);
}
/* NextsimPhysics.cpp:356-369 */
static double sph_ice_dq_dT(double t, double p)
{
    const sphum_coef* k = &SPH_ICE;
    double df_dT = 2 * k->bigC * k->bigB * t;
    double numerator = k->b * k->c * k->d - t * (2 * k->c + t);
    double denominator = k->d * pow(k->c + t, 2);
    double est = sph_est(k, t, 0);
    double f = sph_f(k, t, p);
    double dest_dT = numerator / denominator * est;
    numerator = k->alpha * p * (f * dest_dT + est * df_dT);
    denominator = pow(p - k->beta * est * f, 2);
    return numerator / denominator;
}
/* NextsimPhysics.cpp:383-386 */
static double stefan_boltzmann(double tC) { return ICE_EPSILON * SIGMA_SB * pow(kelvin(tC), 4); }
/* NextsimPhysics.cpp:297-302 */
static double latent_heat_water(double t)
{
    return VAPOUR_LV0 + t * (-2.36418e3 + t * (1.58927 + t * (-6.14342e-2)));
}
/* NextsimPhysics.cpp:304-307 */
static double latent_heat_ice(double t)
{
    return VAPOUR_LV0 + ICE_LF - 240. + t * (-290. + t * (-4.));
}
/* NextsimPhysics.cpp:291-295 */
static double drag_ocean_m(double w) { return 1e-3 * fmax(1.0, fmin(2.0, 0.61 + 0.063 * w)); }

/* LinearFreezing.hpp:30-34, UnescoFreezing.hpp:28-38 */
static double freezing_point(const nsdg_thermo_params* P, double sss)
{
    if (P->freezing_kind == NSDG_FREEZING_UNESCO) {
        const double a0 = -0.0575, a1 = +1.710523e-3, a2 = -2.154996e-4, b = -7.53e-4, p0 = 0;
        return sss * (a0 + a1 * sqrt(sss) + a2 * sss) + b * p0;
    }
    return -WATER_MU * sss;
}

/* SMUIceAlbedo.cpp:19-26, SMU2IceAlbedo.cpp:21-29, CCSMIceAlbedo.cpp:28-36 */
static double ice_albedo(const nsdg_thermo_params* P, double temperature, double snowThickness)
{
    const double ICE_ALBEDO = 0.64, SNOW_ALBEDO = 0.85;
    switch (P->albedo_kind) {
    case NSDG_ALBEDO_SMU:
        if (snowThickness > 0.)
            return SNOW_ALBEDO;
        return ICE_ALBEDO + 0.4 * (1 - ICE_ALBEDO) * P->i0;
    case NSDG_ALBEDO_SMU2:
        if (snowThickness > 0.)
            return fmin(SNOW_ALBEDO, ICE_ALBEDO + (SNOW_ALBEDO - ICE_ALBEDO) * snowThickness / 0.2);
        return ICE_ALBEDO + 0.4 * (1 - ICE_ALBEDO) * P->i0;
    default: {
        const double tLimit = -1.;
        double iceAlbedoT = P->ccsm_ice_albedo - fmax(0., 0.075 * (temperature - tLimit));
        double snowAlbedoT = P->ccsm_snow_albedo - fmax(0., 0.124 * (temperature - tLimit));
        double snowCoverFraction = snowThickness / (snowThickness + 0.02);
        return snowCoverFraction * snowAlbedoT + (1 - snowCoverFraction) * iceAlbedoT;
    }
    }
}

/* One cell, one DevStep::iterate body (core/src/DevStep.cpp:18-21).  state: hice, cice, hsnow,
 * tice0, newice are updated in place.  diag (may be NULL): NSDG_THERMO_NDIAG values. */
void nsdg_oracle_thermo_cell(const nsdg_thermo_params* P, double dt, double* hice_, double* cice_,
    double* hsnow_, double sst, double sss, double* tice0_, double tair, double tdew, double slp,
    double qswIn, double qlwIn, double mld, double snowfall, double wind, double* newice_,
    double* diag)
{
    const double hice = *hice_, cice = *cice_, hsnow = *hsnow_, tice0 = *tice0_;
    double newice = *newice_; /* persistent, SURVEY 8a Q1 */

    /* ---- IPhysics1d::updateDerivedData, IPhysics1d.hpp:33-45 ---- */
    double sphuma = sph(&SPH_WATER, tdew, slp, 0); /* NextsimPhysics.cpp:85-88 */
    double sphumw = sph(&SPH_WATER, sst, slp, sss); /* :90-96 */
    double sphumi = sph(&SPH_ICE, tice0, slp, 0); /* :98-102 */
    double Ra_wet = AIR_RA / (1 - sphuma * (1 - VAPOUR_RA / AIR_RA)); /* :104-109 */
    double rho = slp / (Ra_wet * kelvin(tair));
    double cspec = AIR_CP + sphuma * VAPOUR_CP; /* :111-114 */
    /* PrognosticData.hpp:56,75 */
    const double hiTrue = (cice != 0) ? hice / cice : 0;
    const double hsTrue = (cice != 0) ? hsnow / cice : 0;
    double hs_new = hsTrue;
    double hi_new = hiTrue;

    /* ---- NextsimPhysics::calculate, NextsimPhysics.cpp:116-131 ---- */
    /* massFluxOpenWater :133-137 */
    double evap = P->drag_ocean_q * rho * wind * (sphumw - sphuma);
    /* momentumFluxOpenWater :139-142 */
    double tau = rho * drag_ocean_m(wind);
    /* heatFluxOpenWater :144-162 */
    double Qlhow = evap * latent_heat_water(sst);
    double Qshow = P->drag_ocean_t * rho * cspec * wind * (sst - tair);
    double Qswow = -qswIn * (1 - P->albedo_w);
    double Qlwow = stefan_boltzmann(sst) - qlwIn;
    double Qow = Qlhow + Qshow + Qlwow + Qswow;
    /* massFluxIceAtmosphere :164-168 */
    double subl = P->drag_ice_t * rho * wind * (sphumi - sphuma);
    /* heatFluxIceAtmosphere :170-198 */
    double Qlhi = subl * latent_heat_ice(tice0);
    double dmdot_dT = P->drag_ice_t * rho * wind * sph_ice_dq_dT(tice0, slp);
    double dQlh_dT = latent_heat_ice(tice0) * dmdot_dT;
    double Qshi = P->drag_ice_t * rho * cspec * wind * (tice0 - tair);
    double dQsh_dT = P->drag_ice_t * rho * cspec * wind;
    double albedoValue = ice_albedo(P, tice0, (cice > 0) ? (hsnow / cice) : 0.);
    double Qswi = -qswIn * (1. - P->i0) * (1 - albedoValue);
    double Qlwi = stefan_boltzmann(tice0) - qlwIn;
    double dQlw_dT = 4 / kelvin(tice0) * stefan_boltzmann(tice0);
    double Qia = Qlhi + Qshi + Qlwi + Qswi;
    double dQ_dT = dQlh_dT + dQsh_dT + dQlw_dT;
    /* heatFluxIceOcean :222-226 -> BasicIceOceanHeatFlux.cpp:16-25; ExternalData.hpp:60 */
    const double tf = freezing_point(P, sss);
    const double mlBulkCp = mld * WATER_RHOOCEAN * WATER_CP;
    double Qio = (sst - tf) * mlBulkCp / dt;

    /* ---- massFluxIceOcean :200-220 ---- */
    double hifroms = 0;
    double t_new;
    { /* ThermoIce0::calculate, ThermoIce0.cpp:34-133 */
        const double freezingPointIce = -WATER_MU * ICE_S;
        const double bulkLHFusionSnow = ICE_LF * ICE_RHOSNOW;
        const double bulkLHFusionIce = ICE_LF * ICE_RHO;
        if (hice == 0 || cice == 0) { /* :45-51 */
            hi_new = 0;
            hs_new = 0;
            t_new = freezingPointIce;
        } else {
            double oldIceThickness = hiTrue;
            double k_lSlab = P->ks * ICE_KAPPA / (P->ks * hiTrue + ICE_KAPPA * hsTrue);
            double QIceConduction = k_lSlab * (tf - tice0);
            double remainingFlux = QIceConduction - Qia;
            t_new = tice0 + remainingFlux / (k_lSlab + dQ_dT);
            double meltingLimit = (hsTrue > 0.) ? 0 : freezingPointIce;
            t_new = fmin(meltingLimit, t_new); /* std::min, :66-68 */
            double snowMeltRate = fmin(-remainingFlux, 0.) / bulkLHFusionSnow;
            double snowSublRate = subl / ICE_RHOSNOW;
            hs_new += (snowMeltRate - snowSublRate) * dt;
            double excessIceMelt = fmin(hs_new, 0.) * bulkLHFusionSnow / bulkLHFusionIce;
            hs_new = fmax(hs_new, 0.);
            hs_new += snowfall * dt / ICE_RHOSNOW;
            double iceBottomChange = (QIceConduction - Qio) * dt / bulkLHFusionIce;
            double iceThicknessChange = excessIceMelt + iceBottomChange;
            hi_new += iceThicknessChange;
            double iceDraught = (hi_new * ICE_RHO + hs_new * ICE_RHOSNOW) / WATER_RHOOCEAN;
            if (P->flooding && iceDraught > hi_new) { /* :95-106 */
                double newIce = iceDraught - hi_new;
                hifroms += newIce;
                hi_new = iceDraught;
                hs_new -= newIce * ICE_RHO / ICE_RHOSNOW;
            }
            if (hi_new < P->min_thick) { /* :108-132 */
                hifroms = 0;
                double deltaQio = hi_new * bulkLHFusionIce / dt + hs_new * bulkLHFusionSnow / dt;
                Qio += deltaQio;
                hi_new = 0;
                hs_new = 0;
                t_new = freezingPointIce;
            }
            (void)oldIceThickness; /* topMelt/botMelt scaling (:110-115) has no observable effect */
        }
    }
    { /* newIceFormation, NextsimPhysics.cpp:228-254 */
        double coolingFlux = Qow;
        double deltaTml = -coolingFlux / mlBulkCp * dt;
        double t0 = sst;
        double t1 = t0 + deltaTml;
        if (t1 < tf) {
            double sensibleFlux = (tf - t0) / deltaTml * coolingFlux;
            double latentFlux = coolingFlux - sensibleFlux;
            Qow = sensibleFlux;
            newice = latentFlux * dt * (1 - cice) / (ICE_LF * ICE_RHO);
        }
    }
    double c_new;
    { /* lateralGrowth :262-289; HiblerConcentration.cpp:32-47; updateThickness :257-260 */
        double del_c = 0;
        const double ooh0 = 1. / P->h0;
        del_c += newice * ooh0;
        if (hi_new < hiTrue) {
            double m = 0;
            if (!(cice >= 1)) {
                double del_hi = hi_new - hiTrue;
                m = del_hi * cice * P->phi_m / hiTrue;
            }
            del_c += m;
        }
        c_new = cice + del_c;
        if (c_new >= P->min_conc) {
            hi_new += (newice - hi_new * del_c) / (cice + del_c);
            if (del_c < 0) {
                Qow -= del_c * hs_new * ICE_LF * ICE_RHOSNOW / dt;
            } else {
                hs_new += (0. - hs_new * del_c) / (cice + del_c);
            }
        }
    }
    /* NextsimPhysics.cpp:210-219 */
    if (c_new < P->min_conc || hi_new < P->min_thick) {
        Qow += c_new * ICE_LF * (hi_new * ICE_RHO + hs_new * ICE_RHOSNOW) / dt;
        c_new = 0;
        hi_new = 0;
        hs_new = 0;
    }

    /* ---- PrognosticData::updateAndIntegrate, PrognosticData.cpp:63-71; PhysicsData.hpp:61,66 ---- */
    *hice_ = hi_new * c_new;
    *cice_ = c_new;
    *hsnow_ = hs_new * c_new;
    *tice0_ = t_new;
    *newice_ = newice;
    if (diag) {
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
}

/* SoA sweep: nsteps x DevStep::iterate over n cells. diag: [NSDG_THERMO_NDIAG][n] or NULL. */
void nsdg_oracle_thermo_step(const nsdg_thermo_params* P, double dt, int nsteps, long n,
    double* hice, double* cice, double* hsnow, const double* sst, const double* sss, double* tice0,
    const double* tair, const double* tdew, const double* slp, const double* qsw,
    const double* qlw, const double* mld, const double* snowfall, const double* wind,
    double* newice, double* diag, int nthreads)
{
    if (nthreads < 1)
        nthreads = 1;
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (long k = 0; k < n; ++k) {
        double d[NSDG_THERMO_NDIAG];
        for (int s = 0; s < nsteps; ++s)
            nsdg_oracle_thermo_cell(P, dt, &hice[k], &cice[k], &hsnow[k], sst[k], sss[k],
                &tice0[k], tair[k], tdew[k], slp[k], qsw[k], qlw[k], mld[k], snowfall[k], wind[k],
                &newice[k], diag ? d : NULL);
        if (diag)
            for (int q = 0; q < NSDG_THERMO_NDIAG; ++q)
                diag[(size_t)q * n + k] = d[q];
    }
}

void nsdg_oracle_thermo_default_params(nsdg_thermo_params* p)
{
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
