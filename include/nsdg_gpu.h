/* nsdg_gpu.h -- C ABI of the B200 (sm_100a) hot path of neXtSIM_DG.
 *
 * Plain C, plain pointers and sizes, int status codes; nothing throws across this boundary
 * and no torch/NCCL/CUDA type appears in a signature.  All floating point is FP64.
 * Host arrays are borrowed for the duration of the call only; device state is owned by the
 * context and stays resident in HBM between calls.
 *
 * Which reference interface each group replaces (paths relative to the reference tree):
 *   thermo_*    the body of DevStep::iterate, core/src/DevStep.cpp:14-23, i.e. per cell
 *               IPhysics1d::updateDerivedData (physics/src/modules/include/IPhysics1d.hpp:33-45),
 *               NextsimPhysics::calculate (physics/src/modules/NextsimPhysics.cpp:116-131) incl.
 *               ThermoIce0::calculate (physics/src/modules/ThermoIce0.cpp:34-133) and
 *               PrognosticData::updateAndIntegrate (core/src/PrognosticData.cpp:63-71);
 *               upload/download replace the AoS gather/scatter of core/src/DevGridIO.cpp:101-138,
 *               160-201 (x outer, y inner, tice layer-fastest).
 *   dyn_*       DG transport and the mEVP sub-cycle.  NOT IN THE REFERENCE SNAPSHOT ("dynamics" is
 *               commented out of its build, CMakeLists.txt:43-46): these entry points follow
 *               the discretisation stated in DESIGN.md; parity is against oracle/ only
 *               ("parity unpinned against draenog/nextsimdg").
 *   comm_*      strip decomposition halo exchange (no reference counterpart: the reference is
 *               single-process).
 */
#ifndef NSDG_GPU_H
#define NSDG_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NSDG_ABI_VERSION 3

typedef struct nsdg_ctx nsdg_ctx;

enum nsdg_status {
    NSDG_OK = 0,
    NSDG_EINVAL = -1, /* bad argument / unknown module combination (no fallback) */
    NSDG_ECUDA = -2, /* CUDA runtime error (sticky per context) */
    NSDG_ENOMEM = -3,
    NSDG_ESTATE = -4, /* call order violated (e.g. step before init) */
    NSDG_ECOMM = -5 /* NCCL error */
};

/* ---- module selectors: same names as the [Modules] section values of the reference
 * (core/src/modules/modules.json, physics/src/modules/modules.json) ---- */
enum nsdg_albedo_kind { NSDG_ALBEDO_SMU = 0, NSDG_ALBEDO_SMU2 = 1, NSDG_ALBEDO_CCSM = 2 };
enum nsdg_freezing_kind { NSDG_FREEZING_LINEAR = 0, NSDG_FREEZING_UNESCO = 1 };
enum nsdg_ocean_flux_kind { NSDG_OCEANFLUX_BASIC = 0 };

/* One parameter set per context; the reference keeps these as process-global statics
 * (NextsimPhysics.cpp:26-39,76-82; ThermoIce0.cpp:19-32; HiblerConcentration.cpp:16-30;
 * CCSMIceAlbedo.cpp:16-42).  Defaults = the reference's defaults. */
typedef struct nsdg_thermo_params {
    double drag_ocean_q; /* nextsim_thermo.drag_ocean_q = 1.5e-3 */
    double drag_ocean_t; /* nextsim_thermo.drag_ocean_t = 0.83e-3 */
    double drag_ice_t; /* nextsim_thermo.drag_ice_t   = 1.3e-3 */
    double albedo_w; /* nextsim_thermo.albedoW      = 0.07 */
    double i0; /* nextsim_thermo.I_0          = 0.17 */
    double min_conc; /* nextsim_thermo.min_conc     = 1e-12 */
    double min_thick; /* nextsim_thermo.min_thick    = 0.01 */
    double ks; /* thermoice0.ks               = 0.3096 */
    double h0; /* Hibler.h0                   = 0.25 */
    double phi_m; /* Hibler.phiM                 = 0.5 */
    double ccsm_ice_albedo; /* CCSMIceAlbedo.iceAlbedo     = 0.538 */
    double ccsm_snow_albedo; /* CCSMIceAlbedo.snowAlbedo    = 0.8256 */
    int32_t flooding; /* thermoice0.flooding         = 1 */
    int32_t albedo_kind; /* nsdg_albedo_kind, default SMU (first listed) */
    int32_t freezing_kind; /* nsdg_freezing_kind, default LINEAR */
    int32_t ocean_flux_kind; /* nsdg_ocean_flux_kind */
} nsdg_thermo_params;

/* Thermodynamics field ids for nsdg_thermo_field_ptr */
enum nsdg_thermo_field {
    NSDG_TF_HICE = 0,
    NSDG_TF_CICE,
    NSDG_TF_HSNOW,
    NSDG_TF_SST,
    NSDG_TF_SSS,
    NSDG_TF_TICE0,
    NSDG_TF_TAIR,
    NSDG_TF_TDEW,
    NSDG_TF_SLP,
    NSDG_TF_QSW,
    NSDG_TF_QLW,
    NSDG_TF_MLD,
    NSDG_TF_SNOWFALL,
    NSDG_TF_WIND,
    NSDG_TF_NEWICE, /* hidden persistent state m_newice, NextsimPhysics.hpp:257 */
    NSDG_TF_COUNT
};

/* Per-cell diagnostics of the last step, in this order (13 arrays):
 * rho, sphuma, sphumw, sphumi, cspec, tau   (PhysicsData.hpp:79-85)
 * qia, qio, qow, subl, dqdt, hifroms, tsurf_new (NextsimPhysics.hpp:219-245) */
#define NSDG_THERMO_NDIAG 13

/* ---- dynamics ---- */
enum nsdg_bc_kind {
    NSDG_BC_PERIODIC = 0, /* doubly periodic */
    NSDG_BC_BOX = 1 /* closed box: u=v=0 on the node lines i==0 and j==0 of the periodic
                       topology (== walls of an nx*ny-cell box), no flux through them */
};

typedef struct nsdg_mevp_params {
    double rho_ice; /* 900 */
    double rho_atm; /* 1.3 */
    double rho_ocean; /* 1026 */
    double c_atm; /* 1.2e-3 */
    double c_ocean; /* 5.5e-3 */
    double f_coriolis; /* 1.46e-4 */
    double p_star; /* 27.5e3 */
    double c_star; /* 20 (C in P = P* H exp(-C(1-A))) */
    double ellipse_e; /* 2 */
    double delta_min; /* 2e-9 */
    double alpha; /* 1500 */
    double beta; /* 1500 */
    double h_min; /* lower bound of nodal thickness in the momentum mass, 1e-3 */
    int32_t bc; /* nsdg_bc_kind */
    int32_t _pad;
} nsdg_mevp_params;

/* Dynamics field ids.  DG fields have ncomp in {1,3,6} coefficients per cell, stored
 * component-major: a[c][j][i] (i fastest).  Basis on the reference square, SURVEY 8a' a17:
 *   1, x-1/2, y-1/2, (x-1/2)^2-1/12, (y-1/2)^2-1/12, (x-1/2)(y-1/2).
 * CG1 node (i,j) is the lower-left corner of cell (i,j); node arrays are [j][i], periodic. */
enum nsdg_dyn_field {
    NSDG_DF_H = 0, /* DG, ncomp      ice thickness */
    NSDG_DF_A, /* DG, ncomp      ice concentration */
    NSDG_DF_U, /* node           ice velocity x */
    NSDG_DF_V, /* node           ice velocity y */
    NSDG_DF_S11, /* DG1 (3 comps)  stress */
    NSDG_DF_S12,
    NSDG_DF_S22,
    NSDG_DF_UATM, /* node           10 m wind */
    NSDG_DF_VATM,
    NSDG_DF_UOCN, /* node           ocean surface current */
    NSDG_DF_VOCN,
    NSDG_DF_COUNT
};

typedef struct nsdg_dyn_config {
    int64_t nx; /* cells in x (fast index), global == local */
    int64_t ny_global; /* cells in y over all ranks */
    int64_t ny_local; /* rows of this rank's strip */
    int64_t j0; /* global index of this strip's first row */
    double hx, hy; /* cell size [m] */
    int32_t ncomp; /* 1, 3 or 6: dG0 / dG1 / dG2 tracers */
    int32_t rk_order; /* 0 = match ncomp (1->RK1, 3->Heun RK2, 6->SSP-RK3), else 1,2,3 */
    int32_t limiter; /* 0 off (default), 1 bounds-preserving rescaling of higher moments */
    int32_t flags; /* NSDG_DYN_* bits, 0 = defaults */
} nsdg_dyn_config;

/* use the two-kernel (stress, velocity) sub-iteration instead of the fused kernel */
#define NSDG_DYN_SPLIT_MEVP 1
/* one sub-iteration per launch: do not use the pair-fused (two sub-iterations per launch) kernel */
#define NSDG_DYN_SINGLE_MEVP 2

/* Tuning knobs of the dynamics, per context (nsdg_dyn_set_tuning after nsdg_dyn_init).  0 means
 * "default" in every field, so a zeroed struct is the shipped configuration.  The environment
 * variables named beside each field override it when nsdg_dyn_init runs (A/B runs without
 * touching the caller). */
typedef struct nsdg_dyn_tuning {
    int32_t march_rows; /* rows per band of the marching pair kernel: 0 = chosen per strip
                           (nsdg_mevp_band_rows), > 0 fixed, -1 = tiled pair kernel instead
                           [NSDG_MEVP_MARCH=<rows>, 0 = tiled] */
    int32_t pair; /* two sub-iterations per launch: 0/1 on, -1 off            [NSDG_MEVP_PAIR=0/1] */
    int32_t halo_overlap; /* NCCL halo path: boundary bands first, exchange beside the interior
                             launch: 0/1 on, -1 off                        [NSDG_HALO_OVERLAP=0/1] */
    int32_t halo_peer; /* strips: boundary bands store straight into the neighbour's ghost rows
                          over NVLink peer memory (no pack / NCCL / unpack in the sub-cycle):
                          0/1 on when the peer mapping exists, -1 off         [NSDG_HALO_PEER=0/1] */
    int32_t pair_tile; /* tiled pair kernel: 0 = 16x16, 1 = 32x8, 2 = 32x16, 3 = 16x16 (3 CTAs/SM),
                          4 = 16x8, 5 = 16x12                                [NSDG_PAIR_TILE=WxH] */
    int32_t fused_tile; /* fused kernel: 0 = 32x8, 1 = 64x4, 2 = 32x16, 3 = 64x8, 4 = 16x16
                                                                            [NSDG_FUSED_TILE=WxH] */
    int32_t transport_march; /* transport stage: 0/1 marching kernel for dG2 / dG1, -1 one thread
                                per cell                                 [NSDG_TRANSPORT_MARCH=0/1] */
    int32_t graph; /* CUDA-graph replay of the sub-cycle's launches: 0 = automatic (single rank,
                      grids below 2^20 cells), 1 on, -1 off                    [NSDG_DYN_GRAPH=0/1] */
    int32_t march_ring; /* marching sub-cycle kernel, experiment arms (all bit-identical to the
                           default and measured slower, profiles/r2n_march_ring.txt,
                           r2z_march_prefetch.txt): rows ahead staged in shared memory 1 = by
                           cp.async.bulk, CTA-wide row segments (needs nx % 16 == 0), 2 = by
                           per-lane 8-byte cp.async (single rank); 3 = rows ahead pulled into L2
                           by tensor-box prefetches of the bulk-copy engine (single rank, even
                           nx); 0 / -1 = register pipeline only (default) [NSDG_MEVP_RING=0/1/2/3] */
    int32_t transport_fused; /* two Runge-Kutta stages per launch (transport_pair_march_kernel;
                                limiter off; strips exchange two ghost rows in front of it and
                                have one exchange less per step): the whole Heun step of dG1 (the tracer
                                and its stage buffer swap roles, so nsdg_dyn_field_ptr(H / A)
                                changes with every step), stages 1 + 2 of the SSP-RK3 step of dG2:
                                0 = where the marching stage kernel is the choice (transport_march),
                                1 on any grid, -1 off                  [NSDG_TRANSPORT_FUSED=0/1] */
    int32_t reserved[6]; /* must be 0 */
} nsdg_dyn_tuning;

/* ---- context ---- */
int nsdg_abi_version(void);
int nsdg_create(nsdg_ctx** out, int device);
int nsdg_destroy(nsdg_ctx* ctx);
const char* nsdg_last_error(const nsdg_ctx* ctx);
int nsdg_synchronize(nsdg_ctx* ctx);
/* number of kernels this context has launched since creation (claim for bench "gpu_launches") */
int64_t nsdg_launch_count(const nsdg_ctx* ctx);
/* time the work submitted between the two calls with CUDA events on the context's stream */
int nsdg_timer_start(nsdg_ctx* ctx);
int nsdg_timer_stop(nsdg_ctx* ctx, double* elapsed_ms);

/* per-kernel timing: while enabled every launch is bracketed by a CUDA event pair on the
 * context's stream; nsdg_profile_end resolves them.  Kernel ids 0..N-1, names via
 * nsdg_profile_kernel_name (NULL past the end). */
int nsdg_profile_begin(nsdg_ctx* ctx);
int nsdg_profile_end(nsdg_ctx* ctx);
int nsdg_profile_query(const nsdg_ctx* ctx, int32_t kernel, double* total_ms, int64_t* launches);
const char* nsdg_profile_kernel_name(int32_t kernel);

/* ---- column thermodynamics ---- */
void nsdg_thermo_default_params(nsdg_thermo_params* p);
/* One thermodynamics state and one dynamics state per context.  *_init on a context that
 * already holds a state of that kind fails with NSDG_ESTATE; *_release frees it first. */
int nsdg_thermo_init(nsdg_ctx* ctx, int64_t ncells, int32_t nlayers);
int nsdg_thermo_release(nsdg_ctx* ctx);
int nsdg_thermo_set_params(nsdg_ctx* ctx, const nsdg_thermo_params* p);
/* host -> device; tice is [ncells][nlayers] layer-fastest, only layer 0 is kept */
int nsdg_thermo_upload_prognostic(nsdg_ctx* ctx, const double* hice, const double* cice,
    const double* hsnow, const double* sst, const double* sss, const double* tice);
int nsdg_thermo_upload_forcing(nsdg_ctx* ctx, const double* tair, const double* tdew,
    const double* slp, const double* qsw, const double* qlw, const double* mld,
    const double* snowfall, const double* wind);
/* Forcing ingestion with time interpolation on the device (replaces the
 * constant fill of core/src/include/DummyExternalData.hpp:22-34): two time levels of the 8
 * forcing streams stay resident, nsdg_thermo_blend_forcing sets the active forcing to
 * f0 + w (f1 - f0), 0 <= w <= 1, without touching the host. */
int nsdg_thermo_upload_forcing_level(nsdg_ctx* ctx, int32_t level, const double* tair,
    const double* tdew, const double* slp, const double* qsw, const double* qlw,
    const double* mld, const double* snowfall, const double* wind);
int nsdg_thermo_blend_forcing(nsdg_ctx* ctx, double w);
/* Slab-ocean closure, off by default = the reference's behaviour, which
 * computes Qow and Qio and drops them (core/src/PrognosticData.cpp:63-71, quirk Q5).  When on,
 * every sweep also integrates them into the mixed layer:
 *   sst <- sst - dt ((1 - cice) Qow + cice Qio) / (mld rho_ocean c_p),  cice before the sweep. */
int nsdg_thermo_set_slab_ocean(nsdg_ctx* ctx, int32_t enable);
/* newice is zeroed by init; these two exist for checkpoint/resume of the hidden state */
int nsdg_thermo_upload_newice(nsdg_ctx* ctx, const double* newice);
int nsdg_thermo_download_newice(nsdg_ctx* ctx, double* newice);
/* nsteps sweeps of DevStep::iterate(dt) over all cells, state stays on the device */
int nsdg_thermo_step(nsdg_ctx* ctx, double dt, int32_t nsteps);
/* same, additionally storing the NSDG_THERMO_NDIAG diagnostics of the last sweep */
int nsdg_thermo_step_diag(nsdg_ctx* ctx, double dt, int32_t nsteps);
int nsdg_thermo_download_prognostic(nsdg_ctx* ctx, double* hice, double* cice, double* hsnow,
    double* sst, double* sss, double* tice);
/* diag: [NSDG_THERMO_NDIAG][ncells] */
int nsdg_thermo_download_diagnostics(nsdg_ctx* ctx, double* diag);
/* zero-copy access for callers that already hold device data (bench, coupling) */
int nsdg_thermo_field_ptr(nsdg_ctx* ctx, int32_t field, void** device_ptr);
/* one call, host buffers in and out: upload prognostic+forcing, nsteps sweeps, download */
int nsdg_thermo_run_host(nsdg_ctx* ctx, double dt, int32_t nsteps, double* hice, double* cice,
    double* hsnow, const double* sst, const double* sss, double* tice, const double* tair,
    const double* tdew, const double* slp, const double* qsw, const double* qlw,
    const double* mld, const double* snowfall, const double* wind);

/* ---- dynamics: DG transport + mEVP ---- */
void nsdg_mevp_default_params(nsdg_mevp_params* p);
int nsdg_dyn_init(nsdg_ctx* ctx, const nsdg_dyn_config* cfg);
int nsdg_dyn_release(nsdg_ctx* ctx);
/* t == NULL restores the defaults (plus environment overrides); get returns what is in effect */
int nsdg_dyn_set_tuning(nsdg_ctx* ctx, const nsdg_dyn_tuning* t);
int nsdg_dyn_get_tuning(const nsdg_ctx* ctx, nsdg_dyn_tuning* t);
int nsdg_dyn_set_mevp_params(nsdg_ctx* ctx, const nsdg_mevp_params* p);
/* interior of this rank's strip, host layout [ncomp_of_field][ny_local][nx] */
int nsdg_dyn_upload(nsdg_ctx* ctx, int32_t field, const double* host);
int nsdg_dyn_download(nsdg_ctx* ctx, int32_t field, double* host);
/* a range of interior rows [row0, row0 + nrows) of the strip, host layout [ncomp][nrows][nx]:
 * chunked I/O of strips too large to stage on the host in one piece */
int nsdg_dyn_upload_rows(nsdg_ctx* ctx, int32_t field, int64_t row0, int64_t nrows, const double* host);
int nsdg_dyn_download_rows(nsdg_ctx* ctx, int32_t field, int64_t row0, int64_t nrows, double* host);
int nsdg_dyn_field_ncomp(const nsdg_ctx* ctx, int32_t field);
/* device pointer of component 0, row 0 of the padded array: ny_local+4 rows, of which rows
 * 0,1 and ny_local+2, ny_local+3 are ghost rows of the strip decomposition and rows
 * 2..ny_local+1 the interior; pitch_elems = nx, comp_stride = (ny_local+4)*nx.  The value is valid
 * until the next stepping call: U, V and the stress alternate between two buffers in the
 * sub-cycle, H and A in fused dG1 transport steps (nsdg_dyn_tuning.transport_fused), the forcing
 * fields in deferred host steps. */
int nsdg_dyn_field_ptr(nsdg_ctx* ctx, int32_t field, void** device_ptr);
/* advance tracer `field` (NSDG_DF_H or NSDG_DF_A) by dt with the current U,V */
int nsdg_transport_step(nsdg_ctx* ctx, int32_t field, double dt);
/* advance H and A together (shared velocity reads) */
int nsdg_transport_step_HA(nsdg_ctx* ctx, double dt);
/* nsub mEVP sub-iterations for a momentum step of length dt (U,V,S updated in place) */
int nsdg_mevp_step(nsdg_ctx* ctx, double dt, int32_t nsub);
/* the north-star unit: transport of H,A with the current velocity, then the mEVP sub-cycle */
int nsdg_dyn_step(nsdg_ctx* ctx, double dt, int32_t nsub);
/* One model step with HOST buffers, copies overlapped with compute on separate streams:
 *   in[f]  (f = nsdg_dyn_field, NSDG_DF_COUNT entries, NULL = keep the resident device copy):
 *          U, V, H, A go up first and the transport starts as soon as they have arrived, while
 *          the stress and the forcing are still in flight;
 *   out[f] (NULL = not wanted): H and A come down while the mEVP sub-cycle runs, U, V and the
 *          stress after it.
 * Host layout as nsdg_dyn_upload/download.  Buffers should be page-locked (pinned) for the
 * copies to be asynchronous.  Returns after everything has completed. */
int nsdg_dyn_step_host(nsdg_ctx* ctx, double dt, int32_t nsub, const double* const* in,
    double* const* out);
/* The same step, DEFERRED: returns as soon as everything is enqueued.  A time loop that calls
 * this once per step gets the downloads of step k overlapped with the compute of step k+1 (U, V
 * and the stress leave from a device-side snapshot) and the forcing of step k+1 uploaded into
 * alternate device buffers while step k still computes.  The host buffers of a call (in and
 * out) belong to the library until nsdg_dyn_host_wait returns: use distinct out buffers for
 * consecutive steps, or wait in between.  nsdg_dyn_field_ptr values of the forcing fields change
 * with every deferred call that uploads them.  Call nsdg_dyn_host_wait before using any other
 * entry point on the dynamics state. */
int nsdg_dyn_step_host_async(nsdg_ctx* ctx, double dt, int32_t nsub, const double* const* in,
    double* const* out);
/* blocks until every copy and kernel of earlier nsdg_dyn_step_host_async calls has completed */
int nsdg_dyn_host_wait(nsdg_ctx* ctx);
/* forcing ingestion for the dynamics: two resident time levels of (UATM, VATM, UOCN, VOCN),
 * host layout [ny_local][nx] each; blend sets the active forcing to f0 + w (f1 - f0) */
int nsdg_dyn_upload_forcing_level(nsdg_ctx* ctx, int32_t level, const double* uatm,
    const double* vatm, const double* uocn, const double* vocn);
int nsdg_dyn_blend_forcing(nsdg_ctx* ctx, double w);
/* sum over this rank's cells of the mean coefficient times cell area (mass diagnostic) */
int nsdg_dyn_mass(nsdg_ctx* ctx, int32_t field, double* mass);
/* The same diagnostic for callers that cannot allocate (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Calls no cudaMalloc / cudaFree / allocating library routine and
 * needs no dynamics state: the caller passes the input and the output,
 *   mean      DEVICE pointer, ny rows of nx doubles, `pitch` elements from row to row (>= nx):
 *             e.g. nsdg_dyn_field_ptr(H) + 2 * nx for the interior of a strip, pitch = nx
 *   mass_out  DEVICE pointer to one double: hx hy sum(mean); written by work enqueued on the
 *             context's compute stream (the call does not wait: nsdg_synchronize before reading)
 * Its scratch (one partial sum per row) is a fixed __device__ array inside the library, 8 *
 * NSDG_NOALLOC_MAX_ROWS bytes on every GPU that loads the library, held until the process exits:
 * ny > NSDG_NOALLOC_MAX_ROWS fails with NSDG_EINVAL.  Every call on a GPU shares that array, so
 * calls that OVERLAP on one GPU (other contexts / streams) race; calls on one context are ordered
 * by its stream.  Nothing in it has to start at zero (each call overwrites the entries it reads).
 * The result is reproducible bit for bit (fixed summation order), which nsdg_dyn_mass's atomic
 * accumulation is not; the two agree to rounding.  The first call on each device resolves the
 * array's device address (cudaGetSymbolAddress): make it outside a stream capture. */
#define NSDG_NOALLOC_MAX_ROWS 16384
int nsdg_dyn_mass_noalloc(nsdg_ctx* ctx, const double* mean, int64_t nx, int64_t ny, int64_t pitch,
    double hx, double hy, double* mass_out);
/* rows per band the mEVP marching kernel uses for a strip of nrows x nx cells on a GPU with
 * sm_count SMs (strip != 0: part of a strip decomposition, the halo exchange runs beside the
 * interior bands).  Pure host arithmetic (no GPU needed); the environment variable
 * NSDG_MEVP_MARCH=<rows> overrides it at run time, NSDG_MEVP_MARCH=0 selects the tiled kernel. */
int nsdg_mevp_band_rows(int64_t nrows, int64_t nx, int32_t sm_count, int32_t strip);

/* ---- strip decomposition (one process per GPU) ---- */
#define NSDG_COMM_ID_BYTES 128
int nsdg_comm_get_unique_id(void* id_out /* NSDG_COMM_ID_BYTES */);
int nsdg_comm_init(nsdg_ctx* ctx, const void* id, int32_t rank, int32_t nranks);
int nsdg_comm_rank(const nsdg_ctx* ctx, int32_t* rank, int32_t* nranks);
/* bytes sent per rank by halo exchanges since creation (diagnostic; NCCL messages and peer stores) */
int64_t nsdg_comm_bytes_sent(const nsdg_ctx* ctx);
/* 1 when the sub-cycle's halo goes through NVLink peer stores from inside the marching kernel
 * (neighbours' planes mapped with CUDA IPC at nsdg_dyn_init), 0 when it goes through NCCL */
int nsdg_comm_halo_peer(const nsdg_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* NSDG_GPU_H */
