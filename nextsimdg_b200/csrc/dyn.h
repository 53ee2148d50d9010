// Device-resident state of the dynamics (DG transport + mEVP) and the internal launchers.
//
// HBM layout.  Every array is a stack of planes; a plane is (ny_local + 2*NSDG_GHOST) rows of nx
// doubles: interior rows are [NSDG_GHOST, NSDG_GHOST + ny_local), the NSDG_GHOST rows on either
// side are ghost rows of the strip decomposition (unused when one rank owns the whole periodic
// domain: kernels then wrap row indices instead).  Two ghost rows, because the pair-fused mEVP
// kernel advances two sub-iterations per halo exchange.  DG fields are
// component-major, a[c][row][i], so that each coefficient is its own coalesced stream.  Node
// (CG1) arrays use the same plane shape; node (i,row) is the lower-left corner of cell (i,row).
#pragma once
#include "ctx.h"

constexpr int NSDG_GHOST = 2;

// Ring neighbour whose planes are mapped into this process with CUDA IPC (peer halo, comm.cu)
struct PeerMap {
    double* slab = nullptr; // neighbour's slab (22 planes of ITS plane size, then its flags)
    unsigned* flags = nullptr; // neighbour's flag array [2][nwarps]
    int64_t nyl = 0, plane = 0;
};

struct DynState {
    nsdg_dyn_config cfg;
    nsdg_mevp_params mp;
    nsdg_dyn_tuning tune = {}; // effective knobs (defaults + environment overrides + set_tuning)
    int64_t nx = 0, nyl = 0, rows = 0, plane = 0; // rows = nyl + 2*NSDG_GHOST, plane = rows*nx
    int nc = 0; // tracer components
    int rk = 0; // RK order actually used
    bool wrap = true; // single rank: periodic wrap in y inside the kernels
    double* f[NSDG_DF_COUNT] = {};
    int ncomp[NSDG_DF_COUNT] = {};
    // RK stage buffers for H and A
    double* stage1[2] = {};
    double* stage2[2] = {};
    // per-step mEVP work arrays
    double* pg = nullptr; // HALF the ice strength (P/2) at the 2x2 Gauss points, 4 planes
    double* nd = nullptr; // per-node constants of the sub-cycle, 4 planes: cx0, cy0, mdt, acd
    double* s2[3] = {}; // ping-pong partners of S11, S12, S22 (fused sub-iteration)
    double* u2 = nullptr; // ping-pong partner of U,V
    double* v2 = nullptr;
    double* flevel[2] = {}; // forcing time levels: 4 planes each (UATM, VATM, UOCN, VOCN)
    bool have_flevel[2] = { false, false };
    double* scratch = nullptr; // reductions
    // deferred host steps (nsdg_dyn_step_host_async): the velocity and stress of step k are
    // snapshotted so that they can travel to the host while step k+1 computes; forcing uploads
    // land in the alternate buffers while the current ones are still being read
    double* snap = nullptr; // 11 planes: U, V, S11[3], S12[3], S22[3]
    double* falt[4] = {}; // alternates of UATM, VATM, UOCN, VOCN
    cudaEvent_t ev_done[2] = {}; // compute of host step k finished (k & 1)
    cudaEvent_t ev_ha_down = nullptr, ev_snap_down = nullptr; // downloads of H,A / of the snapshot finished
    int64_t host_calls = 0;
    double* host_stage = nullptr; // pinned staging for upload/download
    size_t host_stage_bytes = 0;
    // halo pack buffers (comm.cu)
    double* send_lo = nullptr; // to rank-1
    double* send_hi = nullptr; // to rank+1
    double* recv_lo = nullptr; // from rank-1
    double* recv_hi = nullptr; // from rank+1
    size_t halo_capacity = 0; // doubles per buffer
    // Strips: the arrays the sub-cycle exchanges (S and its ping-pong partner, u, v and their
    // partners: NSDG_SLAB_PLANES planes) live in ONE allocation, followed by the halo flags, so
    // that a single CUDA IPC handle maps them into the neighbours' processes.  The marching
    // kernel's boundary bands then store their rows straight into the neighbour's ghost rows
    // over NVLink and publish a per-warp sequence number; the neighbour's boundary bands poll it.
    double* slab = nullptr;
    unsigned* flags = nullptr; // [2][nwarps]: [0] written by the rank below, [1] by the rank above
    int nflagw = 0; // warps per band row = ceil(nx / 28)
    unsigned halo_seq = 0; // pairs launched in peer mode (same on every rank)
    PeerMap up, dn;
    void* ipc_up = nullptr; // pointers returned by cudaIpcOpenMemHandle (dn may alias up)
    void* ipc_dn = nullptr;
    bool peer_ok = false; // every rank mapped both neighbours
    // CUDA-graph replay of the launch sequences of one step (single rank; dyn.cu): on small grids
    // a step is launch bound (100 sub-iterations = 50 pair launches of a few microseconds)
    struct Graph {
        uint64_t key[11];
        cudaGraphExec_t exec;
        int64_t launches; // kernels in the graph (nsdg_launch_count keeps counting)
        double* after[14]; // f[U], f[V], u2, v2, f[S11..S22], s2[0..2], f[H], f[A], stage1[0..1]: buffer roles after the body
    };
    std::vector<Graph> graphs;
};
constexpr int NSDG_SLAB_PLANES = 22;

struct PlaneSet { // planes handled together by one halo exchange
    double* p[12];
    int n;
};
// One group of planes of a halo exchange: fill `nlo` lower ghost rows (from the top interior
// rows of the rank below) and `nhi` upper ghost rows (from the bottom interior rows of the rank
// above); 0 <= nlo, nhi <= NSDG_GHOST.
struct HaloSpec {
    PlaneSet planes;
    int nlo, nhi;
};

// transport.cu ------------------------------------------------------------------------------
// dst = a*base + b*(src + dt*L(src)) for nfields tracers sharing the velocity; rows [r0,r1)
int nsdg_transport_stage(nsdg_ctx* ctx, int nfields, double* const* src, double* const* base,
    double* const* dst, double a, double b, double dt, int64_t r0, int64_t r1);
bool nsdg_transport_marching(const nsdg_ctx* ctx);
// two RK stages in one launch: dst = a src + b (k1 + dt L(k1)), k1 = src + dt L(src); dst != src;
// dG1 / dG2; strips: two ghost rows of src and of the velocity on either side current
int nsdg_transport_pair(nsdg_ctx* ctx, int nfields, double* const* src, double* const* dst, double a, double b, double dt,
    int64_t r0, int64_t r1);
int nsdg_transport_limit(nsdg_ctx* ctx, double* phi, double lo, double hi);

// mevp.cu -----------------------------------------------------------------------------------
// pg and the per-node constants from the current H, A, U, V for a momentum step of length dt
int nsdg_mevp_prepare(nsdg_ctx* ctx, double dt);
// split variant: stress in place on cell rows [r0,r1), then velocity on node rows [r0,r1)
int nsdg_mevp_stress(nsdg_ctx* ctx, double dt, const double* u, const double* v, double* const* S,
    int64_t r0, int64_t r1);
int nsdg_mevp_velocity(nsdg_ctx* ctx, double dt, const double* u, const double* v, double* unew,
    double* vnew, double* const* S, int64_t r0, int64_t r1);
// one fused sub-iteration: (u,v,Sold) -> (unew,vnew,Snew) on node rows [r0,r1)
int nsdg_mevp_fused(nsdg_ctx* ctx, double dt, const double* u, const double* v, double* unew,
    double* vnew, double* const* Sold, double* const* Snew, int64_t r0, int64_t r1);
// two sub-iterations in one launch (temporal blocking) on the owned rows [r0,r1); r0 must sit on
// a tile-row boundary (NSDG_GHOST + k * nsdg_mevp_pair_tile_rows(ctx))
// peer_seq != 0 (strips, marching kernel, [r0,r1) = the whole strip): the boundary bands store
// the rows the neighbours need straight into their ghost rows and publish peer_seq; before
// touching their own ghost rows they wait for the neighbours' peer_seq - 1
int nsdg_mevp_pair(nsdg_ctx* ctx, double dt, const double* u, const double* v, double* unew,
    double* vnew, double* const* Sold, double* const* Snew, int64_t r0, int64_t r1,
    cudaStream_t st = nullptr, unsigned peer_seq = 0);
int nsdg_mevp_pair_tile_rows(nsdg_ctx* ctx);
// true when the sub-cycle of this strip exchanges its halo by peer stores (mapping exists, the
// marching kernel is selected and the knob is not off)
bool nsdg_mevp_peer_halo(nsdg_ctx* ctx);
int nsdg_mevp_peer_drain(nsdg_ctx* ctx);

// comm.cu -----------------------------------------------------------------------------------
// Ring halo exchange of up to 3 plane groups in ONE message per neighbour.  No-op on a single
// rank (kernels wrap).
// st: stream to run pack / send+recv / unpack on (nullptr = the compute stream).
int nsdg_halo_exchange(nsdg_ctx* ctx, const HaloSpec* specs, int nspecs, cudaStream_t st = nullptr);
// one group: which = 1 lower ghost row only, 2 upper ghost row only, 3 both (one row each)
int nsdg_halo_exchange(nsdg_ctx* ctx, const PlaneSet& planes, int which);
bool nsdg_comm_active(const nsdg_ctx* ctx);
// after the slab of a strip has been allocated: exchange IPC handles with the ring neighbours,
// map their slabs, agree on peer_ok over all ranks.  Never fails the init for want of peer
// access: peer_ok stays false and the NCCL path is used.
int nsdg_comm_map_peers(nsdg_ctx* ctx);
void nsdg_comm_unmap_peers(nsdg_ctx* ctx);
void nsdg_comm_count_bytes(nsdg_ctx* ctx, int64_t bytes);
// effective knobs -----------------------------------------------------------------------------
void nsdg_dyn_tuning_defaults(nsdg_dyn_tuning* t); // zeros + environment overrides
