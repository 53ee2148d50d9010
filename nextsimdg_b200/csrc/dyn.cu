// Dynamics state management and time-step orchestration (transport RK stages, mEVP sub-cycle,
// halo exchanges of the strip decomposition).  See include/nsdg_gpu.h and dyn.h.
#include "dyn.h"
#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace {

int field_ncomp(const DynState* d, int field)
{
    switch (field) {
    case NSDG_DF_H:
    case NSDG_DF_A:
        return d->nc;
    case NSDG_DF_S11:
    case NSDG_DF_S12:
    case NSDG_DF_S22:
        return 3;
    default:
        return 1;
    }
}

int env_int(const char* name, int dflt)
{
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}
int env_tile(const char* name, const char* const* names, int n)
{
    const char* e = getenv(name);
    if (!e)
        return 0;
    for (int q = 0; q < n; ++q)
        if (!strcmp(e, names[q]))
            return q;
    return 0;
}
// 0/1 environment switch -> tri-state knob (1 on, -1 off), unset -> 0 (default)
int env_switch(const char* name)
{
    const char* e = getenv(name);
    return e ? (atoi(e) != 0 ? 1 : -1) : 0;
}

int alloc_planes(nsdg_ctx* ctx, double** p, int nplanes)
{
    DynState* d = ctx->dyn;
    size_t bytes = sizeof(double) * (size_t)nplanes * d->plane;
    NSDG_CUDA(ctx, cudaMalloc(p, bytes));
    NSDG_CUDA(ctx, cudaMemsetAsync(*p, 0, bytes, ctx->stream));
    return NSDG_OK;
}

__global__ void mass_kernel(const double* mean, int64_t nx, int64_t nyl, double* out)
{
    double s = 0.;
    const int64_t n = nx * nyl;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (int64_t)gridDim.x * blockDim.x)
        s += mean[NSDG_GHOST * nx + q]; // interior starts at row NSDG_GHOST
    for (int off = 16; off > 0; off >>= 1)
        s += __shfl_down_sync(0xffffffffu, s, off);
    __shared__ double part[32];
    if ((threadIdx.x & 31) == 0)
        part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        s = (threadIdx.x < (blockDim.x >> 5)) ? part[threadIdx.x] : 0.;
        for (int off = 16; off > 0; off >>= 1)
            s += __shfl_down_sync(0xffffffffu, s, off);
        if (threadIdx.x == 0)
            atomicAdd(out, s);
    }
}

// interior rows of 4 node planes: out[q] = f0[q] + w (f1[q] - f0[q])   (blockIdx.z = q)
struct Blend4 {
    double* out[4];
};
__global__ void dyn_blend_kernel(const __grid_constant__ Blend4 o, const double* f0, const double* f1, double w,
    int64_t nx, int64_t nyl, int64_t plane)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nx || (int64_t)blockIdx.y >= nyl)
        return;
    const int64_t k = (NSDG_GHOST + (int64_t)blockIdx.y) * nx + i;
    const double a = f0[blockIdx.z * plane + k];
    o.out[blockIdx.z][k] = a + w * (f1[blockIdx.z * plane + k] - a);
}

PlaneSet planes_of(DynState* d, double* base, int ncomp)
{
    PlaneSet ps;
    ps.n = 0;
    for (int c = 0; c < ncomp; ++c)
        ps.p[ps.n++] = base + (size_t)c * d->plane;
    return ps;
}
void append(PlaneSet& ps, DynState* d, double* base, int ncomp)
{
    for (int c = 0; c < ncomp; ++c)
        ps.p[ps.n++] = base + (size_t)c * d->plane;
}

// Two RK stages per launch (transport_pair_march_kernel): no limiter between the stages, and a grid
// on which the marching kernels are the choice anyway (transport_fused = 1 forces it on any grid,
// -1 switches it off).  Strips: the first launch then needs TWO ghost rows of the tracers and of
// the velocity on either side, and the step has one halo exchange less.
bool stages_fused(nsdg_ctx* ctx)
{
    const DynState* d = ctx->dyn;
    const int tf = d->tune.transport_fused;
    return d->nc >= 3 && d->rk == (d->nc == 3 ? 2 : 3) && !d->cfg.limiter && (tf > 0 || (tf == 0 && nsdg_transport_marching(ctx)));
}
// dG1 + Heun: the whole step is one launch
bool heun_fused(nsdg_ctx* ctx) { return ctx->dyn->nc == 3 && stages_fused(ctx); }

// one RK time step for the nfields tracers fld[0..nfields) (d->f[fld[f]] advanced; s1 = the
// stage-1 slots d->stage1, which trade places with the tracers in the fused Heun step)
int transport_fields(nsdg_ctx* ctx, int nfields, const int* fld, double** s1, double** s2,
    const double* lo, const double* hi, double dt)
{
    DynState* d = ctx->dyn;
    const bool multi = nsdg_comm_active(ctx);
    const int64_t r0 = NSDG_GHOST, r1 = NSDG_GHOST + d->nyl;
    int rc;
    double* phi[2] = { d->f[fld[0]], nfields > 1 ? d->f[fld[1]] : nullptr };
    // one message per neighbour: ghost node rows of the velocity and ghost rows of the tracers
    auto halo_uv_phi = [&](int nlo_uv, int nhi_uv, int nphi) -> int {
        if (!multi)
            return NSDG_OK;
        HaloSpec hs[2];
        hs[0].planes.n = 0;
        hs[0].planes.p[hs[0].planes.n++] = d->f[NSDG_DF_U];
        hs[0].planes.p[hs[0].planes.n++] = d->f[NSDG_DF_V];
        hs[0].nlo = nlo_uv;
        hs[0].nhi = nhi_uv;
        hs[1].planes.n = 0;
        for (int f = 0; f < nfields; ++f)
            append(hs[1].planes, d, phi[f], d->nc);
        hs[1].nlo = hs[1].nhi = nphi;
        return nsdg_halo_exchange(ctx, hs, 2);
    };
    auto halo = [&](double** arr) -> int {
        if (!multi)
            return NSDG_OK;
        PlaneSet ps;
        ps.n = 0;
        for (int f = 0; f < nfields; ++f)
            append(ps, d, arr[f], d->nc);
        return nsdg_halo_exchange(ctx, ps, 3);
    };
    auto limit = [&](double** arr) -> int {
        if (!d->cfg.limiter)
            return NSDG_OK;
        for (int f = 0; f < nfields; ++f) {
            int r = nsdg_transport_limit(ctx, arr[f], lo[f], hi[f]);
            if (r)
                return r;
        }
        return NSDG_OK;
    };
    if (stages_fused(ctx)) {
        // stage 1 is computed on one ghost cell row on either side: two ghost rows of the tracers,
        // the ghost node rows 1 and nyl+2, nyl+3 of the velocity
        if ((rc = halo_uv_phi(NSDG_GHOST, NSDG_GHOST, NSDG_GHOST)))
            return rc;
        if (d->nc == 3) { // Heun: the whole step in one launch; it lands in the stage buffer and the two trade places
            if ((rc = nsdg_transport_pair(ctx, nfields, phi, s1, 0.5, 0.5, dt, r0, r1)))
                return rc;
            for (int f = 0; f < nfields; ++f)
                std::swap(d->f[fld[f]], s1[f]);
            return NSDG_OK;
        }
        // SSP-RK3: stages 1 + 2 in one launch (s1 stays in registers), then stage 3
        if ((rc = nsdg_transport_pair(ctx, nfields, phi, s2, 0.75, 0.25, dt, r0, r1)) || (rc = halo(s2))
            || (rc = nsdg_transport_stage(ctx, nfields, s2, phi, phi, 1. / 3., 2. / 3., dt, r0, r1)))
            return rc;
        return NSDG_OK;
    }
    // stage 1: s1 = phi + dt L(phi).  Strips: one message per neighbour carries the upper ghost
    // node row of the velocity (read by the top cell row of every stage) and the tracers' rows.
    if ((rc = halo_uv_phi(0, 1, 1)))
        return rc;
    if ((rc = nsdg_transport_stage(ctx, nfields, phi, nullptr, s1, 0., 1., dt, r0, r1)) || (rc = limit(s1)))
        return rc;
    if (d->rk == 1) {
        // copy the interior of every component (interior rows are contiguous within a plane)
        for (int f = 0; f < nfields; ++f)
            for (int c = 0; c < d->nc; ++c)
                NSDG_CUDA(ctx, cudaMemcpyAsync(phi[f] + (size_t)c * d->plane + NSDG_GHOST * d->nx,
                                   s1[f] + (size_t)c * d->plane + NSDG_GHOST * d->nx,
                                   sizeof(double) * (size_t)d->nx * d->nyl, cudaMemcpyDeviceToDevice,
                                   ctx->stream));
        return NSDG_OK;
    }
    if (d->rk == 2) { // Heun: phi = 1/2 phi + 1/2 (s1 + dt L(s1))
        if ((rc = halo(s1)) || (rc = nsdg_transport_stage(ctx, nfields, s1, phi, phi, 0.5, 0.5, dt, r0, r1))
            || (rc = limit(phi)))
            return rc;
        return NSDG_OK;
    }
    // SSP-RK3: s2 = 3/4 phi + 1/4 (s1 + dt L(s1)); phi = 1/3 phi + 2/3 (s2 + dt L(s2))
    if ((rc = halo(s1)) || (rc = nsdg_transport_stage(ctx, nfields, s1, phi, s2, 0.75, 0.25, dt, r0, r1))
        || (rc = limit(s2)))
        return rc;
    if ((rc = halo(s2))
        || (rc = nsdg_transport_stage(ctx, nfields, s2, phi, phi, 1. / 3., 2. / 3., dt, r0, r1))
        || (rc = limit(phi)))
        return rc;
    return NSDG_OK;
}

int mevp_cycle(nsdg_ctx* ctx, double dt, int nsub)
{
    DynState* d = ctx->dyn;
    const bool multi = nsdg_comm_active(ctx);
    int rc;
    const int G = NSDG_GHOST;
    if (multi) {
        // ghost rows the sub-cycle reads: H, A (strength on the recomputed ghost cells, nodal
        // means), the stress, and every node array (velocity, forcing) -- the pair kernel works
        // on a two-row halo below and above (one row above for cell data)
        HaloSpec hs[3];
        hs[0].planes = planes_of(d, d->f[NSDG_DF_H], d->nc);
        append(hs[0].planes, d, d->f[NSDG_DF_A], d->nc);
        hs[0].nlo = 2;
        hs[0].nhi = 1;
        hs[1].planes = planes_of(d, d->f[NSDG_DF_S11], 3);
        append(hs[1].planes, d, d->f[NSDG_DF_S12], 3);
        append(hs[1].planes, d, d->f[NSDG_DF_S22], 3);
        hs[1].nlo = 2;
        hs[1].nhi = 1;
        hs[2].planes.n = 0;
        for (int q : { NSDG_DF_U, NSDG_DF_V, NSDG_DF_UATM, NSDG_DF_VATM, NSDG_DF_UOCN, NSDG_DF_VOCN })
            hs[2].planes.p[hs[2].planes.n++] = d->f[q];
        hs[2].nlo = hs[2].nhi = 2;
        // one message per neighbour: 54 rows up, 33 down at dG2
        if ((rc = nsdg_halo_exchange(ctx, hs, 3)))
            return rc;
    }
    if ((rc = nsdg_mevp_prepare(ctx, dt)))
        return rc;
    double *u = d->f[NSDG_DF_U], *v = d->f[NSDG_DF_V], *u2 = d->u2, *v2 = d->v2;
    const int64_t s0 = multi ? G - 1 : G, r0 = G, r1 = G + d->nyl;
    const bool fused = !(d->cfg.flags & NSDG_DYN_SPLIT_MEVP);
    // two sub-iterations per launch (and per halo exchange) while at least two remain
    const bool pair = d->tune.pair >= 0 && fused && !(d->cfg.flags & NSDG_DYN_SINGLE_MEVP);
    double* sa[3] = { d->f[NSDG_DF_S11], d->f[NSDG_DF_S12], d->f[NSDG_DF_S22] };
    double* sb[3] = { d->s2[0], d->s2[1], d->s2[2] };
    // Strips: the tile rows next to the ghost rows go first; their halo messages travel on the
    // communication stream while the interior tile rows compute (NSDG_HALO_OVERLAP=0 disables).
    const int ty = nsdg_mevp_pair_tile_rows(ctx);
    const int64_t ntr = (d->nyl + ty - 1) / ty; // tile rows of the strip
    // the two top interior rows are sent up: they must lie inside the last boundary launch
    const int64_t ntop = (d->nyl - (ntr - 1) * ty >= 2) ? 1 : 2;
    // Peer halo: ONE launch per pair; the marching kernel's boundary bands store their rows into
    // the neighbours' ghost rows over NVLink and the neighbours' boundary bands of the next pair
    // poll a per-warp sequence number (mevp.cu).  No pack / NCCL / unpack, no second stream.
    const bool peer = multi && pair && nsdg_mevp_peer_halo(ctx);
    const bool overlap = multi && pair && !peer && d->tune.halo_overlap >= 0 && ntr >= ntop + 2 && ty >= 2;
    bool peer_pending = false; // neighbours may still be storing into this rank's ghost rows
    for (int p = 0; p < nsub; ++p) {
        bool paired = false;
        if (peer_pending && !(pair && p + 1 < nsub)) {
            // the next kernel is not a peer pair (which polls the flags itself)
            if ((rc = nsdg_mevp_peer_drain(ctx)))
                return rc;
            peer_pending = false;
        }
        int64_t rest0 = r0, rest1 = r0; // interior rows that run beside the exchange
        if (pair && p + 1 < nsub) {
            if (peer) {
                if ((rc = nsdg_mevp_pair(ctx, dt, u, v, u2, v2, sa, sb, r0, r1, nullptr, ++d->halo_seq)))
                    return rc;
                // rows this rank stored into its neighbours: S 2 up + 1 down, u, v 2 + 2
                nsdg_comm_count_bytes(ctx, (int64_t)(9 * 3 + 2 * 4) * d->nx * 8);
                peer_pending = true;
            } else if (overlap) {
                // boundary tile rows on the (high-priority) communication stream, behind
                // everything issued so far
                rest0 = r0 + ty;
                rest1 = r0 + (ntr - ntop) * ty;
                NSDG_CUDA(ctx, cudaEventRecord(ctx->ev_halo[0], ctx->stream));
                NSDG_CUDA(ctx, cudaStreamWaitEvent(ctx->comm_stream, ctx->ev_halo[0], 0));
                if ((rc = nsdg_mevp_pair(ctx, dt, u, v, u2, v2, sa, sb, r0, rest0, ctx->comm_stream))
                    || (rc = nsdg_mevp_pair(ctx, dt, u, v, u2, v2, sa, sb, rest1, r1, ctx->comm_stream))
                    || (rc = nsdg_mevp_pair(ctx, dt, u, v, u2, v2, sa, sb, rest0, rest1)))
                    return rc;
            } else if ((rc = nsdg_mevp_pair(ctx, dt, u, v, u2, v2, sa, sb, r0, r1)))
                return rc;
            for (int q = 0; q < 3; ++q)
                std::swap(sa[q], sb[q]);
            ++p;
            paired = true;
        } else if (fused) {
            if ((rc = nsdg_mevp_fused(ctx, dt, u, v, u2, v2, sa, sb, r0, r1)))
                return rc;
            for (int q = 0; q < 3; ++q)
                std::swap(sa[q], sb[q]);
        } else {
            if ((rc = nsdg_mevp_stress(ctx, dt, u, v, sa, s0, r1)))
                return rc;
            if ((rc = nsdg_mevp_velocity(ctx, dt, u, v, u2, v2, sa, r0, r1)))
                return rc;
        }
        if (multi && !(peer && paired)) {
            // after a pair the next pair needs the new stress on its ghost cells as well; a single
            // sub-iteration keeps the first lower ghost row of S current by itself
            HaloSpec hs[2];
            int n = 0;
            if (paired) {
                hs[n].planes = planes_of(d, sa[0], 3);
                append(hs[n].planes, d, sa[1], 3);
                append(hs[n].planes, d, sa[2], 3);
                hs[n].nlo = 2;
                hs[n].nhi = 1;
                ++n;
            }
            hs[n].planes.n = 0;
            hs[n].planes.p[hs[n].planes.n++] = u2;
            hs[n].planes.p[hs[n].planes.n++] = v2;
            hs[n].nlo = hs[n].nhi = paired ? 2 : 1;
            ++n;
            if (rest1 > rest0) {
                // behind the boundary tiles on the communication stream, beside the interior tiles
                if ((rc = nsdg_halo_exchange(ctx, hs, n, ctx->comm_stream)))
                    return rc;
                NSDG_CUDA(ctx, cudaEventRecord(ctx->ev_halo[1], ctx->comm_stream));
                NSDG_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_halo[1], 0));
            } else if ((rc = nsdg_halo_exchange(ctx, hs, n)))
                return rc;
        }
        std::swap(u, u2);
        std::swap(v, v2);
    }
    if (peer_pending && (rc = nsdg_mevp_peer_drain(ctx)))
        return rc;
    d->f[NSDG_DF_U] = u;
    d->f[NSDG_DF_V] = v;
    d->u2 = u2;
    d->v2 = v2;
    d->f[NSDG_DF_S11] = sa[0];
    d->f[NSDG_DF_S12] = sa[1];
    d->f[NSDG_DF_S22] = sa[2];
    for (int q = 0; q < 3; ++q)
        d->s2[q] = sb[q];
    return NSDG_OK;
}

void drop_graphs(DynState* d)
{
    for (auto& g : d->graphs)
        cudaGraphExecDestroy(g.exec);
    d->graphs.clear();
}

// Runs body() -- a sequence of launches on ctx->stream that depends only on (kind, field, dt,
// nsub) and on which physical buffers currently play which role -- either directly or as a
// captured CUDA graph that is replayed on later calls with the same key.  The body also swaps
// host-side buffer roles (ping-pong); those are recorded with the graph and re-applied on replay.
template <class Body> int graphed(nsdg_ctx* ctx, int kind, int field, double dt, int nsub, Body body)
{
    DynState* d = ctx->dyn;
    const bool use = !nsdg_comm_active(ctx) && !ctx->prof.on
        && (d->tune.graph > 0 || (d->tune.graph == 0 && d->nx * d->nyl < ((int64_t)1 << 20)));
    if (!use)
        return body();
    uint64_t dtbits;
    memcpy(&dtbits, &dt, 8);
    const uint64_t key[11] = { (uint64_t)kind, (uint64_t)field, (uint64_t)nsub, dtbits,
        (uint64_t)(uintptr_t)d->f[NSDG_DF_U], (uint64_t)(uintptr_t)d->f[NSDG_DF_S11],
        // the forcing buffers swap with their alternates in deferred host steps
        (uint64_t)(uintptr_t)d->f[NSDG_DF_UATM] ^ ((uint64_t)(uintptr_t)d->f[NSDG_DF_VATM] << 1)
            ^ ((uint64_t)(uintptr_t)d->f[NSDG_DF_UOCN] << 2) ^ ((uint64_t)(uintptr_t)d->f[NSDG_DF_VOCN] << 3),
        // the tracers swap with their stage buffers in fused Heun steps
        (uint64_t)(uintptr_t)d->f[NSDG_DF_H], (uint64_t)(uintptr_t)d->f[NSDG_DF_A], (uint64_t)(uintptr_t)d->stage1[0],
        (uint64_t)(uintptr_t)d->stage1[1] };
    auto roles = [&](double** r) {
        r[0] = d->f[NSDG_DF_U];
        r[1] = d->f[NSDG_DF_V];
        r[2] = d->u2;
        r[3] = d->v2;
        r[4] = d->f[NSDG_DF_S11];
        r[5] = d->f[NSDG_DF_S12];
        r[6] = d->f[NSDG_DF_S22];
        r[7] = d->s2[0];
        r[8] = d->s2[1];
        r[9] = d->s2[2];
        r[10] = d->f[NSDG_DF_H];
        r[11] = d->f[NSDG_DF_A];
        r[12] = d->stage1[0];
        r[13] = d->stage1[1];
    };
    for (auto& g : d->graphs)
        if (!memcmp(g.key, key, sizeof key)) {
            NSDG_CUDA(ctx, cudaGraphLaunch(g.exec, ctx->stream));
            ctx->launches += g.launches;
            d->f[NSDG_DF_U] = g.after[0];
            d->f[NSDG_DF_V] = g.after[1];
            d->u2 = g.after[2];
            d->v2 = g.after[3];
            d->f[NSDG_DF_S11] = g.after[4];
            d->f[NSDG_DF_S12] = g.after[5];
            d->f[NSDG_DF_S22] = g.after[6];
            d->s2[0] = g.after[7];
            d->s2[1] = g.after[8];
            d->s2[2] = g.after[9];
            d->f[NSDG_DF_H] = g.after[10];
            d->f[NSDG_DF_A] = g.after[11];
            d->stage1[0] = g.after[12];
            d->stage1[1] = g.after[13];
            return NSDG_OK;
        }
    const int64_t l0 = ctx->launches;
    NSDG_CUDA(ctx, cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal));
    const int rc = body();
    cudaGraph_t graph = nullptr;
    const cudaError_t e = cudaStreamEndCapture(ctx->stream, &graph);
    if (rc || e != cudaSuccess) {
        if (graph)
            cudaGraphDestroy(graph);
        (void)cudaGetLastError();
        return rc ? rc : nsdg_fail(ctx, NSDG_ECUDA, "graph capture failed: %s", cudaGetErrorString(e));
    }
    DynState::Graph g = {};
    memcpy(g.key, key, sizeof key);
    g.launches = ctx->launches - l0;
    roles(g.after);
    const cudaError_t ei = cudaGraphInstantiate(&g.exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ei != cudaSuccess)
        return nsdg_fail(ctx, NSDG_ECUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(ei));
    if (d->graphs.size() >= 16) // a time loop needs one or two; anything else is churn
        drop_graphs(d);
    d->graphs.push_back(g);
    NSDG_CUDA(ctx, cudaGraphLaunch(g.exec, ctx->stream)); // capture records, it does not run
    return NSDG_OK;
}

} // namespace

void nsdg_dyn_tuning_defaults(nsdg_dyn_tuning* t)
{
    static const char* const pair_tiles[] = { "16x16", "32x8", "32x16", "16x16m3", "16x8", "16x12" };
    static const char* const fused_tiles[] = { "32x8", "64x4", "32x16", "64x8", "16x16" };
    *t = nsdg_dyn_tuning {};
    const int m = env_int("NSDG_MEVP_MARCH", -1);
    t->march_rows = (m < 0 || m > 4096) ? 0 : (m == 0 ? -1 : m);
    t->pair = env_switch("NSDG_MEVP_PAIR");
    t->halo_overlap = env_switch("NSDG_HALO_OVERLAP");
    t->halo_peer = env_switch("NSDG_HALO_PEER");
    t->pair_tile = env_tile("NSDG_PAIR_TILE", pair_tiles, 6);
    t->fused_tile = env_tile("NSDG_FUSED_TILE", fused_tiles, 5);
    t->transport_march = env_switch("NSDG_TRANSPORT_MARCH");
    t->graph = env_switch("NSDG_DYN_GRAPH");
    t->transport_fused = env_switch("NSDG_TRANSPORT_FUSED");
    t->march_ring = env_int("NSDG_MEVP_RING", 0);
    if (t->march_ring < -1 || t->march_ring > 3)
        t->march_ring = 0;
}

void nsdg_dyn_free(nsdg_ctx* ctx)
{
    DynState* d = ctx->dyn;
    if (!d)
        return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    cudaStreamSynchronize(ctx->comm_stream);
    drop_graphs(d);
    if (d->slab) { // the slab owns S, its partners, u, v and their partners
        nsdg_comm_unmap_peers(ctx);
        for (int q : { NSDG_DF_U, NSDG_DF_V, NSDG_DF_S11, NSDG_DF_S12, NSDG_DF_S22 })
            d->f[q] = nullptr;
        d->u2 = d->v2 = nullptr;
        d->s2[0] = d->s2[1] = d->s2[2] = nullptr;
        cudaFree(d->slab);
    }
    for (int q = 0; q < NSDG_DF_COUNT; ++q)
        cudaFree(d->f[q]);
    cudaFree(d->snap);
    for (double* p : d->falt)
        cudaFree(p);
    for (cudaEvent_t e : { d->ev_done[0], d->ev_done[1], d->ev_ha_down, d->ev_snap_down })
        if (e)
            cudaEventDestroy(e);
    for (int f = 0; f < 2; ++f) {
        cudaFree(d->stage1[f]);
        cudaFree(d->stage2[f]);
    }
    for (int q = 0; q < 3; ++q)
        cudaFree(d->s2[q]);
    double* w[] = { d->pg, d->nd, d->u2, d->v2, d->scratch, d->flevel[0], d->flevel[1],
        d->send_lo, d->send_hi, d->recv_lo, d->recv_hi };
    for (double* p : w)
        cudaFree(p);
    if (d->host_stage)
        cudaFreeHost(d->host_stage);
    delete d;
    ctx->dyn = nullptr;
}

extern "C" {

void nsdg_mevp_default_params(nsdg_mevp_params* p)
{
    if (!p)
        return;
    p->rho_ice = 900.;
    p->rho_atm = 1.3;
    p->rho_ocean = 1026.;
    p->c_atm = 1.2e-3;
    p->c_ocean = 5.5e-3;
    p->f_coriolis = 1.46e-4;
    p->p_star = 27.5e3;
    p->c_star = 20.;
    p->ellipse_e = 2.;
    p->delta_min = 2e-9;
    p->alpha = 1500.;
    p->beta = 1500.;
    p->h_min = 1e-3;
    p->bc = NSDG_BC_PERIODIC;
    p->_pad = 0;
}

int nsdg_dyn_init(nsdg_ctx* ctx, const nsdg_dyn_config* cfg)
{
    NSDG_ENTER(ctx);
    NSDG_REQUIRE(ctx, cfg, NSDG_EINVAL, "cfg is null");
    NSDG_REQUIRE(ctx, cfg->nx >= 2 && cfg->ny_local >= 2 && cfg->ny_global >= cfg->ny_local
            && cfg->j0 >= 0 && cfg->j0 + cfg->ny_local <= cfg->ny_global,
        NSDG_EINVAL, "bad grid: nx=%lld ny_local=%lld ny_global=%lld j0=%lld", (long long)cfg->nx,
        (long long)cfg->ny_local, (long long)cfg->ny_global, (long long)cfg->j0);
    NSDG_REQUIRE(ctx, cfg->hx > 0 && cfg->hy > 0, NSDG_EINVAL, "cell sizes must be positive");
    NSDG_REQUIRE(ctx, cfg->ncomp == 1 || cfg->ncomp == 3 || cfg->ncomp == 6, NSDG_EINVAL,
        "ncomp must be 1 (dG0), 3 (dG1) or 6 (dG2), got %d", cfg->ncomp);
    NSDG_REQUIRE(ctx, cfg->rk_order >= 0 && cfg->rk_order <= 3, NSDG_EINVAL, "rk_order in 0..3");
    NSDG_REQUIRE(ctx, (cfg->flags & ~(NSDG_DYN_SPLIT_MEVP | NSDG_DYN_SINGLE_MEVP)) == 0, NSDG_EINVAL,
        "unknown flags 0x%x", cfg->flags);
    // the kernels index every array with 32-bit integers
    NSDG_REQUIRE(ctx, (cfg->ny_local + 2 * NSDG_GHOST) * cfg->nx * (cfg->ncomp > 4 ? cfg->ncomp : 4) < (int64_t)1 << 31,
        NSDG_EINVAL, "strip too large for 32-bit indexing: %lld x %lld cells, %d components",
        (long long)cfg->nx, (long long)cfg->ny_local, cfg->ncomp);
    const bool multi = nsdg_comm_active(ctx);
    NSDG_REQUIRE(ctx, multi || cfg->ny_local == cfg->ny_global, NSDG_EINVAL,
        "a strip (ny_local < ny_global) needs nsdg_comm_init before nsdg_dyn_init");
    NSDG_REQUIRE(ctx, !ctx->dyn, NSDG_ESTATE,
        "this context already holds a dynamics state: call nsdg_dyn_release first");
    DynState* d = new DynState;
    ctx->dyn = d;
    d->cfg = *cfg;
    nsdg_dyn_tuning_defaults(&d->tune);
    nsdg_mevp_default_params(&d->mp);
    d->nx = cfg->nx;
    d->nyl = cfg->ny_local;
    d->rows = d->nyl + 2 * NSDG_GHOST;
    d->plane = d->rows * d->nx;
    d->nc = cfg->ncomp;
    d->rk = cfg->rk_order ? cfg->rk_order : (d->nc == 1 ? 1 : (d->nc == 3 ? 2 : 3));
    d->wrap = !multi;
    int rc;
    const bool fused_bufs = !(cfg->flags & NSDG_DYN_SPLIT_MEVP);
    if (multi) {
        // one allocation (= one IPC handle) for everything the sub-cycle exchanges, then the flags
        d->nflagw = (int)((d->nx + 27) / 28);
        const size_t bytes = sizeof(double) * (size_t)NSDG_SLAB_PLANES * d->plane + sizeof(unsigned) * 2 * (size_t)d->nflagw;
        NSDG_CUDA(ctx, cudaMalloc(&d->slab, bytes));
        NSDG_CUDA(ctx, cudaMemsetAsync(d->slab, 0, bytes, ctx->stream));
        d->flags = reinterpret_cast<unsigned*>(d->slab + (size_t)NSDG_SLAB_PLANES * d->plane);
        double* q = d->slab;
        for (int f : { NSDG_DF_S11, NSDG_DF_S12, NSDG_DF_S22 }) {
            d->f[f] = q;
            q += 3 * d->plane;
        }
        for (int k = 0; k < 3; ++k) {
            if (fused_bufs)
                d->s2[k] = q;
            q += 3 * d->plane;
        }
        d->f[NSDG_DF_U] = q;
        d->f[NSDG_DF_V] = q + d->plane;
        d->u2 = q + 2 * d->plane;
        d->v2 = q + 3 * d->plane;
    }
    for (int q = 0; q < NSDG_DF_COUNT; ++q) {
        d->ncomp[q] = field_ncomp(d, q);
        if (!d->f[q] && (rc = alloc_planes(ctx, &d->f[q], d->ncomp[q])))
            return rc;
    }
    for (int f = 0; f < 2; ++f) {
        if ((rc = alloc_planes(ctx, &d->stage1[f], d->nc)))
            return rc;
        if (d->rk == 3 && (rc = alloc_planes(ctx, &d->stage2[f], d->nc)))
            return rc;
    }
    double** one[] = { &d->u2, &d->v2 };
    for (double** p : one)
        if (!*p && (rc = alloc_planes(ctx, p, 1)))
            return rc;
    if ((rc = alloc_planes(ctx, &d->pg, 4)) || (rc = alloc_planes(ctx, &d->nd, 4)))
        return rc;
    if (fused_bufs)
        for (int q = 0; q < 3; ++q)
            if (!d->s2[q] && (rc = alloc_planes(ctx, &d->s2[q], 3)))
                return rc;
    NSDG_CUDA(ctx, cudaMalloc(&d->scratch, 64));
    if (multi) {
        d->halo_capacity = 64 * (size_t)d->nx;
        double** hb[] = { &d->send_lo, &d->send_hi, &d->recv_lo, &d->recv_hi };
        for (double** p : hb)
            NSDG_CUDA(ctx, cudaMalloc(p, sizeof(double) * d->halo_capacity));
    }
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (multi && (rc = nsdg_comm_map_peers(ctx)))
        return rc;
    return NSDG_OK;
}

int nsdg_dyn_release(nsdg_ctx* ctx)
{
    NSDG_ENTER(ctx);
    nsdg_dyn_free(ctx);
    return NSDG_OK;
}

int nsdg_dyn_set_tuning(nsdg_ctx* ctx, const nsdg_dyn_tuning* t)
{
    NSDG_ENTER(ctx);
    DynState* d = ctx->dyn;
    NSDG_REQUIRE(ctx, d, NSDG_ESTATE, "nsdg_dyn_init has not been called");
    if (!t) {
        nsdg_dyn_tuning_defaults(&d->tune);
        return NSDG_OK;
    }
    NSDG_REQUIRE(ctx, t->march_rows >= -1 && t->march_rows <= 4096, NSDG_EINVAL, "march_rows in -1..4096");
    NSDG_REQUIRE(ctx, t->pair_tile >= 0 && t->pair_tile <= 5 && t->fused_tile >= 0 && t->fused_tile <= 4,
        NSDG_EINVAL, "unknown tile id");
    for (int q : { t->pair, t->halo_overlap, t->halo_peer, t->transport_march, t->graph, t->transport_fused })
        NSDG_REQUIRE(ctx, q >= -1 && q <= 1, NSDG_EINVAL, "switches take -1, 0 or 1");
    NSDG_REQUIRE(ctx, t->march_ring >= -1 && t->march_ring <= 3, NSDG_EINVAL, "march_ring takes -1, 0, 1, 2 or 3");
    for (int q : t->reserved)
        NSDG_REQUIRE(ctx, q == 0, NSDG_EINVAL, "reserved fields must be 0");
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); // knobs never change under work in flight
    drop_graphs(d);
    d->tune = *t;
    return NSDG_OK;
}

int nsdg_dyn_get_tuning(const nsdg_ctx* ctx, nsdg_dyn_tuning* t)
{
    if (!ctx || !ctx->dyn || !t)
        return NSDG_EINVAL;
    *t = ctx->dyn->tune;
    return NSDG_OK;
}

int nsdg_dyn_set_mevp_params(nsdg_ctx* ctx, const nsdg_mevp_params* p)
{
    NSDG_ENTER(ctx);
    NSDG_REQUIRE(ctx, ctx->dyn, NSDG_ESTATE, "nsdg_dyn_init has not been called");
    NSDG_REQUIRE(ctx, p, NSDG_EINVAL, "params is null");
    NSDG_REQUIRE(ctx, p->bc == NSDG_BC_PERIODIC || p->bc == NSDG_BC_BOX, NSDG_EINVAL,
        "unknown boundary condition %d", p->bc);
    NSDG_REQUIRE(ctx, p->alpha > 0 && p->beta >= 0 && p->ellipse_e > 0 && p->delta_min > 0,
        NSDG_EINVAL, "alpha, ellipse_e, delta_min must be positive");
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    drop_graphs(ctx->dyn); // kernel arguments are baked into captured launches
    ctx->dyn->mp = *p;
    return NSDG_OK;
}

int nsdg_dyn_field_ncomp(const nsdg_ctx* ctx, int32_t field)
{
    if (!ctx || !ctx->dyn || field < 0 || field >= NSDG_DF_COUNT)
        return NSDG_EINVAL;
    return ctx->dyn->ncomp[field];
}

int nsdg_dyn_upload(nsdg_ctx* ctx, int32_t field, const double* host)
{
    NSDG_ENTER(ctx);
    DynState* d = ctx->dyn;
    NSDG_REQUIRE(ctx, d, NSDG_ESTATE, "nsdg_dyn_init has not been called");
    NSDG_REQUIRE(ctx, field >= 0 && field < NSDG_DF_COUNT && host, NSDG_EINVAL, "bad field/host");
    const size_t inner = (size_t)d->nx * d->nyl;
    for (int c = 0; c < d->ncomp[field]; ++c)
        NSDG_CUDA(ctx, cudaMemcpyAsync(d->f[field] + (size_t)c * d->plane + NSDG_GHOST * d->nx, host + c * inner,
                           sizeof(double) * inner, cudaMemcpyHostToDevice, ctx->stream));
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return NSDG_OK;
}

int nsdg_dyn_download(nsdg_ctx* ctx, int32_t field, double* host)
{
    NSDG_ENTER(ctx);
    DynState* d = ctx->dyn;
    NSDG_REQUIRE(ctx, d, NSDG_ESTATE, "nsdg_dyn_init has not been called");
    NSDG_REQUIRE(ctx, field >= 0 && field < NSDG_DF_COUNT && host, NSDG_EINVAL, "bad field/host");
    const size_t inner = (size_t)d->nx * d->nyl;
    for (int c = 0; c < d->ncomp[field]; ++c)
        NSDG_CUDA(ctx, cudaMemcpyAsync(host + c * inner, d->f[field] + (size_t)c * d->plane + NSDG_GHOST * d->nx,
                           sizeof(double) * inner, cudaMemcpyDeviceToHost, ctx->stream));
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return NSDG_OK;
}

// rows [row0, row0 + nrows) of the strip's interior; host layout [ncomp][nrows][nx]
int nsdg_dyn_upload_rows(nsdg_ctx* ctx, int32_t field, int64_t row0, int64_t nrows, const double* host)
{
    NSDG_ENTER(ctx);
    DynState* d = ctx->dyn;
    NSDG_REQUIRE(ctx, d, NSDG_ESTATE, "nsdg_dyn_init has not been called");
    NSDG_REQUIRE(ctx, field >= 0 && field < NSDG_DF_COUNT && host, NSDG_EINVAL, "bad field/host");
    NSDG_REQUIRE(ctx, row0 >= 0 && nrows >= 0 && row0 + nrows <= d->nyl, NSDG_EINVAL,
        "rows [%lld, %lld) outside the strip of %lld rows", (long long)row0, (long long)(row0 + nrows), (long long)d->nyl);
    const size_t chunk = (size_t)d->nx * nrows;
    for (int c = 0; c < d->ncomp[field] && chunk; ++c)
        NSDG_CUDA(ctx, cudaMemcpyAsync(d->f[field] + (size_t)c * d->plane + (NSDG_GHOST + row0) * d->nx, host + c * chunk,
                           sizeof(double) * chunk, cudaMemcpyHostToDevice, ctx->stream));
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return NSDG_OK;
}

int nsdg_dyn_download_rows(nsdg_ctx* ctx, int32_t field, int64_t row0, int64_t nrows, double* host)
{
    NSDG_ENTER(ctx);
    DynState* d = ctx->dyn;
    NSDG_REQUIRE(ctx, d, NSDG_ESTATE, "nsdg_dyn_init has not been called");
    NSDG_REQUIRE(ctx, field >= 0 && field < NSDG_DF_COUNT && host, NSDG_EINVAL, "bad field/host");
    NSDG_REQUIRE(ctx, row0 >= 0 && nrows >= 0 && row0 + nrows <= d->nyl, NSDG_EINVAL,
        "rows [%lld, %lld) outside the strip of %lld rows", (long long)row0, (long long)(row0 + nrows), (long long)d->nyl);
    const size_t chunk = (size_t)d->nx * nrows;
    for (int c = 0; c < d->ncomp[field] && chunk; ++c)
        NSDG_CUDA(ctx, cudaMemcpyAsync(host + c * chunk, d->f[field] + (size_t)c * d->plane + (NSDG_GHOST + row0) * d->nx,
                           sizeof(double) * chunk, cudaMemcpyDeviceToHost, ctx->stream));
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return NSDG_OK;
}

int nsdg_dyn_field_ptr(nsdg_ctx* ctx, int32_t field, void** device_ptr)
{
    NSDG_ENTER(ctx);
    NSDG_REQUIRE(ctx, ctx->dyn, NSDG_ESTATE, "nsdg_dyn_init has not been called");
    NSDG_REQUIRE(ctx, field >= 0 && field < NSDG_DF_COUNT && device_ptr, NSDG_EINVAL, "bad field");
    *device_ptr = ctx->dyn->f[field];
    return NSDG_OK;
}

int nsdg_transport_step(nsdg_ctx* ctx, int32_t field, double dt)
{
    NSDG_ENTER(ctx);
    DynState* d = ctx->dyn;
    NSDG_REQUIRE(ctx, d, NSDG_ESTATE, "nsdg_dyn_init has not been called");
    NSDG_REQUIRE(ctx, field == NSDG_DF_H || field == NSDG_DF_A, NSDG_EINVAL,
        "only the tracers H and A can be transported");
    auto body = [&]() {
        const int fld[1] = { field };
        const double lo[1] = { 0. }, hi[1] = { field == NSDG_DF_A ? 1. : 1e300 };
        return transport_fields(ctx, 1, fld, d->stage1, d->stage2, lo, hi, dt);
    };
    if (heun_fused(ctx) && d->tune.graph <= 0) // a single launch: replaying a graph of it only costs host time
        return body();
    return graphed(ctx, 1, field, dt, 0, body);
}

int nsdg_transport_step_HA(nsdg_ctx* ctx, double dt)
{
    NSDG_ENTER(ctx);
    DynState* d = ctx->dyn;
    NSDG_REQUIRE(ctx, d, NSDG_ESTATE, "nsdg_dyn_init has not been called");
    auto body = [&]() {
        const int fld[2] = { NSDG_DF_H, NSDG_DF_A };
        const double lo[2] = { 0., 0. }, hi[2] = { 1e300, 1. };
        return transport_fields(ctx, 2, fld, d->stage1, d->stage2, lo, hi, dt);
    };
    if (heun_fused(ctx) && d->tune.graph <= 0)
        return body();
    return graphed(ctx, 0, 0, dt, 0, body);
}

int nsdg_mevp_step(nsdg_ctx* ctx, double dt, int32_t nsub)
{
    NSDG_ENTER(ctx);
    NSDG_REQUIRE(ctx, ctx->dyn, NSDG_ESTATE, "nsdg_dyn_init has not been called");
    NSDG_REQUIRE(ctx, dt > 0 && nsub >= 0, NSDG_EINVAL, "dt > 0 and nsub >= 0 required");
    return graphed(ctx, 2, 0, dt, nsub, [&]() { return mevp_cycle(ctx, dt, nsub); });
}

int nsdg_dyn_step(nsdg_ctx* ctx, double dt, int32_t nsub)
{
    int rc = nsdg_transport_step_HA(ctx, dt);
    if (rc)
        return rc;
    return nsdg_mevp_step(ctx, dt, nsub);
}

// One model step with host buffers.  deferred = false: returns with every output on the host.
// deferred = true: returns once everything is enqueued; the downloads of step k overlap the
// compute of step k+1 and the forcing of step k+1 goes up while step k still computes.
static int step_host_impl(nsdg_ctx* ctx, double dt, int32_t nsub, const double* const* in,
    double* const* out, bool deferred)
{
    DynState* d = ctx->dyn;
    NSDG_REQUIRE(ctx, d, NSDG_ESTATE, "nsdg_dyn_init has not been called");
    NSDG_REQUIRE(ctx, dt > 0 && nsub >= 0, NSDG_EINVAL, "dt > 0 and nsub >= 0 required");
    const size_t inner = (size_t)d->nx * d->nyl;
    if (!d->ev_done[0]) {
        cudaEvent_t* evs[] = { &d->ev_done[0], &d->ev_done[1], &d->ev_ha_down, &d->ev_snap_down };
        for (cudaEvent_t* e : evs)
            NSDG_CUDA(ctx, cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    }
    const int64_t k = d->host_calls++;
    const int forcing[4] = { NSDG_DF_UATM, NSDG_DF_VATM, NSDG_DF_UOCN, NSDG_DF_VOCN };
    const int snapf[5] = { NSDG_DF_U, NSDG_DF_V, NSDG_DF_S11, NSDG_DF_S12, NSDG_DF_S22 };
    const int snapoff[5] = { 0, 1, 2, 5, 8 };
    int rc;
    if (deferred) {
        bool fresh = false;
        if (out && !d->snap) {
            if ((rc = alloc_planes(ctx, &d->snap, 11)))
                return rc;
            fresh = true;
        }
        for (int q = 0; q < 4; ++q)
            if (in && in[forcing[q]] && !d->falt[q]) {
                if ((rc = alloc_planes(ctx, &d->falt[q], 1)))
                    return rc;
                fresh = true;
            }
        if (fresh) {
            // alloc_planes clears the new buffers on the compute stream; the forcing upload below
            // goes on the copy stream and must not overtake that memset (it would be zeroed)
            NSDG_CUDA(ctx, cudaEventRecord(ctx->ev_pipe[0], ctx->stream));
            NSDG_CUDA(ctx, cudaStreamWaitEvent(ctx->h2d_stream, ctx->ev_pipe[0], 0));
        }
    }
    auto upload = [&](int f, double* devbase) -> cudaError_t {
        for (int c = 0; c < d->ncomp[f]; ++c) {
            cudaError_t e = cudaMemcpyAsync(devbase + (size_t)c * d->plane + NSDG_GHOST * d->nx,
                in[f] + c * inner, sizeof(double) * inner, cudaMemcpyHostToDevice, ctx->h2d_stream);
            if (e != cudaSuccess)
                return e;
        }
        return cudaSuccess;
    };
    auto download = [&](int f, const double* devbase) -> cudaError_t {
        for (int c = 0; c < d->ncomp[f]; ++c) {
            cudaError_t e = cudaMemcpyAsync(out[f] + c * inner,
                devbase + (size_t)c * d->plane + NSDG_GHOST * d->nx, sizeof(double) * inner,
                cudaMemcpyDeviceToHost, ctx->d2h_stream);
            if (e != cudaSuccess)
                return e;
        }
        return cudaSuccess;
    };
    // ---- uploads: what the transport needs first
    const int first[] = { NSDG_DF_U, NSDG_DF_V, NSDG_DF_H, NSDG_DF_A };
    const int second[] = { NSDG_DF_S11, NSDG_DF_S12, NSDG_DF_S22 };
    if (in) {
        bool state_in = false;
        for (int f : first)
            state_in |= in[f] != nullptr;
        for (int f : second)
            state_in |= in[f] != nullptr;
        // state fields are overwritten in place: behind the previous step and its H, A download.
        // Forcing goes to the alternate buffers, last read by the step before the previous one.
        NSDG_CUDA(ctx, cudaStreamWaitEvent(ctx->h2d_stream, d->ev_done[k & 1], 0));
        if (state_in) {
            NSDG_CUDA(ctx, cudaStreamWaitEvent(ctx->h2d_stream, d->ev_done[(k + 1) & 1], 0));
            NSDG_CUDA(ctx, cudaStreamWaitEvent(ctx->h2d_stream, d->ev_ha_down, 0));
        }
        for (int f : first)
            if (in[f])
                NSDG_CUDA(ctx, upload(f, d->f[f]));
        NSDG_CUDA(ctx, cudaEventRecord(ctx->ev_pipe[0], ctx->h2d_stream));
        for (int f : second)
            if (in[f])
                NSDG_CUDA(ctx, upload(f, d->f[f]));
        for (int q = 0; q < 4; ++q) {
            const int f = forcing[q];
            if (!in[f])
                continue;
            if (deferred) {
                NSDG_CUDA(ctx, upload(f, d->falt[q]));
                std::swap(d->f[f], d->falt[q]); // this step reads the new buffer
            } else
                NSDG_CUDA(ctx, upload(f, d->f[f]));
        }
        NSDG_CUDA(ctx, cudaEventRecord(ctx->ev_pipe[1], ctx->h2d_stream));
        NSDG_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_pipe[0], 0));
    }
    // the transport overwrites H, A: behind their previous download
    NSDG_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, d->ev_ha_down, 0));
    if ((rc = nsdg_transport_step_HA(ctx, dt)))
        return rc;
    NSDG_CUDA(ctx, cudaEventRecord(ctx->ev_pipe[2], ctx->stream));
    // ---- H, A are final after the transport: they come down while the sub-cycle runs
    if (out) {
        NSDG_CUDA(ctx, cudaStreamWaitEvent(ctx->d2h_stream, ctx->ev_pipe[2], 0));
        for (int f : { NSDG_DF_H, NSDG_DF_A })
            if (out[f])
                NSDG_CUDA(ctx, download(f, d->f[f]));
        NSDG_CUDA(ctx, cudaEventRecord(d->ev_ha_down, ctx->d2h_stream));
    }
    if (in)
        NSDG_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_pipe[1], 0));
    if ((rc = nsdg_mevp_step(ctx, dt, nsub)))
        return rc;
    if (out && deferred) {
        // snapshot (device to device) once the previous snapshot has left
        NSDG_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, d->ev_snap_down, 0));
        for (int q = 0; q < 5; ++q)
            if (out[snapf[q]])
                NSDG_CUDA(ctx, cudaMemcpyAsync(d->snap + (size_t)snapoff[q] * d->plane, d->f[snapf[q]],
                                   sizeof(double) * (size_t)d->ncomp[snapf[q]] * d->plane,
                                   cudaMemcpyDeviceToDevice, ctx->stream));
    }
    NSDG_CUDA(ctx, cudaEventRecord(ctx->ev_pipe[3], ctx->stream));
    NSDG_CUDA(ctx, cudaEventRecord(d->ev_done[k & 1], ctx->stream));
    if (out) {
        NSDG_CUDA(ctx, cudaStreamWaitEvent(ctx->d2h_stream, ctx->ev_pipe[3], 0));
        for (int q = 0; q < 5; ++q)
            if (out[snapf[q]]) // d->f[] already points at the buffers that hold the new state
                NSDG_CUDA(ctx, download(snapf[q], deferred ? d->snap + (size_t)snapoff[q] * d->plane : d->f[snapf[q]]));
        NSDG_CUDA(ctx, cudaEventRecord(d->ev_snap_down, ctx->d2h_stream));
    }
    return NSDG_OK;
}

int nsdg_dyn_host_wait(nsdg_ctx* ctx)
{
    NSDG_ENTER(ctx);
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->h2d_stream));
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->d2h_stream));
    return NSDG_OK;
}

int nsdg_dyn_step_host(nsdg_ctx* ctx, double dt, int32_t nsub, const double* const* in,
    double* const* out)
{
    NSDG_ENTER(ctx);
    int rc = nsdg_dyn_host_wait(ctx); // deferred steps may still be reading what this one overwrites
    if (!rc)
        rc = step_host_impl(ctx, dt, nsub, in, out, false);
    return rc ? rc : nsdg_dyn_host_wait(ctx);
}

int nsdg_dyn_step_host_async(nsdg_ctx* ctx, double dt, int32_t nsub, const double* const* in,
    double* const* out)
{
    NSDG_ENTER(ctx);
    return step_host_impl(ctx, dt, nsub, in, out, true);
}

int nsdg_dyn_upload_forcing_level(nsdg_ctx* ctx, int32_t level, const double* uatm,
    const double* vatm, const double* uocn, const double* vocn)
{
    NSDG_ENTER(ctx);
    DynState* d = ctx->dyn;
    NSDG_REQUIRE(ctx, d, NSDG_ESTATE, "nsdg_dyn_init has not been called");
    NSDG_REQUIRE(ctx, level == 0 || level == 1, NSDG_EINVAL, "forcing level must be 0 or 1");
    NSDG_REQUIRE(ctx, uatm && vatm && uocn && vocn, NSDG_EINVAL, "null forcing array");
    int rc;
    if (!d->flevel[level] && (rc = alloc_planes(ctx, &d->flevel[level], 4)))
        return rc;
    const double* src[4] = { uatm, vatm, uocn, vocn };
    const size_t inner = (size_t)d->nx * d->nyl;
    for (int q = 0; q < 4; ++q)
        NSDG_CUDA(ctx, cudaMemcpyAsync(d->flevel[level] + (size_t)q * d->plane + NSDG_GHOST * d->nx, src[q],
                           sizeof(double) * inner, cudaMemcpyHostToDevice, ctx->stream));
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    d->have_flevel[level] = true;
    return NSDG_OK;
}

int nsdg_dyn_blend_forcing(nsdg_ctx* ctx, double w)
{
    NSDG_ENTER(ctx);
    DynState* d = ctx->dyn;
    NSDG_REQUIRE(ctx, d, NSDG_ESTATE, "nsdg_dyn_init has not been called");
    NSDG_REQUIRE(ctx, d->have_flevel[0] && d->have_flevel[1], NSDG_ESTATE, "both forcing levels must be uploaded first");
    NSDG_REQUIRE(ctx, w >= 0. && w <= 1., NSDG_EINVAL, "interpolation weight %g outside [0,1]", w);
    Blend4 o = { { d->f[NSDG_DF_UATM], d->f[NSDG_DF_VATM], d->f[NSDG_DF_UOCN], d->f[NSDG_DF_VOCN] } };
    dim3 block(256), grid((unsigned)((d->nx + 255) / 256), (unsigned)d->nyl, 4);
    NSDG_KERNEL(ctx, NSDG_K_BLEND, dyn_blend_kernel<<<grid, block, 0, ctx->stream>>>(o, d->flevel[0], d->flevel[1], w, d->nx, d->nyl, d->plane));
    return NSDG_OK;
}

int nsdg_dyn_mass(nsdg_ctx* ctx, int32_t field, double* mass)
{
    NSDG_ENTER(ctx);
    DynState* d = ctx->dyn;
    NSDG_REQUIRE(ctx, d, NSDG_ESTATE, "nsdg_dyn_init has not been called");
    NSDG_REQUIRE(ctx, (field == NSDG_DF_H || field == NSDG_DF_A) && mass, NSDG_EINVAL, "bad field");
    NSDG_CUDA(ctx, cudaMemsetAsync(d->scratch, 0, sizeof(double), ctx->stream));
    NSDG_KERNEL(ctx, NSDG_K_OTHER, mass_kernel<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(d->f[field], d->nx, d->nyl, d->scratch));
    double s = 0.;
    NSDG_CUDA(ctx, cudaMemcpyAsync(&s, d->scratch, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *mass = s * d->cfg.hx * d->cfg.hy;
    return NSDG_OK;
}

} // extern "C"
