// Strip decomposition over y: one-row halo exchange between ring neighbours with NCCL
// point-to-point (NVLink 5 / NVSwitch).  No counterpart in the reference (single process).
//
// NCCL is bound at run time (dlopen "libnccl.so.2"): inside a torch process this resolves to
// the NCCL torch already loaded, and the library still loads on a CPU-only box.  All planes of
// one exchange are packed into ONE message per neighbour (latency-bound regime: at nx = 16384 a
// row is 128 KiB per plane).
#include "dyn.h"
#include <dlfcn.h>
#include <nccl.h>

struct CommState {
    void* lib = nullptr;
    ncclComm_t comm = nullptr;
    int rank = 0, nranks = 1;
    int64_t bytes_sent = 0;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

namespace {

void* open_nccl(std::string& err)
{
    const char* names[] = { "libnccl.so.2", "libnccl.so" };
    for (const char* n : names) {
        void* h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (h)
            return h;
    }
    const char* e = dlerror(); // one call: dlerror() clears the state it returns
    err = e ? e : "libnccl.so.2 not found";
    return nullptr;
}

// rows to gather into (pack) or scatter from (unpack) a contiguous message buffer, in buffer order
constexpr int MAX_HALO_ROWS = 64;
struct RowList {
    double* plane[MAX_HALO_ROWS];
    int row[MAX_HALO_ROWS];
    int n;
};

// blockIdx.y = entry of the list
__global__ void pack_rows(const __grid_constant__ RowList L, double* buf, int64_t nx)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nx)
        return;
    const int e = blockIdx.y;
    buf[e * nx + i] = L.plane[e][L.row[e] * nx + i];
}
__global__ void unpack_rows(const __grid_constant__ RowList L, const double* buf, int64_t nx)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nx)
        return;
    const int e = blockIdx.y;
    L.plane[e][L.row[e] * nx + i] = buf[e * nx + i];
}

} // namespace

#define NSDG_NCCL(ctx, c, expr)                                                                   \
    do {                                                                                          \
        ncclResult_t _r = (expr);                                                                 \
        if (_r != ncclSuccess)                                                                    \
            return nsdg_fail((ctx), NSDG_ECOMM, "%s failed: %s", #expr,                           \
                (c)->GetErrorString ? (c)->GetErrorString(_r) : "?");                             \
    } while (0)

bool nsdg_comm_active(const nsdg_ctx* ctx) { return ctx->comm && ctx->comm->nranks > 1; }

void nsdg_comm_free(nsdg_ctx* ctx)
{
    CommState* c = ctx->comm;
    if (!c)
        return;
    if (c->comm && c->CommDestroy)
        c->CommDestroy(c->comm);
    delete c; // the NCCL library handle is left open for the life of the process
    ctx->comm = nullptr;
}

int nsdg_halo_exchange(nsdg_ctx* ctx, const HaloSpec* specs, int nspecs, cudaStream_t st)
{
    if (!nsdg_comm_active(ctx) || nspecs == 0)
        return NSDG_OK;
    CommState* c = ctx->comm;
    DynState* d = ctx->dyn;
    const int G = NSDG_GHOST;
    const int nyl = (int)d->nyl;
    // my top interior rows -> lower ghosts of the rank above; my bottom interior rows -> upper
    // ghosts of the rank below.  Sender and receiver enumerate (spec, plane, row) identically.
    RowList send_up = {}, send_dn = {}, recv_lo = {}, recv_hi = {};
    for (int s = 0; s < nspecs; ++s) {
        const HaloSpec& h = specs[s];
        NSDG_REQUIRE(ctx, h.nlo >= 0 && h.nlo <= G && h.nhi >= 0 && h.nhi <= G && h.nlo <= nyl && h.nhi <= nyl,
            NSDG_EINVAL, "halo width out of range");
        for (int q = 0; q < h.planes.n; ++q) {
            for (int r = 0; r < h.nlo; ++r) { // ghost row G-nlo+r <- neighbour's interior row G+nyl-nlo+r
                NSDG_REQUIRE(ctx, send_up.n < MAX_HALO_ROWS, NSDG_EINVAL, "too many rows in one halo exchange");
                send_up.plane[send_up.n] = h.planes.p[q];
                send_up.row[send_up.n++] = G + nyl - h.nlo + r;
                recv_lo.plane[recv_lo.n] = h.planes.p[q];
                recv_lo.row[recv_lo.n++] = G - h.nlo + r;
            }
            for (int r = 0; r < h.nhi; ++r) { // ghost row G+nyl+r <- neighbour's interior row G+r
                NSDG_REQUIRE(ctx, send_dn.n < MAX_HALO_ROWS, NSDG_EINVAL, "too many rows in one halo exchange");
                send_dn.plane[send_dn.n] = h.planes.p[q];
                send_dn.row[send_dn.n++] = G + r;
                recv_hi.plane[recv_hi.n] = h.planes.p[q];
                recv_hi.row[recv_hi.n++] = G + nyl + r;
            }
        }
    }
    const int up = (c->rank + 1) % c->nranks, down = (c->rank + c->nranks - 1) % c->nranks;
    dim3 block(256);
    const unsigned gx = (unsigned)((d->nx + 255) / 256);
    if (!st)
        st = ctx->stream;
    // the per-kernel profiler brackets launches on the compute stream only
#define NSDG_HALO_KERNEL(kid, ...)                                                                \
    do {                                                                                          \
        if (st == ctx->stream) {                                                                  \
            NSDG_KERNEL(ctx, kid, __VA_ARGS__);                                                   \
        } else {                                                                                  \
            __VA_ARGS__;                                                                          \
            NSDG_LAUNCHED(ctx);                                                                   \
        }                                                                                         \
    } while (0)
    if (send_up.n)
        NSDG_HALO_KERNEL(NSDG_K_PACK, pack_rows<<<dim3(gx, send_up.n), block, 0, st>>>(send_up, d->send_hi, d->nx));
    if (send_dn.n)
        NSDG_HALO_KERNEL(NSDG_K_PACK, pack_rows<<<dim3(gx, send_dn.n), block, 0, st>>>(send_dn, d->send_lo, d->nx));
    NSDG_NCCL(ctx, c, c->GroupStart());
    if (send_up.n) {
        const size_t count = (size_t)send_up.n * d->nx;
        NSDG_NCCL(ctx, c, c->Send(d->send_hi, count, ncclFloat64, up, c->comm, st));
        NSDG_NCCL(ctx, c, c->Recv(d->recv_lo, count, ncclFloat64, down, c->comm, st));
        c->bytes_sent += (int64_t)count * 8;
    }
    if (send_dn.n) {
        const size_t count = (size_t)send_dn.n * d->nx;
        NSDG_NCCL(ctx, c, c->Send(d->send_lo, count, ncclFloat64, down, c->comm, st));
        NSDG_NCCL(ctx, c, c->Recv(d->recv_hi, count, ncclFloat64, up, c->comm, st));
        c->bytes_sent += (int64_t)count * 8;
    }
    NSDG_NCCL(ctx, c, c->GroupEnd());
    if (recv_lo.n)
        NSDG_HALO_KERNEL(NSDG_K_UNPACK, unpack_rows<<<dim3(gx, recv_lo.n), block, 0, st>>>(recv_lo, d->recv_lo, d->nx));
    if (recv_hi.n)
        NSDG_HALO_KERNEL(NSDG_K_UNPACK, unpack_rows<<<dim3(gx, recv_hi.n), block, 0, st>>>(recv_hi, d->recv_hi, d->nx));
#undef NSDG_HALO_KERNEL
    return NSDG_OK;
}

// ---- peer halo: map the ring neighbours' slabs with CUDA IPC --------------------------------
namespace {
struct PeerInfo { // what a rank tells its ring neighbours (travels through NCCL as bytes)
    cudaIpcMemHandle_t handle;
    int64_t nyl, plane, nflagw;
    int32_t ok, device;
    char pad[128 - sizeof(cudaIpcMemHandle_t) - 3 * 8 - 2 * 4];
};
static_assert(sizeof(PeerInfo) == 128, "PeerInfo is a 128-byte message");
} // namespace

void nsdg_comm_unmap_peers(nsdg_ctx* ctx)
{
    DynState* d = ctx->dyn;
    if (!d)
        return;
    if (d->ipc_up)
        cudaIpcCloseMemHandle(d->ipc_up);
    if (d->ipc_dn && d->ipc_dn != d->ipc_up)
        cudaIpcCloseMemHandle(d->ipc_dn);
    d->ipc_up = d->ipc_dn = nullptr;
    d->up = d->dn = PeerMap {};
    d->peer_ok = false;
}

int nsdg_comm_map_peers(nsdg_ctx* ctx)
{
    CommState* c = ctx->comm;
    DynState* d = ctx->dyn;
    d->peer_ok = false;
    if (!nsdg_comm_active(ctx) || !d->slab)
        return NSDG_OK;
    const int up = (c->rank + 1) % c->nranks, down = (c->rank + c->nranks - 1) % c->nranks;
    PeerInfo mine = {}, from_up = {}, from_dn = {};
    mine.nyl = d->nyl;
    mine.plane = d->plane;
    mine.nflagw = d->nflagw;
    mine.device = ctx->device;
    mine.ok = cudaIpcGetMemHandle(&mine.handle, d->slab) == cudaSuccess;
    (void)cudaGetLastError();
    PeerInfo* dev = nullptr; // [0] mine, [1] from the rank above, [2] from the rank below
    NSDG_CUDA(ctx, cudaMalloc(&dev, 3 * sizeof(PeerInfo) + 2 * sizeof(int)));
    int* agree = reinterpret_cast<int*>(dev + 3);
    NSDG_CUDA(ctx, cudaMemcpyAsync(dev, &mine, sizeof mine, cudaMemcpyHostToDevice, ctx->stream));
    NSDG_NCCL(ctx, c, c->GroupStart());
    NSDG_NCCL(ctx, c, c->Send(dev, sizeof(PeerInfo), ncclChar, up, c->comm, ctx->stream));
    NSDG_NCCL(ctx, c, c->Recv(dev + 2, sizeof(PeerInfo), ncclChar, down, c->comm, ctx->stream));
    NSDG_NCCL(ctx, c, c->Send(dev, sizeof(PeerInfo), ncclChar, down, c->comm, ctx->stream));
    NSDG_NCCL(ctx, c, c->Recv(dev + 1, sizeof(PeerInfo), ncclChar, up, c->comm, ctx->stream));
    NSDG_NCCL(ctx, c, c->GroupEnd());
    NSDG_CUDA(ctx, cudaMemcpyAsync(&from_up, dev + 1, sizeof from_up, cudaMemcpyDeviceToHost, ctx->stream));
    NSDG_CUDA(ctx, cudaMemcpyAsync(&from_dn, dev + 2, sizeof from_dn, cudaMemcpyDeviceToHost, ctx->stream));
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    int ok = mine.ok && from_up.ok && from_dn.ok && from_up.nflagw == d->nflagw && from_dn.nflagw == d->nflagw;
    auto open = [&](const PeerInfo& pi, void** out) {
        if (cudaIpcOpenMemHandle(out, pi.handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            (void)cudaGetLastError();
            *out = nullptr;
            return false;
        }
        return true;
    };
    if (ok) {
        ok = open(from_up, &d->ipc_up);
        if (ok) {
            if (up == down) // two ranks: both neighbours are the same process, one mapping
                d->ipc_dn = d->ipc_up;
            else
                ok = open(from_dn, &d->ipc_dn);
        }
    }
    // every rank must take the same path: the minimum over all ranks decides
    NSDG_CUDA(ctx, cudaMemcpyAsync(agree, &ok, sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    NSDG_NCCL(ctx, c, c->AllReduce(agree, agree + 1, 1, ncclInt32, ncclMin, c->comm, ctx->stream));
    int all = 0;
    NSDG_CUDA(ctx, cudaMemcpyAsync(&all, agree + 1, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    NSDG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFree(dev);
    if (!all) {
        nsdg_comm_unmap_peers(ctx);
        return NSDG_OK;
    }
    auto fill = [&](PeerMap& m, void* base, const PeerInfo& pi) {
        m.slab = static_cast<double*>(base);
        m.nyl = pi.nyl;
        m.plane = pi.plane;
        m.flags = reinterpret_cast<unsigned*>(m.slab + (size_t)NSDG_SLAB_PLANES * pi.plane);
    };
    fill(d->up, d->ipc_up, from_up);
    fill(d->dn, d->ipc_dn, from_dn);
    d->peer_ok = true;
    return NSDG_OK;
}

void nsdg_comm_count_bytes(nsdg_ctx* ctx, int64_t bytes)
{
    if (ctx->comm)
        ctx->comm->bytes_sent += bytes;
}

int nsdg_halo_exchange(nsdg_ctx* ctx, const PlaneSet& planes, int which)
{
    HaloSpec h;
    h.planes = planes;
    h.nlo = (which & 1) ? 1 : 0;
    h.nhi = (which & 2) ? 1 : 0;
    return nsdg_halo_exchange(ctx, &h, 1);
}

extern "C" {

int nsdg_comm_get_unique_id(void* id_out)
{
    static_assert(sizeof(ncclUniqueId) == NSDG_COMM_ID_BYTES, "unique id size");
    if (!id_out)
        return NSDG_EINVAL;
    std::string err;
    void* lib = open_nccl(err);
    if (!lib)
        return NSDG_ECOMM;
    auto fn = (ncclResult_t(*)(ncclUniqueId*))dlsym(lib, "ncclGetUniqueId");
    if (!fn || fn((ncclUniqueId*)id_out) != ncclSuccess)
        return NSDG_ECOMM;
    return NSDG_OK;
}

int nsdg_comm_init(nsdg_ctx* ctx, const void* id, int32_t rank, int32_t nranks)
{
    NSDG_ENTER(ctx);
    NSDG_REQUIRE(ctx, nranks >= 1 && rank >= 0 && rank < nranks, NSDG_EINVAL, "bad rank %d of %d", rank, nranks);
    NSDG_REQUIRE(ctx, !ctx->dyn, NSDG_ESTATE, "nsdg_comm_init must precede nsdg_dyn_init");
    nsdg_comm_free(ctx);
    CommState* c = new CommState;
    ctx->comm = c;
    c->rank = rank;
    c->nranks = nranks;
    if (nranks == 1)
        return NSDG_OK;
    NSDG_REQUIRE(ctx, id, NSDG_EINVAL, "unique id is null");
    std::string err;
    c->lib = open_nccl(err);
    NSDG_REQUIRE(ctx, c->lib, NSDG_ECOMM, "cannot load NCCL: %s", err.c_str());
#define BIND(field, sym)                                                                          \
    c->field = (decltype(c->field))dlsym(c->lib, sym);                                            \
    NSDG_REQUIRE(ctx, c->field, NSDG_ECOMM, "NCCL symbol %s missing", sym)
    BIND(CommInitRank, "ncclCommInitRank");
    BIND(CommDestroy, "ncclCommDestroy");
    BIND(Send, "ncclSend");
    BIND(Recv, "ncclRecv");
    BIND(AllReduce, "ncclAllReduce");
    BIND(GroupStart, "ncclGroupStart");
    BIND(GroupEnd, "ncclGroupEnd");
    BIND(GetErrorString, "ncclGetErrorString");
#undef BIND
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof uid);
    NSDG_NCCL(ctx, c, c->CommInitRank(&c->comm, nranks, uid, rank));
    return NSDG_OK;
}

int nsdg_comm_rank(const nsdg_ctx* ctx, int32_t* rank, int32_t* nranks)
{
    if (!ctx)
        return NSDG_EINVAL;
    if (rank)
        *rank = ctx->comm ? ctx->comm->rank : 0;
    if (nranks)
        *nranks = ctx->comm ? ctx->comm->nranks : 1;
    return NSDG_OK;
}

int64_t nsdg_comm_bytes_sent(const nsdg_ctx* ctx) { return (ctx && ctx->comm) ? ctx->comm->bytes_sent : 0; }

int nsdg_comm_halo_peer(const nsdg_ctx* ctx)
{
    return (ctx && ctx->dyn && nsdg_mevp_peer_halo(const_cast<nsdg_ctx*>(ctx))) ? 1 : 0;
}

} // extern "C"
