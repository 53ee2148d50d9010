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
    err = dlerror() ? dlerror() : "libnccl.so.2 not found";
    return nullptr;
}

struct PackArgs {
    double* plane[12];
    int n;
    int64_t nx;
    int64_t row_lo, row_hi; // source (pack) or destination (unpack) rows
    double* buf_lo; // pack: rows row_lo -> buf_lo ; unpack: buf_lo -> rows row_lo
    double* buf_hi;
};

// blockIdx.y = plane, blockIdx.z = 0 (lo) / 1 (hi)
__global__ void pack_rows(const __grid_constant__ PackArgs A)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= A.nx)
        return;
    const int p = blockIdx.y;
    if (blockIdx.z == 0) {
        if (A.buf_lo)
            A.buf_lo[p * A.nx + i] = A.plane[p][A.row_lo * A.nx + i];
    } else {
        if (A.buf_hi)
            A.buf_hi[p * A.nx + i] = A.plane[p][A.row_hi * A.nx + i];
    }
}
__global__ void unpack_rows(const __grid_constant__ PackArgs A)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= A.nx)
        return;
    const int p = blockIdx.y;
    if (blockIdx.z == 0) {
        if (A.buf_lo)
            A.plane[p][A.row_lo * A.nx + i] = A.buf_lo[p * A.nx + i];
    } else {
        if (A.buf_hi)
            A.plane[p][A.row_hi * A.nx + i] = A.buf_hi[p * A.nx + i];
    }
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

int nsdg_halo_exchange(nsdg_ctx* ctx, const PlaneSet& planes, int which)
{
    if (!nsdg_comm_active(ctx) || planes.n == 0)
        return NSDG_OK;
    DynState* d = ctx->dyn;
    CommState* c = ctx->comm;
    NSDG_REQUIRE(ctx, (size_t)planes.n * d->nx <= d->halo_capacity, NSDG_EINVAL, "too many planes in one halo exchange");
    const int up = (c->rank + 1) % c->nranks, down = (c->rank + c->nranks - 1) % c->nranks;
    const bool fill_lo = which & 1, fill_hi = which & 2;
    const size_t count = (size_t)planes.n * d->nx;
    PackArgs A;
    for (int q = 0; q < planes.n; ++q)
        A.plane[q] = planes.p[q];
    A.n = planes.n;
    A.nx = d->nx;
    // pack: my top owned row feeds the upper neighbour's lower ghost; my bottom owned row feeds
    // the lower neighbour's upper ghost
    A.row_hi = d->nyl;
    A.row_lo = 1;
    A.buf_hi = fill_lo ? d->send_hi : nullptr;
    A.buf_lo = fill_hi ? d->send_lo : nullptr;
    dim3 block(256), grid((unsigned)((d->nx + 255) / 256), (unsigned)planes.n, 2);
    NSDG_KERNEL(ctx, NSDG_K_PACK, pack_rows<<<grid, block, 0, ctx->stream>>>(A));
    NSDG_NCCL(ctx, c, c->GroupStart());
    if (fill_lo) {
        NSDG_NCCL(ctx, c, c->Send(d->send_hi, count, ncclFloat64, up, c->comm, ctx->stream));
        NSDG_NCCL(ctx, c, c->Recv(d->recv_lo, count, ncclFloat64, down, c->comm, ctx->stream));
        c->bytes_sent += (int64_t)count * 8;
    }
    if (fill_hi) {
        NSDG_NCCL(ctx, c, c->Send(d->send_lo, count, ncclFloat64, down, c->comm, ctx->stream));
        NSDG_NCCL(ctx, c, c->Recv(d->recv_hi, count, ncclFloat64, up, c->comm, ctx->stream));
        c->bytes_sent += (int64_t)count * 8;
    }
    NSDG_NCCL(ctx, c, c->GroupEnd());
    A.row_lo = 0;
    A.row_hi = d->nyl + 1;
    A.buf_lo = fill_lo ? d->recv_lo : nullptr;
    A.buf_hi = fill_hi ? d->recv_hi : nullptr;
    NSDG_KERNEL(ctx, NSDG_K_UNPACK, unpack_rows<<<grid, block, 0, ctx->stream>>>(A));
    return NSDG_OK;
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

} // extern "C"
