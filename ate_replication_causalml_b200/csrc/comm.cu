// comm.cu -- the collectives of a fit shared by several GPUs (no counterpart in
// the single-threaded reference).  One process per GPU; every rank holds a slice of the rows.
//
// NCCL is bound at run time (dlopen of libnccl.so.2: the copy PyTorch already loaded when the caller is
// a torch process, the system one otherwise), so the library itself has no link-time dependency and
// loads on a box without NCCL.  The communicator lives in the ctx and every collective is enqueued on
// the ctx stream: an all-reduce sits between the step kernel and the coordinate-descent kernels of an
// IRLS round without any host synchronisation.  Without a communicator the host hook of
// ate_ctx_set_allreduce is used instead (device -> host -> hook -> device); CPU tests exercise that one.
#include <dlfcn.h>
#include <string.h>

#include <vector>

#include "path.cuh"

namespace {

// the handful of NCCL entry points used, with the ABI of nccl.h (2.x)
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclSum = 0 };
enum { ncclFloat64 = 8 };

struct NcclApi {
    void *handle = nullptr;
    int (*GetUniqueId)(ncclUniqueId *) = nullptr;
    int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    bool ok = false;
};

NcclApi &nccl() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api;
    tried = true;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *nm : names) {
        api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (api.handle) break;
    }
    if (!api.handle) return api;
    api.GetUniqueId = (int (*)(ncclUniqueId *))dlsym(api.handle, "ncclGetUniqueId");
    api.CommInitRank = (int (*)(ncclComm_t *, int, ncclUniqueId, int))dlsym(api.handle, "ncclCommInitRank");
    api.CommDestroy = (int (*)(ncclComm_t))dlsym(api.handle, "ncclCommDestroy");
    api.AllReduce = (int (*)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t))dlsym(api.handle, "ncclAllReduce");
    api.AllGather = (int (*)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t))dlsym(api.handle, "ncclAllGather");
    api.GetErrorString = (const char *(*)(int))dlsym(api.handle, "ncclGetErrorString");
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllReduce && api.AllGather;
    return api;
}

const char *nccl_err(int rc) {
    NcclApi &n = nccl();
    return n.GetErrorString ? n.GetErrorString(rc) : "nccl error";
}

}  // namespace

extern "C" int ate_nccl_unique_id(char *id128) {
    if (!id128) return -1;
    NcclApi &n = nccl();
    if (!n.ok) return -2;
    ncclUniqueId id;
    if (n.GetUniqueId(&id) != ncclSuccess) return -3;
    memcpy(id128, id.internal, 128);
    return 0;
}

extern "C" int ate_ctx_nccl_init(ate_ctx *ctx, const char *id128, int rank, int nranks) {
    if (!ctx) return -1;
    if (!id128 || nranks < 1 || rank < 0 || rank >= nranks) return ate_fail(ctx, -1, "ate_ctx_nccl_init: bad arguments");
    NcclApi &n = nccl();
    if (!n.ok) return ate_fail(ctx, -2, "ate_ctx_nccl_init: libnccl.so.2 cannot be loaded: %s", dlerror());
    ATE_CUDA(ctx, cudaSetDevice(ctx->device));
    if (ctx->nccl_comm) {
        n.CommDestroy((ncclComm_t)ctx->nccl_comm);
        ctx->nccl_comm = nullptr;
    }
    ncclUniqueId id;
    memcpy(id.internal, id128, 128);
    ncclComm_t comm = nullptr;
    const int rc = n.CommInitRank(&comm, nranks, id, rank);
    if (rc != ncclSuccess) return ate_fail(ctx, -3, "ncclCommInitRank: %s", nccl_err(rc));
    ctx->nccl_comm = comm;
    ctx->nccl_rank = rank;
    ctx->nccl_nranks = nranks;
    return 0;
}

extern "C" int ate_ctx_nccl_destroy(ate_ctx *ctx) {
    if (!ctx) return -1;
    if (ctx->nccl_comm) {
        cudaSetDevice(ctx->device);
        cudaStreamSynchronize(ctx->stream);
        nccl().CommDestroy((ncclComm_t)ctx->nccl_comm);
        ctx->nccl_comm = nullptr;
    }
    return 0;
}

int ate_i_allreduce(ate_ctx *ctx, double *dbuf, long long count) {
    if (count <= 0) return 0;
    if (ctx->nccl_comm) {
        const int rc = nccl().AllReduce(dbuf, dbuf, (size_t)count, ncclFloat64, ncclSum, (ncclComm_t)ctx->nccl_comm, ctx->stream);
        if (rc != ncclSuccess) return ate_fail(ctx, -4, "ncclAllReduce: %s", nccl_err(rc));
        ctx->launches++;
        return 0;
    }
    if (!ctx->allreduce) return ate_fail(ctx, -1, "a fit is shared by several ranks but the ctx has neither an NCCL communicator nor an all-reduce hook");
    void *h = nullptr;
    ATE_TRY(ate_cached_pinned(ctx, "comm.host", sizeof(double) * (size_t)count, &h));
    ATE_CUDA(ctx, cudaMemcpyAsync(h, dbuf, sizeof(double) * (size_t)count, cudaMemcpyDeviceToHost, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->allreduce(ctx->allreduce_user, (double *)h, count) != 0) return ate_fail(ctx, -4, "all-reduce hook failed");
    ATE_CUDA(ctx, cudaMemcpyAsync(dbuf, h, sizeof(double) * (size_t)count, cudaMemcpyHostToDevice, ctx->stream));
    ATE_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

// every rank contributes `chunk` doubles at dsend; drecv receives nranks * chunk, rank-major
int ate_i_allgather(ate_ctx *ctx, const double *dsend, double *drecv, long long chunk, int rank, int nranks) {
    if (chunk <= 0) return 0;
    if (ctx->nccl_comm) {
        const int rc = nccl().AllGather(dsend, drecv, (size_t)chunk, ncclFloat64, (ncclComm_t)ctx->nccl_comm, ctx->stream);
        if (rc != ncclSuccess) return ate_fail(ctx, -4, "ncclAllGather: %s", nccl_err(rc));
        ctx->launches++;
        return 0;
    }
    // host hook: zero-fill, place the own slice, sum
    ATE_CUDA(ctx, cudaMemsetAsync(drecv, 0, sizeof(double) * (size_t)chunk * nranks, ctx->stream));
    ATE_CUDA(ctx, cudaMemcpyAsync(drecv + (size_t)rank * chunk, dsend, sizeof(double) * (size_t)chunk, cudaMemcpyDeviceToDevice, ctx->stream));
    return ate_i_allreduce(ctx, drecv, chunk * nranks);
}
