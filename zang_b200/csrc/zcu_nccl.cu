// zcu_nccl.cu — sum of the per-GPU partial mixes: one ncclReduce(sum, f32) over NVLink / NVSwitch.
//
// Voices (or render jobs) are sharded over the GPUs of one box with no data-path exchange; the only
// collective of the whole path is this reduce of the [frames][channels] partial mix.  NCCL is bound at
// run time with dlopen so that the library has no link-time dependency on a particular libnccl (the
// Python harness shares the process with torch's bundled NCCL; a Zig host uses the system one).
// Types, enum values and prototypes come from NCCL's own header (decltype of the declared functions), so the signatures the
// function pointers are called through are the library's, not hand-copied ones; only the binding is deferred to dlopen.
#include <dlfcn.h>
#include <nccl.h>

#include <cstring>

#include "zcu_internal.h"

namespace {

typedef ncclUniqueId ncclUniqueId_t;
typedef ncclComm_t ncclComm_p;
typedef decltype(&ncclGetUniqueId) fn_GetUniqueId;
typedef decltype(&ncclCommInitRank) fn_CommInitRank;
typedef decltype(&ncclCommDestroy) fn_CommDestroy;
typedef decltype(&ncclReduce) fn_Reduce;
typedef decltype(&ncclGetErrorString) fn_GetErrorString;
static_assert(sizeof(ncclUniqueId) == 128, "zcu_nccl_unique_id hands out the 128-byte ncclUniqueId");

struct Nccl {
    void* h = nullptr;
    fn_GetUniqueId GetUniqueId = nullptr;
    fn_CommInitRank CommInitRank = nullptr;
    fn_CommDestroy CommDestroy = nullptr;
    fn_Reduce Reduce = nullptr;
    fn_GetErrorString GetErrorString = nullptr;
    bool load() {
        if (h) return true;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* nme : names) {
            h = dlopen(nme, RTLD_NOW | RTLD_GLOBAL);
            if (h) break;
        }
        if (!h) return false;
        GetUniqueId = (fn_GetUniqueId)dlsym(h, "ncclGetUniqueId");
        CommInitRank = (fn_CommInitRank)dlsym(h, "ncclCommInitRank");
        CommDestroy = (fn_CommDestroy)dlsym(h, "ncclCommDestroy");
        Reduce = (fn_Reduce)dlsym(h, "ncclReduce");
        GetErrorString = (fn_GetErrorString)dlsym(h, "ncclGetErrorString");
        return GetUniqueId && CommInitRank && CommDestroy && Reduce;
    }
};
Nccl g_nccl;
const char* dl_error() {
    const char* e = dlerror();  // NULL when the library loaded but a symbol is missing
    return e ? e : "a required NCCL symbol is missing";
}

}  // namespace

extern "C" {

int zcu_nccl_unique_id(uint8_t id_out[128]) {
    if (!id_out) return ZCU_ERR_INVALID;
    if (!g_nccl.load()) return ZCU_ERR_NCCL;
    ncclUniqueId_t id;
    if (g_nccl.GetUniqueId(&id) != ncclSuccess) return ZCU_ERR_NCCL;
    std::memcpy(id_out, id.internal, 128);
    return ZCU_OK;
}

int zcu_nccl_init(zcu_ctx* ctx, int nranks, int rank, const uint8_t id_in[128]) {
    if (!ctx || !id_in) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, nranks >= 1 && rank >= 0 && rank < nranks, "bad rank / nranks");
    if (!g_nccl.load()) return zcu_fail(ctx, ZCU_ERR_NCCL, "libnccl.so.2 not usable: %s", dl_error());
    ZCU_REQUIRE(ctx, ctx->nccl_comm == nullptr, "zcu_nccl_init called twice (zcu_nccl_finalize first)");
    ncclUniqueId_t id;
    std::memcpy(id.internal, id_in, 128);
    ZCU_CUDA(ctx, cudaSetDevice(ctx->device));
    ncclComm_p comm = nullptr;
    const ncclResult_t rc = g_nccl.CommInitRank(&comm, nranks, id, rank);
    if (rc != ncclSuccess) return zcu_fail(ctx, ZCU_ERR_NCCL, "ncclCommInitRank: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "error");
    ctx->nccl_comm = comm;
    ctx->nccl_rank = rank;
    ctx->nccl_nranks = nranks;
    return ZCU_OK;
}

int zcu_nccl_reduce_mix(zcu_ctx* ctx, zcu_buf* mix, int root) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, mix != nullptr, "null mix");
    ZCU_REQUIRE(ctx, ctx->nccl_comm != nullptr, "zcu_nccl_init has not been called");
    const size_t count = (size_t)mix->frames * mix->voices;
    ZCU_REQUIRE(ctx, root >= 0 && root < ctx->nccl_nranks, "bad root");
    const ncclResult_t rc = g_nccl.Reduce(mix->d, mix->d, count, ncclFloat32, ncclSum, root, (ncclComm_t)ctx->nccl_comm, ctx->stream);
    if (rc != ncclSuccess) return zcu_fail(ctx, ZCU_ERR_NCCL, "ncclReduce: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "error");
    return ZCU_OK;
}

int zcu_nccl_finalize(zcu_ctx* ctx) {
    if (!ctx) return ZCU_ERR_INVALID;
    if (ctx->nccl_comm) {
        cudaStreamSynchronize(ctx->stream);
        g_nccl.CommDestroy((ncclComm_t)ctx->nccl_comm);
        ctx->nccl_comm = nullptr;
    }
    return ZCU_OK;
}

}  // extern "C"
