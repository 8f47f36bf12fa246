// Data-parallel gradient exchange of libdincae_b200: an NCCL communicator owned by the context (one process per GPU, one
// rank per context).  The reference is single-device (src/flux.jl:23-40 trains on one Atype), so this is new capability:
// training shards the mini-batch over ranks, every rank holds a replica, and the flat gradient buffer (+ its loss slot) is
// summed over NVLink/NVSwitch before the identical L2 + ClipGrad + Adam update runs on every replica.
//
// NCCL is bound at run time (dlopen), not at link time: single-GPU users need no NCCL, and a host that already carries one
// (PyTorch, CUDA.jl's NCCL.jl) shares its copy instead of loading a second one.  Order matters for such a host: the dynamic
// loader keeps ONE object per soname, so a host that bundles its own libnccl.so.2 must have loaded it BEFORE the first
// dincae_comm_* call (or point DINCAE_NCCL_LIB at it); otherwise the system copy found here is what the host gets later.
#include "common.cuh"
#include <dlfcn.h>
#include <nccl.h>
#include <mutex>

namespace {

struct NcclApi {
    void* handle = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    decltype(&ncclGetVersion) GetVersion = nullptr;
    std::string path;
};
NcclApi g_nccl;
std::mutex g_nccl_mu;

int32_t nccl_api(NcclApi** out) {
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    if (!g_nccl.handle) {
        const char* env = getenv("DINCAE_NCCL_LIB");
        void* h = nullptr;
        if (env && *env) { h = dlopen(env, RTLD_NOW | RTLD_LOCAL); g_nccl.path = env; }
        if (!h) { h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD); g_nccl.path = "libnccl.so.2 (already loaded by the host process)"; }
        if (!h) { h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL); g_nccl.path = "libnccl.so.2"; }
        if (!h) { h = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL); g_nccl.path = "libnccl.so"; }
        DC_REQUIRE(h != nullptr, DINCAE_EUNSUPPORTED, "NCCL not found (%s); set DINCAE_NCCL_LIB to libnccl.so.2 for multi-GPU training", dlerror());
#define NCCL_SYM(name)                                                                                       \
        g_nccl.name = (decltype(g_nccl.name))dlsym(h, "nccl" #name);                                          \
        DC_REQUIRE(g_nccl.name != nullptr, DINCAE_EUNSUPPORTED, "NCCL library %s lacks nccl" #name, g_nccl.path.c_str());
        NCCL_SYM(GetUniqueId) NCCL_SYM(CommInitRank) NCCL_SYM(CommDestroy) NCCL_SYM(AllReduce) NCCL_SYM(AllGather) NCCL_SYM(GetErrorString) NCCL_SYM(GetVersion)
#undef NCCL_SYM
        g_nccl.handle = h;
    }
    *out = &g_nccl;
    return DINCAE_OK;
}

#define DC_CHECK_NCCL(api, expr)                                                                              \
    do {                                                                                                      \
        ncclResult_t _r = (expr);                                                                             \
        if (_r != ncclSuccess) {                                                                              \
            dincae_set_error("%s failed: %s (%s:%d)", #expr, (api)->GetErrorString(_r), __FILE__, __LINE__);  \
            return DINCAE_ECUDA;                                                                              \
        }                                                                                                     \
    } while (0)

}  // namespace

static_assert(sizeof(ncclUniqueId) == DINCAE_COMM_ID_BYTES, "dincae_b200.h: DINCAE_COMM_ID_BYTES must equal sizeof(ncclUniqueId)");

extern "C" int32_t dincae_comm_unique_id(void* id) {
    DC_REQUIRE(id != nullptr, DINCAE_EINVAL, "dincae_comm_unique_id: id is NULL");
    NcclApi* api;
    DC_TRY(nccl_api(&api));
    DC_CHECK_NCCL(api, api->GetUniqueId((ncclUniqueId*)id));
    return DINCAE_OK;
}

extern "C" int32_t dincae_comm_init_rank(dincae_ctx* ctx, int32_t nranks, int32_t rank, const void* id) {
    DC_REQUIRE(ctx && id, DINCAE_EINVAL, "dincae_comm_init_rank: NULL argument");
    DC_REQUIRE(nranks >= 1 && rank >= 0 && rank < nranks, DINCAE_EINVAL, "dincae_comm_init_rank: rank %d of %d", rank, nranks);
    DC_REQUIRE(ctx->comm == nullptr, DINCAE_EINVAL, "dincae_comm_init_rank: the context already has a communicator");
    NcclApi* api;
    DC_TRY(nccl_api(&api));
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof(uid));
    ncclComm_t comm = nullptr;
    DC_CHECK_NCCL(api, api->CommInitRank(&comm, nranks, uid, rank));
    ctx->comm = comm; ctx->comm_nranks = nranks; ctx->comm_rank = rank;
    DC_CHECK_CUDA(cudaStreamCreateWithFlags(&ctx->comm_stream, cudaStreamNonBlocking));
    DC_CHECK_CUDA(cudaEventCreateWithFlags(&ctx->ev_comm_go, cudaEventDisableTiming));
    DC_CHECK_CUDA(cudaEventCreateWithFlags(&ctx->ev_comm_done, cudaEventDisableTiming));
    return DINCAE_OK;
}

extern "C" int32_t dincae_comm_destroy(dincae_ctx* ctx) {
    if (!ctx || !ctx->comm) return DINCAE_OK;
    NcclApi* api;
    DC_TRY(nccl_api(&api));
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    cudaStreamSynchronize(ctx->comm_stream);
    api->CommDestroy((ncclComm_t)ctx->comm);
    ctx->comm = nullptr; ctx->comm_nranks = 1; ctx->comm_rank = 0;
    cudaStreamDestroy(ctx->comm_stream); ctx->comm_stream = nullptr;
    cudaEventDestroy(ctx->ev_comm_go); cudaEventDestroy(ctx->ev_comm_done);
    ctx->ev_comm_go = ctx->ev_comm_done = nullptr;
    return DINCAE_OK;
}

extern "C" int32_t dincae_comm_info(dincae_ctx* ctx, int32_t* nranks, int32_t* rank, int32_t* nccl_version) {
    DC_REQUIRE(ctx != nullptr, DINCAE_EINVAL, "dincae_comm_info: ctx is NULL");
    if (nranks) *nranks = ctx->comm ? ctx->comm_nranks : 1;
    if (rank) *rank = ctx->comm ? ctx->comm_rank : 0;
    if (nccl_version) {
        *nccl_version = 0;
        if (ctx->comm) {
            NcclApi* api;
            DC_TRY(nccl_api(&api));
            int v = 0;
            api->GetVersion(&v);
            *nccl_version = v;
        }
    }
    return DINCAE_OK;
}

// in-place SUM over all ranks, ordered on `stream`
int32_t dc_allreduce_on(dincae_ctx* ctx, float* buf, int64_t n, cudaStream_t stream) {
    DC_REQUIRE(ctx->comm != nullptr, DINCAE_EINVAL, "no communicator: call dincae_comm_init_rank first");
    if (n == 0) return DINCAE_OK;
    NcclApi* api;
    DC_TRY(nccl_api(&api));
    DC_CHECK_NCCL(api, api->AllReduce(buf, buf, (size_t)n, ncclFloat, ncclSum, (ncclComm_t)ctx->comm, stream));
    ctx->collectives++;
    return DINCAE_OK;
}

// set-up helpers of the peer-memory all-reduce (p2p.cu): exchange of a few bytes per rank and agreement on a flag
int32_t dc_comm_allgather_bytes(dincae_ctx* ctx, const void* d_send, void* d_recv, size_t bytes_per_rank) {
    DC_REQUIRE(ctx->comm != nullptr, DINCAE_EINVAL, "no communicator");
    NcclApi* api;
    DC_TRY(nccl_api(&api));
    DC_CHECK_NCCL(api, api->AllGather(d_send, d_recv, bytes_per_rank, ncclChar, (ncclComm_t)ctx->comm, ctx->stream));
    return DINCAE_OK;
}
int32_t dc_comm_allreduce_min_i32(dincae_ctx* ctx, int32_t* d_val) {
    DC_REQUIRE(ctx->comm != nullptr, DINCAE_EINVAL, "no communicator");
    NcclApi* api;
    DC_TRY(nccl_api(&api));
    DC_CHECK_NCCL(api, api->AllReduce(d_val, d_val, 1, ncclInt32, ncclMin, (ncclComm_t)ctx->comm, ctx->stream));
    return DINCAE_OK;
}

extern "C" int32_t dincae_allreduce(dincae_ctx* ctx, float* dev_buf, int64_t n) {
    DC_REQUIRE(ctx && dev_buf && n >= 0, DINCAE_EINVAL, "dincae_allreduce: bad argument");
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    return dc_allreduce_on(ctx, dev_buf, n, ctx->stream);
}
