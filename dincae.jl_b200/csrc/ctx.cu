// Context, error convention and staging helpers of libdincae_b200.
#include "common.cuh"
#include <stdarg.h>
#include <string.h>
#include <map>
#include <string>

static thread_local char g_err[1024] = "";

void dincae_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* dincae_last_error(void) { return g_err; }
extern "C" int32_t dincae_version(void) { return 100; }

extern "C" int32_t dincae_ctx_create(int32_t device, dincae_ctx** out) {
    DC_REQUIRE(out != nullptr, DINCAE_EINVAL, "dincae_ctx_create: out is NULL");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        dincae_set_error("dincae_ctx_create: no CUDA device (%s); this library has no CPU path", cudaGetErrorString(e));
        return DINCAE_ECUDA;
    }
    DC_REQUIRE(device >= 0 && device < n, DINCAE_EINVAL, "dincae_ctx_create: device %d out of range [0,%d)", device, n);
    DC_CHECK_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    DC_CHECK_CUDA(cudaGetDeviceProperties(&prop, device));
    DC_REQUIRE(prop.major >= 10, DINCAE_EUNSUPPORTED,
               "dincae_ctx_create: device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
    dincae_ctx* c = new dincae_ctx();
    c->device = device;
    c->num_sms = prop.multiProcessorCount;
    DC_CHECK_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    c->own_stream = true;
    *out = c;
    return DINCAE_OK;
}

extern "C" int32_t dincae_ctx_destroy(dincae_ctx* c) {
    if (!c) return DINCAE_OK;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    dincae_comm_destroy(c);
    if (c->scratch) cudaFree(c->scratch);
    if (c->pinned) cudaFreeHost(c->pinned);
    if (c->side) { cudaStreamSynchronize(c->side); cudaStreamDestroy(c->side); c->side = nullptr; }
    if (c->ev_fork) { cudaEventDestroy(c->ev_fork); c->ev_fork = nullptr; }
    if (c->ev_join) { cudaEventDestroy(c->ev_join); c->ev_join = nullptr; }
    if (c->own_stream) cudaStreamDestroy(c->stream);
    delete c;
    return DINCAE_OK;
}

extern "C" int32_t dincae_ctx_set_stream(dincae_ctx* c, void* s) {
    DC_REQUIRE(c != nullptr, DINCAE_EINVAL, "dincae_ctx_set_stream: ctx is NULL");
    DC_CHECK_CUDA(cudaStreamSynchronize(c->stream));
    if (c->own_stream) cudaStreamDestroy(c->stream);
    c->stream = (cudaStream_t)s;
    c->own_stream = false;
    return DINCAE_OK;
}

extern "C" int32_t dincae_sync(dincae_ctx* c) {
    DC_REQUIRE(c != nullptr, DINCAE_EINVAL, "dincae_sync: ctx is NULL");
    DC_CHECK_CUDA(cudaStreamSynchronize(c->stream));
    DC_CHECK_CUDA(cudaGetLastError());
    return DINCAE_OK;
}

extern "C" int64_t dincae_launch_count(dincae_ctx* c) { return c ? c->launches : -1; }

int32_t dc_scratch(dincae_ctx* ctx, size_t bytes, void** out) {
    if (bytes > ctx->scratch_bytes) {
        DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
        if (ctx->scratch) cudaFree(ctx->scratch);
        ctx->scratch = nullptr;
        ctx->scratch_bytes = 0;
        size_t want = bytes + bytes / 4 + 4096;
        cudaError_t e = cudaMalloc(&ctx->scratch, want);
        if (e != cudaSuccess) {
            dincae_set_error("scratch allocation of %zu bytes failed: %s", want, cudaGetErrorString(e));
            return DINCAE_ENOMEM;
        }
        ctx->scratch_bytes = want;
    }
    *out = ctx->scratch;
    return DINCAE_OK;
}

int32_t dc_pinned(dincae_ctx* ctx, size_t bytes, void** out) {
    if (bytes > ctx->pinned_bytes) {
        DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
        if (ctx->pinned) cudaFreeHost(ctx->pinned);
        ctx->pinned = nullptr;
        ctx->pinned_bytes = 0;
        size_t want = bytes < 4096 ? 4096 : bytes;
        DC_CHECK_CUDA(cudaMallocHost(&ctx->pinned, want));
        ctx->pinned_bytes = want;
    }
    *out = ctx->pinned;
    return DINCAE_OK;
}

int32_t dc_stage_in(dincae_ctx* ctx, const void* src, size_t bytes, int32_t loc, DevIn* out) {
    out->owned = nullptr;
    out->ptr = nullptr;
    if (src == nullptr || bytes == 0) return DINCAE_OK;
    if (loc == DINCAE_DEVICE) {
        out->ptr = src;
        return DINCAE_OK;
    }
    cudaError_t e = cudaMallocAsync(&out->owned, bytes, ctx->stream);
    if (e != cudaSuccess) {
        dincae_set_error("staging allocation of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
        return DINCAE_ENOMEM;
    }
    DC_CHECK_CUDA(cudaMemcpyAsync(out->owned, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    out->ptr = out->owned;
    return DINCAE_OK;
}

void dc_stage_free(dincae_ctx* ctx, DevIn* b) {
    if (b->owned) cudaFreeAsync(b->owned, ctx->stream);
    b->owned = nullptr;
}

// ---- per-launch CUDA-event profile ---------------------------------------------------------------
void dc_prof_begin(dincae_ctx* ctx, const char* name) {
    ProfRec r;
    r.name = name;
    r.flops = ctx->cur_flops;
    r.bytes = ctx->cur_bytes;
    cudaEventCreate(&r.a);
    cudaEventCreate(&r.b);
    cudaEventRecord(r.a, ctx->stream);
    ctx->prof_recs.push_back(r);
}
void dc_prof_end(dincae_ctx* ctx) { cudaEventRecord(ctx->prof_recs.back().b, ctx->stream); }

extern "C" int32_t dincae_profile_begin(dincae_ctx* ctx) {
    DC_REQUIRE(ctx != nullptr, DINCAE_EINVAL, "dincae_profile_begin: ctx is NULL");
    DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    for (auto& r : ctx->prof_recs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
    ctx->prof_recs.clear();
    ctx->prof = true;
    return DINCAE_OK;
}

extern "C" int32_t dincae_profile_end(dincae_ctx* ctx, char* json, int64_t cap) {
    DC_REQUIRE(ctx && json && cap > 2, DINCAE_EINVAL, "dincae_profile_end: bad arguments");
    ctx->prof = false;
    DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    struct Agg { int64_t n = 0; double ms = 0, flops = 0, bytes = 0; };
    std::map<std::string, Agg> agg;
    for (auto& r : ctx->prof_recs) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, r.a, r.b);
        Agg& a = agg[r.name];
        a.n++; a.ms += ms; a.flops += r.flops; a.bytes += r.bytes;
        cudaEventDestroy(r.a);
        cudaEventDestroy(r.b);
    }
    ctx->prof_recs.clear();
    std::string out = "[";
    bool first = true;
    for (auto& kv : agg) {
        char buf[512];
        snprintf(buf, sizeof(buf), "%s{\"kernel\": \"%s\", \"launches\": %lld, \"ms\": %.6f, \"flops\": %.6e, \"bytes\": %.6e}",
                 first ? "" : ", ", kv.first.c_str(), (long long)kv.second.n, kv.second.ms, kv.second.flops, kv.second.bytes);
        out += buf;
        first = false;
    }
    out += "]";
    DC_REQUIRE((int64_t)out.size() + 1 <= cap, DINCAE_ESIZE, "dincae_profile_end: buffer too small (%zu needed)", out.size() + 1);
    memcpy(json, out.c_str(), out.size() + 1);
    return DINCAE_OK;
}
