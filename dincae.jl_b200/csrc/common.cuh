// Shared plumbing of libdincae_b200: context, error convention, launch accounting, Philox.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/dincae_b200.h"

struct ProfRec {
    const char* name;
    cudaEvent_t a, b;
    double flops, bytes;
};

struct dincae_ctx {
    // per-launch CUDA-event profile (dincae_profile_begin/_end): off by default
    bool prof = false;
    bool capturing = false;                      // the stream is being captured into the step graph: launches carry no programmatic edge
    std::vector<ProfRec> prof_recs;
    double cur_flops = 0.0, cur_bytes = 0.0;     // algorithmic work of the NEXT launch (set by DC_WORK)
    bool tf32_acts = false;     // element-wise producers of conv inputs round to TF32 on store (set by the model, TF32 mode)
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    // side stream for work that is independent of the main dependency chain (weight gradients during the backward pass);
    // created on first use, fenced against `stream` with the two events
    cudaStream_t side = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    int64_t launches = 0;
    int num_sms = 148;
    // reusable device scratch (grown on demand, never shrunk)
    void* scratch = nullptr;
    size_t scratch_bytes = 0;
    // data-parallel gradient exchange (comm.cu): NCCL communicator of this rank, its stream and the fences against `stream`
    void* comm = nullptr;
    int comm_nranks = 1, comm_rank = 0;
    cudaStream_t comm_stream = nullptr;
    cudaEvent_t ev_comm_go = nullptr, ev_comm_done = nullptr;
    int64_t collectives = 0;
    // pinned staging for small host<->device scalars
    void* pinned = nullptr;
    size_t pinned_bytes = 0;
};

void dincae_set_error(const char* fmt, ...);

#define DC_CHECK_CUDA(expr)                                                                         \
    do {                                                                                            \
        cudaError_t _e = (expr);                                                                    \
        if (_e != cudaSuccess) {                                                                    \
            dincae_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return DINCAE_ECUDA;                                                                    \
        }                                                                                           \
    } while (0)

#define DC_REQUIRE(cond, code, ...)     \
    do {                                \
        if (!(cond)) {                  \
            dincae_set_error(__VA_ARGS__); \
            return (code);              \
        }                               \
    } while (0)

#define DC_TRY(expr)                    \
    do {                                \
        int32_t _s = (expr);            \
        if (_s != DINCAE_OK) return _s; \
    } while (0)

static inline int dc_env_int_early(const char* name, int dflt) { const char* e = getenv(name); return e ? atoi(e) : dflt; }
// every kernel launch of the library goes through this so that gpu_launches is a count, not a guess
// Measurement builds only (-DDINCAE_ABLATE, scripts/ablate.py): DINCAE_ABLATE=name[,name] drops the launches whose kernel expression
// contains one of the names, so that the step-time difference gives the cost a kernel really has inside the overlapped step
// (results are garbage).  The shipped library is built without the flag and has no such switch.
#ifdef DINCAE_ABLATE
static inline bool dc_ablated(const char* name) {
    const char* e = getenv("DINCAE_ABLATE");
    if (!e || !*e) return false;
    std::string list(e);
    size_t pos = 0;
    while (pos <= list.size()) {
        size_t c = list.find(',', pos);
        if (c == std::string::npos) c = list.size();
        if (c > pos && strstr(name, list.substr(pos, c - pos).c_str())) return true;
        pos = c + 1;
    }
    return false;
}
#define DC_ABLATED(kernel) ([] { static const bool s = dc_ablated(#kernel); return s; }())
#else
#define DC_ABLATED(kernel) false
#endif
#define DC_LAUNCH(ctx, kernel, grid, block, smem, ...)                        \
    do {                                                                      \
        if ((ctx)->prof) dc_prof_begin((ctx), #kernel);                       \
        if (!DC_ABLATED(kernel)) kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);      \
        if ((ctx)->prof) dc_prof_end((ctx));                                  \
        (ctx)->cur_flops = 0.0;                                               \
        (ctx)->cur_bytes = 0.0;                                               \
        (ctx)->launches++;                                                    \
    } while (0)
// Programmatic dependent launch: the kernel may start while its predecessor in the stream drains (its CTAs take over SMs as they
// free up and run their prologue), and must execute dc_pdl_wait() before it touches global memory.  Only kernels that contain
// dc_pdl_wait() may be launched through this macro.  Not inside the captured step graph: there plain kernel-to-kernel edges measured
// 6.5 % faster than programmatic ones (1.207 against 1.290 ms per step; launch by launch the attribute is worth 0-8 %).
static inline bool dc_pdl_on() { static const int on = dc_env_int_early("DINCAE_PDL", 1); return on != 0; }
#define DC_LAUNCH_PDL(ctx, kernel, grid, block, smem, ...)                                         \
    do {                                                                                           \
        if ((ctx)->prof) dc_prof_begin((ctx), #kernel);                                            \
        cudaLaunchConfig_t _cfg = {};                                                              \
        _cfg.gridDim = dim3(grid); _cfg.blockDim = dim3(block);                                    \
        _cfg.dynamicSmemBytes = (size_t)(smem); _cfg.stream = (ctx)->stream;                       \
        cudaLaunchAttribute _at[1];                                                                \
        _at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                            \
        _at[0].val.programmaticStreamSerializationAllowed = 1;                                     \
        _cfg.attrs = _at; _cfg.numAttrs = (dc_pdl_on() && !(ctx)->prof && !(ctx)->capturing) ? 1 : 0; \
        if (!DC_ABLATED(kernel)) cudaLaunchKernelEx(&_cfg, kernel, __VA_ARGS__);                   \
        if ((ctx)->prof) dc_prof_end((ctx));                                                       \
        (ctx)->cur_flops = 0.0;                                                                    \
        (ctx)->cur_bytes = 0.0;                                                                    \
        (ctx)->launches++;                                                                         \
    } while (0)
#ifdef __CUDACC__
__device__ __forceinline__ void dc_pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void dc_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#endif
// algorithmic FLOPs / bytes of the next launch (reported by dincae_profile_end, used for rooflines)
#define DC_WORK(ctx, fl, by)        \
    do {                            \
        (ctx)->cur_flops = (fl);    \
        (ctx)->cur_bytes = (by);    \
    } while (0)
void dc_prof_begin(dincae_ctx* ctx, const char* name);
void dc_prof_end(dincae_ctx* ctx);

#define DC_CHECK_LAUNCH()                                                                 \
    do {                                                                                  \
        cudaError_t _e = cudaGetLastError();                                              \
        if (_e != cudaSuccess) {                                                          \
            dincae_set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
            return DINCAE_ECUDA;                                                          \
        }                                                                                 \
    } while (0)

int32_t dc_scratch(dincae_ctx* ctx, size_t bytes, void** out);       // grow-only device scratch
int32_t dc_allreduce_on(dincae_ctx* ctx, float* buf, int64_t n, cudaStream_t stream);   // in-place NCCL SUM ordered on `stream`
int32_t dc_comm_allgather_bytes(dincae_ctx* ctx, const void* d_send, void* d_recv, size_t bytes_per_rank);
int32_t dc_comm_allreduce_min_i32(dincae_ctx* ctx, int32_t* d_val);
int32_t dc_pinned(dincae_ctx* ctx, size_t bytes, void** out);        // grow-only pinned host staging

static inline int64_t dc_cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int dc_env_int(const char* name, int dflt) { const char* e = getenv(name); return e ? atoi(e) : dflt; }
static inline int dc_round_up(int a, int b) { return (a + b - 1) / b * b; }

// RAII-less helper: device buffer that is either the caller's (loc == DEVICE) or a staged copy of host memory
struct DevIn {
    const void* ptr = nullptr;
    void* owned = nullptr;
};
int32_t dc_stage_in(dincae_ctx* ctx, const void* src, size_t bytes, int32_t loc, DevIn* out);
void dc_stage_free(dincae_ctx* ctx, DevIn* b);

#ifdef __CUDACC__
// ---- Philox4x32-10 (counter-based; oracle/philox.py is the CPU twin) ------------------------
struct Philox4 {
    uint32_t v[4];
};
__device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    Philox4 o;
    o.v[0] = c0; o.v[1] = c1; o.v[2] = c2; o.v[3] = c3;
    return o;
}
__device__ __forceinline__ float philox_unit(uint32_t x) { return ((float)(x >> 8) + 0.5f) * 5.9604644775390625e-08f; }
// four N(0,1) draws of counter (c0,c1,c2,c3): Box-Muller on (x0,x1) and (x2,x3)
__device__ __forceinline__ void philox_normal4(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint64_t seed, float z[4]) {
    Philox4 r = philox4x32_10(c0, c1, c2, c3, (uint32_t)seed, (uint32_t)(seed >> 32));
#pragma unroll
    for (int h = 0; h < 2; h++) {
        float u1 = philox_unit(r.v[2 * h]), u2 = philox_unit(r.v[2 * h + 1]);
        float rad = sqrtf(-2.0f * logf(u1));
        float s, c;
        sincosf(6.283185307179586f * u2, &s, &c);
        z[2 * h] = rad * c;
        z[2 * h + 1] = rad * s;
    }
}
#endif
