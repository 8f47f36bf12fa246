// Gradient all-reduce over NVLink peer memory (data-parallel training, one process per GPU).
//
// The gradient buffer is 6 MB: an NCCL all-reduce of that size is latency, not bandwidth (measured +41 / +92 us per 1.34 ms
// step at 2 / 8 GPUs), and it cannot be hidden behind the backward pass here (the persistent conv kernels own every SM).  This
// kernel does the same sum with plain stores into the peers' memory, mapped into this process through CUDA IPC (the phases are
// described at the kernel).  Every element is summed by exactly one rank, in rank order 0..N-1, and the result is what all ranks
// receive: replicas end up with identical bits.  Per rank 2 * (N-1)/N * 6 MB cross NVLink and two flag rounds are paid; no NCCL
// kernel, no proxy thread.  Flags are monotonically increasing epoch tags (nothing to reset) and are read with volatile loads;
// every spin has a time limit: a rank that never arrives makes the others give up and raise an error word instead of hanging the GPU.
#include "model.cuh"

namespace {

constexpr int P2P_MAX_RANKS = 16;
constexpr int P2P_BLOCKS = 64, P2P_THREADS = 512;
// flag words of one rank (uint32 index): [phase 0..2][source rank], then two grid-barrier counters and the error word
constexpr int P2P_F_ARRIVE = 0, P2P_F_COUNT = 3 * P2P_MAX_RANKS, P2P_F_ERROR = P2P_F_COUNT + 2, P2P_F_WORDS = 64;

struct P2PArgs {
    float* g[P2P_MAX_RANKS];            // gradient buffer of every rank (own rank: local pointer)
    float* stage[P2P_MAX_RANKS];        // staging area of every rank: [source rank][slice] contributions pushed by the peers
    uint32_t* flags[P2P_MAX_RANKS];     // flag words of every rank
    int nranks, rank;
    int64_t n4;                         // float4 elements to reduce
    uint32_t tag;                       // epoch tag of this call (3 consecutive values are used)
    long long spin_limit;               // clock64 ticks a flag may be waited for
};

__device__ __forceinline__ void p2p_signal(const P2PArgs& a, int phase, uint32_t tag) {
    // executed by the first nranks threads of ONE block, after that block knows that all of this rank's writes are done
    __threadfence_system();
    if ((int)threadIdx.x < a.nranks) {
        volatile uint32_t* f = a.flags[threadIdx.x] + P2P_F_ARRIVE + phase * P2P_MAX_RANKS + a.rank;
        *f = tag;
    }
}
__device__ __forceinline__ void p2p_wait(const P2PArgs& a, int phase, uint32_t tag) {
    // every block: thread p waits for rank p's word in THIS rank's memory
    if ((int)threadIdx.x < a.nranks) {
        volatile uint32_t* f = a.flags[a.rank] + P2P_F_ARRIVE + phase * P2P_MAX_RANKS + threadIdx.x;
        const long long t0 = clock64();
        while ((int32_t)(*f - tag) < 0) {
            if (clock64() - t0 > a.spin_limit) { *(volatile uint32_t*)(a.flags[a.rank] + P2P_F_ERROR) = 1u + (uint32_t)phase; break; }
        }
    }
    __syncthreads();
    __threadfence_system();
}
// all blocks of this kernel have finished their part (they are co-resident: the grid is smaller than the GPU and nothing else runs
// on the stream); the LAST block to arrive signals the peers
__device__ __forceinline__ void p2p_grid_done_then_signal(const P2PArgs& a, int which, int phase, uint32_t tag) {
    __shared__ int last;
    __threadfence_system();                     // every thread: its writes of this phase are visible to the peers ...
    __syncthreads();                            // ... before this block counts itself as done
    if (threadIdx.x == 0) {
        __threadfence();
        uint32_t* c = a.flags[a.rank] + P2P_F_COUNT + which;
        const uint32_t prev = atomicAdd(c, 1u);
        last = prev == gridDim.x - 1;
        if (last) *c = 0u;                      // ready for the next call (nobody touches it before the next kernel)
    }
    __syncthreads();
    if (last) p2p_signal(a, phase, tag);
}

__device__ __forceinline__ float4 ld_peer(const float4* p) {
    float4 v;
    asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

__device__ __forceinline__ float4 ld_vol(const float4* p) { return ld_peer(p); }

// Push form: remote traffic is posted STORES only (a load over NVLink pays the round trip, a store does not).
//   phase 1  every rank writes its part of slice p into rank p's staging area (slot = source rank), then raises flag 0 at p;
//   phase 2  rank r waits for the N-1 flags, sums slot 0..N-1 (its own part straight from its gradient buffer) in rank order and
//            writes the result into slice r of EVERY rank's gradient buffer, then raises flag 1 everywhere;
//   phase 3  wait for the N-1 flags 1: all slices have landed here, and every peer has finished reading its staging area, so the
//            next call may overwrite it.
__global__ void __launch_bounds__(P2P_THREADS) p2p_allreduce_kernel(const __grid_constant__ P2PArgs a) {
    const int N = a.nranks, r = a.rank;
    const int64_t per = (a.n4 + N - 1) / N;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x, t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    // ---- phase 1: scatter my contributions
    for (int q = 1; q < N; q++) {
        const int p = (r + q) % N;                           // a different first peer on every rank
        const int64_t lo = per * p, hi = lo + per < a.n4 ? lo + per : a.n4;
        const float4* src = reinterpret_cast<const float4*>(a.g[r]);
        float4* dst = reinterpret_cast<float4*>(a.stage[p]) + per * r - lo;
        for (int64_t i = lo + t0; i < hi; i += stride) dst[i] = src[i];
    }
    p2p_grid_done_then_signal(a, 0, 0, a.tag);
    p2p_wait(a, 0, a.tag);
    // ---- phase 2: reduce my slice in rank order, broadcast it
    {
        const int64_t lo = per * r, hi = lo + per < a.n4 ? lo + per : a.n4;
        const float4* st = reinterpret_cast<const float4*>(a.stage[r]);
        for (int64_t i = lo + t0; i < hi; i += stride) {
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int p0 = 0; p0 < N; p0 += 8) {
                float4 v[8];
#pragma unroll
                for (int k = 0; k < 8; k++)
                    if (p0 + k < N) v[k] = (p0 + k == r) ? reinterpret_cast<const float4*>(a.g[r])[i] : ld_vol(st + per * (p0 + k) + (i - lo));
#pragma unroll
                for (int k = 0; k < 8; k++)
                    if (p0 + k < N) { s.x += v[k].x; s.y += v[k].y; s.z += v[k].z; s.w += v[k].w; }
            }
            for (int p = 0; p < N; p++) reinterpret_cast<float4*>(a.g[p])[i] = s;
        }
    }
    p2p_grid_done_then_signal(a, 1, 1, a.tag + 1);
    // ---- phase 3: everybody's slice has arrived (only one block has to see it: the kernel ends after its last block)
    if (blockIdx.x == 0) p2p_wait(a, 1, a.tag + 1);
}

__global__ void p2p_put_handle_kernel(const cudaIpcMemHandle_t h, cudaIpcMemHandle_t* dst) { *dst = h; }

}  // namespace

// Unmap the peers' workspaces.  Collective when the mapping exists: a workspace must not be freed while a peer still has it
// mapped, so every rank closes its mappings and then waits for all the others (a one-word all-reduce) before its caller frees.
void model_p2p_release(dincae_model* M) {
    if (M->p2p_state != 1) return;
    dincae_ctx* ctx = M->ctx;
    cudaStreamSynchronize(ctx->stream);
    for (int p = 0; p < ctx->comm_nranks && p < P2P_MAX_RANKS; p++)
        if (p != ctx->comm_rank && M->p2p_peer_base[p]) { cudaIpcCloseMemHandle(M->p2p_peer_base[p]); M->p2p_peer_base[p] = nullptr; }
    M->p2p_state = -1;
    if (ctx->comm) {
        void* sc = nullptr;
        if (dc_scratch(ctx, 64, &sc) == DINCAE_OK) {
            cudaMemsetAsync(sc, 0, 4, ctx->stream);
            dc_comm_allreduce_min_i32(ctx, (int32_t*)sc);
            cudaStreamSynchronize(ctx->stream);
        }
    }
}

// error word of the peer all-reduce (a flag that never arrived): 0 = fine
int32_t model_p2p_check(dincae_model* M) {
    if (M->p2p_state != 1) return DINCAE_OK;
    uint32_t err = 0;
    DC_CHECK_CUDA(cudaMemcpyAsync(&err, M->p2p_flags + P2P_F_ERROR, 4, cudaMemcpyDeviceToHost, M->ctx->stream));
    DC_CHECK_CUDA(cudaStreamSynchronize(M->ctx->stream));
    DC_REQUIRE(err == 0, DINCAE_ECUDA, "peer-memory all-reduce: a rank did not arrive within the time limit (phase %u); the gradients of this step are invalid", err - 1);
    return DINCAE_OK;
}

// one-time: exchange the IPC handles of the workspaces through the communicator, map the peers', agree that it worked everywhere
static int32_t p2p_setup(dincae_model* M) {
    dincae_ctx* ctx = M->ctx;
    M->p2p_state = -1;
    const int N = ctx->comm_nranks;
    // DINCAE_P2P: 1 = always, 0 = never (NCCL), unset = from 4 ranks on.  Measured per 1.336 ms step: 2 GPUs NCCL +41 us against
    // +47 us here; 8 GPUs NCCL +92 us against +69 us here (NCCL's latency grows with the ring / tree, the two flag rounds do not)
    static const int mode = dc_env_int("DINCAE_P2P", -1);
    const bool on = mode < 0 ? N >= 4 : mode != 0;
    if (!on || N < 2 || N > P2P_MAX_RANKS || M->p2p_flags == nullptr) return DINCAE_OK;
    cudaIpcMemHandle_t mine;
    int32_t ok = cudaIpcGetMemHandle(&mine, M->all_mem) == cudaSuccess ? 1 : 0;
    if (!ok) cudaGetLastError();
    void* sc = nullptr;
    DC_TRY(dc_scratch(ctx, sizeof(cudaIpcMemHandle_t) * (N + 1) + 256, &sc));
    cudaIpcMemHandle_t* d_all = (cudaIpcMemHandle_t*)sc;
    cudaIpcMemHandle_t* d_mine = d_all + N;
    int32_t* d_ok = (int32_t*)(d_mine + 1);
    p2p_put_handle_kernel<<<1, 1, 0, ctx->stream>>>(mine, d_mine);
    DC_CHECK_LAUNCH();
    DC_TRY(dc_comm_allgather_bytes(ctx, d_mine, d_all, sizeof(cudaIpcMemHandle_t)));
    std::vector<cudaIpcMemHandle_t> all(N);
    DC_CHECK_CUDA(cudaMemcpyAsync(all.data(), d_all, sizeof(cudaIpcMemHandle_t) * N, cudaMemcpyDeviceToHost, ctx->stream));
    DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int p = 0; p < N; p++) M->p2p_peer_base[p] = nullptr;
    M->p2p_peer_base[ctx->comm_rank] = M->all_mem;
    for (int p = 0; p < N && ok; p++) {
        if (p == ctx->comm_rank) continue;
        void* ptr = nullptr;
        if (cudaIpcOpenMemHandle(&ptr, all[p], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = 0; break; }
        M->p2p_peer_base[p] = ptr;
    }
    DC_CHECK_CUDA(cudaMemcpyAsync(d_ok, &ok, sizeof(ok), cudaMemcpyHostToDevice, ctx->stream));
    DC_TRY(dc_comm_allreduce_min_i32(ctx, d_ok));
    int32_t all_ok = 0;
    DC_CHECK_CUDA(cudaMemcpyAsync(&all_ok, d_ok, sizeof(all_ok), cudaMemcpyDeviceToHost, ctx->stream));
    DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    if (all_ok) {
        M->p2p_state = 1;
    } else {
        for (int p = 0; p < N; p++)
            if (p != ctx->comm_rank && M->p2p_peer_base[p]) { cudaIpcCloseMemHandle(M->p2p_peer_base[p]); M->p2p_peer_base[p] = nullptr; }
    }
    return DINCAE_OK;
}

int32_t model_allreduce_grads(dincae_model* M) {
    dincae_ctx* ctx = M->ctx;
    DC_REQUIRE(ctx->comm != nullptr, DINCAE_EINVAL, "no communicator: call dincae_comm_init_rank first");
    if (M->p2p_state == 0) DC_TRY(p2p_setup(M));
    if (M->p2p_state != 1) return dc_allreduce_on(ctx, M->g, M->n_int + 1, ctx->stream);
    P2PArgs a{};
    const size_t g_off = (char*)M->g - (char*)M->all_mem, f_off = (char*)M->p2p_flags - (char*)M->all_mem, s_off = (char*)M->p2p_stage - (char*)M->all_mem;
    for (int p = 0; p < ctx->comm_nranks; p++) {
        a.g[p] = (float*)((char*)M->p2p_peer_base[p] + g_off);
        a.flags[p] = (uint32_t*)((char*)M->p2p_peer_base[p] + f_off);
        a.stage[p] = (float*)((char*)M->p2p_peer_base[p] + s_off);
    }
    a.nranks = ctx->comm_nranks; a.rank = ctx->comm_rank;
    a.n4 = (M->n_int + 1 + 3) / 4;                  // the buffer is allocated with four floats of slack
    M->p2p_epoch += 4;
    a.tag = M->p2p_epoch;
    a.spin_limit = 4000000000ll;                    // ~2 s at 1.9 GHz
    DC_WORK(ctx, 0.0, 2.0 * 16.0 * a.n4 * (a.nranks - 1) / a.nranks);       // bytes this rank pushes over NVLink
    DC_LAUNCH(ctx, p2p_allreduce_kernel, P2P_BLOCKS, P2P_THREADS, 0, a);      // 16 / 32 blocks measured slower, 128 / 148 the same
    DC_CHECK_LAUNCH();
    ctx->collectives++;
    return DINCAE_OK;
}
