// Model orchestration of libdincae_b200: StepModel forward (src/model.jl:182-189), loss_function
// (:192-222), reverse pass (replaces Zygote's tape, src/flux.jl:29-31) and the optimiser step (:33).
#include "model.cuh"
#include "k1_assemble.cuh"
#include "k4_interp.cuh"
#include <math.h>
#include <string.h>

namespace {

// valid-pixel counts handed over BY VALUE (kernel parameter): no host->device copy of a stack variable, hence no stream
// synchronisation in the data-parallel step
struct NValidArg { unsigned long long v[16]; };
// the count handed in by the host and, when inv_n is given, its reciprocal in the same launch (see inv_n_kernel)
__global__ void set_nvalid_kernel(NValidArg a, int nout, unsigned long long* __restrict__ n, float* __restrict__ inv_n = nullptr) {
    if ((int)threadIdx.x < nout) {
        n[threadIdx.x] = a.v[threadIdx.x];
        if (inv_n) inv_n[threadIdx.x] = 1.0f / (float)a.v[threadIdx.x];
    }
}
__global__ void inv_n_kernel(const unsigned long long* __restrict__ n, int nout, float* __restrict__ inv_n) {
    int v = threadIdx.x;
    if (v < nout) inv_n[v] = 1.0f / (float)n[v];     // "/ n_noncloud", src/model.jl:86,92 (Inf for n = 0, as the reference)
}
__global__ void store_loss_kernel(const double* __restrict__ acc, float* __restrict__ slot) { slot[0] = (float)acc[0]; }
__global__ void total_loss_kernel(const float* __restrict__ slot, const double* __restrict__ acc, float* __restrict__ out_f, double* __restrict__ out_d) {
    double t = (slot ? (double)slot[0] : acc[0]) + acc[1];
    if (out_f) out_f[0] = (float)t;
    if (out_d) out_d[0] = t;
}
// the per-sample point offsets travel BY VALUE (kernel parameter) like the valid counts: no staging copy of caller memory, hence
// no stream synchronisation in the point-cloud step
constexpr int PT_MAXB = 256;
struct PtOffsets { int64_t v[PT_MAXB + 1]; };
__global__ void plane_of_kernel(const __grid_constant__ PtOffsets offs, int B, int64_t P, int32_t* __restrict__ plane_of) {
    int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (p >= P) return;
    int lo = 0, hi = B;          // largest b with offsets[b] <= p
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (offs.v[mid] <= p) lo = mid; else hi = mid;
    }
    plane_of[p] = lo;
}
__global__ void count_pts_kernel(const float* __restrict__ x2, int64_t P, unsigned long long* __restrict__ n) {
    unsigned int c = 0;
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) c += x2[p] != 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(n, (unsigned long long)c);
}

inline double cflops(const LayerInfo& y, int B, int H, int W) { return 2.0 * 9 * y.Cin * y.Cout * (double)B * H * W; }
// algorithmic HBM bytes of one conv / dgrad / wgrad call: both pixel tensors once + the weights (real channel counts)
inline double cbytes(const LayerInfo& y, int B, int H, int W) { return 4.0 * ((double)(y.Cin + y.Cout) * B * H * W + 9.0 * y.Cin * y.Cout); }
inline bool has_skip(const dincae_model* M, int l) { return l >= 1 && l < M->L && ((M->cfg.skip_mask >> l) & 1u); }
// tensor-core modes, nearest upsampling: the head conv runs fused with its upsample + skip-add producer in exact FP32
// (k2_head_up_conv); U[0] is then only materialised by the backward pass, for the head's weight gradient
inline bool head_fused(const dincae_model* M) {
    return M->wK != nullptr && M->cfg.upsampling == DINCAE_UP_NEAREST && k2_head_up_conv_eligible(M->dec[0][0].Cin, 2 * M->O);
}

// ... and so does its backward pass: data gradient + up-sampling backward in one FP32 kernel straight from the planar loss
// gradient, weight gradient from the half-resolution source (U[0], GD0 and GU[0] are never written)
inline bool head_bwd_fused(const dincae_model* M) { return head_fused(M) && k2_head_bwd_eligible(M->dec[0][0].Cin, 2 * M->O); }

// the producer of conv_dec_l: upsample(+crop) of y_{l+1} = R[l+1] (+ X[l+1] when level l+1 has a skip) or X[L]
int32_t dec_source(dincae_model* M, int c, int l, int B, bool materialise, ConvSrc* s) {
    ChainBufs& cb = M->cb[c];
    const float* src0 = (l + 1 == M->L) ? cb.X[M->L] : cb.R[l + 1];
    const float* src1 = has_skip(M, l + 1) ? cb.X[l + 1] : nullptr;
    const LayerInfo& d = M->dec[c][l];
    const int Cp = d.Cin_p;                          // channels of the low-resolution source tensors
    const int split = d.split_mode == 2;             // 3xTF32 head conv: U[0] carries [hi | lo] = 2*Cp channels
    if (M->cfg.upsampling == DINCAE_UP_NEAREST && !M->wK) {
        *s = ConvSrc{src0, src1, SRC_UP_NEAREST, M->Hl[l + 1], M->Wl[l + 1], Cp};       // FP32 path: fused into the conv producer
    } else {
        // bilinear, or TF32 mode: materialise u_l = crop(upsample(y_{l+1})) once (TF32-rounded), it feeds the TMA conv + wgrad
        if (materialise) {
            if (M->cfg.upsampling == DINCAE_UP_NEAREST) DC_TRY(k2_up_nearest(M->ctx, src0, src1, B, M->Hl[l + 1], M->Wl[l + 1], Cp, M->Hl[l], M->Wl[l], cb.U[l], split));
            else DC_TRY(k2_up_bilinear(M->ctx, src0, src1, B, M->Hl[l + 1], M->Wl[l + 1], Cp, M->Hl[l], M->Wl[l], cb.U[l], split));
        }
        *s = ConvSrc{cb.U[l], nullptr, SRC_DIRECT, M->Hl[l], M->Wl[l], d.Cin_pe};
    }
    return DINCAE_OK;
}

int32_t chain_forward(dincae_model* M, int c, const float* d_xin, int B, bool train) {
    dincae_ctx* ctx = M->ctx;
    ChainBufs& cb = M->cb[c];
    const int L = M->L;
    const float slope = M->cfg.leaky_slope;
    const int split0 = M->enc[c][0].split_mode == 1;      // split first conv: X[0] = [hi | lo] of the real channels
    if (c == 0) DC_TRY(k2_pack(ctx, d_xin, M->cfg.ninput, nullptr, 0, B, M->Hl[0], M->Wl[0], cb.X[0], M->enc[c][0].Cin_pe, split0));
    else DC_TRY(k2_pack(ctx, M->cb[c - 1].RAW, 2 * M->O, d_xin, M->cfg.ninput, B, M->Hl[0], M->Wl[0], cb.X[0], M->enc[c][0].Cin_pe, split0));   // cat(xout,xin), T9
    for (int l = 0; l < L; l++) {
        const LayerInfo& e = M->enc[c][l];
        ConvArgs a{};
        a.in = ConvSrc{cb.X[l], nullptr, SRC_DIRECT, M->Hl[l], M->Wl[l], e.Cin_pe};
        a.B = B; a.H = M->Hl[l]; a.W = M->Wl[l];
        a.wgt = M->w + e.w_off; a.bias = M->w + e.b_off; a.Cout_p = e.Cout_p;
        a.wgtK = M->wK ? M->wK + e.wk_off : nullptr; a.ldk = 9 * e.Cin_pe; a.in_is_tf32 = 1;     // X[0] from pack, X[l>=1] from pool_kernel (both rounded on store)
        a.in_real_C = e.split_mode == 1 ? 2 * e.Cin : e.Cin;                               // channels past this are zero padding
        a.act = ACT_RELU; a.slope = slope; a.addend = nullptr; a.dst = cb.A[l];
        a.work_flops = cflops(e, B, M->Hl[l], M->Wl[l]);
        const bool fuse_pool = cb.MK[l] != nullptr;                               // even H, W on the tcgen05 path: the conv epilogue pools
        if (fuse_pool) {
            // relu(conv) itself is never written: the pooled tensor feeds the next level, the backward pass needs only its sign
            // (one bit per element, and not even that for a pure forward pass)
            a.pool_dst = cb.X[l + 1];
            a.mask_dst = train ? cb.MK[l] : nullptr;
            a.work_bytes = 4.0 * ((double)B * M->Hl[l] * M->Wl[l] * (e.Cin + (train ? e.Cout / 32.0 : 0.0)) + (double)B * M->Hl[l + 1] * M->Wl[l + 1] * e.Cout + 9.0 * e.Cin * e.Cout);
            DC_REQUIRE(k2_conv_fuses_pool(a), DINCAE_EINVAL, "chain_forward: level %d was planned with a fused pool", l);
        } else {
            a.work_bytes = cbytes(e, B, M->Hl[l], M->Wl[l]);
        }
        DC_TRY(k2_conv(ctx, a));                                                   // Conv(..., relu, pad=1), src/model.jl:267
        if (!fuse_pool) DC_TRY(k2_pool(ctx, cb.A[l], B, M->Hl[l], M->Wl[l], e.Cout_p, cb.X[l + 1]));   // MeanPool, :268
    }
    for (int l = L - 1; l >= 0; l--) {
        const LayerInfo& d = M->dec[c][l];
        if (l == 0 && head_fused(M)) {
            const float* src0 = (L == 1) ? cb.X[1] : cb.R[1];
            const float* src1 = has_skip(M, 1) ? cb.X[1] : nullptr;
            DC_TRY(k2_head_up_conv(ctx, src0, src1, B, M->Hl[1], M->Wl[1], d.Cin_p, d.Cin, M->w + d.w_off, d.Cin_p, d.Cout_p, M->w + d.b_off,
                                   2 * M->O, M->Hl[0], M->Wl[0], cb.RAW, cflops(d, B, M->Hl[0], M->Wl[0]),
                                   4.0 * ((double)d.Cin * (has_skip(M, 1) ? 2 : 1) * B * M->Hl[1] * M->Wl[1] + 2.0 * M->O * B * M->Hl[0] * M->Wl[0])));
            continue;                                                              // Upsample + crop + SumSkip + Conv(identity) in one kernel, :270-272
        }
        ConvSrc s;
        DC_TRY(dec_source(M, c, l, B, true, &s));                                  // Upsample + croppadding (+ SumSkip below), :270-271
        if (l > 0) {
            ConvArgs a{};
            a.in = s; a.B = B; a.H = M->Hl[l]; a.W = M->Wl[l];
            a.wgt = M->w + d.w_off; a.bias = M->w + d.b_off; a.Cout_p = d.Cout_p;
            a.wgtK = M->wK ? M->wK + d.wk_off : nullptr; a.ldk = 9 * d.Cin_p; a.in_is_tf32 = M->wK != nullptr;   // U[l] is rounded on store
            a.act = ACT_RELU; a.slope = slope; a.addend = nullptr; a.dst = cb.R[l];
            a.work_flops = cflops(d, B, M->Hl[l], M->Wl[l]);
            a.work_bytes = cbytes(d, B, M->Hl[l], M->Wl[l]);
            DC_TRY(k2_conv(ctx, a));                                               // Conv(..., relu), :272
        } else if (M->wK) {
            // TF32 mode: the 2*nout-channel head conv runs on the tensor cores too (channels padded to 32), planar output
            ConvArgs a{};
            a.in = s; a.B = B; a.H = M->Hl[0]; a.W = M->Wl[0];
            a.wgt = M->w + d.w_off; a.bias = M->w + d.b_off; a.Cout_p = d.Cout_p;
            a.wgtK = M->wK + d.wk_off; a.ldk = 9 * d.Cin_pe; a.in_is_tf32 = 1;
            a.act = ACT_NONE; a.slope = 0.f; a.addend = nullptr; a.dst = nullptr;
            a.planar_out = cb.RAW; a.planar_C = 2 * M->O; a.planar_accumulate = 0; a.planar_fold = d.split_mode == 2;
            a.work_flops = cflops(d, B, M->Hl[0], M->Wl[0]);
            a.work_bytes = cbytes(d, B, M->Hl[0], M->Wl[0]);
            DC_TRY(k2_conv(ctx, a));
        } else {
            ConvSmallArgs a{};
            a.in = s; a.B = B; a.H = M->Hl[0]; a.W = M->Wl[0];
            a.wgt = M->w + d.w_off; a.ldw = d.Cout_p; a.bias = M->w + d.b_off; a.Cout = 2 * M->O; a.accumulate = 0; a.out = cb.RAW;
            a.work_flops = cflops(d, B, M->Hl[0], M->Wl[0]);
            a.work_bytes = cbytes(d, B, M->Hl[0], M->Wl[0]);
            DC_TRY(k2_conv_small(ctx, a, M->nco));                                 // Conv(..., identity) at l == 1, :251-256
        }
    }
    return DINCAE_OK;
}

int32_t ensure_wT(dincae_model* M) {
    if (!M->wT_dirty) return DINCAE_OK;
    DC_TRY(k2_make_wT(M->ctx, M->w, M->wT, M->wK, M->wTK, M->d_wtab, M->n_wlayers, M->n_wT_copy));
    if (M->wK) DC_TRY(k2_make_wK_split(M->ctx, M->w, M->wK, M->split));      // overwrites the forward copies of the 3xTF32 layers
    M->wT_dirty = false;
    return DINCAE_OK;
}

// bias column partials produced by the element-wise backward kernels: a slice of the same arena (TF32 mode only)
float* colpart_alloc(dincae_model* M, int64_t n_elems4, int Cp, int* nblocks) {
    *nblocks = 0;
    if (!M->wK) return nullptr;
    int nb = k2_colpart_blocks(M->ctx, n_elems4, Cp);
    if (nb == 0) return nullptr;
    size_t need = ((size_t)nb * Cp * sizeof(float) + 255) / 256 * 256;
    if (M->wg_off + need > M->wg_scratch_bytes) return nullptr;
    float* p = (float*)((char*)M->wg_scratch + M->wg_off);
    M->wg_off += need;
    *nblocks = nb;
    return p;
}

// The weight gradients hang off the backward chain (nothing downstream reads them before the final reduction), so in TF32 mode
// they are issued on the context's side stream: each one forks after its g tensor has been produced and runs beside the
// data-gradient / level kernels of the main stream; wgrad_join() makes the main stream wait before buffers are reused.
// DINCAE_DP_OVERLAP=1: per-chain gradient buckets all-reduced on the communicator's stream behind the backward pass instead of one
// all-reduce after it.  Off by default because it measured SLOWER (2 GPUs: 1.387 against 1.377 ms per step; with NCCL held to 2 CTAs
// 1.69 ms): the persistent conv kernels are launched with exactly one CTA (pair) per SM, so every SM that a concurrent NCCL kernel
// occupies leaves a CTA waiting for a whole kernel duration.
static bool dp_overlap_on() {
    static const int on = dc_env_int("DINCAE_DP_OVERLAP", 0);
    return on != 0;
}
static bool side_stream_on() {
    static const int on = dc_env_int("DINCAE_SIDE_STREAM", 1);
    return on != 0;
}
static int32_t wgrad_fork(dincae_ctx* ctx) {
    if (!ctx->side) {
        DC_CHECK_CUDA(cudaStreamCreateWithFlags(&ctx->side, cudaStreamNonBlocking));
        DC_CHECK_CUDA(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
        DC_CHECK_CUDA(cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
    }
    DC_CHECK_CUDA(cudaEventRecord(ctx->ev_fork, ctx->stream));
    DC_CHECK_CUDA(cudaStreamWaitEvent(ctx->side, ctx->ev_fork, 0));
    return DINCAE_OK;
}
int32_t wgrad_join(dincae_model* M) {
    dincae_ctx* ctx = M->ctx;
    if (!M->side_pending) return DINCAE_OK;
    DC_CHECK_CUDA(cudaEventRecord(ctx->ev_join, ctx->side));
    DC_CHECK_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_join, 0));
    M->side_pending = false;
    return DINCAE_OK;
}

// weight gradient with its own scratch slice; in TF32 mode the second-stage reductions are queued in M->red
int32_t run_wgrad(dincae_model* M, WgradArgs& wa) {
    if (!M->wK) return k2_wgrad(M->ctx, wa, M->wg_scratch, M->wg_scratch_bytes);
    size_t need = (k2_wgrad_tma_scratch_bytes(M->ctx->num_sms, wa.B, wa.H, wa.W, wa.in.C, wa.Cout_p) + 255) / 256 * 256;
    DC_REQUIRE(M->wg_off + need <= M->wg_scratch_bytes, DINCAE_ESIZE, "weight-gradient scratch arena exhausted");
    wa.defer = &M->red;
    dincae_ctx* ctx = M->ctx;
    const bool side = side_stream_on() && !ctx->prof;          // the per-kernel profile times kernels one after the other
    cudaStream_t main_stream = ctx->stream;
    if (side) {
        DC_TRY(wgrad_fork(ctx));
        ctx->stream = ctx->side;
        M->side_pending = true;
    }
    int32_t st = k2_wgrad(ctx, wa, (char*)M->wg_scratch + M->wg_off, need);
    ctx->stream = main_stream;
    M->wg_off += need;
    return st;
}

int32_t chain_backward(dincae_model* M, int c, int B) {
    dincae_ctx* ctx = M->ctx;
    ChainBufs& cb = M->cb[c];
    const int L = M->L;
    const float slope = M->cfg.leaky_slope;
    const bool nearest = M->cfg.upsampling == DINCAE_UP_NEAREST;
    // ---- decoder, top (l = 0) to bottom
    const bool hbf = head_bwd_fused(M);
    if (hbf) {
        // head layer: weight gradient on the side stream, data gradient + up^T (+ relu' of level 1) on the main stream
        const LayerInfo& d = M->dec[c][0];
        const float* src0 = (L == 1) ? cb.X[1] : cb.R[1];
        const float* src1 = has_skip(M, 1) ? cb.X[1] : nullptr;
        const size_t need = (k2_head_bwd_scratch_bytes(B, M->Hl[1], M->Wl[1], d.Cin, 2 * M->O) + 255) / 256 * 256;
        DC_REQUIRE(M->wg_off + need <= M->wg_scratch_bytes, DINCAE_ESIZE, "weight-gradient scratch arena exhausted");
        const bool side = side_stream_on() && !ctx->prof;
        cudaStream_t main_stream = ctx->stream;
        if (side) {
            DC_TRY(wgrad_fork(ctx));
            ctx->stream = ctx->side;
            M->side_pending = true;
        }
        int32_t st = k2_head_wgrad(ctx, src0, src1, cb.GRAW, B, M->Hl[0], M->Wl[0], M->Hl[1], M->Wl[1], d.Cin_p, d.Cin, 2 * M->O, d.Cin_p, d.Cout_p,
                                   M->g + d.w_off, M->g + d.b_off, (char*)M->wg_scratch + M->wg_off, need);
        ctx->stream = main_stream;
        M->wg_off += need;
        DC_TRY(st);
        DC_TRY(k2_head_bwd_data(ctx, cb.GRAW, B, M->Hl[0], M->Wl[0], M->Hl[1], M->Wl[1], M->w + d.w_off, d.Cin, d.Cin_p, d.Cout_p, 2 * M->O,
                                L == 1 ? nullptr : cb.R[1], slope, d.Cin_p, M->G[1], L == 1 ? nullptr : M->GD[1]));
    } else {
        DC_TRY(k2_pack(ctx, cb.GRAW, 2 * M->O, nullptr, 0, B, M->Hl[0], M->Wl[0], M->GD0, M->nco));
    }
    for (int l = hbf ? 1 : 0; l < L; l++) {
        const LayerInfo& d = M->dec[c][l];
        const float* gd = (l == 0) ? M->GD0 : M->GD[l];
        float* colpart = nullptr;
        int colblocks = 0;
        if (l > 0 && !(hbf && l == 1)) {
            // G[l] = up^T(GU[l-1]); GD[l] = G[l] * relu'(R[l])  (+ per-block column sums of GD[l] = first stage of the bias gradient)
            if (nearest) {
                colpart = colpart_alloc(M, (int64_t)B * M->Hl[l] * M->Wl[l] * (d.Cout_p / 4), d.Cout_p, &colblocks);
                DC_TRY(k2_upbwd(ctx, M->GU[l - 1], B, M->Hl[l - 1], M->Wl[l - 1], d.Cout_p, cb.R[l], slope, M->G[l], M->GD[l], colpart));
            }
            else DC_TRY(k2_up_bilinear_bwd(ctx, M->GU[l - 1], B, M->Hl[l - 1], M->Wl[l - 1], d.Cout_p, M->Hl[l], M->Wl[l], cb.R[l], slope, M->G[l], M->GD[l]));
        }
        ConvSrc s;
        DC_TRY(dec_source(M, c, l, B, l == 0 && head_fused(M), &s));       // fused head: the forward pass did not materialise U[0]
        WgradArgs wa{};
        wa.in = s; wa.B = B; wa.H = M->Hl[l]; wa.W = M->Wl[l]; wa.gout = gd; wa.Cout_p = d.Cout_p;
        wa.dW = d.split_mode ? M->g_exp + d.gexp_off : M->g + d.w_off; wa.db = M->g + d.b_off;     // 3xTF32 layer: expanded gradient, folded after the reductions
        wa.colpart = colpart; wa.colpart_blocks = colblocks;
        wa.work_flops = cflops(d, B, M->Hl[l], M->Wl[l]);
        wa.work_bytes = cbytes(d, B, M->Hl[l], M->Wl[l]);
        wa.tf32 = M->wK != nullptr;
        if (l == 0 && !M->wK) DC_TRY(k2_wgrad_small(ctx, wa, M->nco, M->wg_scratch, M->wg_scratch_bytes));
        else DC_TRY(run_wgrad(M, wa));
        ConvArgs a{};
        a.in = ConvSrc{gd, nullptr, SRC_DIRECT, M->Hl[l], M->Wl[l], d.Cout_p};
        a.B = B; a.H = M->Hl[l]; a.W = M->Wl[l];
        a.wgt = M->wT + d.wT_off; a.bias = nullptr; a.Cout_p = d.Cin_p; a.act = ACT_NONE; a.slope = 0.f; a.addend = nullptr; a.dst = M->GU[l];
        a.wgtK = M->wTK ? M->wTK + d.wtk_off : nullptr; a.ldk = 9 * d.Cout_p; a.in_is_tf32 = 1;   // GD[l>=1] from upbwd, GD0 from pack (both rounded on store)
        a.in_real_C = d.Cout;                                                               // the gradient tensor's channels past Cout are zero
        a.work_flops = cflops(d, B, M->Hl[l], M->Wl[l]);
        a.work_bytes = cbytes(d, B, M->Hl[l], M->Wl[l]);
        DC_TRY(k2_conv(ctx, a));
    }
    if (!(hbf && L == 1)) {   // G[L] = gradient w.r.t. X[L] (the innermost level is the identity, src/model.jl:263-264)
        const LayerInfo& d = M->dec[c][L - 1];
        if (nearest) DC_TRY(k2_upbwd(ctx, M->GU[L - 1], B, M->Hl[L - 1], M->Wl[L - 1], d.Cin_p, nullptr, slope, M->G[L], nullptr));
        else DC_TRY(k2_up_bilinear_bwd(ctx, M->GU[L - 1], B, M->Hl[L - 1], M->Wl[L - 1], d.Cin_p, M->Hl[L], M->Wl[L], nullptr, slope, M->G[L], nullptr));
    }
    // ---- encoder, bottom to top
    const float* xg = M->G[L];
    for (int l = L - 1; l >= 0; l--) {
        const LayerInfo& e = M->enc[c][l];
        int ecolblocks = 0;
        float* ecolpart = colpart_alloc(M, (int64_t)B * M->Hl[l] * M->Wl[l] * (e.Cout_p / 4), e.Cout_p, &ecolblocks);
        if (cb.MK[l]) DC_TRY(k2_poolbwd_mask(ctx, xg, cb.MK[l], B, M->Hl[l], M->Wl[l], e.Cout_p, slope, M->GC[l], ecolpart));
        else DC_TRY(k2_poolbwd(ctx, xg, cb.A[l], B, M->Hl[l], M->Wl[l], e.Cout_p, slope, M->GC[l], ecolpart));
        WgradArgs wa{};
        wa.in = ConvSrc{cb.X[l], nullptr, SRC_DIRECT, M->Hl[l], M->Wl[l], e.Cin_pe};
        wa.B = B; wa.H = M->Hl[l]; wa.W = M->Wl[l]; wa.gout = M->GC[l]; wa.Cout_p = e.Cout_p;
        wa.dW = e.split_mode ? M->g_exp + e.gexp_off : M->g + e.w_off; wa.db = M->g + e.b_off;
        wa.colpart = ecolpart; wa.colpart_blocks = ecolblocks;
        wa.work_flops = cflops(e, B, M->Hl[l], M->Wl[l]);
        wa.work_bytes = cbytes(e, B, M->Hl[l], M->Wl[l]);
        wa.tf32 = M->wK != nullptr;
        DC_TRY(run_wgrad(M, wa));
        if (l >= 1) {
            ConvArgs a{};
            a.in = ConvSrc{M->GC[l], nullptr, SRC_DIRECT, M->Hl[l], M->Wl[l], e.Cout_p};
            a.B = B; a.H = M->Hl[l]; a.W = M->Wl[l];
            a.wgt = M->wT + e.wT_off; a.bias = nullptr; a.Cout_p = e.Cin_p; a.act = ACT_NONE; a.slope = 0.f;
            a.wgtK = M->wTK ? M->wTK + e.wtk_off : nullptr; a.ldk = 9 * e.Cout_p; a.in_is_tf32 = 1;        // GC[l] from poolbwd (rounded)
            a.addend = has_skip(M, l) ? M->G[l] : nullptr;       // SumSkip: the level input also feeds the level output
            a.dst = M->XG[l];
            a.work_flops = cflops(e, B, M->Hl[l], M->Wl[l]);
            a.work_bytes = cbytes(e, B, M->Hl[l], M->Wl[l]);
            DC_TRY(k2_conv(ctx, a));
            xg = M->XG[l];
        } else if (c >= 1 && M->wK) {
            // gradient w.r.t. the first 2*nout input channels of the refinement chain = raw output of chain c-1 (tensor cores, planar +=)
            ConvArgs a{};
            a.in = ConvSrc{M->GC[0], nullptr, SRC_DIRECT, M->Hl[0], M->Wl[0], e.Cout_p};
            a.B = B; a.H = M->Hl[0]; a.W = M->Wl[0];
            a.wgt = M->wT + e.wT_off; a.bias = nullptr; a.Cout_p = e.Cin_p; a.act = ACT_NONE; a.slope = 0.f; a.addend = nullptr; a.dst = nullptr;
            a.wgtK = M->wTK + e.wtk_off; a.ldk = 9 * e.Cout_p; a.in_is_tf32 = 1;
            a.planar_out = M->cb[c - 1].GRAW; a.planar_C = 2 * M->O; a.planar_accumulate = 1;
            a.work_flops = 2.0 * 9 * 2 * M->O * e.Cout * (double)B * M->Hl[0] * M->Wl[0];
            DC_TRY(k2_conv(ctx, a));
        } else if (c >= 1) {
            // gradient w.r.t. the first 2*nout input channels of the refinement chain = raw output of chain c-1
            ConvSmallArgs a{};
            a.in = ConvSrc{M->GC[0], nullptr, SRC_DIRECT, M->Hl[0], M->Wl[0], e.Cout_p};
            a.B = B; a.H = M->Hl[0]; a.W = M->Wl[0];
            a.wgt = M->wT + e.wT_off; a.ldw = e.Cin_p; a.bias = nullptr; a.Cout = 2 * M->O; a.accumulate = 1; a.out = M->cb[c - 1].GRAW;
            a.work_flops = 2.0 * 9 * 2 * M->O * e.Cout * (double)B * M->Hl[0] * M->Wl[0];
            DC_TRY(k2_conv_small(ctx, a, M->nco));
        }
    }
    DC_TRY(wgrad_join(M));       // the next chain (or step) reuses the gradient workspace the side-stream kernels read
    return DINCAE_OK;
}

HeadParams head_params(const dincae_model* M) {
    HeadParams hp;
    hp.gamma = M->cfg.gamma;
    hp.truth_uncertain = M->cfg.truth_uncertain;
    hp.lap = M->cfg.laplacian_penalty;
    hp.lap_err = M->cfg.laplacian_error_penalty;
    return hp;
}

int32_t check_batch(const dincae_model* M, int B, const char* who) {
    DC_REQUIRE(M != nullptr, DINCAE_EINVAL, "%s: model is NULL", who);
    DC_REQUIRE(B >= 1 && B <= M->maxB, DINCAE_ESIZE, "%s: batch %d outside [1, max_batch=%d]", who, B, M->maxB);
    return DINCAE_OK;
}

// bring a (W,H,C,B) host/device tensor to the device
int32_t stage_batch(dincae_model* M, const float* src, int C, int B, int32_t loc, float* staging, const float** out) {
    if (loc == DINCAE_DEVICE) { *out = src; return DINCAE_OK; }
    size_t n = (size_t)M->cfg.W * M->cfg.H * C * B;
    DC_CHECK_CUDA(cudaMemcpyAsync(staging, src, sizeof(float) * n, cudaMemcpyHostToDevice, M->ctx->stream));
    *out = staging;
    return DINCAE_OK;
}

// gridded loss of all chains (+ gradients w.r.t. the raw chain outputs when train)
// 1 / n_valid per output variable (the global count in a data-parallel step, else counted from this batch)
int32_t gridded_counts(dincae_model* M, const float* d_xtrue, int B, const int64_t* n_valid_global) {
    dincae_ctx* ctx = M->ctx;
    const int64_t npix = (int64_t)M->cfg.W * M->cfg.H;
    if (n_valid_global) {
        NValidArg nv{};
        DC_REQUIRE(M->O <= 16, DINCAE_EUNSUPPORTED, "at most 16 output variables");
        for (int v = 0; v < M->O; v++) nv.v[v] = (unsigned long long)n_valid_global[v];
        DC_LAUNCH(ctx, set_nvalid_kernel, 1, 32, 0, nv, M->O, M->d_nvalid, M->d_inv_n);
        DC_CHECK_LAUNCH();
        return DINCAE_OK;
    }
    DC_TRY(k3_count_valid(ctx, d_xtrue, B, M->O, npix, M->d_nvalid));
    DC_LAUNCH(ctx, inv_n_kernel, 1, 32, 0, M->d_nvalid, M->O, M->d_inv_n);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}
int32_t gridded_chain_loss(dincae_model* M, int c, const float* d_xtrue, int B, bool train) {
    dincae_ctx* ctx = M->ctx;
    const int64_t npix = (int64_t)M->cfg.W * M->cfg.H;
    HeadParams hp = head_params(M);
    const bool lap = hp.lap != 0.f || hp.lap_err != 0.f;
    if (lap) DC_TRY(k3_head(ctx, M->cb[c].RAW, B, M->O, npix, hp.gamma, M->MS));
    return k3_loss_grid(ctx, M->cb[c].RAW, d_xtrue, lap ? M->MS : nullptr, B, M->O, M->cfg.W, M->cfg.H, hp, M->d_inv_n,
                        M->cfg.loss_weights[c], M->d_loss, train ? M->cb[c].GRAW : nullptr, M->red_scratch, M->red_scratch_bytes);
}
int32_t gridded_chain_losses(dincae_model* M, const float* d_xtrue, int B, bool train) {
    for (int c = 0; c < M->nchains; c++) DC_TRY(gridded_chain_loss(M, c, d_xtrue, B, train));
    return DINCAE_OK;
}

struct PointTruth {
    const float* pos = nullptr;     // device (P,2)
    const float* x = nullptr;       // device (P,2) column-major
    int32_t* plane_of = nullptr;    // device (P)
    int64_t P = 0;
    DevIn spos, sx;
};

int32_t point_workspace(dincae_model* M, int64_t P) {
    if ((size_t)P <= M->pt_cap) return DINCAE_OK;
    DC_CHECK_CUDA(cudaStreamSynchronize(M->ctx->stream));
    if (M->pt_buf) cudaFree(M->pt_buf);
    M->pt_buf = nullptr;
    M->pt_cap = 0;
    size_t cap = ((size_t)P + (size_t)P / 4 + 1024 + 3) & ~(size_t)3;     // multiple of 4 floats: the int64 offsets behind the float arrays stay 8-byte aligned
    // mv[2P] gathered (m,s2); gms[2P] their gradients; plane_of[P]; offsets
    cudaError_t e = cudaMalloc((void**)&M->pt_buf, sizeof(float) * (5 * cap) + sizeof(int64_t) * (M->maxB + 2));
    if (e != cudaSuccess) { dincae_set_error("point workspace allocation failed: %s", cudaGetErrorString(e)); return DINCAE_ENOMEM; }
    M->pt_cap = cap;
    return DINCAE_OK;
}

int32_t stage_points(dincae_model* M, const float* pos, const float* x, const int64_t* offsets, int B, int32_t loc, PointTruth* pt) {
    dincae_ctx* ctx = M->ctx;
    DC_REQUIRE(offsets != nullptr, DINCAE_EINVAL, "point truth: offsets is NULL (host array of B+1 int64)");
    DC_REQUIRE(offsets[0] == 0, DINCAE_EINVAL, "point truth: offsets[0] must be 0");
    for (int b = 0; b < B; b++) DC_REQUIRE(offsets[b + 1] >= offsets[b], DINCAE_EINVAL, "point truth: offsets not monotone at %d", b);
    pt->P = offsets[B];
    DC_TRY(point_workspace(M, pt->P));
    DC_TRY(dc_stage_in(ctx, pos, sizeof(float) * 2 * pt->P, loc, &pt->spos));
    DC_TRY(dc_stage_in(ctx, x, sizeof(float) * 2 * pt->P, loc, &pt->sx));
    pt->pos = (const float*)pt->spos.ptr;
    pt->x = (const float*)pt->sx.ptr;
    pt->plane_of = (int32_t*)(M->pt_buf + 4 * M->pt_cap);
    DC_REQUIRE(B <= PT_MAXB, DINCAE_ESIZE, "point truth: at most %d samples per mini-batch", PT_MAXB);
    if (pt->P > 0) {
        PtOffsets po;
        for (int b = 0; b <= B; b++) po.v[b] = offsets[b];
        DC_LAUNCH(ctx, plane_of_kernel, (unsigned)dc_cdiv(pt->P, 256), 256, 0, po, B, pt->P, pt->plane_of);
        DC_CHECK_LAUNCH();
    }
    return DINCAE_OK;
}

// point-cloud cost of all chains (src/points.jl:354-426) (+ gradients w.r.t. raw when train)
int32_t points_loss(dincae_model* M, PointTruth& pt, int B, int64_t n_valid_global, bool train) {
    dincae_ctx* ctx = M->ctx;
    DC_REQUIRE(M->O == 1, DINCAE_EUNSUPPORTED, "point loss supports noutput == 1 (src/points.jl:573)");
    const int W = M->cfg.W, H = M->cfg.H;
    const int64_t npix = (int64_t)W * H;
    const int64_t P = pt.P;
    float* mv = M->pt_buf;
    float* gms = M->pt_buf + 2 * M->pt_cap;
    if (n_valid_global >= 0) {
        NValidArg nv{};
        nv.v[0] = (unsigned long long)n_valid_global;
        DC_LAUNCH(ctx, set_nvalid_kernel, 1, 32, 0, nv, 1, M->d_nvalid);
        DC_CHECK_LAUNCH();
    } else {
        DC_CHECK_CUDA(cudaMemsetAsync(M->d_nvalid, 0, sizeof(unsigned long long), ctx->stream));
        if (P > 0) {
            DC_LAUNCH(ctx, count_pts_kernel, 64, 256, 0, pt.x + P, P, M->d_nvalid);
            DC_CHECK_LAUNCH();
        }
    }
    DC_LAUNCH(ctx, inv_n_kernel, 1, 32, 0, M->d_nvalid, 1, M->d_inv_n);
    DC_CHECK_LAUNCH();
    HeadParams hp = head_params(M);
    K4Batch kb{pt.pos, pt.plane_of, nullptr, P, W, H, B, 2 * npix, npix, 2};
    void* sc = nullptr;
    if (train) DC_TRY(dc_scratch(ctx, k4_scatter_scratch_bytes(P), &sc));
    for (int c = 0; c < M->nchains; c++) {
        const float wgt = M->cfg.loss_weights[c];
        DC_TRY(k3_head(ctx, M->cb[c].RAW, B, 1, npix, hp.gamma, M->MS));
        DC_TRY(k4_gather(ctx, kb, M->MS, mv));                                     // interpnd of m and sigma2, src/points.jl:399-402
        DC_TRY(k3_loss_points(ctx, mv, pt.x, P, hp, M->d_inv_n, wgt, M->d_loss, train ? gms : nullptr, M->red_scratch, M->red_scratch_bytes));
        if (train) {
            // pullback of interpnd = interp_adjn (src/points.jl:36-43), both fields in one sorted pass
            DC_TRY(k4_scatter_sorted(ctx, kb, gms, gms + P, M->GMS, sc, ctx->scratch_bytes));
        }
        DC_TRY(k3_head_bwd(ctx, M->cb[c].RAW, M->MS, train ? M->GMS : nullptr, B, 1, W, H, hp, wgt, M->d_loss, train ? M->cb[c].GRAW : nullptr,
                           M->red_scratch, M->red_scratch_bytes));
    }
    return DINCAE_OK;
}

int32_t gridded_loss(dincae_model* M, const float* d_xtrue, int B, const int64_t* n_valid_global, bool train) {
    DC_TRY(gridded_counts(M, d_xtrue, B, n_valid_global));
    return gridded_chain_losses(M, d_xtrue, B, train);
}

int32_t backward_all(dincae_model* M, int B) {
    M->ctx->tf32_acts = M->wK != nullptr;
    DC_TRY(ensure_wT(M));
    M->wg_off = 0;
    M->red.count = 0;
    dincae_ctx* ctx = M->ctx;
    if (!M->dp_overlap) {
        const bool branch = side_stream_on() && !ctx->prof && M->wK != nullptr;
        for (int c = M->nchains - 1; c >= 0; c--) {
            DC_TRY(chain_backward(M, c, B));
            if (c > 0 && branch && M->red.count > 0) {
                // the split-K sums of chain c (an HBM-bound kernel without shared memory) run on the side stream beside the backward
                // pass of chain c-1; that chain's weight gradients queue behind them, its final join covers both
                cudaStream_t main_stream = ctx->stream;
                DC_TRY(wgrad_fork(ctx));
                ctx->stream = ctx->side;
                M->side_pending = true;
                int32_t st = k2_reduce_batched(ctx, M->red);
                ctx->stream = main_stream;
                DC_TRY(st);
            }
        }
        DC_TRY(k2_reduce_batched(ctx, M->red));
        DC_TRY(k2_fold_split_grad(ctx, M->g_exp, M->g, M->split));
        DC_LAUNCH(ctx, store_loss_kernel, 1, 1, 0, M->d_loss, M->g + M->n_int);
        DC_CHECK_LAUNCH();
        return DINCAE_OK;
    }
    // data-parallel step: the bucket of chain c (its parameter range; the last one also carries the loss slot) is finished and
    // handed to NCCL on the communicator's stream as soon as chain c's backward pass ends; chains c-1 .. 0 keep computing
    // (under the per-launch profile the same buckets go out on the main stream: every rank must issue the same collectives)
    cudaStream_t cs = ctx->prof ? ctx->stream : ctx->comm_stream;
    for (int c = M->nchains - 1; c >= 0; c--) {
        DC_TRY(chain_backward(M, c, B));
        DC_TRY(k2_reduce_batched(ctx, M->red));
        DC_TRY(k2_fold_split_grad(ctx, M->g_exp, M->g, M->split_chain[c]));
        int64_t n = M->chain_off[c + 1] - M->chain_off[c];
        if (c == 0) {
            DC_LAUNCH(ctx, store_loss_kernel, 1, 1, 0, M->d_loss, M->g + M->n_int);
            DC_CHECK_LAUNCH();
            if (M->nchains == 1) n += 1;
        }
        if (cs != ctx->stream) {
            DC_CHECK_CUDA(cudaEventRecord(ctx->ev_comm_go, ctx->stream));
            DC_CHECK_CUDA(cudaStreamWaitEvent(cs, ctx->ev_comm_go, 0));
        }
        DC_TRY(dc_allreduce_on(ctx, M->g + M->chain_off[c], n, cs));
        if (c == 0 && M->nchains > 1) DC_TRY(dc_allreduce_on(ctx, M->g + M->n_int, 1, cs));     // the loss slot sits behind the LAST chain's range
    }
    if (cs != ctx->stream) {
        DC_CHECK_CUDA(cudaEventRecord(ctx->ev_comm_done, cs));
        DC_CHECK_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_comm_done, 0));
    }
    return DINCAE_OK;
}

int32_t read_loss(dincae_model* M, bool from_slot, float* out_f, double* out_d) {
    dincae_ctx* ctx = M->ctx;
    void* pin = nullptr;
    DC_TRY(dc_pinned(ctx, 64, &pin));
    float* df = (float*)(M->d_loss + 2);
    double* dd = M->d_loss + 3;
    DC_LAUNCH(ctx, total_loss_kernel, 1, 1, 0, from_slot ? M->g + M->n_int : (const float*)nullptr, M->d_loss, df, dd);
    DC_CHECK_LAUNCH();
    DC_CHECK_CUDA(cudaMemcpyAsync(pin, M->d_loss + 2, 16, cudaMemcpyDeviceToHost, ctx->stream));
    DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    if (out_f) *out_f = *(float*)pin;
    if (out_d) *out_d = *(double*)((char*)pin + 8);
    return model_p2p_check(M);
}

// Flux order <-> internal padded layout (host side)
void flux_to_internal(const dincae_model* M, const float* flux, std::vector<float>& internal) {
    internal.assign((size_t)M->n_int + 1, 0.f);
    for (int c = 0; c < M->nchains; c++)
        for (int kind = 0; kind < 2; kind++)
            for (int l = 0; l < M->L; l++) {
                const LayerInfo& y = kind == 0 ? M->enc[c][l] : M->dec[c][l];
                for (int co = 0; co < y.Cout; co++) {
                    for (int ci = 0; ci < y.Cin; ci++)
                        for (int b = 0; b < 3; b++)
                            for (int a = 0; a < 3; a++) {
                                int ky = 2 - b, kx = 2 - a;      // true convolution -> cross-correlation taps
                                internal[y.w_off + ((size_t)(ky * 3 + kx) * y.Cin_p + ci) * y.Cout_p + co] =
                                    flux[y.flux_w_off + a + 3 * (b + 3 * (ci + (size_t)y.Cin * co))];
                            }
                    internal[y.b_off + co] = flux[y.flux_b_off + co];
                }
            }
}
void internal_to_flux(const dincae_model* M, const std::vector<float>& internal, float* flux) {
    for (int c = 0; c < M->nchains; c++)
        for (int kind = 0; kind < 2; kind++)
            for (int l = 0; l < M->L; l++) {
                const LayerInfo& y = kind == 0 ? M->enc[c][l] : M->dec[c][l];
                for (int co = 0; co < y.Cout; co++) {
                    for (int ci = 0; ci < y.Cin; ci++)
                        for (int b = 0; b < 3; b++)
                            for (int a = 0; a < 3; a++) {
                                int ky = 2 - b, kx = 2 - a;
                                flux[y.flux_w_off + a + 3 * (b + 3 * (ci + (size_t)y.Cin * co))] =
                                    internal[y.w_off + ((size_t)(ky * 3 + kx) * y.Cin_p + ci) * y.Cout_p + co];
                            }
                    flux[y.flux_b_off + co] = internal[y.b_off + co];
                }
            }
}

int32_t upload_internal(dincae_model* M, float* dst, const float* flux) {
    std::vector<float> tmp;
    flux_to_internal(M, flux, tmp);
    DC_CHECK_CUDA(cudaMemcpyAsync(dst, tmp.data(), sizeof(float) * M->n_int, cudaMemcpyHostToDevice, M->ctx->stream));
    DC_CHECK_CUDA(cudaStreamSynchronize(M->ctx->stream));
    return DINCAE_OK;
}
int32_t download_internal(dincae_model* M, const float* src, float* flux) {
    std::vector<float> tmp((size_t)M->n_int + 1);
    DC_CHECK_CUDA(cudaMemcpyAsync(tmp.data(), src, sizeof(float) * M->n_int, cudaMemcpyDeviceToHost, M->ctx->stream));
    DC_CHECK_CUDA(cudaStreamSynchronize(M->ctx->stream));
    internal_to_flux(M, tmp, flux);
    return DINCAE_OK;
}

}  // namespace

int32_t model_forward(dincae_model* M, const float* d_xin, int B, bool train) {
    M->ctx->tf32_acts = M->wK != nullptr;
    if (M->wK) DC_TRY(ensure_wT(M));        // the tensor-core forward reads the K-major TF32 weight copy
    for (int c = 0; c < M->nchains; c++) DC_TRY(chain_forward(M, c, d_xin, B, train));
    return DINCAE_OK;
}

int32_t model_head_last(dincae_model* M, int B, float* d_out) {
    return k3_head(M->ctx, M->cb[M->nchains - 1].RAW, B, M->O, (int64_t)M->cfg.W * M->cfg.H, M->cfg.gamma, d_out);
}

// ---------------------------------------------------------------------------------------------
extern "C" int32_t dincae_model_create(dincae_ctx* ctx, const dincae_config* cfg, dincae_model** out) {
    DC_REQUIRE(ctx && cfg && out, DINCAE_EINVAL, "dincae_model_create: NULL argument");
    DC_REQUIRE(cfg->W >= 2 && cfg->H >= 2 && cfg->ninput >= 1, DINCAE_EINVAL, "dincae_model_create: bad W/H/ninput");
    DC_REQUIRE(cfg->nlevels >= 1 && cfg->nlevels <= DINCAE_MAX_LEVELS, DINCAE_EINVAL, "dincae_model_create: nlevels %d outside [1,%d]", cfg->nlevels, DINCAE_MAX_LEVELS);
    DC_REQUIRE(cfg->nchains >= 1 && cfg->nchains <= DINCAE_MAX_CHAINS, DINCAE_EINVAL, "dincae_model_create: nchains %d outside [1,%d]", cfg->nchains, DINCAE_MAX_CHAINS);
    DC_REQUIRE(cfg->noutput >= 1 && cfg->noutput <= 4, DINCAE_EUNSUPPORTED, "dincae_model_create: noutput %d outside [1,4]", cfg->noutput);
    DC_REQUIRE(cfg->upsampling == DINCAE_UP_NEAREST || cfg->upsampling == DINCAE_UP_BILINEAR, DINCAE_EUNSUPPORTED, "dincae_model_create: unknown upsampling_method %d", cfg->upsampling);
    DC_REQUIRE(!(cfg->skip_mask & 1u), DINCAE_EUNSUPPORTED, "dincae_model_create: a skip connection at level 1 needs ninput == 2*noutput and is not supported");
    DC_REQUIRE(cfg->max_batch >= 1, DINCAE_EINVAL, "dincae_model_create: max_batch < 1");
    DC_REQUIRE(cfg->precision == DINCAE_PREC_FP32 || cfg->precision == DINCAE_PREC_TF32 || cfg->precision == DINCAE_PREC_TF32X1, DINCAE_EUNSUPPORTED, "dincae_model_create: unknown precision %d", cfg->precision);
    for (int l = 0; l < cfg->nlevels; l++) DC_REQUIRE(cfg->nfilter[l] >= 1, DINCAE_EINVAL, "dincae_model_create: nfilter[%d] < 1", l);
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));

    dincae_model* M = new dincae_model();
    M->ctx = ctx; M->cfg = *cfg; M->L = cfg->nlevels; M->nchains = cfg->nchains; M->O = cfg->noutput; M->maxB = cfg->max_batch;
    const bool tc_mode = cfg->precision != DINCAE_PREC_FP32;
    const bool split3 = cfg->precision == DINCAE_PREC_TF32 && 2 * cfg->noutput <= 8;     // 3xTF32 first + head convs
    const int cpad = tc_mode ? 32 : 4;     // tensor-core modes: every conv is tensor-core eligible (32-channel K chunks)
    M->nco = dc_round_up(2 * M->O, cpad);
    const int L = M->L;
    M->Hl[0] = cfg->H; M->Wl[0] = cfg->W;
    for (int l = 0; l < L; l++) { M->Hl[l + 1] = M->Hl[l] / 2 + M->Hl[l] % 2; M->Wl[l + 1] = M->Wl[l] / 2 + M->Wl[l] % 2; }   // sz_small, src/model.jl:258-261

    // ---- topology and parameter layout (Flux.trainables order: enc1..encL, decL..dec1 per chain)
    M->enc.resize(M->nchains); M->dec.resize(M->nchains);
    int64_t off = 0, foff = 0, toff = 0, koff = 0, tkoff = 0, goff = 0;
    M->split.count = 0;
    std::vector<WLayerDesc> wtab;
    for (int c = 0; c < M->nchains; c++) {
        std::vector<int> encf(L + 1), decf(L + 1);
        encf[0] = cfg->ninput + (c > 0 ? 2 * M->O : 0);            // enc_nfilter2[1] += dec_nfilter[1], src/model.jl:333
        decf[0] = 2 * M->O;
        for (int l = 0; l < L; l++) { encf[l + 1] = cfg->nfilter[l]; decf[l + 1] = cfg->nfilter[l]; }
        M->enc[c].resize(L); M->dec[c].resize(L);
        auto fill = [&](LayerInfo& y, int cin, int cout, int split_mode) {
            y.Cin = cin; y.Cout = cout; y.Cin_p = dc_round_up(cin, cpad); y.Cout_p = dc_round_up(cout, cpad);
            y.split_mode = split3 ? split_mode : 0;
            y.Cin_pe = y.split_mode == 1 ? dc_round_up(2 * cin, cpad) : (y.split_mode == 2 ? 2 * y.Cin_p : y.Cin_p);
            y.gexp_off = 0;
            y.Kpad = dc_round_up(9 * y.Cin_p, 16); y.KpadT = dc_round_up(9 * y.Cout_p, 16);
            y.w_off = off; off += (int64_t)y.Kpad * y.Cout_p;
            y.b_off = off; off += y.Cout_p;
            y.flux_w_off = foff; foff += (int64_t)9 * cin * cout;
            y.flux_b_off = foff; foff += cout;
            y.wT_off = toff; toff += (int64_t)y.KpadT * y.Cin_p;
            y.wk_off = koff; koff += (int64_t)9 * y.Cin_pe * y.Cout_p;
            y.wtk_off = tkoff; tkoff += (int64_t)9 * y.Cin_p * y.Cout_p;
            if (y.split_mode) {
                y.gexp_off = goff; goff += (int64_t)9 * y.Cin_pe * y.Cout_p;
                SplitLayerDesc& sd = M->split.d[M->split.count++];
                sd.w_off = y.w_off; sd.wk_off = y.wk_off; sd.gexp_off = y.gexp_off;
                sd.Cin_p = y.Cin_p; sd.Cin_pe = y.Cin_pe; sd.Cout_p = y.Cout_p; sd.mode = y.split_mode; sd.C = cin; sd.n2 = cout;
            }
        };
        for (int l = 0; l < L; l++) fill(M->enc[c][l], encf[l], encf[l + 1], l == 0 ? 1 : 0);
        const bool fused_head = tc_mode && cfg->upsampling == DINCAE_UP_NEAREST && k2_head_up_conv_eligible(decf[1], decf[0]);
        for (int l = L - 1; l >= 0; l--) fill(M->dec[c][l], decf[l + 1], decf[l], (l == 0 && !fused_head) ? 2 : 0);
    }
    for (int c = 0; c < M->nchains; c++) {
        M->chain_off[c] = M->enc[c][0].w_off;
        M->split_chain[c].count = 0;
        for (int i = 0; i < M->split.count; i++) {
            const int64_t wo = M->split.d[i].w_off;
            if (wo >= M->enc[c][0].w_off && (c + 1 == M->nchains || wo < M->enc[c + 1][0].w_off)) M->split_chain[c].d[M->split_chain[c].count++] = M->split.d[i];
        }
    }
    M->chain_off[M->nchains] = off;
    M->n_int = off; M->n_flux = foff; M->n_wT = toff; M->n_wk = koff; M->n_wtk = tkoff; M->n_gexp = goff;
    // table for the transposed copies, in wT order; each entry copies 9*Cin_p*Cout_p values
    {
        // make_wT walks layers by cumulative n, so wT offsets must be visited in the same order as the table
        std::vector<const LayerInfo*> order;
        for (int c = 0; c < M->nchains; c++) {
            for (int l = 0; l < L; l++) order.push_back(&M->enc[c][l]);
            for (int l = L - 1; l >= 0; l--) order.push_back(&M->dec[c][l]);
        }
        int64_t tot = 0;
        for (const LayerInfo* y : order) {
            WLayerDesc d;
            d.w_off = y->w_off; d.wT_off = y->wT_off; d.Cin_p = y->Cin_p; d.Cout_p = y->Cout_p; d.n = (int64_t)9 * y->Cin_p * y->Cout_p; d.wk_off = y->wk_off; d.wtk_off = y->wtk_off;
            wtab.push_back(d);
            tot += d.n;
        }
        M->n_wT_copy = tot;
        M->n_wlayers = (int)wtab.size();
    }

    // ---- workspace plan: one allocation
    struct Want { void** p; size_t bytes; };
    std::vector<Want> wants;
    auto want = [&](void** p, size_t bytes) { wants.push_back(Want{p, (bytes + 255) / 256 * 256}); };
    const size_t Bm = (size_t)M->maxB;
    auto act = [&](int l, int Cp) { return sizeof(float) * Bm * M->Hl[l] * M->Wl[l] * Cp; };
    want((void**)&M->w, sizeof(float) * (M->n_int + 4));
    want((void**)&M->g, sizeof(float) * (M->n_int + 4));
    want((void**)&M->m, sizeof(float) * (M->n_int + 4));
    want((void**)&M->v, sizeof(float) * (M->n_int + 4));
    want((void**)&M->wT, sizeof(float) * (M->n_wT + 4));
    if (tc_mode) {
        want((void**)&M->wK, sizeof(float) * (M->n_wk + 4));
        want((void**)&M->wTK, sizeof(float) * (M->n_wtk + 4));
        if (M->n_gexp) want((void**)&M->g_exp, sizeof(float) * (M->n_gexp + 4));
    }
    want((void**)&M->d_wtab, sizeof(WLayerDesc) * wtab.size());
    const size_t planar_out = sizeof(float) * Bm * 2 * M->O * cfg->W * cfg->H;
    const bool bil = cfg->upsampling == DINCAE_UP_BILINEAR || tc_mode;   // U[l] materialised
    for (int c = 0; c < M->nchains; c++) {
        ChainBufs& cb = M->cb[c];
        memset(&cb, 0, sizeof(cb));
        want((void**)&cb.X[0], act(0, M->enc[c][0].Cin_pe));
        for (int l = 0; l < L; l++) {
            // tensor-core modes, even level size, ReLU-type activation: the epilogue pools and keeps sign bits instead of A[l]
            const bool pool_only = tc_mode && (M->Hl[l] % 2) == 0 && (M->Wl[l] % 2) == 0 && cfg->leaky_slope >= 0.f && cfg->leaky_slope < 1.f &&
                                   dc_env_int("DINCAE_NO_POOL_FUSION", 0) == 0;
            if (pool_only) want((void**)&cb.MK[l], sizeof(uint32_t) * Bm * M->Hl[l] * M->Wl[l] * (M->enc[c][l].Cout_p / 32));
            else want((void**)&cb.A[l], act(l, M->enc[c][l].Cout_p));
            want((void**)&cb.X[l + 1], act(l + 1, M->enc[c][l].Cout_p));
            if (l >= 1) want((void**)&cb.R[l], act(l, M->dec[c][l].Cout_p));
            if (bil) want((void**)&cb.U[l], act(l, M->dec[c][l].Cin_pe));
        }
        want((void**)&cb.RAW, planar_out);
        want((void**)&cb.GRAW, planar_out);
    }
    {
        // gradient workspace: channel counts are the same for every chain except enc[.][0].Cin_p (unused here)
        const std::vector<LayerInfo>& e = M->enc[0];
        const std::vector<LayerInfo>& d = M->dec[0];
        want((void**)&M->GD0, act(0, M->nco));
        for (int l = 0; l < L; l++) {
            want((void**)&M->GU[l], act(l, d[l].Cin_p));
            want((void**)&M->GC[l], act(l, e[l].Cout_p));
            if (l >= 1) {
                want((void**)&M->G[l], act(l, d[l].Cout_p));
                want((void**)&M->GD[l], act(l, d[l].Cout_p));
                want((void**)&M->XG[l], act(l, e[l].Cin_p));
            }
        }
        want((void**)&M->G[L], act(L, d[L - 1].Cin_p));
    }
    want((void**)&M->MS, planar_out);
    want((void**)&M->GMS, planar_out);
    want((void**)&M->p2p_flags, 256);
    want((void**)&M->p2p_stage, sizeof(float) * (M->n_int + 4 + 4 * 16));       // N slots of ceil(n / N) float4 each
    want((void**)&M->d_loss, sizeof(double) * 8);
    want((void**)&M->d_inv_n, sizeof(float) * 16);
    want((void**)&M->d_nvalid, sizeof(unsigned long long) * 16);
    M->red_scratch_bytes = k3_scratch_bytes(0);
    want(&M->red_scratch, M->red_scratch_bytes);
    size_t wgb = 0, wg_sum = 0;
    for (int c = 0; c < M->nchains; c++)
        for (int l = 0; l < L; l++) {
            size_t a = k2_wgrad_scratch_bytes(ctx->num_sms, M->maxB, M->Hl[l], M->Wl[l], M->enc[c][l].Cin_p, M->enc[c][l].Cout_p);
            size_t b = k2_wgrad_scratch_bytes(ctx->num_sms, M->maxB, M->Hl[l], M->Wl[l], M->dec[c][l].Cin_p, M->dec[c][l].Cout_p);
            wgb = a > wgb ? a : wgb;
            wgb = b > wgb ? b : wgb;
            size_t ta = k2_wgrad_tma_scratch_bytes(ctx->num_sms, M->maxB, M->Hl[l], M->Wl[l], M->enc[c][l].Cin_pe, M->enc[c][l].Cout_p);
            size_t tb = k2_wgrad_tma_scratch_bytes(ctx->num_sms, M->maxB, M->Hl[l], M->Wl[l], M->dec[c][l].Cin_pe, M->dec[c][l].Cout_p);
            wg_sum += ((ta + 255) / 256 * 256) + ((tb + 255) / 256 * 256);
            wg_sum += 2 * (((size_t)8 * ctx->num_sms * 512 * sizeof(float) + 255) / 256 * 256);      // bias column partials of both layers
            if (l == 0) wg_sum += (k2_head_bwd_scratch_bytes(M->maxB, M->Hl[1], M->Wl[1], M->dec[c][0].Cin, 2 * M->O) + 255) / 256 * 256;   // fused head
        }
    if (tc_mode && wg_sum > wgb) wgb = wg_sum;    // TF32 mode: one slice per layer (second stages are batched)
    M->wg_scratch_bytes = wgb;
    want(&M->wg_scratch, wgb);
    want((void**)&M->in_xin, sizeof(float) * Bm * cfg->ninput * cfg->W * cfg->H);
    want((void**)&M->in_xtrue, planar_out);
    size_t total = 0;
    for (auto& wnt : wants) total += wnt.bytes;
    cudaError_t e = cudaMalloc(&M->all_mem, total);
    if (e != cudaSuccess) {
        dincae_set_error("dincae_model_create: cannot allocate %.1f MB of workspace for max_batch=%d: %s", total / 1048576.0, M->maxB, cudaGetErrorString(e));
        delete M;
        return DINCAE_ENOMEM;
    }
    DC_CHECK_CUDA(cudaMemsetAsync(M->all_mem, 0, total, ctx->stream));
    {
        char* p = (char*)M->all_mem;
        for (auto& wnt : wants) { *wnt.p = p; p += wnt.bytes; }
    }
    DC_CHECK_CUDA(cudaMemcpyAsync(M->d_wtab, wtab.data(), sizeof(WLayerDesc) * wtab.size(), cudaMemcpyHostToDevice, ctx->stream));
    DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    *out = M;
    return DINCAE_OK;
}

static void step_graphs_release(dincae_model* M);
extern "C" int32_t dincae_model_destroy(dincae_model* M) {
    if (!M) return DINCAE_OK;
    cudaSetDevice(M->ctx->device);
    cudaStreamSynchronize(M->ctx->stream);
    step_graphs_release(M);
    model_p2p_release(M);        // collective when the peers' workspaces are mapped (data-parallel training over peer memory)
    if (M->all_mem) cudaFree(M->all_mem);
    if (M->pt_buf) cudaFree(M->pt_buf);
    if (M->copy_stream) {
        cudaStreamSynchronize(M->copy_stream);
        for (int s = 0; s < 2; s++) { cudaFree(M->up_buf[s]); cudaEventDestroy(M->up_done[s]); cudaEventDestroy(M->slot_free[s]); }
        cudaStreamDestroy(M->copy_stream);
    }
    delete M;
    return DINCAE_OK;
}

extern "C" int32_t dincae_model_num_params(dincae_model* M, int64_t* n) {
    DC_REQUIRE(M && n, DINCAE_EINVAL, "dincae_model_num_params: NULL argument");
    *n = M->n_flux;
    return DINCAE_OK;
}

extern "C" int32_t dincae_model_get_params(dincae_model* M, float* host, int64_t n) {
    DC_REQUIRE(M && host, DINCAE_EINVAL, "dincae_model_get_params: NULL argument");
    DC_REQUIRE(n == M->n_flux, DINCAE_ESIZE, "dincae_model_get_params: n=%lld but the model has %lld parameters", (long long)n, (long long)M->n_flux);
    DC_CHECK_CUDA(cudaSetDevice(M->ctx->device));
    return download_internal(M, M->w, host);
}

extern "C" int32_t dincae_model_set_params(dincae_model* M, const float* host, int64_t n) {
    DC_REQUIRE(M && host, DINCAE_EINVAL, "dincae_model_set_params: NULL argument");
    DC_REQUIRE(n == M->n_flux, DINCAE_ESIZE, "dincae_model_set_params: n=%lld but the model has %lld parameters", (long long)n, (long long)M->n_flux);
    DC_CHECK_CUDA(cudaSetDevice(M->ctx->device));
    M->wT_dirty = true;
    return upload_internal(M, M->w, host);
}

extern "C" int32_t dincae_model_init_glorot(dincae_model* M, uint64_t seed) {
    DC_REQUIRE(M != nullptr, DINCAE_EINVAL, "dincae_model_init_glorot: model is NULL");
    // Flux glorot_uniform: U(-s, s), s = sqrt(6 / (fan_in + fan_out)), fan = 9*cin / 9*cout; biases zero.
    // Host-side splitmix64 stream (one-time set-up, not the hot path).
    std::vector<float> flux((size_t)M->n_flux, 0.f);
    uint64_t st = seed * 0x9E3779B97F4A7C15ull + 0x1234567ull;
    auto next = [&]() {
        st += 0x9E3779B97F4A7C15ull;
        uint64_t z = st;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    };
    for (int c = 0; c < M->nchains; c++)
        for (int kind = 0; kind < 2; kind++)
            for (int l = 0; l < M->L; l++) {
                const LayerInfo& y = kind == 0 ? M->enc[c][l] : M->dec[c][l];
                double s = sqrt(6.0 / (9.0 * y.Cin + 9.0 * y.Cout));
                for (int64_t i = 0; i < (int64_t)9 * y.Cin * y.Cout; i++) {
                    double u = (double)(next() >> 11) * (1.0 / 9007199254740992.0);
                    flux[y.flux_w_off + i] = (float)((2.0 * u - 1.0) * s);
                }
            }
    return dincae_model_set_params(M, flux.data(), M->n_flux);
}

extern "C" int32_t dincae_model_get_opt_state(dincae_model* M, float* host_m, float* host_v, int64_t n, int64_t* t) {
    DC_REQUIRE(M && host_m && host_v && t, DINCAE_EINVAL, "dincae_model_get_opt_state: NULL argument");
    DC_REQUIRE(n == M->n_flux, DINCAE_ESIZE, "dincae_model_get_opt_state: wrong length");
    DC_CHECK_CUDA(cudaSetDevice(M->ctx->device));
    DC_TRY(download_internal(M, M->m, host_m));
    DC_TRY(download_internal(M, M->v, host_v));
    *t = M->t_step;
    return DINCAE_OK;
}

extern "C" int32_t dincae_model_set_opt_state(dincae_model* M, const float* host_m, const float* host_v, int64_t n, int64_t t) {
    DC_REQUIRE(M && host_m && host_v, DINCAE_EINVAL, "dincae_model_set_opt_state: NULL argument");
    DC_REQUIRE(n == M->n_flux && t >= 0, DINCAE_ESIZE, "dincae_model_set_opt_state: wrong length or negative t");
    DC_CHECK_CUDA(cudaSetDevice(M->ctx->device));
    DC_TRY(upload_internal(M, M->m, host_m));
    DC_TRY(upload_internal(M, M->v, host_v));
    M->t_step = t;
    M->bt1 = pow((double)M->cfg.adam_beta1, (double)t);
    M->bt2 = pow((double)M->cfg.adam_beta2, (double)t);
    return DINCAE_OK;
}

extern "C" int32_t dincae_forward(dincae_model* M, const float* xin, int32_t B, float* out, int32_t loc) {
    DC_TRY(check_batch(M, B, "dincae_forward"));
    DC_REQUIRE(xin && out, DINCAE_EINVAL, "dincae_forward: NULL argument");
    dincae_ctx* ctx = M->ctx;
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    const float* dx;
    DC_TRY(stage_batch(M, xin, M->cfg.ninput, B, loc, M->in_xin, &dx));
    DC_TRY(model_forward(M, dx, B, false));
    if (loc == DINCAE_DEVICE) return model_head_last(M, B, out);
    DC_TRY(model_head_last(M, B, M->MS));
    DC_CHECK_CUDA(cudaMemcpyAsync(out, M->MS, sizeof(float) * (size_t)B * 2 * M->O * M->cfg.W * M->cfg.H, cudaMemcpyDeviceToHost, ctx->stream));
    DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    return DINCAE_OK;
}

extern "C" int32_t dincae_count_valid(dincae_model* M, const float* xtrue, int32_t B, int32_t loc, int64_t* n_valid) {
    DC_TRY(check_batch(M, B, "dincae_count_valid"));
    DC_REQUIRE(xtrue && n_valid, DINCAE_EINVAL, "dincae_count_valid: NULL argument");
    dincae_ctx* ctx = M->ctx;
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    const float* dy;
    DC_TRY(stage_batch(M, xtrue, 2 * M->O, B, loc, M->in_xtrue, &dy));
    DC_TRY(k3_count_valid(ctx, dy, B, M->O, (int64_t)M->cfg.W * M->cfg.H, M->d_nvalid));
    unsigned long long tmp[16];
    DC_CHECK_CUDA(cudaMemcpyAsync(tmp, M->d_nvalid, sizeof(unsigned long long) * M->O, cudaMemcpyDeviceToHost, ctx->stream));
    DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int v = 0; v < M->O; v++) n_valid[v] = (int64_t)tmp[v];
    return DINCAE_OK;
}

extern "C" int32_t dincae_loss(dincae_model* M, const float* xin, const float* xtrue, int32_t B, int32_t loc, double* loss) {
    DC_TRY(check_batch(M, B, "dincae_loss"));
    DC_REQUIRE(xin && xtrue && loss, DINCAE_EINVAL, "dincae_loss: NULL argument");
    dincae_ctx* ctx = M->ctx;
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    const float *dx, *dy;
    DC_TRY(stage_batch(M, xin, M->cfg.ninput, B, loc, M->in_xin, &dx));
    DC_TRY(stage_batch(M, xtrue, 2 * M->O, B, loc, M->in_xtrue, &dy));
    DC_CHECK_CUDA(cudaMemsetAsync(M->d_loss, 0, sizeof(double) * 8, ctx->stream));
    DC_TRY(model_forward(M, dx, B, false));
    DC_TRY(gridded_loss(M, dy, B, nullptr, false));
    DC_TRY(k3_reg(ctx, M->w, M->n_int, M->cfg.l1_beta, M->cfg.l2_beta, M->d_loss + 1, M->red_scratch, M->red_scratch_bytes));
    return read_loss(M, false, nullptr, loss);
}

// forward + loss + backward of a gridded step on device buffers.  Two things hang off the main chain on the side stream: the valid
// counts of the truth (count_here: the single-GPU step counts them itself; needed only by the first loss) and the loss of every
// chain but the last, which runs beside the next chain's forward pass (the loss kernels use no shared memory, so they fit on the SMs
// next to the conv CTAs).  The chains' terms are still added to the loss in chain order: results do not change.
// the mini-batch a step assembles for itself from a device-resident dataset (getxy!, src/data.jl:391-517)
struct AsmCall {
    dincae_gridded* ds;
    const int64_t* time_idx;
    const int64_t* imask;
    uint64_t seed, ctr0;
};
static int32_t grad_step_body(dincae_model* M, const float* dx, const float* dy, int B, bool count_here, const AdamParams* ap,
                              const AsmCall* ac = nullptr) {
    dincae_ctx* ctx = M->ctx;
    if (ac) DC_TRY(k1_assemble(ac->ds, ac->time_idx, ac->imask, ac->seed, ac->ctr0, B, const_cast<float*>(dx), const_cast<float*>(dy)));
    DC_CHECK_CUDA(cudaMemsetAsync(M->d_loss, 0, sizeof(double) * 8, ctx->stream));
    HeadParams hp = head_params(M);
    const bool branch = side_stream_on() && !ctx->prof && hp.lap == 0.f && hp.lap_err == 0.f;
    cudaStream_t main_stream = ctx->stream;
    auto on_side = [&](auto&& fn) -> int32_t {
        if (!branch) return fn();
        DC_TRY(wgrad_fork(ctx));
        ctx->stream = ctx->side;
        M->side_pending = true;
        int32_t st = fn();
        ctx->stream = main_stream;
        return st;
    };
    if (count_here) DC_TRY(on_side([&] { return gridded_counts(M, dy, B, nullptr); }));
    ctx->tf32_acts = M->wK != nullptr;
    if (M->wK) DC_TRY(ensure_wT(M));        // the tensor-core forward reads the K-major TF32 weight copy
    for (int c = 0; c < M->nchains; c++) {
        DC_TRY(chain_forward(M, c, dx, B, true));
        if (c + 1 < M->nchains) DC_TRY(on_side([&] { return gridded_chain_loss(M, c, dy, B, true); }));
    }
    DC_TRY(wgrad_join(M));
    DC_TRY(gridded_chain_loss(M, M->nchains - 1, dy, B, true));
    DC_TRY(backward_all(M, B));
    if (ap) {                                  // L1/L2 + ClipGrad + Adam of the same step (single-GPU training)
        DC_CHECK_CUDA(cudaMemsetAsync(M->d_loss + 1, 0, sizeof(double), ctx->stream));
        DC_TRY(k3_adam(ctx, M->w, M->g, M->m, M->v, M->n_int, *ap, M->d_loss + 1, M->red_scratch, M->red_scratch_bytes));
    }
    return DINCAE_OK;
}
static bool step_graph_on() {
    static const int on = dc_env_int("DINCAE_GRAPH", 1);
    return on != 0;
}
static void step_graph_drop(StepGraph& g) {
    if (g.exec) cudaGraphExecDestroy(g.exec);
    if (g.graph) cudaGraphDestroy(g.graph);
    g = StepGraph();
}
static void step_graphs_release(dincae_model* M) {
    for (StepGraph& g : M->step_graph) step_graph_drop(g);
    for (FwdGraph& g : M->fwd_graph) {
        if (g.exec) cudaGraphExecDestroy(g.exec);
        g = FwdGraph();
    }
}
// model(M->in_xin) -> (m, sigma2) in M->MS for reconstruction: the same first-eager / second-captured / then-replayed scheme as the
// training step, keyed on the batch size alone (both buffers belong to the model).  The weights do not change between the batches
// of a reconstruction, so their derived copies are refreshed, if stale, BEFORE the graph.
static int32_t forward_infer_or_graph(dincae_model* M, int B) {
    dincae_ctx* ctx = M->ctx;
    auto body = [&]() -> int32_t {
        DC_TRY(model_forward(M, M->in_xin, B, false));
        return model_head_last(M, B, M->MS);
    };
    const bool capturable = ctx->stream != nullptr && ctx->stream != cudaStreamLegacy && ctx->stream != cudaStreamPerThread;
    if (!step_graph_on() || ctx->prof || M->graph_off || !capturable) return body();
    if (M->wK) DC_TRY(ensure_wT(M));
    FwdGraph* g = nullptr;
    for (FwdGraph& c : M->fwd_graph)
        if (c.B == B) g = &c;
    if (g && g->exec) {
        DC_CHECK_CUDA(cudaGraphLaunch(g->exec, ctx->stream));
        ctx->launches += g->launches;
        return DINCAE_OK;
    }
    if (!g) {
        FwdGraph* slot = &M->fwd_graph[B == M->maxB ? 0 : 1];
        if (slot->exec) cudaGraphExecDestroy(slot->exec);
        *slot = FwdGraph();
        slot->B = B;
        return body();
    }
    const int64_t launches0 = ctx->launches;
    cudaGraph_t graph = nullptr;
    cudaError_t e = cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal);
    int32_t st = DINCAE_OK;
    if (e == cudaSuccess) {
        ctx->capturing = true;
        st = body();
        ctx->capturing = false;
        e = cudaStreamEndCapture(ctx->stream, &graph);
    }
    const int64_t captured = ctx->launches - launches0;
    ctx->launches = launches0;
    if (e == cudaSuccess && st == DINCAE_OK && graph) e = cudaGraphInstantiate(&g->exec, graph, 0);
    if (graph) cudaGraphDestroy(graph);
    if (e != cudaSuccess || st != DINCAE_OK || !g->exec) {
        cudaGetLastError();
        g->exec = nullptr;
        M->graph_off = true;
        return body();
    }
    g->launches = captured;
    DC_CHECK_CUDA(cudaGraphLaunch(g->exec, ctx->stream));
    ctx->launches += g->launches;
    return DINCAE_OK;
}
// Replays the step as a CUDA graph once its key (buffers, batch size) has been seen twice: the first step with a key runs launch by
// launch (and does every lazy set-up: shared-memory attributes, tensor maps, the side stream), the second is captured -- both
// streams, with plain edges instead of programmatic-dependent-launch ones -- and replayed from then on.  The derived weight copies are refreshed
// INSIDE the graph on every replay (after an optimiser step they are always stale; refreshing them when they are not is harmless).
// Not used under the per-launch profile, with bucketed overlap (NCCL calls between the chains), or after a failed capture.
static int32_t grad_step_body_or_graph(dincae_model* M, const float* dx, const float* dy, int B, bool count_here, const AdamParams* ap,
                                       const AsmCall* ac = nullptr) {
    dincae_ctx* ctx = M->ctx;
    const bool capturable = ctx->stream != nullptr && ctx->stream != cudaStreamLegacy && ctx->stream != cudaStreamPerThread;
    if (!step_graph_on() || ctx->prof || M->dp_overlap || M->graph_off || !capturable || (ac && B > k1_assemble_max_single_launch()))
        return grad_step_body(M, dx, dy, B, count_here, ap, ac);
    const bool apply = ap != nullptr;
    const void* ds = ac ? (const void*)ac->ds : nullptr;
    StepGraph* g = nullptr;
    StepGraph* lru = &M->step_graph[0];
    for (StepGraph& c : M->step_graph) {
        if (c.B == B && c.dx == dx && c.dy == dy && c.count_here == count_here && c.apply == apply && c.ds == ds) { g = &c; break; }
        if (c.last_use < lru->last_use) lru = &c;
    }
    M->graph_tick++;
    if (g && g->exec) {
        g->last_use = M->graph_tick;
        if (apply) DC_TRY(k3_update_adam_node(g->exec, g->adam_node, *ap));      // this step's learning rate and bias corrections
        if (ac) DC_TRY(k1_update_assemble_node(g->exec, g->asm_node, ac->ds, ac->time_idx, ac->imask, ac->seed, ac->ctr0, B,
                                               const_cast<float*>(dx), const_cast<float*>(dy)));     // ... and sample indices
        DC_CHECK_CUDA(cudaGraphLaunch(g->exec, ctx->stream));
        ctx->launches += g->launches;
        M->wT_dirty = false;
        return DINCAE_OK;
    }
    if (!g) {                                  // first time: remember the key, run launch by launch
        step_graph_drop(*lru);
        lru->dx = dx; lru->dy = dy; lru->B = B; lru->count_here = count_here; lru->apply = apply; lru->ds = ds; lru->last_use = M->graph_tick;
        return grad_step_body(M, dx, dy, B, count_here, ap, ac);
    }
    g->last_use = M->graph_tick;
    const int64_t launches0 = ctx->launches;
    M->wT_dirty = true;
    cudaGraph_t graph = nullptr;
    cudaError_t e = cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal);
    int32_t st = DINCAE_OK;
    if (e == cudaSuccess) {
        ctx->capturing = true;
        st = grad_step_body(M, dx, dy, B, count_here, ap, ac);
        ctx->capturing = false;
        e = cudaStreamEndCapture(ctx->stream, &graph);
    }
    const int64_t captured = ctx->launches - launches0;
    ctx->launches = launches0;
    if (e == cudaSuccess && st == DINCAE_OK && graph) e = cudaGraphInstantiate(&g->exec, graph, 0);
    g->graph = graph;
    if (e == cudaSuccess && g->exec && apply) {
        g->adam_node = k3_find_adam_node(graph);
        if (!g->adam_node) st = DINCAE_ECUDA;
    }
    if (e == cudaSuccess && g->exec && ac) {
        g->asm_node = k1_find_assemble_node(graph);
        if (!g->asm_node) st = DINCAE_ECUDA;
    }
    if (e != cudaSuccess || st != DINCAE_OK || !g->exec) {
        cudaGetLastError();
        if (g->exec) cudaGraphExecDestroy(g->exec);
        if (g->graph) cudaGraphDestroy(g->graph);
        g->exec = nullptr; g->graph = nullptr; g->adam_node = nullptr; g->asm_node = nullptr;
        M->graph_off = true;
        M->side_pending = false;
        M->wT_dirty = true;
        return grad_step_body(M, dx, dy, B, count_here, ap, ac);
    }
    g->launches = captured;
    DC_CHECK_CUDA(cudaGraphLaunch(g->exec, ctx->stream));
    ctx->launches += g->launches;
    M->wT_dirty = false;
    return DINCAE_OK;
}

extern "C" int32_t dincae_model_step_graph(dincae_model* M, int32_t enable, int32_t* captured) {
    DC_REQUIRE(M != nullptr, DINCAE_EINVAL, "dincae_model_step_graph: model is NULL");
    DC_REQUIRE(enable >= -1 && enable <= 1, DINCAE_EINVAL, "dincae_model_step_graph: enable must be -1 (query), 0 or 1");
    if (enable >= 0) {
        DC_CHECK_CUDA(cudaSetDevice(M->ctx->device));
        DC_CHECK_CUDA(cudaStreamSynchronize(M->ctx->stream));
        if (!enable) step_graphs_release(M);
        M->graph_off = !enable;
    }
    if (captured) {
        *captured = 0;
        for (const StepGraph& g : M->step_graph) *captured += g.exec != nullptr;
    }
    return DINCAE_OK;
}

extern "C" int32_t dincae_grad_step(dincae_model* M, const float* xin, const float* xtrue, int32_t B,
                                    const int64_t* n_valid_global, int32_t loc) {
    DC_TRY(check_batch(M, B, "dincae_grad_step"));
    DC_REQUIRE(xin && xtrue, DINCAE_EINVAL, "dincae_grad_step: NULL argument");
    dincae_ctx* ctx = M->ctx;
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    const float *dx, *dy;
    DC_TRY(stage_batch(M, xin, M->cfg.ninput, B, loc, M->in_xin, &dx));
    DC_TRY(stage_batch(M, xtrue, 2 * M->O, B, loc, M->in_xtrue, &dy));
    if (n_valid_global) DC_TRY(gridded_counts(M, dy, B, n_valid_global));      // per-step host values: stay outside the captured part
    return grad_step_body_or_graph(M, dx, dy, B, n_valid_global == nullptr, nullptr);
}

extern "C" int32_t dincae_grad_buffer(dincae_model* M, float** dev_ptr, int64_t* n) {
    DC_REQUIRE(M && dev_ptr && n, DINCAE_EINVAL, "dincae_grad_buffer: NULL argument");
    *dev_ptr = M->g;
    *n = M->n_int + 1;
    return DINCAE_OK;
}

// optimiser step t+1: advances the step count and the running beta powers
static AdamParams adam_params_next(dincae_model* M, float lr) {
    M->t_step += 1;
    M->bt1 *= (double)M->cfg.adam_beta1;
    M->bt2 *= (double)M->cfg.adam_beta2;
    AdamParams p;
    p.lr = lr; p.beta1 = M->cfg.adam_beta1; p.beta2 = M->cfg.adam_beta2; p.eps = M->cfg.adam_eps;
    p.clip = M->cfg.clip_grad; p.l1 = M->cfg.l1_beta; p.l2 = M->cfg.l2_beta;
    p.inv_bc1 = (float)(1.0 / (1.0 - M->bt1));
    p.inv_bc2 = (float)(1.0 / (1.0 - M->bt2));
    return p;
}
// single-GPU training step on device buffers: gradient step and optimiser update as ONE captured graph
static int32_t train_step_device(dincae_model* M, const float* dx, const float* dy, int B, float lr) {
    const AdamParams p = adam_params_next(M, lr);
    int32_t st = grad_step_body_or_graph(M, dx, dy, B, true, &p);
    M->wT_dirty = true;
    return st;
}

extern "C" int32_t dincae_apply_step(dincae_model* M, float lr, float* loss) {
    DC_REQUIRE(M != nullptr, DINCAE_EINVAL, "dincae_apply_step: model is NULL");
    dincae_ctx* ctx = M->ctx;
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    const AdamParams p = adam_params_next(M, lr);
    DC_CHECK_CUDA(cudaMemsetAsync(M->d_loss + 1, 0, sizeof(double), ctx->stream));
    DC_TRY(k3_adam(ctx, M->w, M->g, M->m, M->v, M->n_int, p, M->d_loss + 1, M->red_scratch, M->red_scratch_bytes));
    M->wT_dirty = true;
    if (loss) return read_loss(M, true, loss, nullptr);
    return DINCAE_OK;
}

extern "C" int32_t dincae_train_step(dincae_model* M, const float* xin, const float* xtrue, int32_t B, float lr,
                                     int32_t loc, float* loss) {
    DC_TRY(check_batch(M, B, "dincae_train_step"));
    DC_REQUIRE(xin && xtrue, DINCAE_EINVAL, "dincae_train_step: NULL argument");
    dincae_ctx* ctx = M->ctx;
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    const float *dx, *dy;
    DC_TRY(stage_batch(M, xin, M->cfg.ninput, B, loc, M->in_xin, &dx));
    DC_TRY(stage_batch(M, xtrue, 2 * M->O, B, loc, M->in_xtrue, &dy));
    DC_TRY(train_step_device(M, dx, dy, B, lr));
    if (loss) return read_loss(M, true, loss, nullptr);
    return DINCAE_OK;
}

extern "C" int32_t dincae_allreduce_grads(dincae_model* M) {
    DC_REQUIRE(M != nullptr, DINCAE_EINVAL, "dincae_allreduce_grads: model is NULL");
    DC_CHECK_CUDA(cudaSetDevice(M->ctx->device));
    return model_allreduce_grads(M);
}

extern "C" int32_t dincae_train_step_dp(dincae_model* M, const float* xin, const float* xtrue, int32_t B, const int64_t* n_valid_global,
                                        float lr, int32_t loc, float* loss) {
    DC_REQUIRE(M != nullptr, DINCAE_EINVAL, "dincae_train_step_dp: model is NULL");
    DC_REQUIRE(M->ctx->comm != nullptr, DINCAE_EINVAL, "dincae_train_step_dp: no communicator (dincae_comm_init_rank)");
    DC_REQUIRE(n_valid_global != nullptr, DINCAE_EINVAL, "dincae_train_step_dp: n_valid_global is NULL (the gradient sums must be normalised by the GLOBAL count)");
    M->dp_overlap = dp_overlap_on();
    int32_t st = dincae_grad_step(M, xin, xtrue, B, n_valid_global, loc);
    const bool bucketed = M->dp_overlap;
    M->dp_overlap = false;
    DC_TRY(st);
    if (!bucketed) DC_TRY(dincae_allreduce_grads(M));
    return dincae_apply_step(M, lr, loss);
}

static int32_t check_dataset(const dincae_model* M, const dincae_gridded* ds, const char* who) {
    DC_REQUIRE(ds != nullptr, DINCAE_EINVAL, "%s: dataset is NULL", who);
    DC_REQUIRE(ds->ctx == M->ctx, DINCAE_EINVAL, "%s: dataset and model belong to different contexts", who);
    DC_REQUIRE(ds->W == M->cfg.W && ds->H == M->cfg.H && 4 + 2 * ds->T * ds->V == M->cfg.ninput && ds->nout == M->O, DINCAE_EINVAL,
               "%s: dataset (W=%d,H=%d,channels=%d,nout=%d) does not match the model", who, ds->W, ds->H, 4 + 2 * ds->T * ds->V, ds->nout);
    return DINCAE_OK;
}

extern "C" int32_t dincae_train_step_gridded(dincae_model* M, dincae_gridded* ds, const int64_t* time_idx, const int64_t* imask, uint64_t seed,
                                             uint64_t sample_ctr0, int32_t B, float lr, float* loss) {
    DC_TRY(check_batch(M, B, "dincae_train_step_gridded"));
    DC_TRY(check_dataset(M, ds, "dincae_train_step_gridded"));
    DC_REQUIRE(time_idx != nullptr, DINCAE_EINVAL, "dincae_train_step_gridded: time_idx is NULL");
    DC_TRY(k1_check_indices(ds, time_idx, imask, B));        // a caller's bad index must not look like a failed capture
    DC_CHECK_CUDA(cudaSetDevice(M->ctx->device));
    const AsmCall ac{ds, time_idx, imask, seed, sample_ctr0};
    const int64_t t0 = M->t_step; const double b1 = M->bt1, b2 = M->bt2;
    const AdamParams p = adam_params_next(M, lr);
    int32_t st = grad_step_body_or_graph(M, M->in_xin, M->in_xtrue, B, true, &p, &ac);
    if (st != DINCAE_OK) { M->t_step = t0; M->bt1 = b1; M->bt2 = b2; return st; }     // e.g. an index out of range: nothing was updated
    M->wT_dirty = true;
    if (loss) return read_loss(M, true, loss, nullptr);
    return DINCAE_OK;
}

extern "C" int32_t dincae_train_step_dp_gridded(dincae_model* M, dincae_gridded* ds, const int64_t* time_idx, const int64_t* imask, uint64_t seed,
                                                uint64_t sample_ctr0, int32_t B, const int64_t* n_valid_global, float lr, float* loss) {
    DC_TRY(check_batch(M, B, "dincae_train_step_dp_gridded"));
    DC_TRY(check_dataset(M, ds, "dincae_train_step_dp_gridded"));
    DC_REQUIRE(time_idx != nullptr, DINCAE_EINVAL, "dincae_train_step_dp_gridded: time_idx is NULL");
    DC_REQUIRE(M->ctx->comm != nullptr, DINCAE_EINVAL, "dincae_train_step_dp_gridded: no communicator (dincae_comm_init_rank)");
    DC_REQUIRE(n_valid_global != nullptr, DINCAE_EINVAL, "dincae_train_step_dp_gridded: n_valid_global is NULL (the gradient sums must be normalised by the GLOBAL count)");
    DC_TRY(k1_check_indices(ds, time_idx, imask, B));
    DC_CHECK_CUDA(cudaSetDevice(M->ctx->device));
    const AsmCall ac{ds, time_idx, imask, seed, sample_ctr0};
    M->dp_overlap = dp_overlap_on();
    int32_t st = gridded_counts(M, M->in_xtrue, B, n_valid_global);
    if (st == DINCAE_OK) st = grad_step_body_or_graph(M, M->in_xin, M->in_xtrue, B, false, nullptr, &ac);
    const bool bucketed = M->dp_overlap;
    M->dp_overlap = false;
    DC_TRY(st);
    if (!bucketed) DC_TRY(dincae_allreduce_grads(M));
    return dincae_apply_step(M, lr, loss);
}

extern "C" int32_t dincae_replica_divergence(dincae_model* M, float* max_abs_diff) {
    DC_REQUIRE(M && max_abs_diff, DINCAE_EINVAL, "dincae_replica_divergence: NULL argument");
    dincae_ctx* ctx = M->ctx;
    DC_REQUIRE(ctx->comm != nullptr, DINCAE_EINVAL, "dincae_replica_divergence: no communicator");
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    DC_TRY(wgrad_join(M));
    // the gradient workspace is free between steps: tmp = (rank == 0 ? w : 0), SUM over ranks = w_0, then max |w - w_0|
    float* tmp = M->g;
    if (ctx->comm_rank == 0) DC_CHECK_CUDA(cudaMemcpyAsync(tmp, M->w, sizeof(float) * M->n_int, cudaMemcpyDeviceToDevice, ctx->stream));
    else DC_CHECK_CUDA(cudaMemsetAsync(tmp, 0, sizeof(float) * M->n_int, ctx->stream));
    DC_TRY(dc_allreduce_on(ctx, tmp, M->n_int, ctx->stream));
    DC_TRY(k3_max_abs_diff(ctx, M->w, tmp, M->n_int, (float*)M->d_inv_n));
    DC_TRY(dc_allreduce_on(ctx, (float*)M->d_inv_n, 1, ctx->stream));       // SUM of non-negative per-rank maxima: 0 iff every rank is 0
    void* pin = nullptr;
    DC_TRY(dc_pinned(ctx, 64, &pin));
    DC_CHECK_CUDA(cudaMemcpyAsync(pin, M->d_inv_n, 4, cudaMemcpyDeviceToHost, ctx->stream));
    DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    *max_abs_diff = *(float*)pin;
    return DINCAE_OK;
}

// ---- one-batch-ahead prefetch (PrefetchDataIter, src/data.jl:872-904) ------------------------------------------
extern "C" int32_t dincae_batch_upload(dincae_model* M, int32_t slot, const float* xin_host, const float* xtrue_host, int32_t B) {
    DC_TRY(check_batch(M, B, "dincae_batch_upload"));
    DC_REQUIRE(slot == 0 || slot == 1, DINCAE_EINVAL, "dincae_batch_upload: slot must be 0 or 1");
    DC_REQUIRE(xin_host && xtrue_host, DINCAE_EINVAL, "dincae_batch_upload: NULL argument");
    dincae_ctx* ctx = M->ctx;
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    const size_t npix = (size_t)M->cfg.W * M->cfg.H;
    const size_t nx = npix * M->cfg.ninput * M->maxB, ny = npix * 2 * M->O * M->maxB;
    if (!M->copy_stream) {
        // allocate into locals and publish only when everything succeeded: a failed allocation must not leave a half-built prefetch state
        cudaStream_t cs = nullptr;
        float* buf[2] = {nullptr, nullptr};
        cudaEvent_t done[2] = {nullptr, nullptr}, fre[2] = {nullptr, nullptr};
        cudaError_t e = cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking);
        for (int s = 0; s < 2 && e == cudaSuccess; s++) {
            e = cudaMalloc((void**)&buf[s], sizeof(float) * (nx + ny));
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&done[s], cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&fre[s], cudaEventDisableTiming);
        }
        if (e != cudaSuccess) {
            for (int s = 0; s < 2; s++) { if (buf[s]) cudaFree(buf[s]); if (done[s]) cudaEventDestroy(done[s]); if (fre[s]) cudaEventDestroy(fre[s]); }
            if (cs) cudaStreamDestroy(cs);
            dincae_set_error("dincae_batch_upload: cannot set up the prefetch slots: %s", cudaGetErrorString(e));
            return e == cudaErrorMemoryAllocation ? DINCAE_ENOMEM : DINCAE_ECUDA;
        }
        M->copy_stream = cs;
        for (int s = 0; s < 2; s++) { M->up_buf[s] = buf[s]; M->up_done[s] = done[s]; M->slot_free[s] = fre[s]; }
    }
    if (M->slot_used[slot]) DC_CHECK_CUDA(cudaStreamWaitEvent(M->copy_stream, M->slot_free[slot], 0));   // the step that read this slot has finished
    DC_CHECK_CUDA(cudaMemcpyAsync(M->up_buf[slot], xin_host, sizeof(float) * npix * M->cfg.ninput * B, cudaMemcpyHostToDevice, M->copy_stream));
    DC_CHECK_CUDA(cudaMemcpyAsync(M->up_buf[slot] + nx, xtrue_host, sizeof(float) * npix * 2 * M->O * B, cudaMemcpyHostToDevice, M->copy_stream));
    DC_CHECK_CUDA(cudaEventRecord(M->up_done[slot], M->copy_stream));
    return DINCAE_OK;
}

extern "C" int32_t dincae_train_step_slot(dincae_model* M, int32_t slot, int32_t B, float lr, float* loss) {
    DC_TRY(check_batch(M, B, "dincae_train_step_slot"));
    DC_REQUIRE((slot == 0 || slot == 1) && M->copy_stream, DINCAE_EINVAL, "dincae_train_step_slot: nothing was uploaded to slot %d", slot);
    dincae_ctx* ctx = M->ctx;
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    const size_t nx = (size_t)M->cfg.W * M->cfg.H * M->cfg.ninput * M->maxB;
    DC_CHECK_CUDA(cudaStreamWaitEvent(ctx->stream, M->up_done[slot], 0));
    DC_TRY(train_step_device(M, M->up_buf[slot], M->up_buf[slot] + nx, B, lr));
    DC_CHECK_CUDA(cudaEventRecord(M->slot_free[slot], ctx->stream));
    M->slot_used[slot] = true;
    if (loss) return read_loss(M, true, loss, nullptr);
    return DINCAE_OK;
}

extern "C" int32_t dincae_grad_step_slot(dincae_model* M, int32_t slot, int32_t B, const int64_t* n_valid_global) {
    DC_TRY(check_batch(M, B, "dincae_grad_step_slot"));
    DC_REQUIRE((slot == 0 || slot == 1) && M->copy_stream, DINCAE_EINVAL, "dincae_grad_step_slot: nothing was uploaded to slot %d", slot);
    dincae_ctx* ctx = M->ctx;
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    const size_t nx = (size_t)M->cfg.W * M->cfg.H * M->cfg.ninput * M->maxB;
    DC_CHECK_CUDA(cudaStreamWaitEvent(ctx->stream, M->up_done[slot], 0));
    DC_TRY(dincae_grad_step(M, M->up_buf[slot], M->up_buf[slot] + nx, B, n_valid_global, DINCAE_DEVICE));
    DC_CHECK_CUDA(cudaEventRecord(M->slot_free[slot], ctx->stream));
    M->slot_used[slot] = true;
    return DINCAE_OK;
}

extern "C" int32_t dincae_train_step_dp_slot(dincae_model* M, int32_t slot, int32_t B, const int64_t* n_valid_global, float lr, float* loss) {
    DC_REQUIRE(M != nullptr, DINCAE_EINVAL, "dincae_train_step_dp_slot: model is NULL");
    DC_REQUIRE(M->ctx->comm != nullptr, DINCAE_EINVAL, "dincae_train_step_dp_slot: no communicator (dincae_comm_init_rank)");
    DC_REQUIRE(n_valid_global != nullptr, DINCAE_EINVAL, "dincae_train_step_dp_slot: n_valid_global is NULL");
    M->dp_overlap = dp_overlap_on();
    int32_t st = dincae_grad_step_slot(M, slot, B, n_valid_global);
    const bool bucketed = M->dp_overlap;
    M->dp_overlap = false;
    DC_TRY(st);
    if (!bucketed) DC_TRY(dincae_allreduce_grads(M));
    return dincae_apply_step(M, lr, loss);
}

extern "C" int32_t dincae_loss_points(dincae_model* M, const float* xin, const float* pos, const float* x,
                                      const int64_t* offsets, int32_t B, int32_t loc, double* loss) {
    DC_TRY(check_batch(M, B, "dincae_loss_points"));
    DC_REQUIRE(xin && loss, DINCAE_EINVAL, "dincae_loss_points: NULL argument");
    dincae_ctx* ctx = M->ctx;
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    const float* dx;
    DC_TRY(stage_batch(M, xin, M->cfg.ninput, B, loc == DINCAE_HOST ? DINCAE_HOST : DINCAE_DEVICE, M->in_xin, &dx));
    PointTruth pt;
    DC_TRY(stage_points(M, pos, x, offsets, B, loc == DINCAE_DEVICE ? DINCAE_DEVICE : DINCAE_HOST, &pt));
    DC_CHECK_CUDA(cudaMemsetAsync(M->d_loss, 0, sizeof(double) * 8, ctx->stream));
    int32_t st = model_forward(M, dx, B, false);
    if (st == DINCAE_OK) st = points_loss(M, pt, B, -1, false);
    dc_stage_free(ctx, &pt.spos);
    dc_stage_free(ctx, &pt.sx);
    DC_TRY(st);
    DC_TRY(k3_reg(ctx, M->w, M->n_int, M->cfg.l1_beta, M->cfg.l2_beta, M->d_loss + 1, M->red_scratch, M->red_scratch_bytes));
    return read_loss(M, false, nullptr, loss);
}

extern "C" int32_t dincae_grad_step_points(dincae_model* M, const float* xin, const float* pos, const float* x,
                                           const int64_t* offsets, int32_t B, int64_t n_valid_global, int32_t loc) {
    DC_TRY(check_batch(M, B, "dincae_grad_step_points"));
    DC_REQUIRE(xin != nullptr, DINCAE_EINVAL, "dincae_grad_step_points: NULL argument");
    dincae_ctx* ctx = M->ctx;
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    const float* dx;
    DC_TRY(stage_batch(M, xin, M->cfg.ninput, B, loc == DINCAE_HOST ? DINCAE_HOST : DINCAE_DEVICE, M->in_xin, &dx));
    PointTruth pt;
    DC_TRY(stage_points(M, pos, x, offsets, B, loc == DINCAE_DEVICE ? DINCAE_DEVICE : DINCAE_HOST, &pt));
    DC_CHECK_CUDA(cudaMemsetAsync(M->d_loss, 0, sizeof(double) * 8, ctx->stream));
    int32_t st = model_forward(M, dx, B, true);
    if (st == DINCAE_OK) st = points_loss(M, pt, B, n_valid_global, true);
    dc_stage_free(ctx, &pt.spos);
    dc_stage_free(ctx, &pt.sx);
    DC_TRY(st);
    return backward_all(M, B);
}

extern "C" int32_t dincae_train_step_points(dincae_model* M, const float* xin, const float* pos, const float* x,
                                            const int64_t* offsets, int32_t B, float lr, int32_t loc, float* loss) {
    DC_TRY(dincae_grad_step_points(M, xin, pos, x, offsets, B, -1, loc));
    return dincae_apply_step(M, lr, loss);
}

extern "C" int32_t dincae_reconstruct_range(dincae_model* M, dincae_gridded* ds, int64_t t0, int64_t t1,
                                            int32_t count, float fill_value, float* m_out, float* sigma_out, int32_t loc) {
    DC_REQUIRE(M && ds && m_out && sigma_out, DINCAE_EINVAL, "dincae_reconstruct_range: NULL argument");
    DC_REQUIRE(0 <= t0 && t0 <= t1 && t1 <= ds->N, DINCAE_EINVAL, "dincae_reconstruct_range: bad range [%lld,%lld) of %lld", (long long)t0, (long long)t1, (long long)ds->N);
    DC_REQUIRE(count >= 1, DINCAE_EINVAL, "dincae_reconstruct_range: count must be >= 1 (ds.attrib[\"count\"], src/data.jl:621-624)");
    DC_REQUIRE(ds->W == M->cfg.W && ds->H == M->cfg.H && 4 + 2 * ds->T * ds->V == M->cfg.ninput && ds->nout == M->O, DINCAE_EINVAL,
               "dincae_reconstruct_range: dataset (W=%d,H=%d,channels=%d,nout=%d) does not match the model", ds->W, ds->H, 4 + 2 * ds->T * ds->V, ds->nout);
    dincae_ctx* ctx = M->ctx;
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    const int64_t npix = (int64_t)ds->W * ds->H, nsteps = t1 - t0;
    if (nsteps == 0) return DINCAE_OK;
    float *dm = m_out, *dsg = sigma_out;
    const size_t nres = (size_t)npix * nsteps * M->O;
    if (loc == DINCAE_HOST) {
        DC_CHECK_CUDA(cudaMallocAsync((void**)&dm, sizeof(float) * 2 * nres, ctx->stream));
        dsg = dm + nres;
        if (count > 1) {
            DC_CHECK_CUDA(cudaMemcpyAsync(dm, m_out, sizeof(float) * nres, cudaMemcpyHostToDevice, ctx->stream));
            DC_CHECK_CUDA(cudaMemcpyAsync(dsg, sigma_out, sizeof(float) * nres, cudaMemcpyHostToDevice, ctx->stream));
        }
    }
    std::vector<int64_t> idx(M->maxB);
    for (int64_t s = 0; s < nsteps; s += M->maxB) {
        int B = (int)(nsteps - s < M->maxB ? nsteps - s : M->maxB);
        for (int b = 0; b < B; b++) idx[b] = t0 + s + b;          // DataBatches(train=false): sequential, no perturbation
        DC_TRY(k1_assemble(ds, idx.data(), nullptr, 0, 0, B, M->in_xin, M->in_xtrue));
        DC_TRY(forward_infer_or_graph(M, B));
        DC_TRY(k3_savesample(ctx, M->MS, B, M->O, npix, ds->d_mean_out, ds->d_landmask, count, fill_value,
                             dm + (size_t)s * npix, dsg + (size_t)s * npix, npix * nsteps));
    }
    if (loc == DINCAE_HOST) {
        DC_CHECK_CUDA(cudaMemcpyAsync(m_out, dm, sizeof(float) * nres, cudaMemcpyDeviceToHost, ctx->stream));
        DC_CHECK_CUDA(cudaMemcpyAsync(sigma_out, dsg, sizeof(float) * nres, cudaMemcpyDeviceToHost, ctx->stream));
        DC_CHECK_CUDA(cudaFreeAsync(dm, ctx->stream));
        DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return DINCAE_OK;
}
