// K2: 3x3 convolution family (forward, data-gradient, weight-gradient) and the element-wise
// pool / upsample / crop / skip operators of the U-Net level (recmodel4, src/model.jl:242-284).
//
// Internal activation layout: NHWC Float32, channels padded to a multiple of 4 ("Cp"), i.e.
// t[((b*H + y)*W + x)*Cp + c] with x = lon index (Julia dim 1) and y = lat index (Julia dim 2).
// Internal weight layout of one Conv: w[(tap*Cin_p + ci)*Cout_p + co], tap = ky*3+kx of the
// cross-correlation out[y,x] = sum w[tap] * in[y+ky-1, x+kx-1]; Flux's true convolution kernel
// W[a,b,ci,co] (src/model.jl:267) maps to tap (ky,kx) = (2-b, 2-a)  (SURVEY T8).
#pragma once
#include "common.cuh"

enum { SRC_DIRECT = 0, SRC_UP_NEAREST = 1 };
enum { ACT_NONE = 0, ACT_RELU = 1 };

// Producer description: the conv input at output-resolution coordinates (b,y,x,c).
//   SRC_DIRECT     : src0[b,y,x,c]                       (+ src1[b,y,x,c] if given)
//   SRC_UP_NEAREST : src0[b,y>>1,x>>1,c] (+ src1[...])   = Upsample(:nearest,scale=2) then croppadding
//                    (src/flux.jl:4, src/model.jl:270-271) and SumSkip of the level below (:277-279)
struct ConvSrc {
    const float* src0;
    const float* src1;
    int mode;
    int Hs, Ws;   // spatial size of the source tensors
    int C;        // padded channel count of the source tensors
};

struct ConvArgs {
    ConvSrc in;
    int B, H, W;          // conv input == output spatial size
    const float* wgt;     // [Kpad][Cout_p], Kpad = round_up(9*Cin_p, 16), zero padded
    const float* wgtK;    // tensor-core path: K-major copy [Cout_p][ldk] (values pre-rounded to TF32), or NULL
    int ldk;              // 9*Cin_p
    int in_is_tf32;       // the input tensor already holds TF32-representable values (TMA path needs no producer math)
    int in_real_C;        // 0 or the number of leading input channels that can be non-zero (the rest is zero padding): the persistent
                          // tensor-core kernel then skips the K = 8 steps that would only multiply zeros
    const float* bias;    // [Cout_p] or NULL
    int Cout_p;
    int act;
    float slope;          // leaky slope (0 = ReLU)
    const float* addend;  // optional [B,H,W,Cout_p], added AFTER the activation
    float* dst;           // [B,H,W,Cout_p] (may be NULL when planar_out is given)
    float* planar_out;    // optional planar (Julia WHCN) output [B][planar_C][H][W] of the first planar_C channels (TMA kernel only)
    int planar_C;
    int planar_accumulate;   // 1: planar_out += value
    int planar_fold;         // 1: planar channel k = accumulator column k + column (8 + k): the head conv's low-order weight
                             // columns (3xTF32 split, see k2_make_wK_split) are summed in the epilogue; needs planar_C <= 8
    float* pool_dst;         // optional: MeanPool((2,2)) of the output, only when k2_conv_fuses_pool(args) is true
    uint32_t* mask_dst;      // pool-only mode (dst == NULL, pool_dst set): the activation itself is never written, only one bit per
                             // element (output > 0 = what the backward pass needs of it), word [pixel][Cout_p / 32]
    double work_flops;    // algorithmic FLOPs (2*9*Cin*Cout*B*H*W with the REAL channel counts), for the profile
    double work_bytes;    // algorithmic HBM bytes (input + output + weights, REAL channel counts)
};

// deferred second-stage reductions: every weight gradient appends its split-K partial sums and its bias column partials here;
// ONE batched kernel at the end of the backward pass finishes them all (deterministic fixed-order sums)
struct RedEntry {
    const float* part;   // [splits][n]
    float* dst;          // [n]
    int64_t n;
    int splits;
};
constexpr int RED_MAX = 48;
struct RedList {
    RedEntry e[RED_MAX];
    int blk_start[RED_MAX + 1];
    int count;
};
int32_t k2_reduce_batched(dincae_ctx* ctx, RedList& list);

struct WgradArgs {
    ConvSrc in;
    int B, H, W;
    const float* gout;    // [B,H,W,Cout_p] gradient w.r.t. the conv output (pre-activation)
    int Cout_p;
    float* dW;            // [9*Cin_p][Cout_p]
    float* db;            // [Cout_p]
    float* partial;       // split-K scratch
    int splits;
    int64_t pix_per_split;
    double work_flops;
    double work_bytes;
    int tf32;             // operands are TF32-representable: the TMA/tcgen05 kernel may be used
    RedList* defer;       // if set, the second-stage reductions are queued instead of launched
    const float* colpart; // optional precomputed column partials of gout [colpart_blocks][Cout_p] (bias gradient)
    int colpart_blocks;
};

// small-Cout forward (Cout_p in {4,8}): NHWC in -> planar out[b][co][y][x] (co < Cout real)
struct ConvSmallArgs {
    ConvSrc in;
    int B, H, W;
    const float* wgt;     // element (k, j) at wgt[k*ldw + j], k = tap*Cin_p + ci, j < NCO
    int ldw;
    const float* bias;    // or NULL
    int Cout;             // real output channels written
    int accumulate;       // 0: out = v ; 1: out += v
    float* out;           // planar [B][Cout][H][W]
    double work_flops;
    double work_bytes;
};

int32_t k2_conv(dincae_ctx* ctx, const ConvArgs& a);        // dispatches to the tcgen05 kernel when eligible
bool k2_conv_tma_eligible(const ConvArgs& a);
bool k2_conv_fuses_pool(const ConvArgs& a);            // k2_conv(a) can also write MeanPool((2,2)) of its output to a.pool_dst
int32_t k2_conv_tma(dincae_ctx* ctx, const ConvArgs& a);
size_t k2_wgrad_scratch_bytes(int num_sms, int B, int H, int W, int Cin_p, int Cout_p);
int32_t k2_wgrad(dincae_ctx* ctx, WgradArgs a, void* scratch, size_t scratch_bytes);   // dispatches to the tcgen05 kernel when eligible
bool k2_wgrad_tma_eligible(const WgradArgs& a);
size_t k2_wgrad_tma_scratch_bytes(int num_sms, int B, int H, int W, int Cin_p, int Cout_p);
int32_t k2_wgrad_tma(dincae_ctx* ctx, const WgradArgs& a, void* scratch, size_t scratch_bytes);
int32_t k2_conv_small(dincae_ctx* ctx, const ConvSmallArgs& a, int nco);
// weight gradient for gout with NCO (4 or 8) padded channels
int32_t k2_wgrad_small(dincae_ctx* ctx, WgradArgs a, int nco, void* scratch, size_t scratch_bytes);

// Head conv fused with nearest upsample (+ crop + skip add) of its half-resolution source, exact FP32 (CUDA cores), planar output
// out[b][o][y][x], o < nout2 = 2*noutput (src/model.jl:270-272, src/flux.jl:4); src NHWC with Cp channels of which Cin are real
bool k2_head_up_conv_eligible(int Cin, int nout2);
int32_t k2_head_up_conv(dincae_ctx* ctx, const float* src0, const float* src1, int B, int Hs, int Ws, int Cp, int Cin,
                        const float* wgt, int Cin_p, int Cout_p, const float* bias, int nout2, int H, int W, float* out,
                        double work_flops, double work_bytes);
// Backward of the fused head, exact FP32 from the planar gradient g = d loss / d raw (nout2 channels at (H, W)):
//   k2_head_bwd_data : G = up^T(conv^T(g)) at (Hs, Ws) with Cp channels, GD = G * act'(R) (TF32-rounded; R, G or GD may be NULL)
//   k2_head_wgrad    : dW (padded layout, padding zeroed) and db of the head conv from the half-resolution source src0 (+ src1);
//                      scratch holds the per-block partials (k2_head_bwd_scratch_bytes), summed in block order
bool k2_head_bwd_eligible(int Cin, int nout2);
size_t k2_head_bwd_scratch_bytes(int B, int Hs, int Ws, int Cin, int nout2);
int32_t k2_head_bwd_data(dincae_ctx* ctx, const float* g, int B, int H, int W, int Hs, int Ws, const float* wgt, int Cin, int Cin_p, int Cout_p,
                         int nout2, const float* R, float slope, int Cp, float* G, float* GD);
int32_t k2_head_wgrad(dincae_ctx* ctx, const float* src0, const float* src1, const float* g, int B, int H, int W, int Hs, int Ws, int Cp, int Cin,
                      int nout2, int Cin_p, int Cout_p, float* dW, float* db, void* scratch, size_t scratch_bytes);
// planar (Julia WHCN) -> NHWC padded; channels [0,C0) from src0, [C0,C0+C1) from src1 (cat(xout,xin,dims=3), src/model.jl:186)
// split != 0 (tensor-core mode, first conv of a chain): the C = C0 + C1 real channels are written twice as
// [tf32(a) | tf32(a - tf32(a))], so that the conv multiplies the unrounded activation (a_hi + a_lo) * tf32(w); the matching
// weight rows are laid out by k2_make_wK_split (Cp >= 2*C)
int32_t k2_pack(dincae_ctx* ctx, const float* src0, int C0, const float* src1, int C1, int B, int H, int W, float* dst, int Cp, int split = 0);
// MeanPool((2,2), pad=odd): count-include-pad (src/model.jl:268)
int32_t k2_pool(dincae_ctx* ctx, const float* a, int B, int H, int W, int Cp, float* x);
// backward of nearest-upsample + crop (+ activation mask of r): G = up^T(gu); gd = G * act'(r)
// colpart (optional): [k2_colpart_blocks][Cp] per-block column sums of gd (the bias gradient's first stage, fused here)
int k2_colpart_blocks(dincae_ctx* ctx, int64_t n_elems4, int Cp);
int32_t k2_upbwd(dincae_ctx* ctx, const float* gu, int B, int H, int W, int Cp, const float* r, float slope, float* G, float* gd, float* colpart = nullptr);
// backward of mean-pool + activation mask: gc = 0.25 * xg[(y+oh)>>1,(x+ow)>>1] * act'(a)
int32_t k2_poolbwd(dincae_ctx* ctx, const float* xg, const float* a, int B, int H, int W, int Cp, float slope, float* gc, float* colpart = nullptr);
// the same with the activation's sign bits (written by the pool-only conv epilogue) in place of the activation
int32_t k2_poolbwd_mask(dincae_ctx* ctx, const float* xg, const uint32_t* mask, int B, int H, int W, int Cp, float slope, float* gc, float* colpart = nullptr);
// bilinear (align_corners=true) upsample+crop of (src0 + src1) and its transpose
// nearest upsample x2 + crop of (src0 + src1), materialised (TF32 mode: feeds the TMA kernels)
// split != 0: dst has 2*Cp channels per pixel, [tf32(v) | tf32(v - tf32(v))] (input of the 3xTF32 head conv)
int32_t k2_up_nearest(dincae_ctx* ctx, const float* src0, const float* src1, int B, int Hs, int Ws, int Cp, int H, int W, float* dst, int split = 0);
int32_t k2_up_bilinear(dincae_ctx* ctx, const float* src0, const float* src1, int B, int Hs, int Ws, int Cp, int H, int W, float* dst, int split = 0);
int32_t k2_up_bilinear_bwd(dincae_ctx* ctx, const float* gu, int B, int H, int W, int Cp, int Hs, int Ws, const float* r, float slope, float* G, float* gd);

// table-driven weight transpose+flip for the data-gradient convs
struct WLayerDesc {
    int64_t w_off, wT_off;
    int Cin_p, Cout_p;
    int64_t n;   // 9*Cin_p*Cout_p
    int64_t wk_off;   // offset of this layer in the forward K-major copy wK
    int64_t wtk_off;  // offset of this layer in the data-gradient K-major copy wTK
};
// wT: transposed + tap-flipped [9*Cout_p][Cin_p] (FP32 path).  If wK/wTK are given also the K-major TF32-rounded copies:
//   wK [Cout_p][9*Cin_p] for the forward conv, wTK [Cin_p][9*Cout_p] (tap-flipped) for the data gradient.
// Split-operand layers (DINCAE_PREC_TF32: the first conv and the head conv of every chain, the thin full-resolution layers whose
// operand rounding dominates the output error): K-major forward weights wK [Cout_p][9*Cin_pe] in the expanded operand layout,
// and the fold of the expanded weight gradient back onto the master layout.
//   mode 1 (first conv, C real input channels; 2xTF32 = exact activations): k-slots [0,C) = hi(w), [C,2C) = hi(w), rest 0
//   mode 2 (head conv, P0 = Cin_p, n2 = real output channels): rows co < n2: slots [0,P0) = hi(w), [P0,2P0) = hi(w);
//           rows 8 <= co < 8 + n2: slots [0,P0) = lo(w[., co - 8]); everything else 0 (n2 <= 8)
//   hi(w) = tf32(w), lo(w) = tf32(w - hi(w)).
struct SplitLayerDesc {
    int64_t w_off;       // master weights [9*Cin_p][Cout_p] (and the gradient slot of the same layout)
    int64_t wk_off;      // expanded K-major copy [Cout_p][9*Cin_pe]
    int64_t gexp_off;    // expanded weight gradient [9*Cin_pe][Cout_p] in the temporary buffer
    int Cin_p, Cin_pe, Cout_p;
    int mode, C, n2;
};
constexpr int SPLIT_MAX = 2 * DINCAE_MAX_CHAINS;
struct SplitTable { SplitLayerDesc d[SPLIT_MAX]; int count; };
int32_t k2_make_wK_split(dincae_ctx* ctx, const float* w, float* wK, const SplitTable& tab);
// g[(tap*Cin_p + ci)*Cout_p + co] = gexp[(tap*Cin_pe + ci)*Cout_p + co] + gexp[(tap*Cin_pe + lo + ci)*Cout_p + co]  (lo = C | Cin_p),
// zero for the padding channels: the hi and lo activation slots multiply the same weight
int32_t k2_fold_split_grad(dincae_ctx* ctx, const float* gexp, float* g, const SplitTable& tab);
int32_t k2_make_wT(dincae_ctx* ctx, const float* w, float* wT, float* wK, float* wTK, const WLayerDesc* d_table, int nlayers, int64_t total);
