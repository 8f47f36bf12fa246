// K3: (m, sigma^2) head fused with the masked Gaussian NLL (src/model.jl:7-148), Laplacian penalty
// (:44-52), L1/L2 regularisation (:210-219) and the ClipGrad+Adam update (src/flux.jl:10-14,:33).
// All HBM-bound element-wise / reduction kernels over planar (Julia WHCN) tensors.
#pragma once
#include "common.cuh"

struct HeadParams {
    float gamma;
    int truth_uncertain;
    float lap, lap_err;      // laplacian_penalty, laplacian_error_penalty
};

// n_valid[v] += #(xtrue[:,:,2v+1,:] != 0)   (n_noncloud, src/model.jl:61,114) -- integer, exact
int32_t k3_count_valid(dincae_ctx* ctx, const float* xtrue, int B, int nout, int64_t npix, unsigned long long* n_valid);
// ms = transform_msigma2(raw)  (src/model.jl:16-41); raw, ms planar (W,H,2*nout,B)
int32_t k3_head(dincae_ctx* ctx, const float* raw, int B, int nout, int64_t npix, float gamma, float* ms);
// gridded loss of one chain + gradient w.r.t. raw:
//   loss_acc[0] += weight * costfun(head(raw), xtrue);  graw = weight * d cost / d raw     (graw may be NULL)
// inv_n[v] = 1 / n_noncloud of the (global) mini-batch.  ms = head(raw) must be given when a Laplacian penalty is set.
int32_t k3_loss_grid(dincae_ctx* ctx, const float* raw, const float* xtrue, const float* ms, int B, int nout, int W, int H,
                     const HeadParams& hp, const float* inv_n, float weight, double* loss_acc, float* graw,
                     void* scratch, size_t scratch_bytes);
// point loss pieces (src/points.jl:354-426): per-observation NLL terms from gathered (m, s2)
//   gms[0:P] = d cost/d m_p, gms[P:2P] = d cost/d s2_p ; loss_acc += weight*cost
int32_t k3_loss_points(dincae_ctx* ctx, const float* m_s2_at_pts, const float* x, int64_t P, const HeadParams& hp,
                       const float* inv_n, float weight, double* loss_acc, float* gms, void* scratch, size_t scratch_bytes);
// chain rule through the head (+ Laplacian terms): graw = weight*(d/d raw)[ <g_ms, head(raw)> + lap penalties ]
//   g_ms planar like ms, already multiplied by weight; loss_acc += weight * laplacian penalties
int32_t k3_head_bwd(dincae_ctx* ctx, const float* raw, const float* ms, const float* g_ms, int B, int nout, int W, int H,
                    const HeadParams& hp, float weight, double* loss_acc, float* graw, void* scratch, size_t scratch_bytes);

struct AdamParams {
    float lr, beta1, beta2, eps, clip, l1, l2;
    float inv_bc1, inv_bc2;        // 1/(1-beta1^t), 1/(1-beta2^t)
};
// fused: g += 2*l2*w + l1*sign(w); clamp; Adam; also reg_acc[0] += l1*sum|w| + l2*sum w^2 (pre-update weights)
int32_t k3_adam(dincae_ctx* ctx, float* w, const float* g, float* m, float* v, int64_t n, const AdamParams& p,
                double* reg_acc, void* scratch, size_t scratch_bytes);
// CUDA-graph support: the (single) node of the optimiser kernel in a captured step, and the per-step update of its AdamParams
cudaGraphNode_t k3_find_adam_node(cudaGraph_t graph);
int32_t k3_update_adam_node(cudaGraphExec_t exec, cudaGraphNode_t node, const AdamParams& p);
// reg_acc[0] += l1*sum|w| + l2*sum w^2 only (dincae_loss)
// out[0] = max_i |a[i] - b[i]| (integer atomicMax on the bit pattern of non-negative floats: order-independent)
int32_t k3_max_abs_diff(dincae_ctx* ctx, const float* a, const float* b, int64_t n, float* out);
int32_t k3_reg(dincae_ctx* ctx, const float* w, int64_t n, float l1, float l2, double* reg_acc, void* scratch, size_t scratch_bytes);
size_t k3_scratch_bytes(int64_t n_elems);
// savesample arithmetic (src/data.jl:615-619,:653-675) fused: ms planar (W,H,2*nout,B) -> running mean into m_out/sigma_out
int32_t k3_savesample(dincae_ctx* ctx, const float* ms, int B, int nout, int64_t npix, const float* meandata_out /*(W,H,nout)*/,
                      const uint8_t* landmask, int count, float fill_value, float* m_out, float* sigma_out, int64_t var_stride);
