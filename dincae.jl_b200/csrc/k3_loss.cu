// K3: head + masked Gaussian NLL + Laplacian + regularisers + ClipGrad/Adam (see k3_loss.cuh).
// Derivatives follow SURVEY.md Appendix A, derived from src/model.jl:16-29 and :79-92.
#include "k3_loss.cuh"

namespace {

constexpr int RED_THREADS = 256;
constexpr int RED_MAX_BLOCKS = 1024;

__device__ __forceinline__ double block_sum(double v) {
    __shared__ double sh[RED_THREADS / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    double s = 0.0;
    if (threadIdx.x == 0)
        for (int i = 0; i < RED_THREADS / 32; i++) s += sh[i];
    return s;   // valid on thread 0
}

__global__ void final_sum_kernel(const double* __restrict__ part, int n, double* __restrict__ acc) {
    __shared__ double sh[256];
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) s += part[i];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) acc[0] += sh[0];
}

__global__ void count_valid_kernel(const float* __restrict__ xtrue, int B, int nout, int64_t npix, unsigned long long* __restrict__ n_valid) {
    // grid: (blocks, nout)
    int v = blockIdx.y;
    int64_t total = (int64_t)B * npix;
    unsigned int c = 0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t b = i / npix, p = i - b * npix;
        c += (xtrue[(b * 2 * nout + 2 * v + 1) * npix + p] != 0.f) ? 1u : 0u;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&n_valid[v], (unsigned long long)c);   // integer atomics: exact, order-free
}

struct HeadVals {
    float p, q, s2, m;
    bool open;     // d q / d b != 0
};
__device__ __forceinline__ HeadVals head_eval(float a, float l, float gamma) {
    HeadVals h;
    h.p = expf(fminf(l, gamma));            // invσ² = exp(min(x2, γ)), src/model.jl:21
    h.q = fmaxf(h.p, 1e-3f);                // sinv clamp, src/model.jl:7-10
    h.s2 = 1.0f / h.q;
    h.m = a * h.s2;                         // m = x1 .* σ², src/model.jl:23
    h.open = (l < gamma) && (h.p > 1e-3f);
    return h;
}

__global__ void head_kernel(const float* __restrict__ raw, int64_t nvb, int64_t npix, float gamma, float* __restrict__ ms) {
    int64_t total = nvb * npix;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t vb = i / npix, p = i - vb * npix;
        size_t o = (size_t)(2 * vb) * npix + p;
        HeadVals h = head_eval(raw[o], raw[o + npix], gamma);
        ms[o] = h.m;
        ms[o + npix] = h.s2;
    }
}

// NLL terms of one observation given (m, s2) and the target pair (xt1, xt2); returns cost term, sets gm, gs (unnormalised by n)
__device__ __forceinline__ float nll_terms(float m, float s2, float xt1, float xt2, int truth_uncertain, float& gm, float& gs) {
    if (xt2 == 0.f) { gm = 0.f; gs = 0.f; return 0.f; }       // mask_noncloud, src/model.jl:114
    float s2t = 1.0f / fmaxf(xt2, 1e-3f);                         // :109
    float mt = xt1 * s2t;                                         // :110
    float d = m - mt;
    float inv = 1.0f / s2;
    float num = truth_uncertain ? (d * d + s2t) : d * d;          // :83 / :90
    float cost = (truth_uncertain ? logf(s2 / s2t) : logf(s2)) + num * inv;   // :86 / :92
    gm = 2.f * d * inv;
    gs = inv - num * inv * inv;
    return cost;
}

// chain rule through the head: (gm, gs) -> (d/da, d/dl)
__device__ __forceinline__ void head_chain(float a, const HeadVals& h, float gm, float gs, float& ga, float& gl) {
    ga = gm * h.s2;
    float gs2 = gs + gm * a;              // m = a*s2
    float gq = -gs2 * h.s2 * h.s2;        // s2 = 1/q
    gl = h.open ? gq * h.p : 0.f;
}

template <bool FUSE_BWD>
__global__ void __launch_bounds__(RED_THREADS) loss_grid_kernel(const float* __restrict__ raw, const float* __restrict__ xtrue, int B, int nout,
                                                               int64_t npix, HeadParams hp, const float* __restrict__ inv_n, float weight,
                                                               double* __restrict__ part, float* __restrict__ gout) {
    int64_t total = (int64_t)B * nout * npix;
    double acc = 0.0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t vb = i / npix, p = i - vb * npix;
        int v = (int)(vb % nout);
        size_t o = (size_t)(2 * vb) * npix + p;
        float a = raw[o], l = raw[o + npix];
        HeadVals h = head_eval(a, l, hp.gamma);
        float gm, gs;
        float c = nll_terms(h.m, h.s2, xtrue[o], xtrue[o + npix], hp.truth_uncertain, gm, gs);
        float sc = weight * inv_n[v];
        acc += (double)c * (double)sc;
        if (gout) {
            gm *= sc; gs *= sc;
            if (FUSE_BWD) {
                float ga, gl;
                head_chain(a, h, gm, gs, ga, gl);
                gout[o] = ga;
                gout[o + npix] = gl;
            } else {
                gout[o] = gm;
                gout[o + npix] = gs;
            }
        }
    }
    double s = block_sum(acc);
    if (threadIdx.x == 0) part[blockIdx.x] = s;
}

__global__ void __launch_bounds__(RED_THREADS) loss_points_kernel(const float* __restrict__ mv, const float* __restrict__ x, int64_t P, HeadParams hp,
                                                                 const float* __restrict__ inv_n, float weight, double* __restrict__ part,
                                                                 float* __restrict__ gms) {
    double acc = 0.0;
    float sc = weight * inv_n[0];
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
        float gm, gs;
        float c = nll_terms(mv[p], mv[P + p], x[p], x[P + p], hp.truth_uncertain, gm, gs);
        acc += (double)c * (double)sc;
        if (gms) { gms[p] = gm * sc; gms[P + p] = gs * sc; }
    }
    double s = block_sum(acc);
    if (threadIdx.x == 0) part[blockIdx.x] = s;
}

// second difference along one axis and its adjoint applied to itself: returns [D^T D u]_i and adds (D u)_i^2 to pen
__device__ __forceinline__ float lap_axis(const float* __restrict__ u, int i, int n, int64_t stride, float& pen) {
    auto dd = [&](int k) -> float {     // (D u)_k for 1 <= k <= n-2, else 0
        if (k < 1 || k > n - 2) return 0.f;
        return u[(int64_t)(k - 1 - i) * stride] - 2.f * u[(int64_t)(k - i) * stride] + u[(int64_t)(k + 1 - i) * stride];
    };
    float d0 = dd(i);
    pen += d0 * d0;
    return dd(i - 1) - 2.f * d0 + dd(i + 1);
}

__global__ void __launch_bounds__(RED_THREADS) head_bwd_kernel(const float* __restrict__ raw, const float* __restrict__ ms, const float* g_ms,
                                                              int B, int nout, int W, int H, HeadParams hp, float weight,
                                                              double* __restrict__ part, float* graw) {
    int64_t npix = (int64_t)W * H;
    int64_t total = (int64_t)B * nout * npix;
    double acc = 0.0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t vb = i / npix, p = i - vb * npix;
        int y = (int)(p / W), x = (int)(p - (int64_t)y * W);
        size_t o = (size_t)(2 * vb) * npix + p;
        float a = raw[o], l = raw[o + npix];
        HeadVals h = head_eval(a, l, hp.gamma);
        float gm = g_ms ? g_ms[o] : 0.f, gs = g_ms ? g_ms[o + npix] : 0.f;
        if (hp.lap != 0.f) {        // sum_laplacian_penalty(laplacian_penalty, m_rec), src/model.jl:44-52,:99
            float pen = 0.f;
            float g = lap_axis(ms + o, x, W, 1, pen) + lap_axis(ms + o, y, H, W, pen);
            gm += weight * 2.f * hp.lap * g;
            acc += (double)weight * (double)hp.lap * (double)pen;
        }
        if (hp.lap_err != 0.f) {    // ... (laplacian_error_penalty, σ2_rec), src/model.jl:100
            float pen = 0.f;
            float g = lap_axis(ms + o + npix, x, W, 1, pen) + lap_axis(ms + o + npix, y, H, W, pen);
            gs += weight * 2.f * hp.lap_err * g;
            acc += (double)weight * (double)hp.lap_err * (double)pen;
        }
        if (graw) {
            float ga, gl;
            head_chain(a, h, gm, gs, ga, gl);
            graw[o] = ga;
            graw[o + npix] = gl;
        }
    }
    double s = block_sum(acc);
    if (threadIdx.x == 0) part[blockIdx.x] = s;
}

template <bool UPDATE>
__global__ void __launch_bounds__(RED_THREADS) adam_kernel(float* __restrict__ w, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                                                          int64_t n, AdamParams p, double* __restrict__ part) {
    double acc = 0.0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float wi = w[i];
        if (p.l1 != 0.f) acc += (double)p.l1 * fabs((double)wi);
        if (p.l2 != 0.f) acc += (double)p.l2 * (double)wi * (double)wi;
        if (UPDATE) {
            float gi = g[i];
            if (p.l2 != 0.f) gi += 2.f * p.l2 * wi;                       // d/dw beta*sum(abs2,w), src/model.jl:215-217
            if (p.l1 != 0.f) gi += p.l1 * (wi > 0.f ? 1.f : (wi < 0.f ? -1.f : 0.f));
            if (p.clip > 0.f) gi = fminf(fmaxf(gi, -p.clip), p.clip);    // ClipGrad(delta): element clamp
            float mi = p.beta1 * m[i] + (1.f - p.beta1) * gi;
            float vi = p.beta2 * v[i] + (1.f - p.beta2) * gi * gi;
            m[i] = mi;
            v[i] = vi;
            w[i] = wi - p.lr * (mi * p.inv_bc1) / (sqrtf(vi * p.inv_bc2) + p.eps);
        }
    }
    double s = block_sum(acc);
    if (threadIdx.x == 0) part[blockIdx.x] = s;
}

__global__ void savesample_kernel(const float* __restrict__ ms, int B, int nout, int64_t npix, const float* __restrict__ mean,
                                  const uint8_t* __restrict__ landmask, int count, float fill, float* __restrict__ m_out,
                                  float* __restrict__ s_out, int64_t var_stride) {
    int64_t total = (int64_t)B * nout * npix;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t vb = i / npix, p = i - vb * npix;
        int v = (int)(vb % nout);
        int64_t b = vb / nout;
        size_t o = (size_t)(2 * vb) * npix + p;
        float rec = ms[o] + mean[(size_t)v * npix + p];       // recdata = batch_m_rec .+ meandata, src/data.jl:661
        float sig = sqrtf(ms[o + npix]);                      // :662
        if (rec != rec) sig = nanf("");                       // :664
        if (landmask && !landmask[p]) { rec = nanf(""); sig = rec; }   // :668-669
        size_t q = (size_t)v * var_stride + (size_t)b * npix + p;
        float om = count > 1 ? m_out[q] : 0.f, os = count > 1 ? s_out[q] : 0.f;
        float nm = ((float)(count - 1) * om + rec) / (float)count;      // accumulate!, :615-619
        float ns = ((float)(count - 1) * os + sig) / (float)count;
        m_out[q] = (nm != nm) ? fill : nm;
        s_out[q] = (ns != ns) ? fill : ns;
    }
}

int red_blocks(dincae_ctx* ctx, int64_t n) {
    int64_t b = dc_cdiv(n, RED_THREADS * 4);
    int64_t cap = 8 * (int64_t)ctx->num_sms;
    if (cap > RED_MAX_BLOCKS) cap = RED_MAX_BLOCKS;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

}  // namespace

size_t k3_scratch_bytes(int64_t) { return RED_MAX_BLOCKS * sizeof(double); }

int32_t k3_count_valid(dincae_ctx* ctx, const float* xtrue, int B, int nout, int64_t npix, unsigned long long* n_valid) {
    DC_CHECK_CUDA(cudaMemsetAsync(n_valid, 0, sizeof(unsigned long long) * nout, ctx->stream));
    if (B == 0) return DINCAE_OK;
    dim3 grid(red_blocks(ctx, (int64_t)B * npix), nout);
    DC_LAUNCH(ctx, count_valid_kernel, grid, 256, 0, xtrue, B, nout, npix, n_valid);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

int32_t k3_head(dincae_ctx* ctx, const float* raw, int B, int nout, int64_t npix, float gamma, float* ms) {
    int64_t n = (int64_t)B * nout * npix;
    if (n == 0) return DINCAE_OK;
    DC_WORK(ctx, 0.0, 16.0 * B * nout * npix);
    DC_LAUNCH(ctx, head_kernel, (unsigned)dc_cdiv(n, 256), 256, 0, raw, (int64_t)B * nout, npix, gamma, ms);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

int32_t k3_loss_grid(dincae_ctx* ctx, const float* raw, const float* xtrue, const float* ms, int B, int nout, int W, int H,
                     const HeadParams& hp, const float* inv_n, float weight, double* loss_acc, float* graw,
                     void* scratch, size_t scratch_bytes) {
    int64_t npix = (int64_t)W * H;
    int64_t n = (int64_t)B * nout * npix;
    if (n == 0) return DINCAE_OK;
    DC_REQUIRE(scratch_bytes >= k3_scratch_bytes(n), DINCAE_ESIZE, "k3_loss_grid: scratch too small");
    double* part = (double*)scratch;
    int nb = red_blocks(ctx, n);
    bool lap = hp.lap != 0.f || hp.lap_err != 0.f;
    // algorithmic bytes: raw (m, log 1/s2) + truth (2 channels) read, gradient (2 channels) written when training
    const double lbytes = 4.0 * n * (4 + (graw ? 2 : 0));
    if (!lap) {
        DC_WORK(ctx, 0.0, lbytes);
        DC_LAUNCH(ctx, loss_grid_kernel<true>, nb, RED_THREADS, 0, raw, xtrue, B, nout, npix, hp, inv_n, weight, part, graw);
        DC_CHECK_LAUNCH();
        DC_LAUNCH(ctx, final_sum_kernel, 1, 256, 0, part, nb, loss_acc);
        DC_CHECK_LAUNCH();
        return DINCAE_OK;
    }
    DC_REQUIRE(ms != nullptr, DINCAE_EINVAL, "k3_loss_grid: Laplacian penalty needs head(raw)");
    // (gm, gs) into graw's storage first, then the Laplacian + chain rule in place
    DC_WORK(ctx, 0.0, lbytes);
    DC_LAUNCH(ctx, loss_grid_kernel<false>, nb, RED_THREADS, 0, raw, xtrue, B, nout, npix, hp, inv_n, weight, part, graw);
    DC_CHECK_LAUNCH();
    DC_LAUNCH(ctx, final_sum_kernel, 1, 256, 0, part, nb, loss_acc);
    DC_CHECK_LAUNCH();
    return k3_head_bwd(ctx, raw, ms, graw, B, nout, W, H, hp, weight, loss_acc, graw, scratch, scratch_bytes);
}

int32_t k3_loss_points(dincae_ctx* ctx, const float* mv, const float* x, int64_t P, const HeadParams& hp, const float* inv_n,
                       float weight, double* loss_acc, float* gms, void* scratch, size_t scratch_bytes) {
    if (P == 0) return DINCAE_OK;
    DC_REQUIRE(scratch_bytes >= k3_scratch_bytes(P), DINCAE_ESIZE, "k3_loss_points: scratch too small");
    double* part = (double*)scratch;
    int nb = red_blocks(ctx, P);
    DC_LAUNCH(ctx, loss_points_kernel, nb, RED_THREADS, 0, mv, x, P, hp, inv_n, weight, part, gms);
    DC_CHECK_LAUNCH();
    DC_LAUNCH(ctx, final_sum_kernel, 1, 256, 0, part, nb, loss_acc);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

int32_t k3_head_bwd(dincae_ctx* ctx, const float* raw, const float* ms, const float* g_ms, int B, int nout, int W, int H,
                    const HeadParams& hp, float weight, double* loss_acc, float* graw, void* scratch, size_t scratch_bytes) {
    int64_t n = (int64_t)B * nout * W * H;
    if (n == 0) return DINCAE_OK;
    DC_REQUIRE(scratch_bytes >= k3_scratch_bytes(n), DINCAE_ESIZE, "k3_head_bwd: scratch too small");
    double* part = (double*)scratch;
    int nb = red_blocks(ctx, n);
    // in-place use (g_ms == graw) is safe: each thread reads its own two elements before writing them
    DC_WORK(ctx, 0.0, 4.0 * n * (2 + 2 + (g_ms ? 2 : 0) + (graw ? 2 : 0)));       // raw, ms (with stencil re-reads served by L2), g_ms in, graw out
    DC_LAUNCH(ctx, head_bwd_kernel, nb, RED_THREADS, 0, raw, ms, g_ms, B, nout, W, H, hp, weight, part, graw);
    DC_CHECK_LAUNCH();
    DC_LAUNCH(ctx, final_sum_kernel, 1, 256, 0, part, nb, loss_acc);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

int32_t k3_adam(dincae_ctx* ctx, float* w, const float* g, float* m, float* v, int64_t n, const AdamParams& p,
                double* reg_acc, void* scratch, size_t scratch_bytes) {
    if (n == 0) return DINCAE_OK;
    DC_REQUIRE(scratch_bytes >= k3_scratch_bytes(n), DINCAE_ESIZE, "k3_adam: scratch too small");
    double* part = (double*)scratch;
    int nb = red_blocks(ctx, n);
    DC_WORK(ctx, 0.0, 28.0 * n);              // w, m, v read + written, g read
    DC_LAUNCH(ctx, adam_kernel<true>, nb, RED_THREADS, 0, w, g, m, v, n, p, part);
    DC_CHECK_LAUNCH();
    if (reg_acc && (p.l1 != 0.f || p.l2 != 0.f)) {
        DC_LAUNCH(ctx, final_sum_kernel, 1, 256, 0, part, nb, reg_acc);
        DC_CHECK_LAUNCH();
    }
    return DINCAE_OK;
}

// ---- CUDA-graph support: the optimiser kernel's AdamParams argument (learning rate, bias corrections) changes every step ------
cudaGraphNode_t k3_find_adam_node(cudaGraph_t graph) {
    size_t n = 0;
    if (cudaGraphGetNodes(graph, nullptr, &n) != cudaSuccess || n == 0) return nullptr;
    std::vector<cudaGraphNode_t> nodes(n);
    if (cudaGraphGetNodes(graph, nodes.data(), &n) != cudaSuccess) return nullptr;
    cudaGraphNode_t found = nullptr;
    for (size_t i = 0; i < n; i++) {
        cudaGraphNodeType ty;
        if (cudaGraphNodeGetType(nodes[i], &ty) != cudaSuccess || ty != cudaGraphNodeTypeKernel) continue;
        cudaKernelNodeParams kp;
        if (cudaGraphKernelNodeGetParams(nodes[i], &kp) != cudaSuccess) continue;
        if (kp.func == (void*)adam_kernel<true>) {
            if (found) return nullptr;               // more than one: not a step graph this code knows
            found = nodes[i];
        }
    }
    return found;
}
int32_t k3_update_adam_node(cudaGraphExec_t exec, cudaGraphNode_t node, const AdamParams& p) {
    cudaKernelNodeParams kp;
    DC_CHECK_CUDA(cudaGraphKernelNodeGetParams(node, &kp));
    void* args[7];
    for (int i = 0; i < 7; i++) args[i] = kp.kernelParams[i];     // (w, g, m, v, n, p, part): everything but p stays as captured
    AdamParams copy = p;
    args[5] = &copy;
    kp.kernelParams = args;
    kp.extra = nullptr;
    DC_CHECK_CUDA(cudaGraphExecKernelNodeSetParams(exec, node, &kp));
    return DINCAE_OK;
}

namespace {
__global__ void __launch_bounds__(256) max_abs_diff_kernel(const float* __restrict__ a, const float* __restrict__ b, int64_t n, unsigned int* __restrict__ out) {
    float mx = 0.f;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float d = fabsf(a[i] - b[i]);
        mx = (d > mx || d != d) ? (d != d ? __int_as_float(0x7f800000) : d) : mx;      // NaN counts as +inf
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_down_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) atomicMax(out, __float_as_uint(mx));
}
}  // namespace
int32_t k3_max_abs_diff(dincae_ctx* ctx, const float* a, const float* b, int64_t n, float* out) {
    DC_CHECK_CUDA(cudaMemsetAsync(out, 0, sizeof(float), ctx->stream));
    if (n == 0) return DINCAE_OK;
    DC_LAUNCH(ctx, max_abs_diff_kernel, 2 * ctx->num_sms, 256, 0, a, b, n, (unsigned int*)out);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

int32_t k3_reg(dincae_ctx* ctx, const float* w, int64_t n, float l1, float l2, double* reg_acc, void* scratch, size_t scratch_bytes) {
    if (n == 0 || (l1 == 0.f && l2 == 0.f)) return DINCAE_OK;
    DC_REQUIRE(scratch_bytes >= k3_scratch_bytes(n), DINCAE_ESIZE, "k3_reg: scratch too small");
    double* part = (double*)scratch;
    int nb = red_blocks(ctx, n);
    AdamParams p{};
    p.l1 = l1;
    p.l2 = l2;
    DC_LAUNCH(ctx, adam_kernel<false>, nb, RED_THREADS, 0, const_cast<float*>(w), nullptr, nullptr, nullptr, n, p, part);
    DC_CHECK_LAUNCH();
    DC_LAUNCH(ctx, final_sum_kernel, 1, 256, 0, part, nb, reg_acc);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

int32_t k3_savesample(dincae_ctx* ctx, const float* ms, int B, int nout, int64_t npix, const float* meandata_out,
                      const uint8_t* landmask, int count, float fill_value, float* m_out, float* sigma_out, int64_t var_stride) {
    int64_t n = (int64_t)B * nout * npix;
    if (n == 0) return DINCAE_OK;
    DC_WORK(ctx, 0.0, (8.0 + (count > 1 ? 16.0 : 8.0)) * B * nout * npix);     // (m, s2) in; running mean read (count > 1) + written
    DC_LAUNCH(ctx, savesample_kernel, (unsigned)dc_cdiv(n, 256), 256, 0, ms, B, nout, npix, meandata_out, landmask, count, fill_value,
              m_out, sigma_out, var_stride);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}


// ---- 2-D vector (HF-radar radial currents) cost, src/vector2.jl:13-100 ------------------------------------------------
namespace {
// xrec (W,H,5,B) = (u', v', L11, L21, L22): P = L L^T is the error covariance of (u, v) = P (u', v')  (vector2_covariance / _mean,
// :13-51); every site s sees the projection on its bearing (:27-44) m = sin(t) u + cos(t) v, s2 = max(p' P p, 1e-4) (:78) and
// is scored against xtrue (W,H,2*nsites,B) = (m_true / s2_true, 1 / s2_true) per site with costfun_single (src/model.jl:54-100).
// One thread = one pixel of one sample: the five inputs are read once, the nsites terms and their gradient summed in site order.
__global__ void __launch_bounds__(RED_THREADS) vector2_count_kernel(const float* __restrict__ xtrue, int64_t npix, int nsites, int B,
                                                                    unsigned long long* __restrict__ n_valid) {
    const int64_t total = (int64_t)B * nsites * npix;
    unsigned int c = 0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t sb = i / npix, p = i - sb * npix;          // sb = b * nsites + s
        c += (xtrue[(2 * sb + 1) * npix + p] != 0.f) ? 1u : 0u;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(n_valid, (unsigned long long)c);   // integer atomics: exact, order-free
}

__global__ void __launch_bounds__(RED_THREADS) vector2_loss_kernel(const float* __restrict__ xrec, const float* __restrict__ xtrue,
                                                                   const float* __restrict__ dir, int64_t npix, int nsites, int B,
                                                                   int dir_has_batch, int truth_uncertain,
                                                                   const unsigned long long* __restrict__ n_valid, double* __restrict__ part,
                                                                   float* __restrict__ grad) {
    const int64_t total = (int64_t)B * npix;
    const float inv_n = 1.0f / (float)n_valid[0];
    double acc = 0.0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t b = i / npix, p = i - b * npix;
        const float* xr = xrec + (b * 5) * npix + p;
        const float x1 = xr[0], x2 = xr[npix], l11 = xr[2 * npix], l21 = xr[3 * npix], l22 = xr[4 * npix];
        const float p11 = l11 * l11, p12 = l11 * l21, p22 = l21 * l21 + l22 * l22;                  // :19-21
        const float u = x1 * p11 + x2 * p12, v = x1 * p12 + x2 * p22;                                // :47-48
        float gx1 = 0.f, gx2 = 0.f, gp11 = 0.f, gp12 = 0.f, gp22 = 0.f;
        for (int s = 0; s < nsites; s++) {
            const float d = dir[((dir_has_batch ? b * nsites : 0) + s) * npix + p];
            const float t = d * 3.14159274f / 180.f;                                                  // directionobs * pi / 180 in Float32, :28
            const float sn = sinf(t), cs = cosf(t);
            const float m = sn * u + cs * v;                                                          // :36-37
            const float s2raw = sn * sn * p11 + 2.f * sn * cs * p12 + cs * cs * p22;                  // :39-41
            const float s2 = fmaxf(s2raw, 1e-4f);                                                     // :78
            const float* xt = xtrue + ((b * nsites + s) * 2) * npix + p;
            const float inv_true = xt[npix];
            if (inv_true == 0.f) continue;                                                            // mask_noncloud, :90
            const float s2t = 1.0f / fmaxf(inv_true, 1e-3f);                                          // sinv, src/model.jl:7-10
            const float mt = xt[0] * s2t;                                                             // :86
            const float df = m - mt;
            float term, num;
            if (truth_uncertain) { term = logf(s2 / s2t) + (df * df + s2t) / s2; num = df * df + s2t; }   // src/model.jl:80-88
            else { term = logf(s2) + df * df / s2; num = df * df; }                                       // :90-94
            acc += (double)term;
            if (grad) {
                const float gm = 2.f * df / s2;
                const float gs = s2raw > 1e-4f ? (1.f / s2 - num / (s2 * s2)) : 0.f;
                // m = sn u + cs v with u, v linear in (x1, x2) and in P
                gx1 += gm * (sn * p11 + cs * p12);
                gx2 += gm * (sn * p12 + cs * p22);
                gp11 += gm * sn * x1 + gs * sn * sn;
                gp12 += gm * (sn * x2 + cs * x1) + gs * 2.f * sn * cs;
                gp22 += gm * cs * x2 + gs * cs * cs;
            }
        }
        if (grad) {
            float* g = grad + (b * 5) * npix + p;
            g[0] = gx1 * inv_n;
            g[npix] = gx2 * inv_n;
            g[2 * npix] = (2.f * l11 * gp11 + l21 * gp12) * inv_n;         // P11 = L11^2, P12 = L11 L21
            g[3 * npix] = (l11 * gp12 + 2.f * l21 * gp22) * inv_n;         // P12, P22 = L21^2 + L22^2
            g[4 * npix] = (2.f * l22 * gp22) * inv_n;
        }
    }
    const double sblk = block_sum(acc);
    if (threadIdx.x == 0) part[blockIdx.x] = sblk;
}
__global__ void vector2_finish_kernel(const double* __restrict__ sum, const unsigned long long* __restrict__ n_valid, double* __restrict__ out) {
    out[0] = sum[0] / (double)n_valid[0];
}
}  // namespace

extern "C" int32_t dincae_vector2_loss(dincae_ctx* ctx, const float* xrec, const float* xtrue, const float* directionobs, int32_t W, int32_t H,
                                       int32_t nsites, int32_t B, int32_t dir_has_batch, int32_t truth_uncertain, int32_t loc, double* loss,
                                       float* grad) {
    DC_REQUIRE(ctx && xrec && xtrue && directionobs && loss, DINCAE_EINVAL, "dincae_vector2_loss: NULL argument");
    DC_REQUIRE(W >= 1 && H >= 1 && nsites >= 1 && B >= 1, DINCAE_EINVAL, "dincae_vector2_loss: bad sizes W=%d H=%d nsites=%d B=%d", W, H, nsites, B);
    DC_REQUIRE(loc == DINCAE_HOST || loc == DINCAE_DEVICE, DINCAE_EINVAL, "dincae_vector2_loss: loc must be DINCAE_HOST or DINCAE_DEVICE");
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    const int64_t npix = (int64_t)W * H;
    const size_t nrec = sizeof(float) * 5 * npix * B, ntrue = sizeof(float) * 2 * nsites * npix * B;
    const size_t ndir = sizeof(float) * nsites * npix * (dir_has_batch ? B : 1);
    DevIn drec, dtrue, ddir;
    DC_TRY(dc_stage_in(ctx, xrec, nrec, loc, &drec));
    DC_TRY(dc_stage_in(ctx, xtrue, ntrue, loc, &dtrue));
    DC_TRY(dc_stage_in(ctx, directionobs, ndir, loc, &ddir));
    // scratch: block partials, [sum, cost] doubles, the valid count; + the gradient when it goes back to the host
    const size_t head = (RED_MAX_BLOCKS + 4) * sizeof(double);
    void* sc = nullptr;
    DC_TRY(dc_scratch(ctx, head + ((grad && loc == DINCAE_HOST) ? nrec : 0), &sc));
    double* part = (double*)sc;
    double* sums = part + RED_MAX_BLOCKS;                                   // [0] = sum of terms, [1] = cost
    unsigned long long* nv = (unsigned long long*)(sums + 2);
    float* dgrad = grad ? (loc == DINCAE_HOST ? (float*)((char*)sc + head) : grad) : nullptr;
    DC_CHECK_CUDA(cudaMemsetAsync(sums, 0, 4 * sizeof(double), ctx->stream));
    const int nb = red_blocks(ctx, (int64_t)B * npix);
    DC_LAUNCH(ctx, vector2_count_kernel, red_blocks(ctx, (int64_t)B * nsites * npix), RED_THREADS, 0, (const float*)dtrue.ptr, npix, nsites, B, nv);
    DC_CHECK_LAUNCH();
    DC_WORK(ctx, 0.0, (double)nrec * (grad ? 2 : 1) + ntrue + (double)sizeof(float) * nsites * npix * B);
    DC_LAUNCH(ctx, vector2_loss_kernel, nb, RED_THREADS, 0, (const float*)drec.ptr, (const float*)dtrue.ptr, (const float*)ddir.ptr, npix, nsites, B,
              dir_has_batch ? 1 : 0, truth_uncertain ? 1 : 0, nv, part, dgrad);
    DC_CHECK_LAUNCH();
    DC_LAUNCH(ctx, final_sum_kernel, 1, 256, 0, part, nb, sums);
    DC_CHECK_LAUNCH();
    DC_LAUNCH(ctx, vector2_finish_kernel, 1, 1, 0, sums, nv, sums + 1);
    DC_CHECK_LAUNCH();
    void* pin = nullptr;
    DC_TRY(dc_pinned(ctx, 64, &pin));
    DC_CHECK_CUDA(cudaMemcpyAsync(pin, sums + 1, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (grad && loc == DINCAE_HOST) DC_CHECK_CUDA(cudaMemcpyAsync(grad, dgrad, nrec, cudaMemcpyDeviceToHost, ctx->stream));
    DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    dc_stage_free(ctx, &drec);
    dc_stage_free(ctx, &dtrue);
    dc_stage_free(ctx, &ddir);
    *loss = *(double*)pin;
    DC_REQUIRE(*loss == *loss, DINCAE_EINVAL, "cost function value is NaN");     // src/vector2.jl:95-97
    return DINCAE_OK;
}
