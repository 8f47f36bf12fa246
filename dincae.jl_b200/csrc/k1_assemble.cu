// K1: gridded dataset precompute + batch assembly kernels (HBM-bound streaming; see k1_assemble.cuh).
//
// HBM layout: the cube is stored once as d = data - mean with NaN marking missing values; the
// reference's two channels x[...,1] = d/sigma^2 (NaN->0) and x[...,2] = (1-isnan)/sigma^2
// (src/data.jl:313-327) are derived on the fly, halving the resident footprint.
// Algorithmic bytes per sample: read 4*T*V*W*H (+ 4*W*H mask step) ; write 4*(4+2TV+2O)*W*H.
#include "k1_assemble.cuh"

namespace {

// mean over time ignoring NaN, sequential Float32 sum like sum(...,dims=4)./sum(.!isnan,dims=4) (src/data.jl:270)
__global__ void mean_kernel(const float* __restrict__ data, int64_t VHW, int64_t N, float* __restrict__ mean) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= VHW) return;
    float s = 0.f, c = 0.f;
    for (int64_t n = 0; n < N; n++) {
        float v = data[n * VHW + i];
        if (v == v) { s += v; c += 1.f; }
    }
    mean[i] = s / c;     // 0/0 = NaN where never observed, as in the reference
}

__global__ void subtract_mean_kernel(float* __restrict__ data, int64_t VHW, int64_t total, const float* __restrict__ mean) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x)
        data[i] = data[i] - mean[i % VHW];      // data = data_full .- meandata, src/data.jl:299
}

__global__ void valid_count_kernel(const float* __restrict__ data, const float* __restrict__ err2, int64_t HW, int V, int v,
                                   unsigned long long* __restrict__ counts) {
    // block n counts the targets of variable v at time step n whose weight channel (1 - isnan) / err^2 is non-zero
    int64_t n = blockIdx.x;
    const float* p = data + (n * V + v) * HW;
    const float* q = err2 ? err2 + (n * V + v) * HW : nullptr;
    unsigned int c = 0;
    for (int64_t i = threadIdx.x; i < HW; i += blockDim.x) c += (p[i] == p[i] && (!q || 1.0f / q[i] != 0.f)) ? 1u : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&counts[n], (unsigned long long)c);
}

constexpr int ASM_MAXB = 128;
struct AsmArgs {
    const float* data;
    const float* err2;       // per-pixel error variance cube or NULL (then params[v] applies to the whole variable)
    const float* lon;
    const float* lat;
    const float* tcos;
    const float* tsin;
    const float* params;     // [V] obs_err_std^2 then [V] jitter_std
    const int* outvars;
    int tidx[ASM_MAXB];      // time_idx of this launch's samples (kernel parameters: no staging copy, no sync)
    int imk[ASM_MAXB];       // imask
    int has_mask;
    int W, H, V, T, nout, B;
    int64_t N;
    uint64_t seed, ctr0;
    float* xin;              // planar [B][4+2TV][H][W]
    float* xtrue;            // planar [B][2*nout][H][W]
};

// one thread = 4 consecutive pixels of one sample (one Philox call gives their 4 normals)
__global__ void __launch_bounds__(256) assemble_kernel(AsmArgs a) {
    const int64_t HW = (int64_t)a.W * a.H;
    const int64_t groups = (HW + 3) / 4;
    int64_t gi = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (gi >= groups * a.B) return;
    const int b = (int)(gi / groups);
    const int64_t g = gi - (int64_t)b * groups;
    const int64_t p0 = g * 4;
    const int np = (int)((HW - p0) < 4 ? (HW - p0) : 4);
    const int C = 4 + 2 * a.T * a.V;
    const int64_t ind = a.tidx[b];
    const int half = (a.T - 1) / 2, centraln = (a.T + 1) / 2 - 1;
    const bool perturb = a.has_mask != 0;
    const int64_t im = perturb ? a.imk[b] : 0;
    const bool vec = (HW & 3) == 0;
    float* xo = a.xin + (size_t)b * C * HW + p0;

    auto store = [&](float* dst, const float v[4]) {
        if (vec) *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
        else for (int e = 0; e < np; e++) dst[e] = v[e];
    };
    auto load = [&](const float* src, float v[4]) {
        if (vec) { float4 t = __ldg(reinterpret_cast<const float4*>(src)); v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w; }
        else for (int e = 0; e < 4; e++) v[e] = e < np ? __ldg(src + e) : 0.f;
    };

    float v4[4];
    // metadata channels: lon, lat, cos, sin (src/data.jl:409-419)
    for (int e = 0; e < 4; e++) { int64_t p = p0 + e; v4[e] = e < np ? a.lon[p % a.W] : 0.f; }
    store(xo, v4);
    for (int e = 0; e < 4; e++) { int64_t p = p0 + e; v4[e] = e < np ? a.lat[p / a.W] : 0.f; }
    store(xo + HW, v4);
    float tc = a.tcos[ind], ts = a.tsin[ind];
    for (int e = 0; e < 4; e++) v4[e] = tc;
    store(xo + 2 * HW, v4);
    for (int e = 0; e < 4; e++) v4[e] = ts;
    store(xo + 3 * HW, v4);

    // transplanted cloud mask: missingmask[:,:,imask] = isnan of variable 1 at step imask (src/data.jl:91,471-488)
    bool mm[4] = {false, false, false, false};
    if (perturb) {
        float t4[4];
        load(a.data + ((size_t)im * a.V + 0) * HW + p0, t4);
        for (int e = 0; e < 4; e++) mm[e] = t4[e] != t4[e];
    }
    int off = 4;
    for (int v = 0; v < a.V; v++) {
        const float s2 = a.params[v], jit = a.params[a.V + v];
        for (int k = 0; k < a.T; k++) {
            int64_t n = ind - half + k;                       // nrange = clamp(ind-h:ind+h, 1, N), src/data.jl:405
            n = n < 0 ? 0 : (n > a.N - 1 ? a.N - 1 : n);
            float d4[4], x1[4], x2[4], e4[4] = {s2, s2, s2, s2};
            load(a.data + ((size_t)n * a.V + v) * HW + p0, d4);
            if (a.err2) load(a.err2 + ((size_t)n * a.V + v) * HW + p0, e4);      // Inf where the error field is missing: both channels 0
            for (int e = 0; e < 4; e++) {
                bool nan = d4[e] != d4[e];
                x1[e] = (nan ? 0.f : d4[e]) / e4[e];
                x2[e] = nan ? 0.f : 1.0f / e4[e];
            }
            if (perturb) {
                if (k == centraln)
                    for (int e = 0; e < 4; e++)
                        if (mm[e]) { x1[e] = 0.f; x2[e] = 0.f; }      // zero FIRST (all channels of the centre step)
                float z[4];
                uint64_t sc = a.ctr0 + (uint64_t)b;
                philox_normal4((uint32_t)g, (uint32_t)(v * a.T + k), (uint32_t)sc, (uint32_t)(sc >> 32) & 0x7FFFFFFFu, a.seed, z);
                for (int e = 0; e < 4; e++) x1[e] += jit * z[e];       // noise AFTER, on every pixel
            }
            store(xo + (size_t)off * HW, x1);
            store(xo + (size_t)(off + 1) * HW, x2);
            off += 2;
        }
    }
    // xtrue: untouched (x1, x2) of the output variables at `ind` (src/data.jl:504-514)
    float* yo = a.xtrue + (size_t)b * 2 * a.nout * HW + p0;
    for (int o = 0; o < a.nout; o++) {
        int v = a.outvars[o];
        const float s2 = a.params[v];
        float d4[4], x1[4], x2[4], e4[4] = {s2, s2, s2, s2};
        load(a.data + ((size_t)ind * a.V + v) * HW + p0, d4);
        if (a.err2) load(a.err2 + ((size_t)ind * a.V + v) * HW + p0, e4);
        for (int e = 0; e < 4; e++) {
            bool nan = d4[e] != d4[e];
            x1[e] = (nan ? 0.f : d4[e]) / e4[e];
            x2[e] = nan ? 0.f : 1.0f / e4[e];
        }
        store(yo + (size_t)(2 * o) * HW, x1);
        store(yo + (size_t)(2 * o + 1) * HW, x2);
    }
}

}  // namespace

int32_t k1_check_indices(const dincae_gridded* ds, const int64_t* h_time_idx, const int64_t* h_imask, int B) {
    for (int b = 0; b < B; b++) {
        DC_REQUIRE(h_time_idx[b] >= 0 && h_time_idx[b] < ds->N, DINCAE_EINVAL, "assemble: time_idx[%d]=%lld out of [0,%lld)", b,
                   (long long)h_time_idx[b], (long long)ds->N);
        if (h_imask)
            DC_REQUIRE(h_imask[b] >= 0 && h_imask[b] < ds->N, DINCAE_EINVAL, "assemble: imask[%d]=%lld out of range", b, (long long)h_imask[b]);
    }
    DC_REQUIRE(ds->N < 0x7FFFFFFF, DINCAE_ESIZE, "assemble: more than 2^31 time steps");
    return DINCAE_OK;
}

// validates the indices of samples [b0, b0 + nb) and fills the kernel arguments of their launch
static int32_t asm_fill(dincae_gridded* ds, const int64_t* h_time_idx, const int64_t* h_imask, uint64_t seed, uint64_t ctr0,
                        int b0, int nb, float* d_xin, float* d_xtrue, AsmArgs& a) {
    DC_TRY(k1_check_indices(ds, h_time_idx + b0, h_imask ? h_imask + b0 : nullptr, nb));
    const size_t HW = (size_t)ds->W * ds->H;
    const int C = 4 + 2 * ds->T * ds->V;
    a.data = ds->d_data; a.err2 = ds->d_err2; a.lon = ds->d_lon; a.lat = ds->d_lat; a.tcos = ds->d_tcos; a.tsin = ds->d_tsin;
    a.params = ds->d_params; a.outvars = ds->d_outvars;
    for (int b = 0; b < nb; b++) { a.tidx[b] = (int)h_time_idx[b0 + b]; a.imk[b] = h_imask ? (int)h_imask[b0 + b] : 0; }
    a.has_mask = h_imask != nullptr;
    a.W = ds->W; a.H = ds->H; a.V = ds->V; a.T = ds->T; a.nout = ds->nout; a.B = nb; a.N = ds->N;
    a.seed = seed; a.ctr0 = ctr0 + (uint64_t)b0;
    a.xin = d_xin + (size_t)b0 * C * HW;
    a.xtrue = d_xtrue + (size_t)b0 * 2 * ds->nout * HW;
    return DINCAE_OK;
}

int32_t k1_assemble(dincae_gridded* ds, const int64_t* h_time_idx, const int64_t* h_imask, uint64_t seed, uint64_t ctr0,
                    int B, float* d_xin, float* d_xtrue) {
    dincae_ctx* ctx = ds->ctx;
    if (B == 0) return DINCAE_OK;
    const size_t HW = (size_t)ds->W * ds->H;
    const int C = 4 + 2 * ds->T * ds->V;
    for (int b0 = 0; b0 < B; b0 += ASM_MAXB) {
        int nb = B - b0 < ASM_MAXB ? B - b0 : ASM_MAXB;
        AsmArgs a;
        DC_TRY(asm_fill(ds, h_time_idx, h_imask, seed, ctr0, b0, nb, d_xin, d_xtrue, a));
        int64_t groups = ((int64_t)HW + 3) / 4 * nb;
        // algorithmic bytes per sample: T*V cube planes (+ the transplanted mask plane) read, C input + 2*nout target planes written
        DC_WORK(ctx, 0.0, 4.0 * HW * nb * ((double)ds->T * ds->V + (h_imask ? 1 : 0) + C + 2 * ds->nout));
        DC_LAUNCH(ctx, assemble_kernel, (unsigned)dc_cdiv(groups, 256), 256, 0, a);
    }
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

// ---- CUDA-graph support (dincae_train_step_gridded): the sample indices of the assembly launch change every step ---------------
int k1_assemble_max_single_launch() { return ASM_MAXB; }
cudaGraphNode_t k1_find_assemble_node(cudaGraph_t graph) {
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
        if (kp.func == (void*)assemble_kernel) {
            if (found) return nullptr;
            found = nodes[i];
        }
    }
    return found;
}
int32_t k1_update_assemble_node(cudaGraphExec_t exec, cudaGraphNode_t node, dincae_gridded* ds, const int64_t* h_time_idx, const int64_t* h_imask,
                                uint64_t seed, uint64_t ctr0, int B, float* d_xin, float* d_xtrue) {
    DC_REQUIRE(B >= 1 && B <= ASM_MAXB, DINCAE_ESIZE, "assemble node: batch %d outside [1,%d]", B, ASM_MAXB);
    AsmArgs a;
    DC_TRY(asm_fill(ds, h_time_idx, h_imask, seed, ctr0, 0, B, d_xin, d_xtrue, a));
    cudaKernelNodeParams kp;
    DC_CHECK_CUDA(cudaGraphKernelNodeGetParams(node, &kp));      // grid and block stay as captured (same B, same grid size)
    void* args[1] = {&a};
    kp.kernelParams = args;
    kp.extra = nullptr;
    DC_CHECK_CUDA(cudaGraphExecKernelNodeSetParams(exec, node, &kp));
    return DINCAE_OK;
}

// static per-step valid counts of every output variable (n_noncloud without a collective)
static int32_t compute_valid_counts(dincae_gridded* ds) {
    dincae_ctx* ctx = ds->ctx;
    const int64_t HW = (int64_t)ds->W * ds->H, N = ds->N;
    unsigned long long* d_counts = nullptr;
    DC_CHECK_CUDA(cudaMalloc((void**)&d_counts, sizeof(unsigned long long) * N));
    ds->valid_counts.resize(ds->nout);
    std::vector<unsigned long long> hc(N);
    for (int o = 0; o < ds->nout; o++) {
        DC_CHECK_CUDA(cudaMemsetAsync(d_counts, 0, sizeof(unsigned long long) * N, ctx->stream));
        DC_LAUNCH(ctx, valid_count_kernel, (unsigned)N, 256, 0, ds->d_data, ds->d_err2, HW, ds->V, ds->outvars[o], d_counts);
        DC_CHECK_LAUNCH();
        DC_CHECK_CUDA(cudaMemcpyAsync(hc.data(), d_counts, sizeof(unsigned long long) * N, cudaMemcpyDeviceToHost, ctx->stream));
        DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
        ds->valid_counts[o].assign(hc.begin(), hc.end());
    }
    cudaFree(d_counts);
    return DINCAE_OK;
}

extern "C" int32_t dincae_gridded_create(dincae_ctx* ctx, const float* data_full, int32_t W, int32_t H, int32_t V, int64_t N,
                                         const float* obs_err_std, const float* jitter_std, const uint8_t* isoutput,
                                         int32_t remove_mean, const float* mean_data, const uint8_t* landmask,
                                         const float* lon_scaled, const float* lat_scaled,
                                         const float* time_cos, const float* time_sin, int32_t ntime_win,
                                         int32_t loc, dincae_gridded** out) {
    DC_REQUIRE(ctx && data_full && out && lon_scaled && lat_scaled && time_cos && time_sin, DINCAE_EINVAL, "dincae_gridded_create: NULL argument");
    DC_REQUIRE(W > 0 && H > 0 && V > 0 && N > 0, DINCAE_EINVAL, "dincae_gridded_create: bad sizes");
    DC_REQUIRE(ntime_win >= 1 && (ntime_win & 1), DINCAE_EINVAL, "dincae_gridded_create: ntime_win must be odd (src/data.jl:396)");
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    dincae_gridded* ds = new dincae_gridded();
    ds->ctx = ctx; ds->W = W; ds->H = H; ds->V = V; ds->N = N; ds->T = ntime_win;
    const int64_t HW = (int64_t)W * H, VHW = HW * V, total = VHW * N;
    for (int v = 0; v < V; v++)
        if (isoutput ? isoutput[v] != 0 : v == 0) ds->outvars.push_back(v);     // isoutput default: first variable (src/data.jl:79)
    ds->nout = (int)ds->outvars.size();
    DC_REQUIRE(ds->nout >= 1, DINCAE_EINVAL, "dincae_gridded_create: no output variable");
    cudaError_t e = cudaMalloc((void**)&ds->d_data, sizeof(float) * total);
    if (e != cudaSuccess) { dincae_set_error("dincae_gridded_create: cannot allocate %lld bytes for the cube: %s", (long long)total * 4, cudaGetErrorString(e)); delete ds; return DINCAE_ENOMEM; }
    DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_data, data_full, sizeof(float) * total,
                                  loc == DINCAE_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, ctx->stream));
    DC_CHECK_CUDA(cudaMalloc((void**)&ds->d_mean, sizeof(float) * VHW));
    DC_CHECK_CUDA(cudaMalloc((void**)&ds->d_mean_out, sizeof(float) * HW * ds->nout));
    DC_CHECK_CUDA(cudaMalloc((void**)&ds->d_lon, sizeof(float) * W));
    DC_CHECK_CUDA(cudaMalloc((void**)&ds->d_lat, sizeof(float) * H));
    DC_CHECK_CUDA(cudaMalloc((void**)&ds->d_tcos, sizeof(float) * N));
    DC_CHECK_CUDA(cudaMalloc((void**)&ds->d_tsin, sizeof(float) * N));
    DC_CHECK_CUDA(cudaMalloc((void**)&ds->d_params, sizeof(float) * 2 * V));
    DC_CHECK_CUDA(cudaMalloc((void**)&ds->d_outvars, sizeof(int) * ds->nout));
    DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_lon, lon_scaled, sizeof(float) * W, cudaMemcpyHostToDevice, ctx->stream));
    DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_lat, lat_scaled, sizeof(float) * H, cudaMemcpyHostToDevice, ctx->stream));
    DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_tcos, time_cos, sizeof(float) * N, cudaMemcpyHostToDevice, ctx->stream));
    DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_tsin, time_sin, sizeof(float) * N, cudaMemcpyHostToDevice, ctx->stream));
    std::vector<float> params(2 * V);
    for (int v = 0; v < V; v++) {
        float s = obs_err_std ? obs_err_std[v] : 1.0f;
        params[v] = s * s;                                       // x ./= obs_err_std^2, src/data.jl:321-322
        params[V + v] = jitter_std ? jitter_std[v] : 0.05f;      // default jitter_std, src/data.jl:82
    }
    DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_params, params.data(), sizeof(float) * 2 * V, cudaMemcpyHostToDevice, ctx->stream));
    DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_outvars, ds->outvars.data(), sizeof(int) * ds->nout, cudaMemcpyHostToDevice, ctx->stream));
    if (landmask) {
        DC_CHECK_CUDA(cudaMalloc((void**)&ds->d_landmask, HW));
        DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_landmask, landmask, HW, cudaMemcpyHostToDevice, ctx->stream));
    }
    if (remove_mean) {
        if (mean_data) {
            DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_mean, mean_data, sizeof(float) * VHW, cudaMemcpyHostToDevice, ctx->stream));
        } else {
            DC_LAUNCH(ctx, mean_kernel, (unsigned)dc_cdiv(VHW, 256), 256, 0, ds->d_data, VHW, N, ds->d_mean);
            DC_CHECK_LAUNCH();
        }
        int nb = (int)(dc_cdiv(total, 256) < 65535 ? dc_cdiv(total, 256) : 65535);
        DC_LAUNCH(ctx, subtract_mean_kernel, nb, 256, 0, ds->d_data, VHW, total, ds->d_mean);
        DC_CHECK_LAUNCH();
    } else {
        DC_CHECK_CUDA(cudaMemsetAsync(ds->d_mean, 0, sizeof(float) * VHW, ctx->stream));   // zeros, src/data.jl:275
    }
    for (int o = 0; o < ds->nout; o++)
        DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_mean_out + (size_t)o * HW, ds->d_mean + (size_t)ds->outvars[o] * HW, sizeof(float) * HW,
                                      cudaMemcpyDeviceToDevice, ctx->stream));
    DC_TRY(compute_valid_counts(ds));
    DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    *out = ds;
    return DINCAE_OK;
}

extern "C" int32_t dincae_gridded_destroy(dincae_gridded* ds) {
    if (!ds) return DINCAE_OK;
    cudaSetDevice(ds->ctx->device);
    cudaStreamSynchronize(ds->ctx->stream);
    void* ptrs[] = {ds->d_data, ds->d_err2, ds->d_mean, ds->d_mean_out, ds->d_lon, ds->d_lat, ds->d_tcos, ds->d_tsin, ds->d_landmask,
                    ds->d_params, ds->d_outvars, ds->d_idx};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    delete ds;
    return DINCAE_OK;
}

namespace {
__global__ void square_kernel(float* __restrict__ x, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) x[i] = x[i] * x[i];
}
}  // namespace

extern "C" int32_t dincae_gridded_set_error(dincae_gridded* ds, const float* obs_err_std, int32_t loc) {
    DC_REQUIRE(ds && obs_err_std, DINCAE_EINVAL, "dincae_gridded_set_error: NULL argument");
    dincae_ctx* ctx = ds->ctx;
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    const int64_t total = (int64_t)ds->W * ds->H * ds->V * ds->N;
    if (!ds->d_err2) {
        cudaError_t e = cudaMalloc((void**)&ds->d_err2, sizeof(float) * total);
        if (e != cudaSuccess) { dincae_set_error("dincae_gridded_set_error: cannot allocate %lld bytes: %s", (long long)total * 4, cudaGetErrorString(e)); ds->d_err2 = nullptr; return DINCAE_ENOMEM; }
    }
    DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_err2, obs_err_std, sizeof(float) * total, loc == DINCAE_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, ctx->stream));
    int nb = (int)(dc_cdiv(total, 256) < 65535 ? dc_cdiv(total, 256) : 65535);
    DC_LAUNCH(ctx, square_kernel, nb, 256, 0, ds->d_err2, total);          // x ./= obs_err_std.^2, src/data.jl:324-325
    DC_CHECK_LAUNCH();
    DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    return compute_valid_counts(ds);       // a target whose error is infinite carries no weight: it no longer counts
}

extern "C" int32_t dincae_gridded_get_mean(dincae_gridded* ds, float* host_mean) {
    DC_REQUIRE(ds && host_mean, DINCAE_EINVAL, "dincae_gridded_get_mean: NULL argument");
    DC_CHECK_CUDA(cudaSetDevice(ds->ctx->device));
    DC_CHECK_CUDA(cudaMemcpyAsync(host_mean, ds->d_mean, sizeof(float) * (size_t)ds->W * ds->H * ds->V, cudaMemcpyDeviceToHost, ds->ctx->stream));
    DC_CHECK_CUDA(cudaStreamSynchronize(ds->ctx->stream));
    return DINCAE_OK;
}

extern "C" int32_t dincae_gridded_valid_counts(dincae_gridded* ds, int32_t v, int64_t* host_counts) {
    DC_REQUIRE(ds && host_counts, DINCAE_EINVAL, "dincae_gridded_valid_counts: NULL argument");
    DC_REQUIRE(v >= 0 && v < ds->nout, DINCAE_EINVAL, "dincae_gridded_valid_counts: output variable %d out of range", v);
    for (int64_t n = 0; n < ds->N; n++) host_counts[n] = ds->valid_counts[v][n];
    return DINCAE_OK;
}

extern "C" int32_t dincae_gridded_assemble(dincae_gridded* ds, const int64_t* time_idx, const int64_t* imask,
                                           uint64_t seed, uint64_t sample_ctr0, int32_t B,
                                           float* xin, float* xtrue, int32_t loc) {
    DC_REQUIRE(ds && time_idx && xin && xtrue, DINCAE_EINVAL, "dincae_gridded_assemble: NULL argument");
    DC_REQUIRE(B >= 0, DINCAE_EINVAL, "dincae_gridded_assemble: B < 0");
    dincae_ctx* ctx = ds->ctx;
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    if (loc == DINCAE_DEVICE) return k1_assemble(ds, time_idx, imask, seed, sample_ctr0, B, xin, xtrue);
    const size_t HW = (size_t)ds->W * ds->H;
    const size_t nx = HW * (4 + 2 * ds->T * ds->V) * B, ny = HW * 2 * ds->nout * B;
    float *dx = nullptr, *dy = nullptr;
    DC_CHECK_CUDA(cudaMallocAsync((void**)&dx, sizeof(float) * (nx + ny), ctx->stream));
    dy = dx + nx;
    int32_t st = k1_assemble(ds, time_idx, imask, seed, sample_ctr0, B, dx, dy);
    if (st == DINCAE_OK) {
        DC_CHECK_CUDA(cudaMemcpyAsync(xin, dx, sizeof(float) * nx, cudaMemcpyDeviceToHost, ctx->stream));
        DC_CHECK_CUDA(cudaMemcpyAsync(xtrue, dy, sizeof(float) * ny, cudaMemcpyDeviceToHost, ctx->stream));
    }
    cudaFreeAsync(dx, ctx->stream);
    DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    return st;
}
