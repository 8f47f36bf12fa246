// Point-cloud dataset and gridding assembly (PointCloud src/points.jl:137-227, posindices! :98-134,
// pertpos :229-235, getxy! :237-313).  The reference grids on the CPU (interp_adjn! on Arrays inside
// getxy!); here the whole mini-batch is gridded by ONE sorted segmented scatter on the device.
#include "common.cuh"
#include "k4_interp.cuh"
#include <vector>

struct dincae_points {
    dincae_ctx* ctx = nullptr;
    int W = 0, H = 0, T = 0, naux = 0;
    int64_t N = 0, Ptot = 0;
    float jit_x = 0.f, jit_y = 0.f;
    float* d_pos = nullptr;     // (Ptot,2)
    float* d_x = nullptr;       // (Ptot,2) column-major
    float* d_aux = nullptr;     // [N][naux][H][W]
    float* d_lon = nullptr;
    float* d_lat = nullptr;
    float* d_cos = nullptr;
    float* d_sin = nullptr;
    std::vector<int64_t> offsets;   // host (N+1)
    // per-call workspace (grown on demand)
    char* d_work = nullptr;
    size_t work_bytes = 0;
};

namespace {

__global__ void posindices_kernel(const float* __restrict__ lon, const float* __restrict__ lat, int64_t P,
                                  double gx0, double gx1, int W, double gy0, double gy1, int H,
                                  float* __restrict__ pos, uint8_t* __restrict__ inside) {
    int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (p >= P) return;
    // pos = length(grid) * (coord - grid[1]) / (grid[end] - grid[1]) + 1   (length, not length-1: SURVEY T15)
    double px = __dadd_rn(__ddiv_rn(__dmul_rn((double)W, __dsub_rn((double)lon[p], gx0)), __dsub_rn(gx1, gx0)), 1.0);
    double py = __dadd_rn(__ddiv_rn(__dmul_rn((double)H, __dsub_rn((double)lat[p], gy0)), __dsub_rn(gy1, gy0)), 1.0);
    inside[p] = (1.0 <= px && px < (double)W && 1.0 <= py && py < (double)H) ? 1 : 0;   // src/points.jl:117
    pos[2 * p] = (float)px;
    pos[2 * p + 1] = (float)py;
}

struct FillArgs {
    const float* lon; const float* lat; const float* dcos; const float* dsin; const float* aux;
    const int* tidx;        // [B] device
    int W, H, naux, T, B;
    int64_t N;
    float* xin;
};
// meta + aux channels, zero the 2T gridded channels (src/points.jl:242-259; xin = zeros, :326)
__global__ void points_fill_kernel(FillArgs a) {
    const int64_t HW = (int64_t)a.W * a.H;
    const int C = 4 + a.naux + 2 * a.T;
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= (int64_t)a.B * C * HW) return;
    int64_t p = i % HW;
    int c = (int)((i / HW) % C);
    int b = (int)(i / (HW * C));
    int ind = a.tidx[b];
    float v = 0.f;
    if (c == 0) v = a.lon[p % a.W];
    else if (c == 1) v = a.lat[p / a.W];
    else if (c == 2) v = a.dcos[ind];
    else if (c == 3) v = a.dsin[ind];
    else if (c < 4 + a.naux) v = a.aux[((size_t)ind * a.naux + (c - 4)) * HW + p];
    a.xin[i] = v;
}

struct GenArgs {
    const float* pos; const float* x;     // dataset arrays
    int64_t Ptot;
    const int64_t* seg_src;   // [B*T] first dataset point of segment (b,k)
    const int64_t* seg_dst;   // [B*T+1] first batch point of segment
    const int64_t* keep_off;  // [B] offset of sample b's keep flags (or NULL)
    const uint8_t* keep;      // concatenated keep flags of the centre steps (or NULL)
    int nseg, T, W, H, perturb;
    float jit_x, jit_y;
    uint64_t seed, ctr0;
    float* out_pos; float* out_v0; float* out_v1; int32_t* out_plane; uint8_t* out_keep;
};
__global__ void points_gen_kernel(GenArgs a, int64_t Pb) {
    int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (q >= Pb) return;
    int lo = 0, hi = a.nseg;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (a.seg_dst[mid] <= q) lo = mid; else hi = mid;
    }
    const int seg = lo, b = seg / a.T, k = seg % a.T;
    const int64_t local = q - a.seg_dst[seg];
    const int64_t src = a.seg_src[seg] + local;
    float px = a.pos[2 * src], py = a.pos[2 * src + 1];
    if (a.perturb) {    // pertpos: clamp(p + jitter_std_pos*randn, 1, sz-1), src/points.jl:229-235
        float z[4];
        uint64_t sc = a.ctr0 + (uint64_t)b;
        philox_normal4((uint32_t)(local >> 1), (uint32_t)k, (uint32_t)sc, ((uint32_t)(sc >> 32)) | 0x80000000u, a.seed, z);
        float zx = (local & 1) ? z[2] : z[0], zy = (local & 1) ? z[3] : z[1];
        px = fminf(fmaxf(__fadd_rn(px, __fmul_rn(a.jit_x, zx)), 1.0f), (float)(a.W - 1));
        py = fminf(fmaxf(__fadd_rn(py, __fmul_rn(a.jit_y, zy)), 1.0f), (float)(a.H - 1));
    }
    a.out_pos[2 * q] = px;
    a.out_pos[2 * q + 1] = py;
    a.out_v0[q] = a.x[src];
    a.out_v1[q] = a.x[a.Ptot + src];
    a.out_plane[q] = seg;
    uint8_t kp = 1;
    if (a.keep && k == (a.T + 1) / 2 - 1) kp = a.keep[a.keep_off[b] + local];   // track dropout on the centre step only, :284-300
    a.out_keep[q] = kp;
}

}  // namespace

extern "C" int32_t dincae_posindices(dincae_ctx* ctx, const float* lon, const float* lat, int64_t P,
                                     double gx0, double gx1, int32_t W, double gy0, double gy1, int32_t H,
                                     float* pos, uint8_t* inside, int32_t loc) {
    DC_REQUIRE(ctx && (P == 0 || (lon && lat && pos && inside)), DINCAE_EINVAL, "dincae_posindices: NULL argument");
    DC_REQUIRE(W >= 2 && H >= 2 && gx1 != gx0 && gy1 != gy0, DINCAE_EINVAL, "dincae_posindices: degenerate grid");
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    if (P == 0) return DINCAE_OK;
    DevIn dlon, dlat;
    DC_TRY(dc_stage_in(ctx, lon, sizeof(float) * P, loc, &dlon));
    DC_TRY(dc_stage_in(ctx, lat, sizeof(float) * P, loc, &dlat));
    float* dpos = pos;
    uint8_t* dins = inside;
    if (loc == DINCAE_HOST) {
        DC_CHECK_CUDA(cudaMallocAsync((void**)&dpos, sizeof(float) * 2 * P + P, ctx->stream));
        dins = (uint8_t*)(dpos + 2 * P);
    }
    DC_LAUNCH(ctx, posindices_kernel, (unsigned)dc_cdiv(P, 256), 256, 0, (const float*)dlon.ptr, (const float*)dlat.ptr, P, gx0, gx1, W, gy0, gy1, H, dpos, dins);
    DC_CHECK_LAUNCH();
    if (loc == DINCAE_HOST) {
        DC_CHECK_CUDA(cudaMemcpyAsync(pos, dpos, sizeof(float) * 2 * P, cudaMemcpyDeviceToHost, ctx->stream));
        DC_CHECK_CUDA(cudaMemcpyAsync(inside, dins, P, cudaMemcpyDeviceToHost, ctx->stream));
        DC_CHECK_CUDA(cudaFreeAsync(dpos, ctx->stream));
        DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    dc_stage_free(ctx, &dlon);
    dc_stage_free(ctx, &dlat);
    return DINCAE_OK;
}

extern "C" int32_t dincae_points_create(dincae_ctx* ctx, int32_t W, int32_t H, int64_t N, const float* pos, const float* x,
                                        const int64_t* offsets, const float* aux, int32_t naux,
                                        const float* lon_scaled, const float* lat_scaled,
                                        const float* doy_cos, const float* doy_sin, int32_t ntime_win,
                                        float jitter_std_x, float jitter_std_y, dincae_points** out) {
    DC_REQUIRE(ctx && offsets && lon_scaled && lat_scaled && doy_cos && doy_sin && out, DINCAE_EINVAL, "dincae_points_create: NULL argument");
    DC_REQUIRE(W >= 2 && H >= 2 && N >= 1 && naux >= 0, DINCAE_EINVAL, "dincae_points_create: bad sizes");
    DC_REQUIRE(ntime_win >= 1 && (ntime_win & 1), DINCAE_EINVAL, "dincae_points_create: ntime_win must be odd");
    DC_REQUIRE(naux == 0 || aux, DINCAE_EINVAL, "dincae_points_create: naux > 0 but aux is NULL");
    DC_REQUIRE(offsets[0] == 0, DINCAE_EINVAL, "dincae_points_create: offsets[0] must be 0");
    for (int64_t n = 0; n < N; n++) DC_REQUIRE(offsets[n + 1] >= offsets[n], DINCAE_EINVAL, "dincae_points_create: offsets not monotone");
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    dincae_points* ds = new dincae_points();
    ds->ctx = ctx; ds->W = W; ds->H = H; ds->N = N; ds->T = ntime_win; ds->naux = naux;
    ds->jit_x = jitter_std_x; ds->jit_y = jitter_std_y;
    ds->offsets.assign(offsets, offsets + N + 1);
    ds->Ptot = offsets[N];
    DC_REQUIRE(ds->Ptot == 0 || (pos && x), DINCAE_EINVAL, "dincae_points_create: pos/x is NULL");
    const size_t HW = (size_t)W * H;
    size_t np = ds->Ptot > 0 ? ds->Ptot : 1;
    DC_CHECK_CUDA(cudaMalloc((void**)&ds->d_pos, sizeof(float) * 2 * np));
    DC_CHECK_CUDA(cudaMalloc((void**)&ds->d_x, sizeof(float) * 2 * np));
    DC_CHECK_CUDA(cudaMalloc((void**)&ds->d_lon, sizeof(float) * W));
    DC_CHECK_CUDA(cudaMalloc((void**)&ds->d_lat, sizeof(float) * H));
    DC_CHECK_CUDA(cudaMalloc((void**)&ds->d_cos, sizeof(float) * N));
    DC_CHECK_CUDA(cudaMalloc((void**)&ds->d_sin, sizeof(float) * N));
    if (ds->Ptot > 0) {
        DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_pos, pos, sizeof(float) * 2 * ds->Ptot, cudaMemcpyHostToDevice, ctx->stream));
        DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_x, x, sizeof(float) * 2 * ds->Ptot, cudaMemcpyHostToDevice, ctx->stream));
    }
    DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_lon, lon_scaled, sizeof(float) * W, cudaMemcpyHostToDevice, ctx->stream));
    DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_lat, lat_scaled, sizeof(float) * H, cudaMemcpyHostToDevice, ctx->stream));
    DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_cos, doy_cos, sizeof(float) * N, cudaMemcpyHostToDevice, ctx->stream));
    DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_sin, doy_sin, sizeof(float) * N, cudaMemcpyHostToDevice, ctx->stream));
    if (naux > 0) {
        DC_CHECK_CUDA(cudaMalloc((void**)&ds->d_aux, sizeof(float) * HW * naux * N));
        DC_CHECK_CUDA(cudaMemcpyAsync(ds->d_aux, aux, sizeof(float) * HW * naux * N, cudaMemcpyHostToDevice, ctx->stream));
    }
    DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    *out = ds;
    return DINCAE_OK;
}

extern "C" int32_t dincae_points_destroy(dincae_points* ds) {
    if (!ds) return DINCAE_OK;
    cudaSetDevice(ds->ctx->device);
    cudaStreamSynchronize(ds->ctx->stream);
    void* ptrs[] = {ds->d_pos, ds->d_x, ds->d_aux, ds->d_lon, ds->d_lat, ds->d_cos, ds->d_sin, ds->d_work};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    delete ds;
    return DINCAE_OK;
}

extern "C" int32_t dincae_points_assemble(dincae_points* ds, const int64_t* time_idx, int32_t B, int32_t perturb,
                                          uint64_t seed, uint64_t sample_ctr0, const uint8_t* keep, const int64_t* keep_offsets,
                                          float* xin, int32_t loc) {
    DC_REQUIRE(ds && time_idx && xin, DINCAE_EINVAL, "dincae_points_assemble: NULL argument");
    DC_REQUIRE(B >= 1, DINCAE_EINVAL, "dincae_points_assemble: B < 1");
    DC_REQUIRE(!keep || keep_offsets, DINCAE_EINVAL, "dincae_points_assemble: keep given without keep_offsets");
    dincae_ctx* ctx = ds->ctx;
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    const int T = ds->T, half = (T - 1) / 2, nseg = B * T;
    const int64_t HW = (int64_t)ds->W * ds->H;
    const int C = 4 + ds->naux + 2 * T;
    // host-side segment table (integer work on B*T entries)
    std::vector<int64_t> seg_src(nseg), seg_dst(nseg + 1), koff(B + 1, 0);
    std::vector<int> tidx(B);
    int64_t Pb = 0, nkeep = 0;
    for (int b = 0; b < B; b++) {
        int64_t ind = time_idx[b];
        DC_REQUIRE(ind >= 0 && ind < ds->N, DINCAE_EINVAL, "dincae_points_assemble: time_idx[%d]=%lld out of range", b, (long long)ind);
        tidx[b] = (int)ind;
        for (int k = 0; k < T; k++) {
            int64_t n = ind - half + k;                            // nrange, src/points.jl:269
            n = n < 0 ? 0 : (n > ds->N - 1 ? ds->N - 1 : n);
            seg_src[b * T + k] = ds->offsets[n];
            seg_dst[b * T + k] = Pb;
            Pb += ds->offsets[n + 1] - ds->offsets[n];
        }
        int64_t pc = ds->offsets[ind + 1] - ds->offsets[ind];
        if (keep) {
            DC_REQUIRE(keep_offsets[b + 1] - keep_offsets[b] == pc, DINCAE_EINVAL,
                       "dincae_points_assemble: sample %d has %lld keep flags for %lld centre-step points", b,
                       (long long)(keep_offsets[b + 1] - keep_offsets[b]), (long long)pc);
            koff[b] = keep_offsets[b] - keep_offsets[0];
            nkeep += pc;
        }
    }
    seg_dst[nseg] = Pb;
    // workspace: [seg_src | seg_dst | koff | tidx | keep | out_pos | v0 | v1 | plane | keepflag | sort scratch]
    auto al = [](size_t v) { return (v + 255) / 256 * 256; };
    size_t o_src = 0, o_dst = o_src + al(8 * nseg), o_koff = o_dst + al(8 * (nseg + 1)), o_tidx = o_koff + al(8 * (B + 1)),
           o_keep = o_tidx + al(4 * B), o_pos = o_keep + al(nkeep + 1), o_v0 = o_pos + al(8 * (Pb + 1)), o_v1 = o_v0 + al(4 * (Pb + 1)),
           o_plane = o_v1 + al(4 * (Pb + 1)), o_kf = o_plane + al(4 * (Pb + 1)), o_sort = o_kf + al(Pb + 1);
    size_t sort_bytes = k4_scatter_scratch_bytes(Pb);
    size_t need = o_sort + sort_bytes;
    if (need > ds->work_bytes) {
        DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
        if (ds->d_work) cudaFree(ds->d_work);
        ds->d_work = nullptr;
        ds->work_bytes = 0;
        size_t want = need + need / 4;
        cudaError_t e = cudaMalloc((void**)&ds->d_work, want);
        if (e != cudaSuccess) { dincae_set_error("dincae_points_assemble: workspace of %zu bytes: %s", want, cudaGetErrorString(e)); return DINCAE_ENOMEM; }
        ds->work_bytes = want;
    }
    char* wk = ds->d_work;
    DC_CHECK_CUDA(cudaMemcpyAsync(wk + o_src, seg_src.data(), 8 * nseg, cudaMemcpyHostToDevice, ctx->stream));
    DC_CHECK_CUDA(cudaMemcpyAsync(wk + o_dst, seg_dst.data(), 8 * (nseg + 1), cudaMemcpyHostToDevice, ctx->stream));
    DC_CHECK_CUDA(cudaMemcpyAsync(wk + o_koff, koff.data(), 8 * (B + 1), cudaMemcpyHostToDevice, ctx->stream));
    DC_CHECK_CUDA(cudaMemcpyAsync(wk + o_tidx, tidx.data(), 4 * B, cudaMemcpyHostToDevice, ctx->stream));
    if (keep && nkeep > 0)
        DC_CHECK_CUDA(cudaMemcpyAsync(wk + o_keep, keep + keep_offsets[0], nkeep, cudaMemcpyHostToDevice, ctx->stream));
    float* dxin = xin;
    const size_t nx = (size_t)B * C * HW;
    if (loc == DINCAE_HOST) DC_CHECK_CUDA(cudaMallocAsync((void**)&dxin, sizeof(float) * nx, ctx->stream));
    FillArgs fa{ds->d_lon, ds->d_lat, ds->d_cos, ds->d_sin, ds->d_aux, (const int*)(wk + o_tidx), ds->W, ds->H, ds->naux, T, B, ds->N, dxin};
    DC_LAUNCH(ctx, points_fill_kernel, (unsigned)dc_cdiv((int64_t)nx, 256), 256, 0, fa);
    DC_CHECK_LAUNCH();
    if (Pb > 0) {
        GenArgs ga;
        ga.pos = ds->d_pos; ga.x = ds->d_x; ga.Ptot = ds->Ptot;
        ga.seg_src = (const int64_t*)(wk + o_src); ga.seg_dst = (const int64_t*)(wk + o_dst);
        ga.keep_off = keep ? (const int64_t*)(wk + o_koff) : nullptr;
        ga.keep = keep ? (const uint8_t*)(wk + o_keep) : nullptr;
        ga.nseg = nseg; ga.T = T; ga.W = ds->W; ga.H = ds->H; ga.perturb = perturb;
        ga.jit_x = ds->jit_x; ga.jit_y = ds->jit_y; ga.seed = seed; ga.ctr0 = sample_ctr0;
        ga.out_pos = (float*)(wk + o_pos); ga.out_v0 = (float*)(wk + o_v0); ga.out_v1 = (float*)(wk + o_v1);
        ga.out_plane = (int32_t*)(wk + o_plane); ga.out_keep = (uint8_t*)(wk + o_kf);
        DC_LAUNCH(ctx, points_gen_kernel, (unsigned)dc_cdiv(Pb, 256), 256, 0, ga, Pb);
        DC_CHECK_LAUNCH();
        // plane (b,k) -> channels 4+naux+2k, 4+naux+2k+1 of sample b (offset = 5+naux+2(localn-1), src/points.jl:275)
        K4Batch kb{ga.out_pos, ga.out_plane, ga.out_keep, Pb, ds->W, ds->H, nseg, 2 * HW, HW, 2};
        kb.inner = T;
        kb.outer_stride = (int64_t)C * HW;
        kb.zero_out = 0;
        DC_TRY(k4_scatter_sorted(ctx, kb, ga.out_v0, ga.out_v1, dxin + (size_t)(4 + ds->naux) * HW, wk + o_sort, sort_bytes));
    }
    if (loc == DINCAE_HOST) {
        DC_CHECK_CUDA(cudaMemcpyAsync(xin, dxin, sizeof(float) * nx, cudaMemcpyDeviceToHost, ctx->stream));
        DC_CHECK_CUDA(cudaFreeAsync(dxin, ctx->stream));
        DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return DINCAE_OK;
}
