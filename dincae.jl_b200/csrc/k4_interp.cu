// K4: bilinear gather + deterministic sorted segmented scatter-add (see k4_interp.cuh).
//
// Adjoint design (HBM/latency-bound integer work, no tensor cores):
//   1. every point emits 4 entries (one per corner) keyed by plane*W*H + cell;
//   2. a stable LSD radix sort (8-bit digits, warp match ranking) orders the entries by key, so
//      inside one cell they stay in (point, corner) order = the order of the reference's
//      sequential CPU loop (src/points.jl:55-68);
//   3. the first entry of every segment sums its segment left to right with un-fused
//      multiply/add and writes the cell once.  No float atomics => bit-reproducible, and
//      bit-identical to the sequential CPU loop.
#include "k4_interp.cuh"
#include <mutex>

namespace {

constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 8;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;
constexpr uint32_t KEY_INVALID = 0xFFFFFFFFu;

__device__ __forceinline__ void corner_weights(float px, float py, int& ix, int& iy, float& cx, float& cy) {
    float fx = floorf(px), fy = floorf(py);
    ix = (int)fx;   // unsafe_trunc(Int32, floor(p)), ext/CUDAExt.jl:15
    iy = (int)fy;
    cx = __fsub_rn(px, fx);
    cy = __fsub_rn(py, fy);
}

__device__ __forceinline__ float corner_coeff(int c, float cx, float cy) {
    // CartesianIndices((0:1,0:1)) order: c = oi + 2*oj
    float wx = (c & 1) ? cx : __fsub_rn(1.0f, cx);
    float wy = (c & 2) ? cy : __fsub_rn(1.0f, cy);
    return __fmul_rn(wx, wy);
}

__global__ void k4_make_keys(K4Batch b, uint32_t* __restrict__ keys, uint32_t* __restrict__ idx, int64_t E) {
    int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (e >= E) return;
    int64_t p = e >> 2;
    int c = (int)(e & 3);
    float px = b.pos[2 * p], py = b.pos[2 * p + 1];
    int ix, iy;
    float cx, cy;
    corner_weights(px, py, ix, iy, cx, cy);
    int x = ix + (c & 1), y = iy + ((c >> 1) & 1);       // 1-based
    bool ok = (x >= 1) && (x <= b.W) && (y >= 1) && (y <= b.H) && (px == px) && (py == py);
    if (b.keep && !b.keep[p]) ok = false;
    uint32_t plane = b.plane_of ? (uint32_t)b.plane_of[p] : 0u;
    uint32_t key = ok ? (plane * (uint32_t)(b.W * b.H) + (uint32_t)((x - 1) + b.W * (y - 1))) : KEY_INVALID;
    keys[e] = key;
    idx[e] = (uint32_t)e;
}

__global__ void __launch_bounds__(SORT_THREADS) radix_hist(const uint32_t* __restrict__ keys, int64_t E, int shift,
                                                           uint32_t* __restrict__ hist, int nblk) {
    __shared__ uint32_t sh[256];
    sh[threadIdx.x] = 0;
    __syncthreads();
    int64_t base = (int64_t)blockIdx.x * SORT_TILE;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        int64_t e = base + i * SORT_THREADS + threadIdx.x;
        if (e < E) atomicAdd(&sh[(keys[e] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[(size_t)threadIdx.x * nblk + blockIdx.x] = sh[threadIdx.x];
}

// Exclusive scan of the n = 256 * nblk digit counts in two levels: every block scans one chunk of SCAN_CHUNK values in place
// (coalesced loads, warp shuffles) and leaves the chunk total; one block then scans the chunk totals.  radix_scatter adds the
// chunk offset when it reads its 256 bases, so no third pass over the counts is needed.
constexpr int SCAN_CHUNK = 4096;            // 1024 threads x 4 values
__global__ void __launch_bounds__(1024) radix_scan_chunks(uint32_t* __restrict__ data, int64_t n, uint32_t* __restrict__ totals) {
    __shared__ uint32_t wsum[32];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int64_t base = (int64_t)blockIdx.x * SCAN_CHUNK + (int64_t)t * 4;
    uint32_t v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) v[k] = base + k < n ? data[base + k] : 0u;
    const uint32_t mine = v[0] + v[1] + v[2] + v[3];
    uint32_t inc = mine;                                           // inclusive scan of the per-thread sums inside the warp
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, inc, off);
        if (lane >= off) inc += o;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = wsum[lane], winc = w;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, winc, off);
            if (lane >= off) winc += o;
        }
        wsum[lane] = winc - w;                                     // exclusive offset of each warp
        if (lane == 31) totals[blockIdx.x] = winc;
    }
    __syncthreads();
    uint32_t run = wsum[warp] + inc - mine;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        if (base + k < n) data[base + k] = run;
        run += v[k];
    }
}
// exclusive scan of the (few) chunk totals by one block
__global__ void __launch_bounds__(1024) radix_scan_totals(uint32_t* __restrict__ totals, int nchunks) {
    __shared__ uint32_t part[1024];
    const int t = threadIdx.x;
    const int per = (nchunks + 1023) / 1024;
    const int lo = t * per, hi = lo + per < nchunks ? lo + per : nchunks;
    uint32_t s = 0;
    for (int i = lo; i < hi; i++) s += totals[i];
    part[t] = s;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {
        uint32_t v = (t >= off) ? part[t - off] : 0;
        __syncthreads();
        part[t] += v;
        __syncthreads();
    }
    uint32_t run = part[t] - s;
    for (int i = lo; i < hi; i++) {
        uint32_t v = totals[i];
        totals[i] = run;
        run += v;
    }
}

__global__ void __launch_bounds__(SORT_THREADS) radix_scatter(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ idx_in,
                                                              uint32_t* __restrict__ keys_out, uint32_t* __restrict__ idx_out,
                                                              int64_t E, int shift, const uint32_t* __restrict__ hist, int nblk,
                                                              const uint32_t* __restrict__ chunk_off) {
    __shared__ uint32_t wcount[SORT_THREADS / 32][256];
    __shared__ uint32_t gbase[256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < (SORT_THREADS / 32) * 256; i += SORT_THREADS) (&wcount[0][0])[i] = 0;
    __syncthreads();
    int64_t base = (int64_t)blockIdx.x * SORT_TILE + (int64_t)warp * (32 * SORT_ITEMS);
    uint32_t key[SORT_ITEMS], val[SORT_ITEMS], rank[SORT_ITEMS];
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        int64_t e = base + i * 32 + lane;
        bool valid = e < E;
        key[i] = valid ? keys_in[e] : KEY_INVALID;
        val[i] = valid ? idx_in[e] : 0u;
        uint32_t d = (key[i] >> shift) & 255u;
        uint32_t mask = __match_any_sync(0xffffffffu, d);
        uint32_t before = __popc(mask & ((1u << lane) - 1u));
        uint32_t prev = wcount[warp][d];
        __syncwarp();
        if (before == 0) wcount[warp][d] = prev + __popc(mask);
        __syncwarp();
        rank[i] = prev + before;
    }
    __syncthreads();
    {
        int d = threadIdx.x;   // SORT_THREADS == 256 digits
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < SORT_THREADS / 32; w++) {
            uint32_t c = wcount[w][d];
            wcount[w][d] = run;
            run += c;
        }
        const size_t hi_ = (size_t)d * nblk + blockIdx.x;
        gbase[d] = hist[hi_] + chunk_off[hi_ / SCAN_CHUNK];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        int64_t e = base + i * 32 + lane;
        if (e < E) {
            uint32_t d = (key[i] >> shift) & 255u;
            uint32_t dst = gbase[d] + wcount[warp][d] + rank[i];
            keys_out[dst] = key[i];
            idx_out[dst] = val[i];
        }
    }
}

template <int NF>
__global__ void k4_segment_reduce(K4Batch b, const uint32_t* __restrict__ keys, const uint32_t* __restrict__ idx, int64_t E,
                                  const float* __restrict__ vals0, const float* __restrict__ vals1, float* __restrict__ out) {
    int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (e >= E) return;
    uint32_t key = keys[e];
    if (key == KEY_INVALID) return;
    if (e > 0 && keys[e - 1] == key) return;      // not a segment head
    float s0 = 0.f, s1 = 0.f;
    for (int64_t j = e; j < E && keys[j] == key; j++) {
        uint32_t ent = idx[j];
        int64_t p = ent >> 2;
        int c = ent & 3;
        float px = b.pos[2 * p], py = b.pos[2 * p + 1];
        int ix, iy;
        float cx, cy;
        corner_weights(px, py, ix, iy, cx, cy);
        float cc = corner_coeff(c, cx, cy);
        s0 = __fadd_rn(s0, __fmul_rn(vals0[p], cc));           // A2[p2...] += values[i] * cc, src/points.jl:66
        if (NF == 2) s1 = __fadd_rn(s1, __fmul_rn(vals1[p], cc));
    }
    uint32_t ncell = (uint32_t)(b.W * b.H);
    uint32_t plane = key / ncell, cell = key - plane * ncell;
    float* o = out + (b.inner > 0 ? (size_t)(plane / b.inner) * b.outer_stride + (size_t)(plane % b.inner) * b.plane_stride
                                  : (size_t)plane * b.plane_stride) + cell;
    o[0] = s0;
    if (NF == 2) o[b.field_stride] = s1;
}

__global__ void k4_gather_kernel(K4Batch b, const float* __restrict__ grids, float* __restrict__ out) {
    int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (p >= b.P) return;
    float px = b.pos[2 * p], py = b.pos[2 * p + 1];
    int ix, iy;
    float cx, cy;
    corner_weights(px, py, ix, iy, cx, cy);
    const size_t plane = b.plane_of ? b.plane_of[p] : 0;
    const float* g = grids + (b.inner > 0 ? (plane / b.inner) * b.outer_stride + (plane % b.inner) * b.plane_stride : plane * b.plane_stride);
    for (int f = 0; f < b.nfields; f++) {
        float v = 0.f;
#pragma unroll
        for (int c = 0; c < 4; c++) {
            int x = ix + (c & 1), y = iy + ((c >> 1) & 1);
            if (x >= 1 && x <= b.W && y >= 1 && y <= b.H)
                v = __fadd_rn(v, __fmul_rn(__ldg(g + f * b.field_stride + (x - 1) + (size_t)b.W * (y - 1)), corner_coeff(c, cx, cy)));
        }
        out[(size_t)f * b.P + p] = v;    // vec[i] += A[p2...] * cc, src/points.jl:20 (vec zero initially)
    }
}

__global__ void k4_cells_kernel(const float* __restrict__ pos, int64_t P, int32_t* __restrict__ cells) {
    int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (p >= P) return;
    cells[2 * p] = (int)floorf(pos[2 * p]);
    cells[2 * p + 1] = (int)floorf(pos[2 * p + 1]);
}

struct SortLayout {
    int64_t E;
    int nblk;
    size_t off_keys[2], off_idx[2], off_hist, off_tot, total;
    int nchunks;
};
SortLayout sort_layout(int64_t P) {
    SortLayout L;
    L.E = 4 * P;
    L.nblk = (int)dc_cdiv(L.E > 0 ? L.E : 1, SORT_TILE);
    size_t a = ((size_t)L.E * 4 + 255) / 256 * 256;
    size_t o = 0;
    for (int i = 0; i < 2; i++) { L.off_keys[i] = o; o += a; }
    for (int i = 0; i < 2; i++) { L.off_idx[i] = o; o += a; }
    L.off_hist = o;
    o += ((size_t)256 * L.nblk * 4 + 255) / 256 * 256;
    L.nchunks = (int)dc_cdiv((int64_t)256 * L.nblk, SCAN_CHUNK);
    L.off_tot = o;
    o += ((size_t)L.nchunks * 4 + 255) / 256 * 256;
    L.total = o;
    return L;
}

}  // namespace

size_t k4_scatter_scratch_bytes(int64_t P) { return sort_layout(P).total; }

int32_t k4_scatter_sorted(dincae_ctx* ctx, const K4Batch& b, const float* vals0, const float* vals1, float* out,
                          void* scratch, size_t scratch_bytes) {
    DC_REQUIRE(b.nfields == 1 || b.nfields == 2, DINCAE_EINVAL, "k4_scatter_sorted: nfields must be 1 or 2");
    DC_REQUIRE((int64_t)b.nplanes * b.W * b.H < (int64_t)KEY_INVALID, DINCAE_ESIZE,
               "k4_scatter_sorted: planes*W*H = %lld does not fit a 32-bit key", (long long)b.nplanes * b.W * b.H);
    DC_REQUIRE(4 * b.P < (int64_t)0x7FFFFFFF, DINCAE_ESIZE, "k4_scatter_sorted: too many points (%lld)", (long long)b.P);
    // zero the output grids (interp_adjn! starts with A2 .= 0, src/points.jl:53)
    if (b.zero_out) {
        if (b.inner <= 0 && b.plane_stride == (int64_t)b.nfields * b.W * b.H && b.field_stride == (int64_t)b.W * b.H) {
            DC_CHECK_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * (size_t)b.nplanes * b.plane_stride, ctx->stream));
        } else {
            for (int q = 0; q < b.nplanes; q++)
                for (int f = 0; f < b.nfields; f++) {
                    size_t o = b.inner > 0 ? (size_t)(q / b.inner) * b.outer_stride + (size_t)(q % b.inner) * b.plane_stride : (size_t)q * b.plane_stride;
                    DC_CHECK_CUDA(cudaMemsetAsync(out + o + (size_t)f * b.field_stride, 0, sizeof(float) * (size_t)b.W * b.H, ctx->stream));
                }
        }
    }
    if (b.P == 0) return DINCAE_OK;
    SortLayout L = sort_layout(b.P);
    DC_REQUIRE(scratch_bytes >= L.total, DINCAE_ESIZE, "k4_scatter_sorted: scratch too small (%zu < %zu)", scratch_bytes, L.total);
    char* sc = (char*)scratch;
    uint32_t* keys[2] = {(uint32_t*)(sc + L.off_keys[0]), (uint32_t*)(sc + L.off_keys[1])};
    uint32_t* idx[2] = {(uint32_t*)(sc + L.off_idx[0]), (uint32_t*)(sc + L.off_idx[1])};
    uint32_t* hist = (uint32_t*)(sc + L.off_hist);
    uint32_t* tot = (uint32_t*)(sc + L.off_tot);
    DC_LAUNCH(ctx, k4_make_keys, (unsigned)dc_cdiv(L.E, 256), 256, 0, b, keys[0], idx[0], L.E);
    // KEY_INVALID has all bits set, so it sorts last as long as every pass that could separate it is run:
    // sort on the bits of (nkeys) where nkeys = planes*W*H; invalid keys differ from valid ones in a sorted bit
    // only if we include one bit above the largest valid key.
    uint64_t nkeys = (uint64_t)b.nplanes * b.W * b.H;     // valid keys are < nkeys
    int bits = 1;
    while ((1ull << bits) <= nkeys) bits++;                // valid keys fit in `bits`-1.. and 1<<(bits-1) <= nkeys
    int passes = (bits + 7) / 8;
    if (passes > 4) passes = 4;
    int cur = 0;
    for (int p = 0; p < passes; p++) {
        DC_LAUNCH(ctx, radix_hist, L.nblk, SORT_THREADS, 0, keys[cur], L.E, 8 * p, hist, L.nblk);
        DC_LAUNCH(ctx, radix_scan_chunks, L.nchunks, 1024, 0, hist, (int64_t)256 * L.nblk, tot);
        DC_LAUNCH(ctx, radix_scan_totals, 1, 1024, 0, tot, L.nchunks);
        DC_LAUNCH(ctx, radix_scatter, L.nblk, SORT_THREADS, 0, keys[cur], idx[cur], keys[cur ^ 1], idx[cur ^ 1], L.E, 8 * p, hist, L.nblk, tot);
        cur ^= 1;
    }
    if (b.nfields == 1)
        DC_LAUNCH(ctx, k4_segment_reduce<1>, (unsigned)dc_cdiv(L.E, 256), 256, 0, b, keys[cur], idx[cur], L.E, vals0, vals1, out);
    else
        DC_LAUNCH(ctx, k4_segment_reduce<2>, (unsigned)dc_cdiv(L.E, 256), 256, 0, b, keys[cur], idx[cur], L.E, vals0, vals1, out);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

int32_t k4_gather(dincae_ctx* ctx, const K4Batch& b, const float* grids, float* out) {
    if (b.P == 0) return DINCAE_OK;
    DC_LAUNCH(ctx, k4_gather_kernel, (unsigned)dc_cdiv(b.P, 256), 256, 0, b, grids, out);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

// ---------------------------------------------------------------------------------------------
extern "C" int32_t dincae_interp(dincae_ctx* ctx, const float* pos, int64_t P, const float* A, int32_t W, int32_t H,
                                 float* out, int32_t loc) {
    DC_REQUIRE(ctx && (P == 0 || (pos && A && out)), DINCAE_EINVAL, "dincae_interp: NULL argument");
    DC_REQUIRE(P >= 0 && W >= 2 && H >= 2, DINCAE_EINVAL, "dincae_interp: bad sizes P=%lld W=%d H=%d", (long long)P, W, H);
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    if (P == 0) return DINCAE_OK;
    DevIn dpos, dA;
    DC_TRY(dc_stage_in(ctx, pos, sizeof(float) * 2 * P, loc, &dpos));
    DC_TRY(dc_stage_in(ctx, A, sizeof(float) * (size_t)W * H, loc, &dA));
    float* dout = out;
    if (loc == DINCAE_HOST) DC_CHECK_CUDA(cudaMallocAsync((void**)&dout, sizeof(float) * P, ctx->stream));
    K4Batch b{(const float*)dpos.ptr, nullptr, nullptr, P, W, H, 1, (int64_t)W * H, (int64_t)W * H, 1};
    int32_t st = k4_gather(ctx, b, (const float*)dA.ptr, dout);
    if (st == DINCAE_OK && loc == DINCAE_HOST) {
        DC_CHECK_CUDA(cudaMemcpyAsync(out, dout, sizeof(float) * P, cudaMemcpyDeviceToHost, ctx->stream));
        DC_CHECK_CUDA(cudaFreeAsync(dout, ctx->stream));
        DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    dc_stage_free(ctx, &dpos);
    dc_stage_free(ctx, &dA);
    return st;
}

extern "C" int32_t dincae_interp_adjoint(dincae_ctx* ctx, const float* pos, const float* values, int64_t P,
                                         int32_t W, int32_t H, float* out, int32_t loc) {
    DC_REQUIRE(ctx && out && (P == 0 || (pos && values)), DINCAE_EINVAL, "dincae_interp_adjoint: NULL argument");
    DC_REQUIRE(P >= 0 && W >= 2 && H >= 2, DINCAE_EINVAL, "dincae_interp_adjoint: bad sizes P=%lld W=%d H=%d", (long long)P, W, H);
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    DevIn dpos, dval;
    DC_TRY(dc_stage_in(ctx, pos, sizeof(float) * 2 * P, loc, &dpos));
    DC_TRY(dc_stage_in(ctx, values, sizeof(float) * P, loc, &dval));
    float* dout = out;
    if (loc == DINCAE_HOST) DC_CHECK_CUDA(cudaMallocAsync((void**)&dout, sizeof(float) * (size_t)W * H, ctx->stream));
    void* scratch = nullptr;
    size_t need = k4_scatter_scratch_bytes(P);
    DC_TRY(dc_scratch(ctx, need, &scratch));
    K4Batch b{(const float*)dpos.ptr, nullptr, nullptr, P, W, H, 1, (int64_t)W * H, (int64_t)W * H, 1};
    int32_t st = k4_scatter_sorted(ctx, b, (const float*)dval.ptr, nullptr, dout, scratch, ctx->scratch_bytes);
    if (st == DINCAE_OK && loc == DINCAE_HOST) {
        DC_CHECK_CUDA(cudaMemcpyAsync(out, dout, sizeof(float) * (size_t)W * H, cudaMemcpyDeviceToHost, ctx->stream));
        DC_CHECK_CUDA(cudaFreeAsync(dout, ctx->stream));
        DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    dc_stage_free(ctx, &dpos);
    dc_stage_free(ctx, &dval);
    return st;
}

extern "C" int32_t dincae_interp_cells(dincae_ctx* ctx, const float* pos, int64_t P, int32_t* cells, int32_t loc) {
    DC_REQUIRE(ctx && (P == 0 || (pos && cells)), DINCAE_EINVAL, "dincae_interp_cells: NULL argument");
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    if (P == 0) return DINCAE_OK;
    DevIn dpos;
    DC_TRY(dc_stage_in(ctx, pos, sizeof(float) * 2 * P, loc, &dpos));
    int32_t* dc = cells;
    if (loc == DINCAE_HOST) DC_CHECK_CUDA(cudaMallocAsync((void**)&dc, sizeof(int32_t) * 2 * P, ctx->stream));
    DC_LAUNCH(ctx, k4_cells_kernel, (unsigned)dc_cdiv(P, 256), 256, 0, (const float*)dpos.ptr, P, dc);
    DC_CHECK_LAUNCH();
    if (loc == DINCAE_HOST) {
        DC_CHECK_CUDA(cudaMemcpyAsync(cells, dc, sizeof(int32_t) * 2 * P, cudaMemcpyDeviceToHost, ctx->stream));
        DC_CHECK_CUDA(cudaFreeAsync(dc, ctx->stream));
        DC_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    dc_stage_free(ctx, &dpos);
    return DINCAE_OK;
}

// ---------------------------------------------------------------------------------------------
// Adjoint scatter-add for callers that must not allocate device memory.
//
// Who needs this: a host program whose GPU memory is owned by somebody else's pool.  The typical case
// is the Julia side of DINCAE itself: CUDA.jl's caching allocator may already hold nearly the whole
// device, so a foreign cudaMalloc issued from inside a ccall can fail with out-of-memory (or force the
// pool to trim) exactly when the pullback of interpnd runs in the middle of a training step.  The same
// holds for any embedding application that gives a plug-in a fixed memory budget of zero.
//
// The sort buffers and digit histograms that dincae_interp_adjoint takes from the context's grow-only
// scratch live here in a fixed array inside the library image instead.  Every GPU that loads the
// library carries the array (NOALLOC_SCRATCH_BYTES, about 4.1 MiB) until the process exits.
namespace {
constexpr int64_t NOALLOC_MAX_POINTS = 65536;
constexpr size_t noalloc_align(size_t v) { return (v + 255) / 256 * 256; }
constexpr size_t NOALLOC_ENTRIES = 4 * (size_t)NOALLOC_MAX_POINTS;
constexpr size_t NOALLOC_NBLK = (NOALLOC_ENTRIES + SORT_TILE - 1) / SORT_TILE;
constexpr size_t NOALLOC_NCHUNKS = (256 * NOALLOC_NBLK + SCAN_CHUNK - 1) / SCAN_CHUNK;
constexpr size_t NOALLOC_SCRATCH_BYTES = 4 * noalloc_align(NOALLOC_ENTRIES * 4) + noalloc_align(256 * NOALLOC_NBLK * 4) + noalloc_align(NOALLOC_NCHUNKS * 4);
}  // namespace

static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mutex;
void* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};

// In host code `g_noalloc_scratch` is only a handle for the runtime, not an address a kernel can use:
// the device address has to be asked for, once per device, while that device is current.
int32_t noalloc_scratch_address(int device, void** out) {
    DC_REQUIRE(device >= 0 && device < NOALLOC_MAX_DEVICES, DINCAE_EINVAL, "dincae_interp_adjoint_noalloc: device %d out of range", device);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_ptr[device] == nullptr) {
        void* p = nullptr;
        DC_CHECK_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        DC_REQUIRE(p != nullptr, DINCAE_ECUDA, "dincae_interp_adjoint_noalloc: cudaGetSymbolAddress returned NULL on device %d", device);
        g_noalloc_ptr[device] = p;
    }
    *out = g_noalloc_ptr[device];
    return DINCAE_OK;
}
}  // namespace

extern "C" int32_t dincae_interp_adjoint_noalloc(dincae_ctx* ctx, const float* d_pos, const float* d_values, int64_t P,
                                                 int32_t W, int32_t H, float* d_out) {
    DC_REQUIRE(ctx && d_out && (P == 0 || (d_pos && d_values)), DINCAE_EINVAL, "dincae_interp_adjoint_noalloc: NULL argument");
    DC_REQUIRE(P >= 0 && W >= 2 && H >= 2, DINCAE_EINVAL, "dincae_interp_adjoint_noalloc: bad sizes P=%lld W=%d H=%d", (long long)P, W, H);
    DC_REQUIRE(P <= NOALLOC_MAX_POINTS, DINCAE_ESIZE,
               "dincae_interp_adjoint_noalloc: %lld points exceed the built-in scratch (at most %lld points per call); "
               "split the call or use dincae_interp_adjoint, which allocates",
               (long long)P, (long long)NOALLOC_MAX_POINTS);
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    void* scratch = nullptr;
    DC_TRY(noalloc_scratch_address(ctx->device, &scratch));
    DC_REQUIRE(k4_scatter_scratch_bytes(P) <= NOALLOC_SCRATCH_BYTES, DINCAE_ESIZE, "dincae_interp_adjoint_noalloc: internal scratch layout mismatch");
    // The array still holds whatever the previous call left in it.  Nothing in it has to start at zero:
    // keys and indices are rewritten for all 4*P entries and every histogram cell is overwritten by its
    // block before the scan reads it; the only zero-initialised memory is the caller's grid, cleared below
    // by k4_scatter_sorted.  The per-launch event profile creates CUDA events, so it is bypassed here.
    const bool prof = ctx->prof;
    ctx->prof = false;
    K4Batch b{d_pos, nullptr, nullptr, P, W, H, 1, (int64_t)W * H, (int64_t)W * H, 1};
    int32_t st = k4_scatter_sorted(ctx, b, d_values, nullptr, d_out, scratch, NOALLOC_SCRATCH_BYTES);
    ctx->prof = prof;
    return st;
}
