// K2 (FP32 CUDA-core path): implicit-GEMM 3x3 convolution, its weight gradient and the
// element-wise level operators.  This is the exact-arithmetic path (DINCAE_PREC_FP32) and the
// path for the HBM-bound edge convolutions (Cin <= 16 or Cout = 2*noutput); the tensor-core
// (tcgen05) implicit GEMMs live in k2_conv_tma.cu.
#include "k2_conv.cuh"

namespace {

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

__device__ __forceinline__ float4 src_load4(const ConvSrc& s, int b, int y, int x, int c) {
    if (s.mode == SRC_UP_NEAREST) { y >>= 1; x >>= 1; }
    size_t off = (((size_t)b * s.Hs + y) * s.Ws + x) * s.C + c;
    float4 v = ldg4(s.src0 + off);
    if (s.src1) {
        float4 w = ldg4(s.src1 + off);
        v.x += w.x; v.y += w.y; v.z += w.z; v.w += w.w;
    }
    return v;
}
__device__ __forceinline__ float src_load1(const ConvSrc& s, int b, int y, int x, int c) {
    if (s.mode == SRC_UP_NEAREST) { y >>= 1; x >>= 1; }
    size_t off = (((size_t)b * s.Hs + y) * s.Ws + x) * s.C + c;
    float v = __ldg(s.src0 + off);
    if (s.src1) v += __ldg(s.src1 + off);
    return v;
}

__device__ __forceinline__ float act_apply(float v, int act, float slope) { return (act == ACT_RELU && v < 0.f) ? v * slope : v; }
__device__ __forceinline__ float act_grad(float a, float slope) { return a > 0.f ? 1.f : slope; }
__device__ __forceinline__ float round_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}
__device__ __forceinline__ float4 round_tf32(float4 v, int on) {
    return on ? make_float4(round_tf32(v.x), round_tf32(v.y), round_tf32(v.z), round_tf32(v.w)) : v;
}

// ---------------------------------------------------------------------------------------------
// generic forward: tile 128 pixels x BN couts, K chunk 16, 256 threads, 8 x (BN/16) per thread
template <int BN>
__global__ void __launch_bounds__(256) conv3x3_simt(ConvArgs a, int Mtot, int nchunks) {
    constexpr int BM = 128, BK = 16, TM = 8, TN = BN / 16;
    __shared__ __align__(16) float As[BK][BM + 4];
    __shared__ __align__(16) float Bs[BK][BN + 4];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    const int Cin_p = a.in.C, C4 = Cin_p >> 2;
    const int HW = a.H * a.W;

    int pb[2], py[2], px[2];
    bool pv[2];
#pragma unroll
    for (int j = 0; j < 2; j++) {
        int m = m0 + (tid >> 2) + 64 * j;
        pv[j] = m < Mtot;
        int mm = pv[j] ? m : 0;
        pb[j] = mm / HW;
        int r = mm - pb[j] * HW;
        py[j] = r / a.W;
        px[j] = r - py[j] * a.W;
    }
    int tap = 0, c4 = tid & 3;      // K position of this thread's float4 in the current chunk
    while (c4 >= C4) { c4 -= C4; tap++; }

    constexpr int NB4 = BK * BN / 4;
    float4 ra[2], rb;
    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; i++)
#pragma unroll
        for (int j = 0; j < TN; j++) acc[i][j] = 0.f;

    auto load_chunk = [&](int kc) {
        int ky = tap / 3, kx = tap - ky * 3;
#pragma unroll
        for (int j = 0; j < 2; j++) {
            int yy = py[j] + ky - 1, xx = px[j] + kx - 1;
            bool ok = pv[j] && tap < 9 && yy >= 0 && yy < a.H && xx >= 0 && xx < a.W;
            ra[j] = ok ? src_load4(a.in, pb[j], yy, xx, c4 * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (tid < NB4) {
            int k = tid / (BN / 4), n4 = tid - k * (BN / 4);
            int n = n0 + n4 * 4;
            rb = (n < a.Cout_p) ? ldg4(a.wgt + (size_t)(kc * BK + k) * a.Cout_p + n) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        c4 += 4;
        while (c4 >= C4) { c4 -= C4; tap++; }
    };

    load_chunk(0);
    for (int kc = 0; kc < nchunks; kc++) {
        {
            int kr = (tid & 3) * 4;
#pragma unroll
            for (int j = 0; j < 2; j++) {
                int pl = (tid >> 2) + 64 * j;
                As[kr + 0][pl] = ra[j].x; As[kr + 1][pl] = ra[j].y; As[kr + 2][pl] = ra[j].z; As[kr + 3][pl] = ra[j].w;
            }
            if (tid < NB4) {
                int k = tid / (BN / 4), n4 = tid - k * (BN / 4);
                *reinterpret_cast<float4*>(&Bs[k][n4 * 4]) = rb;
            }
        }
        __syncthreads();
        if (kc + 1 < nchunks) load_chunk(kc + 1);
#pragma unroll
        for (int k = 0; k < BK; k++) {
            float av[TM], bv[TN];
            float4 a0 = *reinterpret_cast<const float4*>(&As[k][ty * TM]);
            float4 a1 = *reinterpret_cast<const float4*>(&As[k][ty * TM + 4]);
            av[0] = a0.x; av[1] = a0.y; av[2] = a0.z; av[3] = a0.w; av[4] = a1.x; av[5] = a1.y; av[6] = a1.z; av[7] = a1.w;
            if (TN == 4) {
                float4 b0 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
                bv[0] = b0.x; bv[1] = b0.y; bv[2] = b0.z; bv[3] = b0.w;
            } else {
                float2 b0 = *reinterpret_cast<const float2*>(&Bs[k][tx * 2]);
                bv[0] = b0.x; bv[1] = b0.y;
            }
#pragma unroll
            for (int i = 0; i < TM; i++)
#pragma unroll
                for (int j = 0; j < TN; j++) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }

    const int n = n0 + tx * TN;
    if (n >= a.Cout_p) return;
    float bias[TN];
#pragma unroll
    for (int j = 0; j < TN; j++) bias[j] = a.bias ? __ldg(a.bias + n + j) : 0.f;
#pragma unroll
    for (int i = 0; i < TM; i++) {
        int m = m0 + ty * TM + i;
        if (m >= Mtot) continue;
        float o[TN];
#pragma unroll
        for (int j = 0; j < TN; j++) o[j] = act_apply(acc[i][j] + bias[j], a.act, a.slope);
        size_t off = (size_t)m * a.Cout_p + n;
        if (a.addend) {
#pragma unroll
            for (int j = 0; j < TN; j++) o[j] += __ldg(a.addend + off + j);
        }
        if (TN == 4) *reinterpret_cast<float4*>(a.dst + off) = make_float4(o[0], o[1], o[2], o[3]);
        else *reinterpret_cast<float2*>(a.dst + off) = make_float2(o[0], o[1]);
    }
}

// ---------------------------------------------------------------------------------------------
// weight gradient: dW[k][n] = sum_m im2col(in)[m][k] * gout[m][n];  tile 64 x 64, pixel chunk 16,
// split over pixel ranges (blockIdx.z) with a deterministic second-stage reduction.
__global__ void __launch_bounds__(256) wgrad_simt(WgradArgs a, int Mtot, int Krows, int Kt, int Nt) {
    constexpr int BP = 16;
    __shared__ __align__(16) float As[BP][64 + 4];
    __shared__ __align__(16) float Gs[BP][64 + 4];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int k0 = blockIdx.x * 64, n0 = blockIdx.y * 64;
    const int Cin_p = a.in.C;
    const int HW = a.H * a.W;
    int64_t mlo = (int64_t)blockIdx.z * a.pix_per_split;
    int64_t mhi = mlo + a.pix_per_split;
    if (mhi > Mtot) mhi = Mtot;

    // loader roles: pixel pl = tid>>4 of the chunk, float4 column q = tid&15
    const int pl = tid >> 4, q = tid & 15;
    const int kk = k0 + q * 4;
    const bool kvalid = kk < Krows;
    const int tap = kvalid ? kk / Cin_p : 0, ci = kvalid ? kk - tap * Cin_p : 0;
    const int ky = tap / 3, kx = tap - ky * 3;
    const int nn = n0 + q * 4;
    const bool nvalid = nn < a.Cout_p;

    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.f;
    float bsum = 0.f;

    for (int64_t mc = mlo; mc < mhi; mc += BP) {
        int64_t m = mc + pl;
        float4 va = make_float4(0.f, 0.f, 0.f, 0.f), vg = va;
        if (m < mhi) {
            int b = (int)(m / HW);
            int r = (int)(m - (int64_t)b * HW);
            int y = r / a.W, x = r - y * a.W;
            int yy = y + ky - 1, xx = x + kx - 1;
            if (kvalid && yy >= 0 && yy < a.H && xx >= 0 && xx < a.W) va = src_load4(a.in, b, yy, xx, ci);
            if (nvalid) vg = ldg4(a.gout + (size_t)m * a.Cout_p + nn);
        }
        __syncthreads();
        *reinterpret_cast<float4*>(&As[pl][q * 4]) = va;
        *reinterpret_cast<float4*>(&Gs[pl][q * 4]) = vg;
        __syncthreads();
#pragma unroll
        for (int p = 0; p < BP; p++) {
            float4 av = *reinterpret_cast<const float4*>(&As[p][ty * 4]);
            float4 gv = *reinterpret_cast<const float4*>(&Gs[p][tx * 4]);
            float aa[4] = {av.x, av.y, av.z, av.w}, gg[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) acc[i][j] = fmaf(aa[i], gg[j], acc[i][j]);
        }
        if (blockIdx.x == 0 && tid < 64) {
#pragma unroll
            for (int p = 0; p < BP; p++) bsum += Gs[p][tid];
        }
    }
    // write (partial) result
    float* dst = (a.splits == 1) ? a.dW : a.partial + (size_t)blockIdx.z * ((size_t)Kt * 64 * Nt * 64 + (size_t)Nt * 64);
    const int ld = (a.splits == 1) ? a.Cout_p : Nt * 64;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int k = k0 + ty * 4 + i;
        int n = n0 + tx * 4;
        if (a.splits == 1) {
            if (k < Krows && n < a.Cout_p) *reinterpret_cast<float4*>(dst + (size_t)k * ld + n) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        } else {
            *reinterpret_cast<float4*>(dst + (size_t)k * ld + n) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        }
    }
    if (blockIdx.x == 0 && tid < 64) {
        int n = n0 + tid;
        if (a.splits == 1) {
            if (n < a.Cout_p) a.db[n] = bsum;
        } else {
            dst[(size_t)Kt * 64 * ld + n] = bsum;
        }
    }
}

__global__ void wgrad_reduce(const float* __restrict__ partial, int splits, size_t stride, int ld, int Krows, int Cout_p,
                             float* __restrict__ dW, float* __restrict__ db, size_t bias_off) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int64_t nW = (int64_t)Krows * Cout_p;
    if (i < nW) {
        int k = (int)(i / Cout_p), n = (int)(i - (int64_t)k * Cout_p);
        float s = 0.f;
        for (int z = 0; z < splits; z++) s += partial[(size_t)z * stride + (size_t)k * ld + n];
        dW[i] = s;
    } else if (i < nW + Cout_p) {
        int n = (int)(i - nW);
        float s = 0.f;
        for (int z = 0; z < splits; z++) s += partial[(size_t)z * stride + bias_off + n];
        db[n] = s;
    }
}

// ---------------------------------------------------------------------------------------------
// small-Cout forward: one thread per pixel, all NCO outputs; weights in shared memory
template <int NCO>
__global__ void __launch_bounds__(256) conv3x3_small(ConvSmallArgs a, int Mtot, int Krows) {
    extern __shared__ float ws[];   // [Krows][NCO]
    for (int i = threadIdx.x; i < Krows * NCO; i += blockDim.x) {
        int k = i / NCO, j = i - k * NCO;
        ws[i] = __ldg(a.wgt + (size_t)k * a.ldw + j);
    }
    __syncthreads();
    int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= Mtot) return;
    const int HW = a.H * a.W, Cin_p = a.in.C;
    int b = m / HW, r = m - b * HW, y = r / a.W, x = r - y * a.W;
    float acc[NCO];
#pragma unroll
    for (int j = 0; j < NCO; j++) acc[j] = 0.f;
    for (int tap = 0; tap < 9; tap++) {
        int ky = tap / 3, kx = tap - ky * 3;
        int yy = y + ky - 1, xx = x + kx - 1;
        if (yy < 0 || yy >= a.H || xx < 0 || xx >= a.W) continue;
        const float* w = ws + (size_t)tap * Cin_p * NCO;
        for (int c = 0; c < Cin_p; c += 4) {
            float4 v = src_load4(a.in, b, yy, xx, c);
            float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int e = 0; e < 4; e++)
#pragma unroll
                for (int j = 0; j < NCO; j++) acc[j] = fmaf(vv[e], w[(c + e) * NCO + j], acc[j]);
        }
    }
#pragma unroll
    for (int j = 0; j < NCO; j++) {
        if (j < a.Cout) {
            float v = acc[j] + (a.bias ? __ldg(a.bias + j) : 0.f);
            size_t off = ((size_t)b * a.Cout + j) * HW + r;
            a.out[off] = a.accumulate ? a.out[off] + v : v;
        }
    }
}

// small-Cout weight gradient: gout has NCO padded channels.  Thread = one K row; block = pixel range.
template <int NCO>
__global__ void __launch_bounds__(256) wgrad_small(WgradArgs a, int Mtot, int Krows) {
    const int HW = a.H * a.W, Cin_p = a.in.C;
    int64_t mlo = (int64_t)blockIdx.x * a.pix_per_split, mhi = mlo + a.pix_per_split;
    if (mhi > Mtot) mhi = Mtot;
    float* dst = a.partial + (size_t)blockIdx.x * (size_t)(Krows + 1) * NCO;
    for (int k = threadIdx.x; k <= Krows; k += blockDim.x) {
        const bool ones = (k == Krows);     // virtual row of ones = bias gradient
        int tap = ones ? 4 : k / Cin_p, ci = ones ? 0 : k - tap * Cin_p;
        int ky = tap / 3, kx = tap - ky * 3;
        float acc[NCO];
#pragma unroll
        for (int j = 0; j < NCO; j++) acc[j] = 0.f;
        int b = (int)(mlo / HW);
        int r = (int)(mlo - (int64_t)b * HW);
        int y = r / a.W, x = r - y * a.W;
        for (int64_t m = mlo; m < mhi; m++) {
            int yy = y + ky - 1, xx = x + kx - 1;
            if (yy >= 0 && yy < a.H && xx >= 0 && xx < a.W) {
                float v = ones ? 1.f : src_load1(a.in, b, yy, xx, ci);
                const float* g = a.gout + (size_t)m * NCO;
                if (NCO == 4) {
                    float4 gv = ldg4(g);
                    acc[0] = fmaf(v, gv.x, acc[0]); acc[1] = fmaf(v, gv.y, acc[1]); acc[2] = fmaf(v, gv.z, acc[2]); acc[3] = fmaf(v, gv.w, acc[3]);
                } else {
#pragma unroll
                    for (int j = 0; j < NCO; j++) acc[j] = fmaf(v, __ldg(g + j), acc[j]);
                }
            }
            if (++x == a.W) { x = 0; if (++y == a.H) { y = 0; b++; } }
        }
#pragma unroll
        for (int j = 0; j < NCO; j++) dst[(size_t)k * NCO + j] = acc[j];
    }
}

__global__ void wgrad_small_reduce(const float* __restrict__ partial, int splits, int Krows, int nco, int Cout_p,
                                   float* __restrict__ dW, float* __restrict__ db) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (Krows + 1) * nco) return;
    int k = i / nco, j = i - k * nco;
    float s = 0.f;
    for (int z = 0; z < splits; z++) s += partial[(size_t)z * (Krows + 1) * nco + i];
    if (j < Cout_p) {
        if (k < Krows) dW[(size_t)k * Cout_p + j] = s;
        else db[j] = s;
    }
}

// ---------------------------------------------------------------------------------------------
// planar (WHCN) -> NHWC: a warp moves 32 consecutive pixels; reads are coalesced along pixels (one channel plane at a time),
// a padded smem tile transposes, writes are coalesced along channels (4 pixels x 128 B per store instruction)
__global__ void __launch_bounds__(256) pack_planar_kernel(const float* __restrict__ s0, int C0, const float* __restrict__ s1, int C1,
                                                          int B, int HW, float* __restrict__ dst, int Cp, int Cw, int rnd, int split) {
    dc_pdl_launch_dependents();
    dc_pdl_wait();
    __shared__ float tile[8][32][33];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t total = (int64_t)B * HW;
    const int64_t m0 = ((int64_t)blockIdx.x * 8 + warp) * 32;
    if (m0 >= total) return;
    const int64_t m = m0 + lane;
    const bool pv = m < total;
    const int b = pv ? (int)(m / HW) : 0, r = pv ? (int)(m - (int64_t)b * HW) : 0;
    const int Ct = C0 + C1;
    // per-pixel base pointers: the channel loop only adds multiples of the plane size (no 64-bit multiplies per load)
    const float* p0 = s0 + ((size_t)b * C0) * HW + r;
    const float* p1 = s1 ? s1 + ((size_t)b * C1) * HW + r : nullptr;
    const bool one_pass = split && 2 * Ct <= 32;            // [hi | lo] of every channel fits one 32-channel group: load each value once
    for (int cg = 0; cg < Cw; cg += 32) {                  // channels >= Cw are zero padding and stay as the buffer was created (zeros)
        const int nc = Cw - cg < 32 ? Cw - cg : 32;
        if (one_pass) {
            for (int c8 = 0; c8 < Ct; c8 += 8) {
                float v[8];
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const int cc = c8 + k;
                    v[k] = 0.f;
                    if (pv && cc < Ct) v[k] = cc < C0 ? __ldg(p0 + (size_t)cc * HW) : __ldg(p1 + (size_t)(cc - C0) * HW);
                }
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const int cc = c8 + k;
                    if (cc < Ct) {
                        const float hi = round_tf32(v[k]);
                        tile[warp][lane][cc] = hi;
                        tile[warp][lane][Ct + cc] = round_tf32(v[k] - hi);
                    }
                }
            }
            for (int cc = 2 * Ct; cc < nc; cc++) tile[warp][lane][cc] = 0.f;
        } else
        for (int c8 = 0; c8 < nc; c8 += 8) {               // eight independent loads in flight per thread, then the shared stores
            float v[8];
#pragma unroll
            for (int k = 0; k < 8; k++) {
                int cc = cg + c8 + k;
                if (split && cc >= Ct) cc = cc < 2 * Ct ? cc - Ct : Ct;         // [hi | lo | zero padding]
                v[k] = 0.f;
                if (pv) {
                    if (cc < C0) v[k] = __ldg(p0 + (size_t)cc * HW);
                    else if (cc < Ct) v[k] = __ldg(p1 + (size_t)(cc - C0) * HW);
                }
            }
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const float hi = rnd ? round_tf32(v[k]) : v[k];
                const bool lo = split && cg + c8 + k >= Ct;
                tile[warp][lane][c8 + k] = lo ? round_tf32(v[k] - hi) : hi;
            }
        }
        __syncwarp();
        const int chunks = nc >> 2;                         // float4 chunks per pixel row in this group
        for (int i = lane; i < 32 * chunks; i += 32) {
            const int px = i / chunks, ch = i - px * chunks;
            if (m0 + px < total) {
                float4 o = make_float4(tile[warp][px][4 * ch], tile[warp][px][4 * ch + 1], tile[warp][px][4 * ch + 2], tile[warp][px][4 * ch + 3]);
                *reinterpret_cast<float4*>(dst + (size_t)(m0 + px) * Cp + cg + 4 * ch) = o;
            }
        }
        __syncwarp();
    }
}

// The element-wise level kernels below share one indexing scheme: a block walks whole pixel rows (row = b * H + y, one 32-bit
// division per row), a thread owns the float4 column group c4 = t % CW for good and steps through the row PL = 256 / CW
// pixels at a time.  No per-element 64-bit div/mod, so the kernels are limited by HBM rather than by index arithmetic.
struct RowThread {
    int CW, PL, c4, lane;
    __device__ __forceinline__ RowThread(int C4) {
        CW = C4 < 256 ? C4 : 256;
        PL = 256 / CW;
        c4 = (int)threadIdx.x % CW;
        lane = (int)threadIdx.x / CW;       // >= PL for the (256 - PL * CW) threads without work
    }
};

__global__ void __launch_bounds__(256) pool_kernel(const float* __restrict__ a, int B, int H, int W, int Cp, int H2, int W2, float* __restrict__ x, int rnd) {
    dc_pdl_launch_dependents();
    dc_pdl_wait();
    const int C4 = Cp >> 2;
    const RowThread rt(C4);
    if (rt.lane >= rt.PL) return;
    const int oh = H & 1, ow = W & 1;
    for (int row = blockIdx.x; row < B * H2; row += gridDim.x) {
        const int b = row / H2, y2 = row - b * H2;
        const int ya = 2 * y2 - oh, yb = ya + 1;                        // ya may be -1 (odd H: the pad row), yb < H always
        const float* ra = a + ((size_t)b * H + (ya < 0 ? 0 : ya)) * W * Cp;
        const float* rb = a + ((size_t)b * H + yb) * W * Cp;
        float* ro = x + (size_t)row * W2 * Cp;
        for (int c4 = rt.c4; c4 < C4; c4 += rt.CW)
            for (int x2 = rt.lane; x2 < W2; x2 += rt.PL) {
                const int xa = 2 * x2 - ow, xb = xa + 1;
                float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
                // same accumulation order as the reference loop: (dy,dx) = (0,0), (0,1), (1,0), (1,1)
                if (ya >= 0 && xa >= 0) { float4 v = ldg4(ra + (size_t)xa * Cp + c4 * 4); s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; }
                if (ya >= 0) { float4 v = ldg4(ra + (size_t)xb * Cp + c4 * 4); s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; }
                if (xa >= 0) { float4 v = ldg4(rb + (size_t)xa * Cp + c4 * 4); s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; }
                { float4 v = ldg4(rb + (size_t)xb * Cp + c4 * 4); s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; }
                s.x *= 0.25f; s.y *= 0.25f; s.z *= 0.25f; s.w *= 0.25f;
                *reinterpret_cast<float4*>(ro + (size_t)x2 * Cp + c4 * 4) = round_tf32(s, rnd);
            }
    }
}

// block-level column partial sums of the float4 each thread accumulated for ITS column group (requires 256 % C4 == 0 and a
// grid stride that is a multiple of C4, so a thread always sees the same columns); fixed summation order -> deterministic
__device__ __forceinline__ void block_colpart(float4 acc, int C4, int Cp, float* __restrict__ colpart) {
    __shared__ float4 shp[256];
    shp[threadIdx.x] = acc;
    __syncthreads();
    if ((int)threadIdx.x < C4) {
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int k = threadIdx.x; k < 256; k += C4) { float4 v = shp[k]; t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w; }
        reinterpret_cast<float4*>(colpart + (size_t)blockIdx.x * Cp)[threadIdx.x] = t;
    }
}

constexpr int LV_U = 1;      // cells per thread and iteration of the 4-load adjoint kernel
constexpr int LV_V = 1;      // ... of the 1- and 2-load kernels
// gu at (H,W); outputs at (Hs,Ws) = (H/2 + H%2, W/2 + W%2)
__global__ void __launch_bounds__(256, 8) upbwd_kernel(const float* __restrict__ gu, int B, int H, int W, int Cp, int Hs, int Ws,
                                                    const float* __restrict__ r, float slope, float* __restrict__ G, float* __restrict__ gd, int rnd,
                                                    float* __restrict__ colpart) {
    dc_pdl_launch_dependents();
    dc_pdl_wait();
    const int C4 = Cp >> 2;
    const RowThread rt(C4);
    float4 cacc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (rt.lane < rt.PL) {
        for (int row = blockIdx.x; row < B * Hs; row += gridDim.x) {
            const int b = row / Hs, ys = row - b * Hs;
            const int ya = 2 * ys, yb = ya + 1;
            const float* ra = gu + ((size_t)b * H + ya) * W * Cp;
            const float* rb = gu + ((size_t)b * H + (yb < H ? yb : ya)) * W * Cp;
            const size_t ro = (size_t)row * Ws * Cp;
            // LV_U cells per thread and iteration, every load issued before the first use: these kernels are bound by the number
            // of memory round trips a thread makes one after the other, not by bandwidth
            for (int c4 = rt.c4; c4 < C4; c4 += rt.CW)
                for (int x0 = rt.lane; x0 < Ws; x0 += LV_U * rt.PL) {
                    float4 va[LV_U], vb[LV_U], vc[LV_U], vd[LV_U], rv[LV_U];
                    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                    for (int k = 0; k < LV_U; k++) {
                        const int xs = x0 + k * rt.PL, xa = 2 * xs, xb = xa + 1;
                        const bool in = xs < Ws;
                        va[k] = in ? ldg4(ra + (size_t)xa * Cp + c4 * 4) : z4;
                        vb[k] = in && xb < W ? ldg4(ra + (size_t)xb * Cp + c4 * 4) : z4;
                        vc[k] = in && yb < H ? ldg4(rb + (size_t)xa * Cp + c4 * 4) : z4;
                        vd[k] = in && yb < H && xb < W ? ldg4(rb + (size_t)xb * Cp + c4 * 4) : z4;
                        rv[k] = in && gd && r ? ldg4(r + ro + (size_t)xs * Cp + c4 * 4) : z4;
                    }
#pragma unroll
                    for (int k = 0; k < LV_U; k++) {
                        const int xs = x0 + k * rt.PL, xb = 2 * xs + 1;
                        if (xs >= Ws) break;
                        float4 s = va[k];
                        if (xb < W) { s.x += vb[k].x; s.y += vb[k].y; s.z += vb[k].z; s.w += vb[k].w; }
                        if (yb < H) {
                            s.x += vc[k].x; s.y += vc[k].y; s.z += vc[k].z; s.w += vc[k].w;
                            if (xb < W) { s.x += vd[k].x; s.y += vd[k].y; s.z += vd[k].z; s.w += vd[k].w; }
                        }
                        const size_t off = ro + (size_t)xs * Cp + c4 * 4;
                        if (G) *reinterpret_cast<float4*>(G + off) = s;
                        if (gd) {
                            if (r) {
                                s.x *= act_grad(rv[k].x, slope); s.y *= act_grad(rv[k].y, slope);
                                s.z *= act_grad(rv[k].z, slope); s.w *= act_grad(rv[k].w, slope);
                            }
                            s = round_tf32(s, rnd);
                            *reinterpret_cast<float4*>(gd + off) = s;
                            cacc.x += s.x; cacc.y += s.y; cacc.z += s.z; cacc.w += s.w;
                        }
                    }
                }
        }
    }
    if (colpart) block_colpart(cacc, C4, Cp, colpart);
}

// act'(a) comes either from the stored activation `a` or, when the conv epilogue kept only its sign bits, from `mask`
// (word [pixel][Cp / 32], bit c % 32 set <=> a[pixel][c] > 0)
__global__ void __launch_bounds__(256, 8) poolbwd_kernel(const float* __restrict__ xg, const float* __restrict__ a, const uint32_t* __restrict__ mask,
                                                      int B, int H, int W, int Cp,
                                                      int H2, int W2, float slope, float* __restrict__ gc, int rnd, float* __restrict__ colpart) {
    dc_pdl_launch_dependents();
    dc_pdl_wait();
    const int C4 = Cp >> 2;
    const RowThread rt(C4);
    float4 cacc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (rt.lane < rt.PL) {
        const int oh = H & 1, ow = W & 1;
        for (int row = blockIdx.x; row < B * H; row += gridDim.x) {
            const int b = row / H, y = row - b * H;
            const float* rx = xg + ((size_t)b * H2 + ((y + oh) >> 1)) * W2 * Cp;
            const size_t ro = (size_t)row * W * Cp;
            for (int c4 = rt.c4; c4 < C4; c4 += rt.CW)
                for (int x0 = rt.lane; x0 < W; x0 += LV_V * rt.PL) {
                    float4 gv[LV_V], av[LV_V];
                    uint32_t bits[LV_V];
                    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                    for (int k = 0; k < LV_V; k++) {                       // all loads of LV_V cells first (see upbwd_kernel)
                        const int x = x0 + k * rt.PL;
                        const bool in = x < W;
                        gv[k] = in ? ldg4(rx + (size_t)((x + ow) >> 1) * Cp + c4 * 4) : z4;
                        if (mask) bits[k] = in ? __ldg(mask + ((size_t)row * W + x) * (Cp >> 5) + (c4 >> 3)) >> ((c4 & 7) * 4) : 0u;
                        else av[k] = in ? ldg4(a + ro + (size_t)x * Cp + c4 * 4) : z4;
                    }
#pragma unroll
                    for (int k = 0; k < LV_V; k++) {
                        const int x = x0 + k * rt.PL;
                        if (x >= W) break;
                        float4 g = gv[k];
                        if (mask) {
                            g.x *= 0.25f * ((bits[k] & 1u) ? 1.f : slope); g.y *= 0.25f * ((bits[k] & 2u) ? 1.f : slope);
                            g.z *= 0.25f * ((bits[k] & 4u) ? 1.f : slope); g.w *= 0.25f * ((bits[k] & 8u) ? 1.f : slope);
                        } else {
                            g.x *= 0.25f * act_grad(av[k].x, slope); g.y *= 0.25f * act_grad(av[k].y, slope);
                            g.z *= 0.25f * act_grad(av[k].z, slope); g.w *= 0.25f * act_grad(av[k].w, slope);
                        }
                        g = round_tf32(g, rnd);
                        *reinterpret_cast<float4*>(gc + ro + (size_t)x * Cp + c4 * 4) = g;
                        cacc.x += g.x; cacc.y += g.y; cacc.z += g.z; cacc.w += g.w;
                    }
                }
        }
    }
    if (colpart) block_colpart(cacc, C4, Cp, colpart);
}

// bilinear, align_corners = true (NNlib.upsample_bilinear default), full size (2Hs,2Ws) cropped to (H,W)
__device__ __forceinline__ void bil_coord(int o, int nin, int nout_full, int& i0, int& i1, float& f) {
    float s = (nout_full > 1) ? (float)o * (float)(nin - 1) / (float)(nout_full - 1) : 0.f;
    i0 = (int)floorf(s);
    if (i0 > nin - 1) i0 = nin - 1;
    i1 = i0 + 1 < nin ? i0 + 1 : nin - 1;
    f = s - (float)i0;
}
__global__ void up_bilinear_kernel(const float* __restrict__ s0, const float* __restrict__ s1, int B, int Hs, int Ws, int Cp,
                                   int H, int W, float* __restrict__ dst, int rnd, int split) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int C4 = Cp >> 2;
    int64_t total = (int64_t)B * H * W * C4;
    if (i >= total) return;
    int c4 = (int)(i % C4);
    int64_t p = i / C4;
    int x = (int)(p % W);
    p /= W;
    int y = (int)(p % H);
    int b = (int)(p / H);
    int x0, x1, y0, y1;
    float fx, fy;
    bil_coord(x, Ws, 2 * Ws, x0, x1, fx);
    bil_coord(y, Hs, 2 * Hs, y0, y1, fy);
    auto L = [&](int yy, int xx) {
        size_t off = (((size_t)b * Hs + yy) * Ws + xx) * Cp + c4 * 4;
        float4 v = ldg4(s0 + off);
        if (s1) { float4 w = ldg4(s1 + off); v.x += w.x; v.y += w.y; v.z += w.z; v.w += w.w; }
        return v;
    };
    float4 v00 = L(y0, x0), v01 = L(y0, x1), v10 = L(y1, x0), v11 = L(y1, x1);
    float w00 = (1 - fy) * (1 - fx), w01 = (1 - fy) * fx, w10 = fy * (1 - fx), w11 = fy * fx;
    float4 o;
    o.x = w00 * v00.x + w01 * v01.x + w10 * v10.x + w11 * v11.x;
    o.y = w00 * v00.y + w01 * v01.y + w10 * v10.y + w11 * v11.y;
    o.z = w00 * v00.z + w01 * v01.z + w10 * v10.z + w11 * v11.z;
    o.w = w00 * v00.w + w01 * v01.w + w10 * v10.w + w11 * v11.w;
    if (!split) {
        *reinterpret_cast<float4*>(dst + (((size_t)b * H + y) * W + x) * Cp + c4 * 4) = round_tf32(o, rnd);
    } else {
        const float4 hi = round_tf32(o, 1);
        float* d = dst + (((size_t)b * H + y) * W + x) * (2 * Cp) + c4 * 4;
        *reinterpret_cast<float4*>(d) = hi;
        *reinterpret_cast<float4*>(d + Cp) = round_tf32(make_float4(o.x - hi.x, o.y - hi.y, o.z - hi.z, o.w - hi.w), 1);
    }
}
// one thread per SOURCE cell: the (optionally summed) value is read once and written to its 2 x 2 (cropped) destination cells
__global__ void __launch_bounds__(256, 8) up_nearest_kernel(const float* __restrict__ s0, const float* __restrict__ s1, int B, int Hs, int Ws, int Cp,
                                                         int H, int W, float* __restrict__ dst, int rnd, int split) {
    dc_pdl_launch_dependents();
    dc_pdl_wait();
    const int C4 = Cp >> 2;
    const RowThread rt(C4);
    if (rt.lane >= rt.PL) return;
    const int Cd = split ? 2 * Cp : Cp;             // destination channels per pixel
    for (int row = blockIdx.x; row < B * Hs; row += gridDim.x) {
        const int b = row / Hs, ys = row - b * Hs;
        const int ya = 2 * ys, yb = ya + 1;
        if (ya >= H) continue;
        const size_t rs = (size_t)row * Ws * Cp;
        float* da = dst + ((size_t)b * H + ya) * W * Cd;
        float* db = yb < H ? dst + ((size_t)b * H + yb) * W * Cd : nullptr;
        for (int c4 = rt.c4; c4 < C4; c4 += rt.CW)
          for (int x0 = rt.lane; x0 < Ws; x0 += LV_V * rt.PL) {
            float4 v0[LV_V], v1[LV_V];
            const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int k = 0; k < LV_V; k++) {                               // all loads of LV_V cells first (see upbwd_kernel)
                const int xs = x0 + k * rt.PL;
                const bool in = xs < Ws && 2 * xs < W;
                v0[k] = in ? ldg4(s0 + rs + (size_t)xs * Cp + c4 * 4) : z4;
                v1[k] = in && s1 ? ldg4(s1 + rs + (size_t)xs * Cp + c4 * 4) : z4;
            }
#pragma unroll
            for (int k = 0; k < LV_V; k++) {
                const int xs = x0 + k * rt.PL;
                const int xa = 2 * xs, xb = xa + 1;
                if (xs >= Ws || xa >= W) break;
                float4 v = v0[k];
                if (s1) { v.x += v1[k].x; v.y += v1[k].y; v.z += v1[k].z; v.w += v1[k].w; }
                const float4 full = v;
                v = round_tf32(v, rnd);
                *reinterpret_cast<float4*>(da + (size_t)xa * Cd + c4 * 4) = v;
                if (xb < W) *reinterpret_cast<float4*>(da + (size_t)xb * Cd + c4 * 4) = v;
                if (db) {
                    *reinterpret_cast<float4*>(db + (size_t)xa * Cd + c4 * 4) = v;
                    if (xb < W) *reinterpret_cast<float4*>(db + (size_t)xb * Cd + c4 * 4) = v;
                }
                if (split) {                         // low-order part behind the Cp high-order channels
                    const float4 lo = round_tf32(make_float4(full.x - v.x, full.y - v.y, full.z - v.z, full.w - v.w), 1);
                    *reinterpret_cast<float4*>(da + (size_t)xa * Cd + Cp + c4 * 4) = lo;
                    if (xb < W) *reinterpret_cast<float4*>(da + (size_t)xb * Cd + Cp + c4 * 4) = lo;
                    if (db) {
                        *reinterpret_cast<float4*>(db + (size_t)xa * Cd + Cp + c4 * 4) = lo;
                        if (xb < W) *reinterpret_cast<float4*>(db + (size_t)xb * Cd + Cp + c4 * 4) = lo;
                    }
                }
            }
          }
    }
}
// transpose: each low-res cell gathers from the high-res cells whose stencil touches it (deterministic)
__global__ void up_bilinear_bwd_kernel(const float* __restrict__ gu, int B, int H, int W, int Cp, int Hs, int Ws,
                                       const float* __restrict__ r, float slope, float* __restrict__ G, float* __restrict__ gd, int rnd) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int C4 = Cp >> 2;
    int64_t total = (int64_t)B * Hs * Ws * C4;
    if (i >= total) return;
    int c4 = (int)(i % C4);
    int64_t p = i / C4;
    int xs = (int)(p % Ws);
    p /= Ws;
    int ys = (int)(p % Hs);
    int b = (int)(p / Hs);
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    // candidate high-res rows/cols: source coordinate within (ys-1, ys+1)  =>  o in about [2ys-3, 2ys+3]
    for (int y = max(0, 2 * ys - 3); y <= min(H - 1, 2 * ys + 3); y++) {
        int y0, y1; float fy;
        bil_coord(y, Hs, 2 * Hs, y0, y1, fy);
        float wy = (y0 == ys ? (1 - fy) : 0.f) + (y1 == ys ? fy : 0.f);
        if (wy == 0.f) continue;
        for (int x = max(0, 2 * xs - 3); x <= min(W - 1, 2 * xs + 3); x++) {
            int x0, x1; float fx;
            bil_coord(x, Ws, 2 * Ws, x0, x1, fx);
            float wx = (x0 == xs ? (1 - fx) : 0.f) + (x1 == xs ? fx : 0.f);
            if (wx == 0.f) continue;
            float4 v = ldg4(gu + (((size_t)b * H + y) * W + x) * Cp + c4 * 4);
            float w = wy * wx;
            s.x += w * v.x; s.y += w * v.y; s.z += w * v.z; s.w += w * v.w;
        }
    }
    size_t off = (((size_t)b * Hs + ys) * Ws + xs) * Cp + c4 * 4;
    if (G) *reinterpret_cast<float4*>(G + off) = s;
    if (gd) {
        if (r) {
            float4 rv = ldg4(r + off);
            s.x *= act_grad(rv.x, slope); s.y *= act_grad(rv.y, slope); s.z *= act_grad(rv.z, slope); s.w *= act_grad(rv.w, slope);
        }
        *reinterpret_cast<float4*>(gd + off) = round_tf32(s, rnd);
    }
}

// derived weight copies, one layer per blockIdx.y.  Pass A walks the wT index (ci fastest): wT and wK are written coalesced,
// the strided reads of w hit L2.  Pass B walks the wTK index (co fastest): read and write are both coalesced.
__global__ void __launch_bounds__(256) make_wT_kernel(const float* __restrict__ w, float* __restrict__ wT, float* __restrict__ wK, float* __restrict__ wTK,
                                                      const WLayerDesc* __restrict__ tab, int nlayers) {
    const WLayerDesc L = tab[blockIdx.y];
    const unsigned Cin_p = (unsigned)L.Cin_p, Cout_p = (unsigned)L.Cout_p, n = (unsigned)L.n;
    const float* wl = w + L.w_off;
    for (unsigned j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        // j = (tapT*Cout_p + co)*Cin_p + ci
        const unsigned ci = j % Cin_p, t = j / Cin_p, co = t % Cout_p, tapT = t / Cout_p, tap = 8u - tapT;
        const float v = wl[((size_t)tap * Cin_p + ci) * Cout_p + co];
        wT[L.wT_off + j] = v;
        if (wK) wK[L.wk_off + (size_t)co * (9 * Cin_p) + (size_t)tap * Cin_p + ci] = round_tf32(v);
    }
    if (wTK) {
        for (unsigned j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
            // j = (tap*Cin_p + ci)*Cout_p + co  (the layout of w itself)
            const unsigned co = j % Cout_p, t = j / Cout_p, ci = t % Cin_p, tap = t / Cin_p, tapT = 8u - tap;
            wTK[L.wtk_off + (size_t)ci * (9 * Cout_p) + (size_t)tapT * Cout_p + co] = round_tf32(wl[j]);
        }
    }
}

// TF32 mode (all channel counts are multiples of 32): one block = one 32 (ci) x 32 (co) tile of one (layer, tap).  The tile is
// read once, coalesced along co; wTK (co fastest) is written straight from it, wT and wK (ci fastest) from its shared-memory
// transpose: every global access is a full 128-byte line.
__global__ void __launch_bounds__(256) make_wT_tiled_kernel(const float* __restrict__ w, float* __restrict__ wT, float* __restrict__ wK,
                                                            float* __restrict__ wTK, const WLayerDesc* __restrict__ tab, int nlayers) {
    __shared__ float tile[32][33];
    int l = 0, rem = (int)blockIdx.x;
    for (; l < nlayers - 1; l++) {
        const int nt = 9 * (tab[l].Cin_p >> 5) * (tab[l].Cout_p >> 5);
        if (rem < nt) break;
        rem -= nt;
    }
    const WLayerDesc L = tab[l];
    const int Cin_p = L.Cin_p, Cout_p = L.Cout_p, nco = Cout_p >> 5, nci = Cin_p >> 5;
    const int tap = rem / (nci * nco), r2 = rem - tap * (nci * nco), ci0 = (r2 / nco) * 32, co0 = (r2 - (r2 / nco) * nco) * 32;
    const int tapT = 8 - tap;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;        // 8 rows per pass
    const float* wl = w + L.w_off;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int ci = ci0 + ty + 8 * k, co = co0 + tx;
        const float v = wl[((size_t)tap * Cin_p + ci) * Cout_p + co];
        tile[ty + 8 * k][tx] = v;
        wTK[L.wtk_off + (size_t)ci * (9 * Cout_p) + (size_t)tapT * Cout_p + co] = round_tf32(v);
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int co = co0 + ty + 8 * k, ci = ci0 + tx;
        const float v = tile[tx][ty + 8 * k];
        wT[L.wT_off + ((size_t)tapT * Cout_p + co) * Cin_p + ci] = v;
        wK[L.wk_off + (size_t)co * (9 * Cin_p) + (size_t)tap * Cin_p + ci] = round_tf32(v);
    }
}


// ---------------------------------------------------------------------------------------------
// Head conv fused with its producer (src/model.jl:270-272 at l = 1 + src/flux.jl:4): raw = Conv3x3(crop(Upsample2x_nearest(s0 + s1)))
// with NO = 2*noutput output channels, identity activation, planar (WHCN) output.  Exact FP32 on the CUDA cores in every
// precision mode: the layer has 2*noutput columns (a 32-column MMA would be 94 % padding) and its rounding error reaches
// (m, sigma) undamped, so this is where FP32 operands buy the most accuracy.
// One thread = TWO neighbouring LOW-resolution cells = the 2 x 4 output pixels above them.  The (8+2) x (64+2) source tile
// (s0 + s1) sits channel-major in shared memory; per input channel a thread reads its 3 x 4 source values as six 8-byte loads
// and the weights as broadcast vector loads, so the inner loop is FMA-bound, not load-bound.
//   PHASE = true (even H and W): an output pixel of phase (p, q) = (y & 1, x & 1) only sees 2 x 2 DISTINCT source cells, so the
//     nine taps are pre-summed into 4 phases x 4 weights when the block stages them (2.25x fewer FMAs);
//   PHASE = false (odd H or W): the nine taps are applied one by one and a tap that lands on the cropped row / column
//     (output row 2*ys+1 == H) reads zero, as the conv padding does.
constexpr int HU_TW = 64, HU_TH = 8;                        // low-resolution cells per block (x, y); 256 threads x 2 cells
constexpr int HU_RS = HU_TW + 2;                            // row stride of the tile (even: 8-byte loads stay aligned)
constexpr int HU_SP = HU_RS * (HU_TH + 2) + 2;              // channel stride (even)
template <int NO, bool PHASE>
__global__ void __launch_bounds__(256, 2) head_up_conv_kernel(const float* __restrict__ s0, const float* __restrict__ s1, int B, int Hs, int Ws, int Cp, int Cin,
                                                              const float* __restrict__ wgt, int Cin_p, int Cout_p, const float* __restrict__ bias,
                                                              int H, int W, float* __restrict__ out) {
    extern __shared__ __align__(16) float hu_smem[];
    constexpr int NW = (PHASE ? 16 : 9) * NO;                // staged weights per input channel
    float* S = hu_smem;                                      // [Cin][HU_SP]
    float* wS = hu_smem + (((size_t)Cin * HU_SP + 3) & ~(size_t)3);      // [Cin][NW], 16-byte aligned
    dc_pdl_launch_dependents();
    dc_pdl_wait();
    const int tilesx = (Ws + HU_TW - 1) / HU_TW, tilesy = (Hs + HU_TH - 1) / HU_TH;
    int bid = blockIdx.x;
    const int b = bid / (tilesx * tilesy);
    bid -= b * tilesx * tilesy;
    const int ty0 = (bid / tilesx) * HU_TH, tx0 = (bid - (bid / tilesx) * tilesx) * HU_TW;
    const int t = threadIdx.x;
    for (int i = t; i < Cin * NW; i += 256) {
        const int o = i % NO, r = i / NO, e = r % (PHASE ? 16 : 9), ci = r / (PHASE ? 16 : 9);
        const float* wc = wgt + (size_t)ci * Cout_p + o;     // element (tap, ci, o) at wc[tap * Cin_p * Cout_p]
        const size_t ts = (size_t)Cin_p * Cout_p;
        float v;
        if (PHASE) {
            // e = ((p*2 + q)*2 + i)*2 + j: source cell (p + i, q + j) of the 3 x 3 neighbourhood; the taps dy that reach it from an
            // output row of phase p are {0} / {1,2} for p = 0 and {0,1} / {2} for p = 1 (same for columns)
            const int j = e & 1, ii = (e >> 1) & 1, q = (e >> 2) & 1, pp = e >> 3;
            const int dy0 = pp == 0 ? (ii == 0 ? 0 : 1) : (ii == 0 ? 0 : 2), dy1 = pp == 0 ? (ii == 0 ? 0 : 2) : (ii == 0 ? 1 : 2);
            const int dx0 = q == 0 ? (j == 0 ? 0 : 1) : (j == 0 ? 0 : 2), dx1 = q == 0 ? (j == 0 ? 0 : 2) : (j == 0 ? 1 : 2);
            v = 0.f;
            for (int dy = dy0; dy <= dy1; dy++)
                for (int dx = dx0; dx <= dx1; dx++) v += __ldg(wc + (size_t)(dy * 3 + dx) * ts);
        } else {
            v = __ldg(wc + (size_t)e * ts);
        }
        wS[i] = v;
    }
    {   // tile (with a one-cell halo, zero outside the source image), transposed to channel-major.  Eight (pairs of) 16-byte loads
        // are in flight per thread before the first shared-memory store: one load per round trip would make the fill a chain of
        // ~40 memory latencies per thread, which is longer than the arithmetic of the whole block
        const int C4 = (Cin + 3) >> 2;
        constexpr int NP = HU_RS * (HU_TH + 2);
        const int nitems = NP * C4;
        for (int i0 = t; i0 < nitems; i0 += 8 * 256) {
            float4 v[8], u[8];
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const int i = i0 + k * 256;
                v[k] = make_float4(0.f, 0.f, 0.f, 0.f); u[k] = v[k];
                if (i < nitems) {
                    const int c4 = i % C4, pix = i / C4, py = pix / HU_RS, px = pix - py * HU_RS;
                    const int ys = ty0 + py - 1, xs = tx0 + px - 1;
                    if (ys >= 0 && ys < Hs && xs >= 0 && xs < Ws) {
                        const size_t off = (((size_t)b * Hs + ys) * Ws + xs) * Cp + 4 * c4;
                        v[k] = ldg4(s0 + off);
                        if (s1) u[k] = ldg4(s1 + off);
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const int i = i0 + k * 256;
                if (i < nitems) {
                    const int c4 = i % C4, pix = i / C4, c = 4 * c4;
                    S[(size_t)c * HU_SP + pix] = v[k].x + u[k].x;
                    if (c + 1 < Cin) S[(size_t)(c + 1) * HU_SP + pix] = v[k].y + u[k].y;
                    if (c + 2 < Cin) S[(size_t)(c + 2) * HU_SP + pix] = v[k].z + u[k].z;
                    if (c + 3 < Cin) S[(size_t)(c + 3) * HU_SP + pix] = v[k].w + u[k].w;
                }
            }
        }
    }
    __syncthreads();
    const int ly = t >> 5, lx = t & 31;
    const int ys = ty0 + ly, xs = tx0 + 2 * lx;              // this thread's cells: (ys, xs) and (ys, xs + 1)
    float acc[2][2][2][NO];                                  // [cell][p][q][o]
#pragma unroll
    for (int n = 0; n < 2; n++)
#pragma unroll
        for (int p = 0; p < 2; p++)
#pragma unroll
            for (int q = 0; q < 2; q++)
#pragma unroll
                for (int o = 0; o < NO; o++) acc[n][p][q][o] = 0.f;
    const float* Sp = S + ly * HU_RS + 2 * lx;
    const bool rowB = 2 * ys + 1 < H;                        // output row 2*ys+1 exists (not cropped)
    const bool colB0 = 2 * xs + 1 < W, colB1 = 2 * xs + 3 < W;
    for (int ci = 0; ci < Cin; ci++) {
        float v[3][4];                                       // source cells (ys-1+a, xs-1+c), c = 0..3 covers both cells
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const float2 lo = *reinterpret_cast<const float2*>(Sp + (size_t)ci * HU_SP + a * HU_RS);
            const float2 hi = *reinterpret_cast<const float2*>(Sp + (size_t)ci * HU_SP + a * HU_RS + 2);
            v[a][0] = lo.x; v[a][1] = lo.y; v[a][2] = hi.x; v[a][3] = hi.y;
        }
        float w[NW];
        if (NW % 4 == 0) {
#pragma unroll
            for (int e = 0; e < NW / 4; e++) {
                const float4 t4 = *reinterpret_cast<const float4*>(wS + (size_t)ci * NW + 4 * e);
                w[4 * e] = t4.x; w[4 * e + 1] = t4.y; w[4 * e + 2] = t4.z; w[4 * e + 3] = t4.w;
            }
        } else {
#pragma unroll
            for (int e = 0; e < NW / 2; e++) {
                const float2 t2 = *reinterpret_cast<const float2*>(wS + (size_t)ci * NW + 2 * e);
                w[2 * e] = t2.x; w[2 * e + 1] = t2.y;
            }
        }
        if (PHASE) {
#pragma unroll
            for (int n = 0; n < 2; n++)
#pragma unroll
                for (int p = 0; p < 2; p++)
#pragma unroll
                    for (int q = 0; q < 2; q++)
#pragma unroll
                        for (int i = 0; i < 2; i++)
#pragma unroll
                            for (int j = 0; j < 2; j++) {
                                const float x = v[p + i][n + q + j];
#pragma unroll
                                for (int o = 0; o < NO; o++) acc[n][p][q][o] = fmaf(x, w[((((p * 2 + q) * 2 + i) * 2 + j)) * NO + o], acc[n][p][q][o]);
                            }
        } else {
#pragma unroll
            for (int n = 0; n < 2; n++) {
                const bool colB = n == 0 ? colB0 : colB1;
#pragma unroll
                for (int dy = 0; dy < 3; dy++)
#pragma unroll
                    for (int dx = 0; dx < 3; dx++)
#pragma unroll
                        for (int p = 0; p < 2; p++)
#pragma unroll
                            for (int q = 0; q < 2; q++) {
                                // output row 2*ys+p, tap dy-1 -> high-res row 2*ys+p+dy-1 -> source offset floor((p+dy-1)/2) in {-1,0,+1};
                                // (p,dy) = (0,2) lands on row 2*ys+1, which may be cropped (then it is padding: zero)
                                const int a = (p + dy + 1) >> 1, c = (q + dx + 1) >> 1;
                                float x = v[a][n + c];
                                if (p == 0 && dy == 2 && !rowB) x = 0.f;
                                if (q == 0 && dx == 2 && !colB) x = 0.f;
#pragma unroll
                                for (int o = 0; o < NO; o++) acc[n][p][q][o] = fmaf(x, w[(dy * 3 + dx) * NO + o], acc[n][p][q][o]);
                            }
            }
        }
    }
    if (ys >= Hs) return;
#pragma unroll
    for (int o = 0; o < NO; o++) {
        const float bv = bias ? __ldg(bias + o) : 0.f;
#pragma unroll
        for (int p = 0; p < 2; p++) {
            const int y = 2 * ys + p;
            if (y >= H) continue;
            float* dst = out + (((size_t)b * NO + o) * H + y) * W + 2 * xs;
#pragma unroll
            for (int n = 0; n < 2; n++) {
                if (xs + n >= Ws) continue;
                if (PHASE) {                                  // even W: both columns exist, 8-byte aligned
                    *reinterpret_cast<float2*>(dst + 2 * n) = make_float2(acc[n][p][0][o] + bv, acc[n][p][1][o] + bv);
                } else {
                    dst[2 * n] = acc[n][p][0][o] + bv;
                    if (2 * (xs + n) + 1 < W) dst[2 * n + 1] = acc[n][p][1][o] + bv;
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Backward of the fused head (adjoint of head_up_conv_kernel), exact FP32 on the CUDA cores, straight from the planar gradient
// g = d loss / d raw (NO channels at full resolution):
//
// head_bwd_data_kernel: G[i,j,ci] = sum over the (up to) 2 x 2 output pixels (y,x) above low-resolution cell (i,j) and the nine
//   taps of g[y-(dy-1), x-(dx-1), co] * w[tap][ci][co]  = data gradient of the conv followed by the adjoint of nearest up-sampling
//   + crop; GD = G * act'(R) (rounded to TF32: it feeds the tensor-core convs of level 1).  HB_LPC lanes share one cell: each holds the
//   4 x 4 x NO window of g that reaches the cell in registers and accumulates its 32 / HB_LPC channels of a chunk against weights staged
//   in shared memory ([combination][ci], 16-byte broadcast loads).  Replaces: pack(g) + data-gradient conv at full resolution +
//   the 51 MB tensor between them + up-sampling backward.
//   PHASE = true (even H, W): the taps that pair a window pixel with the cell are pre-summed (16 x NO weights per ci);
//   PHASE = false: pixel by pixel with bounds (the cropped row / column of an odd size does not exist).
constexpr int HB_LPC = 1;                                    // lanes that share one cell (each owns 32 / HB_LPC channels of a chunk)
constexpr int HB_THREADS = 128;
constexpr int HB_ROWS = HB_THREADS / HB_LPC / 32;            // block tile: HB_ROWS x 32 cells
template <int NO, bool PHASE>
__global__ void __launch_bounds__(HB_THREADS, 768 / HB_THREADS) head_bwd_data_kernel(const float* __restrict__ g, int B, int H, int W, int Hs, int Ws,
                                                                       const float* __restrict__ wgt, int Cin, int Cin_p, int Cout_p,
                                                                       const float* __restrict__ R, float slope, int Cp,
                                                                       float* __restrict__ G, float* __restrict__ GD, int rnd) {
    extern __shared__ __align__(16) float hb_smem[];
    constexpr int NC = (PHASE ? 16 : 9) * NO;                // weight rows (combinations) per input channel
    constexpr int NQ = 8 / HB_LPC;                           // 16-byte channel groups per lane and chunk: groups sub, sub + HB_LPC, ...
    const int Cr = (Cin + 31) & ~31;                         // row length: whole 32-channel chunks (zero padded)
    dc_pdl_launch_dependents();
    dc_pdl_wait();
    const int t = threadIdx.x;
    for (int i = t; i < NC * Cr; i += HB_THREADS) {
        const int ci = i % Cr, e = i / Cr, co = e % NO, k = e / NO;
        float v = 0.f;
        if (ci < Cin) {
            const float* wc = wgt + (size_t)ci * Cout_p + co;
            const size_t ts = (size_t)Cin_p * Cout_p;
            if (PHASE) {
                // window pixel (a, b) (row 2i-1+a, column 2j-1+b) reaches output pixel (2i+p, 2j+q) through tap dy = p - a + 2
                const int a = k >> 2, b = k & 3;
                for (int pp = 0; pp < 2; pp++)
                    for (int q = 0; q < 2; q++) {
                        const int dy = pp - a + 2, dx = q - b + 2;
                        if (dy >= 0 && dy < 3 && dx >= 0 && dx < 3) v += __ldg(wc + (size_t)(dy * 3 + dx) * ts);
                    }
            } else {
                v = __ldg(wc + (size_t)k * ts);
            }
        }
        hb_smem[i] = v;
    }
    __syncthreads();
    // HB_LPC neighbouring lanes share a cell: together they store whole 64-byte runs of G / GD, and the grid holds HB_LPC times the
    // threads of a thread-per-cell layout (the kernel is bound by memory latency, not by arithmetic)
    const int tilesx = (Ws + 31) / 32, tilesy = (Hs + HB_ROWS - 1) / HB_ROWS;
    int bid = blockIdx.x;
    const int b = bid / (tilesx * tilesy);
    bid -= b * tilesx * tilesy;
    const int sub = t % HB_LPC, ct = t / HB_LPC;
    const int ci_ = (bid / tilesx) * HB_ROWS + (ct >> 5), cj = (bid - (bid / tilesx) * tilesx) * 32 + (ct & 31);
    if (ci_ >= Hs || cj >= Ws) return;
    float gw[4][4][NO];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const int y = 2 * ci_ - 1 + a, x = 2 * cj - 1 + c;
            const bool in = y >= 0 && y < H && x >= 0 && x < W;
#pragma unroll
            for (int o = 0; o < NO; o++) gw[a][c][o] = in ? __ldg(g + (((size_t)b * NO + o) * H + y) * W + x) : 0.f;
        }
    const bool rowB = 2 * ci_ + 1 < H, colB = 2 * cj + 1 < W;
    const size_t cell = ((size_t)b * Hs + ci_) * Ws + cj;
    for (int c0 = 0; c0 < Cp; c0 += 32) {
        const size_t off = cell * Cp + c0 + 4 * sub;
        // act'(R) of this lane's channels as sign bits, fetched BEFORE the arithmetic so that the latency hides behind it
        uint32_t rbits = 0xffffffffu;
        if (GD && R) {
            rbits = 0u;
#pragma unroll
            for (int k = 0; k < NQ; k++) {
                const float4 rv = ldg4(R + off + 4 * HB_LPC * k);
                rbits |= (rv.x > 0.f ? 1u : 0u) << (4 * k) | (rv.y > 0.f ? 1u : 0u) << (4 * k + 1) | (rv.z > 0.f ? 1u : 0u) << (4 * k + 2) |
                         (rv.w > 0.f ? 1u : 0u) << (4 * k + 3);
            }
        }
        float acc[4 * NQ];
#pragma unroll
        for (int k = 0; k < 4 * NQ; k++) acc[k] = 0.f;
        if (c0 < Cr) {
            if (PHASE) {
#pragma unroll
                for (int a = 0; a < 4; a++)
#pragma unroll
                    for (int c = 0; c < 4; c++)
#pragma unroll
                        for (int o = 0; o < NO; o++) {
                            const float x = gw[a][c][o];
                            const float4* wr = reinterpret_cast<const float4*>(hb_smem + (size_t)((a * 4 + c) * NO + o) * Cr + c0) + sub;
#pragma unroll
                            for (int k = 0; k < NQ; k++) {
                                const float4 w4 = wr[HB_LPC * k];
                                acc[4 * k] = fmaf(x, w4.x, acc[4 * k]); acc[4 * k + 1] = fmaf(x, w4.y, acc[4 * k + 1]);
                                acc[4 * k + 2] = fmaf(x, w4.z, acc[4 * k + 2]); acc[4 * k + 3] = fmaf(x, w4.w, acc[4 * k + 3]);
                            }
                        }
            } else {
#pragma unroll
                for (int pp = 0; pp < 2; pp++)
#pragma unroll
                    for (int q = 0; q < 2; q++) {
                        if ((pp == 1 && !rowB) || (q == 1 && !colB)) continue;      // that output pixel was cropped
#pragma unroll
                        for (int dy = 0; dy < 3; dy++)
#pragma unroll
                            for (int dx = 0; dx < 3; dx++)
#pragma unroll
                                for (int o = 0; o < NO; o++) {
                                    const float x = gw[pp - dy + 2][q - dx + 2][o];
                                    const float4* wr = reinterpret_cast<const float4*>(hb_smem + (size_t)((dy * 3 + dx) * NO + o) * Cr + c0) + sub;
#pragma unroll
                                    for (int k = 0; k < NQ; k++) {
                                        const float4 w4 = wr[HB_LPC * k];
                                        acc[4 * k] = fmaf(x, w4.x, acc[4 * k]); acc[4 * k + 1] = fmaf(x, w4.y, acc[4 * k + 1]);
                                        acc[4 * k + 2] = fmaf(x, w4.z, acc[4 * k + 2]); acc[4 * k + 3] = fmaf(x, w4.w, acc[4 * k + 3]);
                                    }
                                }
                    }
            }
        }
#pragma unroll
        for (int k = 0; k < NQ; k++) {
            float4 v = make_float4(acc[4 * k], acc[4 * k + 1], acc[4 * k + 2], acc[4 * k + 3]);
            if (G) *reinterpret_cast<float4*>(G + off + 4 * HB_LPC * k) = v;
            if (GD) {
                const uint32_t m = rbits >> (4 * k);
                v.x *= (m & 1u) ? 1.f : slope; v.y *= (m & 2u) ? 1.f : slope; v.z *= (m & 4u) ? 1.f : slope; v.w *= (m & 8u) ? 1.f : slope;
                *reinterpret_cast<float4*>(GD + off + 4 * HB_LPC * k) = round_tf32(v, rnd);
            }
        }
    }
}

// head_wgrad_kernel: dW[tap][ci][co] = sum over pixels of U[y+dy-1, x+dx-1, ci] * g[y, x, co] with U = crop(upsample(s0 + s1)) never
//   materialised: for a low-resolution cell the source value multiplies Gsum[tap][co] = the sum of g over the (up to four) pixels
//   whose tap lands in that cell, so dW = sum over cells of S[cell][ci] x Gsum[cell][tap][co] -- an outer-product reduction with 4x
//   fewer products than the pixel form.  One block = 8 x 32 cells: thread t < 256 prepares Gsum (and the bias partials) of cell t,
//   the tile of S is staged pixel-major (no transposition), then thread (ci, tap) of each 32-channel chunk walks the 256 cells.
//   Block partials [block][9*Cin*NO + NO] are summed in block order by head_wgrad_reduce_kernel: deterministic.
template <int NO>
__global__ void __launch_bounds__(288) head_wgrad_kernel(const float* __restrict__ s0, const float* __restrict__ s1, const float* __restrict__ g,
                                                          int B, int H, int W, int Hs, int Ws, int Cp, int Cin, float* __restrict__ part) {
    extern __shared__ __align__(16) float hw_smem[];
    constexpr int NG = 9 * NO, NGP = (NG + 3) & ~3;          // Gsum row (padded to 16 bytes)
    const int Cr = (Cin + 31) & ~31;
    float* S = hw_smem;                                      // [256 cells][Cr]
    float* Gs = hw_smem + 256 * Cr;                          // [256 cells][NGP]
    float* bs = Gs + 256 * NGP;                              // [256 cells][NO] bias partials
    const int t = threadIdx.x;
    const int tilesx = (Ws + 31) / 32, tilesy = (Hs + 7) / 8;
    int bid = blockIdx.x;
    const int b = bid / (tilesx * tilesy);
    bid -= b * tilesx * tilesy;
    const int i0 = (bid / tilesx) * 8, j0 = (bid - (bid / tilesx) * tilesx) * 32;
    // ---- S tile: 256 cells x Cr channels, float4 along channels
    {
        const int C4 = Cr >> 2;
        for (int i = t; i < 256 * C4; i += 288) {
            const int c4 = i % C4, cl = i / C4, ci_ = i0 + (cl >> 5), cj = j0 + (cl & 31);
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ci_ < Hs && cj < Ws && 4 * c4 < Cp) {
                const size_t off = (((size_t)b * Hs + ci_) * Ws + cj) * Cp + 4 * c4;
                v = ldg4(s0 + off);
                if (s1) { const float4 u = ldg4(s1 + off); v.x += u.x; v.y += u.y; v.z += u.z; v.w += u.w; }
            }
            *reinterpret_cast<float4*>(S + (size_t)cl * Cr + 4 * c4) = v;
        }
    }
    // ---- Gsum and bias partials of every cell
    if (t < 256) {
        const int ci_ = i0 + (t >> 5), cj = j0 + (t & 31);
        float gw[4][4][NO];
#pragma unroll
        for (int a = 0; a < 4; a++)
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const int y = 2 * ci_ - 1 + a, x = 2 * cj - 1 + c;
                const bool in = ci_ < Hs && cj < Ws && y >= 0 && y < H && x >= 0 && x < W;
#pragma unroll
                for (int o = 0; o < NO; o++) gw[a][c][o] = in ? __ldg(g + (((size_t)b * NO + o) * H + y) * W + x) : 0.f;
            }
        // pixel (y, x) = window (a, c) meets this cell through tap dy iff Y = y + dy - 1 is an existing row of U with Y >> 1 == cell row:
        // Y in {2i, 2i+1} (2i+1 only if it is < H)  =>  a = Y - (dy - 1) - (2i - 1) = (Y - 2i) - dy + 2
#pragma unroll
        for (int dy = 0; dy < 3; dy++)
#pragma unroll
            for (int dx = 0; dx < 3; dx++)
#pragma unroll
                for (int o = 0; o < NO; o++) {
                    float sum = 0.f;
#pragma unroll
                    for (int pp = 0; pp < 2; pp++)
#pragma unroll
                        for (int q = 0; q < 2; q++) {
                            const bool ok = (pp == 0 || 2 * ci_ + 1 < H) && (q == 0 || 2 * cj + 1 < W);
                            sum += ok ? gw[pp - dy + 2][q - dx + 2][o] : 0.f;
                        }
                    Gs[(size_t)t * NGP + (dy * 3 + dx) * NO + o] = sum;
                }
#pragma unroll
        for (int o = 0; o < NO; o++) bs[t * NO + o] = gw[1][1][o] + gw[1][2][o] + gw[2][1][o] + gw[2][2][o];     // the cell's own pixels (zero where outside)
    }
    __syncthreads();
    float* pb = part + (size_t)blockIdx.x * (9 * Cin * NO + NO);
    const int lane_c = t & 31, tap = t >> 5;                 // 288 threads = 32 channels x 9 taps
    for (int c0 = 0; c0 < Cr; c0 += 32) {
        float acc[NO];
#pragma unroll
        for (int o = 0; o < NO; o++) acc[o] = 0.f;
#pragma unroll 4
        for (int cl = 0; cl < 256; cl++) {
            const float sv = S[(size_t)cl * Cr + c0 + lane_c];
#pragma unroll
            for (int o = 0; o < NO; o++) acc[o] = fmaf(sv, Gs[(size_t)cl * NGP + tap * NO + o], acc[o]);
        }
        if (c0 + lane_c < Cin) {
#pragma unroll
            for (int o = 0; o < NO; o++) pb[((size_t)tap * Cin + c0 + lane_c) * NO + o] = acc[o];
        }
    }
    if (t < NO) {
        float sum = 0.f;
        for (int cl = 0; cl < 256; cl++) sum += bs[cl * NO + t];
        pb[9 * Cin * NO + t] = sum;
    }
}

// second stage: one warp per gradient value sums its column of the block partials (lane k takes blocks k, k+32, ... in order, then a
// fixed shuffle tree: deterministic) and writes it at its place in the padded layout [9*Cin_p][Cout_p] / db (zeroed beforehand)
__global__ void __launch_bounds__(256) head_wgrad_reduce_kernel(const float* __restrict__ part, int nblocks, int Cin, int NO, int Cin_p, int Cout_p,
                                                                float* __restrict__ dW, float* __restrict__ db) {
    const int rowlen = 9 * Cin * NO + NO;
    const int v = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (v >= rowlen) return;
    float s = 0.f;
    for (int k = lane; k < nblocks; k += 32) s += part[(size_t)k * rowlen + v];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if (lane == 0) {
        if (v < 9 * Cin * NO) {
            const int co = v % NO, r = v / NO, ci = r % Cin, tap = r / Cin;
            dW[((size_t)tap * Cin_p + ci) * Cout_p + co] = s;
        } else {
            db[v - 9 * Cin * NO] = s;
        }
    }
}

// expanded K-major forward weights of the 3xTF32 layers (layout: SplitLayerDesc in k2_conv.cuh); one layer per blockIdx.y
__global__ void __launch_bounds__(256) make_wK_split_kernel(const float* __restrict__ w, float* __restrict__ wK, const __grid_constant__ SplitTable tab) {
    const SplitLayerDesc L = tab.d[blockIdx.y];
    const int ldk = 9 * L.Cin_pe;
    const int64_t n = (int64_t)L.Cout_p * ldk;
    const float* wl = w + L.w_off;
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
        const int co = (int)(j / ldk), k = (int)(j - (int64_t)co * ldk), tap = k / L.Cin_pe, s = k - tap * L.Cin_pe;
        int ci = -1, cs = co, lo = 0;                   // source channel / output column of the master weight, low-order part?
        if (L.mode == 1) {
            if (s < 2 * L.C) ci = s % L.C;
        } else {
            if (co < L.n2) { if (s < 2 * L.Cin_p) ci = s % L.Cin_p; }
            else if (co >= 8 && co < 8 + L.n2 && s < L.Cin_p) { ci = s; cs = co - 8; lo = 1; }
        }
        float v = 0.f;
        if (ci >= 0) {
            const float x = wl[((size_t)tap * L.Cin_p + ci) * L.Cout_p + cs];
            const float hi = round_tf32(x);
            v = lo ? round_tf32(x - hi) : hi;
        }
        wK[L.wk_off + j] = v;
    }
}
__global__ void __launch_bounds__(256) fold_split_grad_kernel(const float* __restrict__ gexp, float* __restrict__ g, const __grid_constant__ SplitTable tab) {
    const SplitLayerDesc L = tab.d[blockIdx.y];
    const int64_t n = (int64_t)9 * L.Cin_p * L.Cout_p;
    const int nhi = L.mode == 1 ? L.C : L.Cin_p, lo_off = nhi;
    const float* ge = gexp + L.gexp_off;
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
        const int co = (int)(j % L.Cout_p);
        const int64_t t = j / L.Cout_p;
        const int ci = (int)(t % L.Cin_p), tap = (int)(t / L.Cin_p);
        float v = 0.f;
        if (ci < nhi) v = ge[((size_t)tap * L.Cin_pe + ci) * L.Cout_p + co] + ge[((size_t)tap * L.Cin_pe + lo_off + ci) * L.Cout_p + co];
        g[L.w_off + j] = v;
    }
}

}  // namespace

int32_t k2_make_wK_split(dincae_ctx* ctx, const float* w, float* wK, const SplitTable& tab) {
    if (tab.count == 0) return DINCAE_OK;
    DC_LAUNCH(ctx, make_wK_split_kernel, dim3(32u, (unsigned)tab.count), 256, 0, w, wK, tab);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}
int32_t k2_fold_split_grad(dincae_ctx* ctx, const float* gexp, float* g, const SplitTable& tab) {
    if (tab.count == 0) return DINCAE_OK;
    DC_LAUNCH(ctx, fold_split_grad_kernel, dim3(32u, (unsigned)tab.count), 256, 0, gexp, g, tab);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

static size_t head_up_smem_bytes(int Cin, int nout2) {
    return ((((size_t)Cin * HU_SP + 3) & ~(size_t)3) + (size_t)Cin * 16 * nout2) * sizeof(float);
}
bool k2_head_up_conv_eligible(int Cin, int nout2) {
    return (nout2 == 2 || nout2 == 4 || nout2 == 6 || nout2 == 8) && head_up_smem_bytes(Cin, nout2) <= 200 * 1024;
}
int32_t k2_head_up_conv(dincae_ctx* ctx, const float* src0, const float* src1, int B, int Hs, int Ws, int Cp, int Cin,
                        const float* wgt, int Cin_p, int Cout_p, const float* bias, int nout2, int H, int W, float* out,
                        double work_flops, double work_bytes) {
    if ((int64_t)B * H * W == 0) return DINCAE_OK;
    DC_REQUIRE(k2_head_up_conv_eligible(Cin, nout2), DINCAE_EUNSUPPORTED, "k2_head_up_conv: %d input / %d output channels not supported", Cin, nout2);
    DC_REQUIRE(Hs == H / 2 + H % 2 && Ws == W / 2 + W % 2, DINCAE_EINVAL, "k2_head_up_conv: source is not the half-resolution level");
    const int smem = (int)head_up_smem_bytes(Cin, nout2);
    const unsigned grid = (unsigned)(B * ((Ws + HU_TW - 1) / HU_TW) * ((Hs + HU_TH - 1) / HU_TH));
    const bool phase = (H % 2) == 0 && (W % 2) == 0;
    DC_WORK(ctx, work_flops, work_bytes);
#define HU_LAUNCH(NO_, PH_)                                                                                                 \
    do {                                                                                                                    \
        static int attr = 0;                                                                                                \
        auto head_up_conv = head_up_conv_kernel<NO_, PH_>;                                                                  \
        if (smem > attr) { DC_CHECK_CUDA(cudaFuncSetAttribute(head_up_conv, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); attr = smem; } \
        DC_LAUNCH_PDL(ctx, head_up_conv, grid, 256, smem, src0, src1, B, Hs, Ws, Cp, Cin, wgt, Cin_p, Cout_p, bias, H, W, out); \
    } while (0)
#define HU_PICK(NO_) do { if (phase) HU_LAUNCH(NO_, true); else HU_LAUNCH(NO_, false); } while (0)
    if (nout2 == 2) HU_PICK(2);
    else if (nout2 == 4) HU_PICK(4);
    else if (nout2 == 6) HU_PICK(6);
    else HU_PICK(8);
#undef HU_PICK
#undef HU_LAUNCH
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

size_t k2_head_bwd_scratch_bytes(int B, int Hs, int Ws, int Cin, int nout2) {
    const size_t nblocks = (size_t)B * ((Ws + 31) / 32) * ((Hs + 7) / 8);
    return nblocks * (size_t)(9 * Cin * nout2 + nout2) * sizeof(float);
}
bool k2_head_bwd_eligible(int Cin, int nout2) {
    const int Cr = (Cin + 31) & ~31;
    return (nout2 == 2 || nout2 == 4) && (size_t)16 * nout2 * Cr * sizeof(float) <= 96 * 1024 &&
           (size_t)(256 * Cr + 256 * ((9 * nout2 + 3) & ~3) + 256 * nout2) * sizeof(float) <= 200 * 1024;
}
int32_t k2_head_bwd_data(dincae_ctx* ctx, const float* g, int B, int H, int W, int Hs, int Ws, const float* wgt, int Cin, int Cin_p, int Cout_p,
                         int nout2, const float* R, float slope, int Cp, float* G, float* GD) {
    if ((int64_t)B * Hs * Ws == 0) return DINCAE_OK;
    DC_REQUIRE(k2_head_bwd_eligible(Cin, nout2), DINCAE_EUNSUPPORTED, "k2_head_bwd_data: %d input / %d output channels not supported", Cin, nout2);
    DC_REQUIRE((Cp % 32) == 0, DINCAE_EINVAL, "k2_head_bwd_data: %d channels are not whole 32-channel chunks", Cp);
    const bool phase = (H % 2) == 0 && (W % 2) == 0;
    const int Cr = (Cin + 31) & ~31;
    const int smem = (int)((size_t)(phase ? 16 : 9) * nout2 * Cr * sizeof(float));
    const unsigned grid = (unsigned)(B * ((Ws + 31) / 32) * ((Hs + HB_ROWS - 1) / HB_ROWS));
    DC_WORK(ctx, 2.0 * 9 * Cin * nout2 * (double)B * H * W, 4.0 * ((double)nout2 * B * H * W + (double)Cp * B * Hs * Ws * ((G ? 1 : 0) + (GD ? 1 : 0) + (R ? 1 : 0))));
#define HB_LAUNCH(NO_, PH_)                                                                                                  \
    do {                                                                                                                     \
        static int attr = 0;                                                                                                 \
        auto head_bwd_data = head_bwd_data_kernel<NO_, PH_>;                                                                 \
        if (smem > attr) { DC_CHECK_CUDA(cudaFuncSetAttribute(head_bwd_data, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); attr = smem; } \
        DC_LAUNCH_PDL(ctx, head_bwd_data, grid, HB_THREADS, smem, g, B, H, W, Hs, Ws, wgt, Cin, Cin_p, Cout_p, R, slope, Cp, G, GD, (int)ctx->tf32_acts); \
    } while (0)
    if (nout2 == 2) { if (phase) HB_LAUNCH(2, true); else HB_LAUNCH(2, false); }
    else { if (phase) HB_LAUNCH(4, true); else HB_LAUNCH(4, false); }
#undef HB_LAUNCH
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}
int32_t k2_head_wgrad(dincae_ctx* ctx, const float* src0, const float* src1, const float* g, int B, int H, int W, int Hs, int Ws, int Cp, int Cin,
                      int nout2, int Cin_p, int Cout_p, float* dW, float* db, void* scratch, size_t scratch_bytes) {
    if ((int64_t)B * Hs * Ws == 0) return DINCAE_OK;
    DC_REQUIRE(k2_head_bwd_eligible(Cin, nout2), DINCAE_EUNSUPPORTED, "k2_head_wgrad: %d input / %d output channels not supported", Cin, nout2);
    DC_REQUIRE(scratch_bytes >= k2_head_bwd_scratch_bytes(B, Hs, Ws, Cin, nout2), DINCAE_ESIZE, "k2_head_wgrad: scratch too small");
    const int Cr = (Cin + 31) & ~31;
    const int smem = (int)((size_t)(256 * Cr + 256 * ((9 * nout2 + 3) & ~3) + 256 * nout2) * sizeof(float));
    const unsigned grid = (unsigned)(B * ((Ws + 31) / 32) * ((Hs + 7) / 8));
    DC_WORK(ctx, 2.0 * 9 * Cin * nout2 * (double)B * H * W, 4.0 * ((double)nout2 * B * H * W + (double)Cin * B * Hs * Ws * (src1 ? 2 : 1)));
#define HW_LAUNCH(NO_)                                                                                                       \
    do {                                                                                                                     \
        static int attr = 0;                                                                                                 \
        auto head_wgrad = head_wgrad_kernel<NO_>;                                                                            \
        if (smem > attr) { DC_CHECK_CUDA(cudaFuncSetAttribute(head_wgrad, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); attr = smem; } \
        DC_LAUNCH(ctx, head_wgrad, grid, 288, smem, src0, src1, g, B, H, W, Hs, Ws, Cp, Cin, (float*)scratch);               \
    } while (0)
    if (nout2 == 2) HW_LAUNCH(2); else HW_LAUNCH(4);
#undef HW_LAUNCH
    DC_CHECK_LAUNCH();
    DC_CHECK_CUDA(cudaMemsetAsync(dW, 0, sizeof(float) * 9 * (size_t)Cin_p * Cout_p, ctx->stream));        // the padding entries stay exactly zero
    DC_CHECK_CUDA(cudaMemsetAsync(db, 0, sizeof(float) * Cout_p, ctx->stream));
    DC_LAUNCH(ctx, head_wgrad_reduce_kernel, (unsigned)dc_cdiv(9 * Cin * nout2 + nout2, 8), 256, 0, (const float*)scratch, (int)grid, Cin, nout2, Cin_p, Cout_p, dW, db);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

// ---------------------------------------------------------------------------------------------
int32_t k2_conv(dincae_ctx* ctx, const ConvArgs& a) {
    int Mtot = a.B * a.H * a.W;
    if (Mtot == 0) return DINCAE_OK;
    if (k2_conv_tma_eligible(a)) return k2_conv_tma(ctx, a);
    DC_REQUIRE(a.planar_out == nullptr, DINCAE_EUNSUPPORTED, "k2_conv: planar output needs the TMA tensor-core kernel (driver without cuTensorMapEncodeTiled?)");
    int Kpad = dc_round_up(9 * a.in.C, 16);
    int nchunks = Kpad / 16;
    DC_WORK(ctx, a.work_flops > 0 ? a.work_flops : 2.0 * 9 * a.in.C * a.Cout_p * (double)Mtot, a.work_bytes);
    if (a.Cout_p <= 32) {
        dim3 grid((unsigned)dc_cdiv(Mtot, 128), (unsigned)dc_cdiv(a.Cout_p, 32));
        DC_LAUNCH(ctx, conv3x3_simt<32>, grid, 256, 0, a, Mtot, nchunks);
    } else {
        dim3 grid((unsigned)dc_cdiv(Mtot, 128), (unsigned)dc_cdiv(a.Cout_p, 64));
        DC_LAUNCH(ctx, conv3x3_simt<64>, grid, 256, 0, a, Mtot, nchunks);
    }
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

static void wgrad_plan(int num_sms, int Mtot, int Krows, int Cout_p, int& Kt, int& Nt, int& splits, int64_t& pps) {
    Kt = (int)dc_cdiv(Krows, 64);
    Nt = (int)dc_cdiv(Cout_p, 64);
    int tiles = Kt * Nt;
    splits = (4 * num_sms + tiles - 1) / tiles;
    int maxs = (int)dc_cdiv(Mtot, 256);
    if (splits > maxs) splits = maxs;
    if (splits < 1) splits = 1;
    pps = dc_cdiv(dc_cdiv(Mtot, splits), 16) * 16;
    splits = (int)dc_cdiv(Mtot, pps);
}

size_t k2_wgrad_scratch_bytes(int num_sms, int B, int H, int W, int Cin_p, int Cout_p) {
    int Kt, Nt, splits;
    int64_t pps;
    wgrad_plan(num_sms, B * H * W, 9 * Cin_p, Cout_p, Kt, Nt, splits, pps);
    size_t big = (size_t)splits * ((size_t)Kt * 64 * Nt * 64 + (size_t)Nt * 64) * sizeof(float);
    // small path: one block per pixel range
    size_t small = (size_t)(8 * num_sms) * (size_t)(9 * Cin_p + 1) * 8 * sizeof(float);
    return big > small ? big : small;
}

int32_t k2_wgrad(dincae_ctx* ctx, WgradArgs a, void* scratch, size_t scratch_bytes) {
    if (k2_wgrad_tma_eligible(a)) return k2_wgrad_tma(ctx, a, scratch, scratch_bytes);
    int Mtot = a.B * a.H * a.W;
    int Krows = 9 * a.in.C;
    int Kt, Nt, splits;
    int64_t pps;
    wgrad_plan(ctx->num_sms, Mtot, Krows, a.Cout_p, Kt, Nt, splits, pps);
    a.splits = splits;
    a.pix_per_split = pps;
    a.partial = (float*)scratch;
    size_t stride = (size_t)Kt * 64 * Nt * 64 + (size_t)Nt * 64;
    if (splits > 1)
        DC_REQUIRE(scratch_bytes >= stride * splits * sizeof(float), DINCAE_ESIZE, "k2_wgrad: scratch too small");
    dim3 grid(Kt, Nt, splits);
    DC_WORK(ctx, a.work_flops > 0 ? a.work_flops : 2.0 * Krows * a.Cout_p * (double)Mtot, a.work_bytes);
    DC_LAUNCH(ctx, wgrad_simt, grid, 256, 0, a, Mtot, Krows, Kt, Nt);
    DC_CHECK_LAUNCH();
    if (splits > 1) {
        int64_t n = (int64_t)Krows * a.Cout_p + a.Cout_p;
        DC_LAUNCH(ctx, wgrad_reduce, (unsigned)dc_cdiv(n, 256), 256, 0, a.partial, splits, stride, Nt * 64, Krows, a.Cout_p, a.dW, a.db,
                  (size_t)Kt * 64 * Nt * 64);
        DC_CHECK_LAUNCH();
    }
    return DINCAE_OK;
}

int32_t k2_conv_small(dincae_ctx* ctx, const ConvSmallArgs& a, int nco) {
    int Mtot = a.B * a.H * a.W;
    if (Mtot == 0) return DINCAE_OK;
    int Krows = 9 * a.in.C;
    size_t smem = (size_t)Krows * nco * sizeof(float);
    DC_REQUIRE(smem <= 200 * 1024, DINCAE_EUNSUPPORTED, "k2_conv_small: %d input channels do not fit shared memory", a.in.C);
    DC_WORK(ctx, a.work_flops > 0 ? a.work_flops : 2.0 * Krows * a.Cout * (double)Mtot, a.work_bytes);
    if (nco == 4) {
        if (smem > 48 * 1024) DC_CHECK_CUDA(cudaFuncSetAttribute(conv3x3_small<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        DC_LAUNCH(ctx, conv3x3_small<4>, (unsigned)dc_cdiv(Mtot, 256), 256, smem, a, Mtot, Krows);
    } else if (nco == 8) {
        if (smem > 48 * 1024) DC_CHECK_CUDA(cudaFuncSetAttribute(conv3x3_small<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        DC_LAUNCH(ctx, conv3x3_small<8>, (unsigned)dc_cdiv(Mtot, 256), 256, smem, a, Mtot, Krows);
    } else {
        DC_REQUIRE(false, DINCAE_EUNSUPPORTED, "k2_conv_small: nco=%d", nco);
    }
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

int32_t k2_wgrad_small(dincae_ctx* ctx, WgradArgs a, int nco, void* scratch, size_t scratch_bytes) {
    int Mtot = a.B * a.H * a.W;
    int Krows = 9 * a.in.C;
    int splits = 8 * ctx->num_sms;
    int64_t pps = dc_cdiv(Mtot, splits);
    if (pps < 64) pps = 64;
    splits = (int)dc_cdiv(Mtot, pps);
    DC_REQUIRE(scratch_bytes >= (size_t)splits * (Krows + 1) * nco * sizeof(float), DINCAE_ESIZE, "k2_wgrad_small: scratch too small");
    a.partial = (float*)scratch;
    a.pix_per_split = pps;
    a.splits = splits;
    int threads = Krows + 1 < 256 ? dc_round_up(Krows + 1, 32) : 256;
    DC_WORK(ctx, a.work_flops > 0 ? a.work_flops : 2.0 * Krows * a.Cout_p * (double)Mtot, a.work_bytes);
    if (nco == 4) DC_LAUNCH(ctx, wgrad_small<4>, splits, threads, 0, a, Mtot, Krows);
    else if (nco == 8) DC_LAUNCH(ctx, wgrad_small<8>, splits, threads, 0, a, Mtot, Krows);
    else DC_REQUIRE(false, DINCAE_EUNSUPPORTED, "k2_wgrad_small: nco=%d", nco);
    DC_CHECK_LAUNCH();
    DC_LAUNCH(ctx, wgrad_small_reduce, (unsigned)dc_cdiv((Krows + 1) * nco, 256), 256, 0, a.partial, splits, Krows, nco, a.Cout_p, a.dW, a.db);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

// grid of the row-walking level kernels: one block per pixel row, capped at 8 resident blocks per SM
static int k2_row_blocks(dincae_ctx* ctx, int64_t rows, int64_t n_elems4) {
    (void)n_elems4;
    int64_t cap = 8 * (int64_t)ctx->num_sms;
    return (int)(rows < cap ? rows : cap);
}

int32_t k2_pack(dincae_ctx* ctx, const float* src0, int C0, const float* src1, int C1, int B, int H, int W, float* dst, int Cp, int split) {
    int64_t M = (int64_t)B * H * W;
    if (M == 0) return DINCAE_OK;
    DC_REQUIRE(!split || (ctx->tf32_acts && 2 * (C0 + C1) <= Cp), DINCAE_EINVAL, "k2_pack: split layout needs TF32 mode and Cp >= 2*C");
    int Cw = (int)dc_round_up((split ? 2 : 1) * (C0 + C1), 8);                  // whole 32-byte sectors; the rest of the row is never written (zero since allocation)
    if (Cw > Cp) Cw = Cp;
    DC_WORK(ctx, 0.0, 4.0 * M * ((C0 + C1) + Cw));                 // real planes in, written channels out
    DC_LAUNCH_PDL(ctx, pack_planar_kernel, (unsigned)dc_cdiv(M, 256), 256, 0, src0, C0, src1, C1, B, H * W, dst, Cp, Cw, (int)ctx->tf32_acts, split);   // 8 warps x 32 pixels per block
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

int32_t k2_pool(dincae_ctx* ctx, const float* a, int B, int H, int W, int Cp, float* x) {
    int H2 = H / 2 + H % 2, W2 = W / 2 + W % 2;
    int64_t n = (int64_t)B * H2 * W2 * (Cp / 4);
    if (n == 0) return DINCAE_OK;
    DC_WORK(ctx, 0.0, 4.0 * Cp * ((double)B * H * W + (double)B * H2 * W2));
    DC_LAUNCH_PDL(ctx, pool_kernel, (unsigned)k2_row_blocks(ctx, (int64_t)B * H2, n), 256, 0, a, B, H, W, Cp, H2, W2, x, (int)ctx->tf32_acts);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

// number of blocks the producers use when they also emit bias column partials (0 = not supported for this channel count)
int k2_colpart_blocks(dincae_ctx* ctx, int64_t n_elems4, int Cp) {
    int C4 = Cp >> 2;
    if (C4 < 1 || C4 > 256 || (256 % C4) != 0) return 0;
    int64_t nb = dc_cdiv(n_elems4, 256);
    int64_t cap = 8 * (int64_t)ctx->num_sms;
    return (int)(nb < cap ? nb : cap);
}

int32_t k2_upbwd(dincae_ctx* ctx, const float* gu, int B, int H, int W, int Cp, const float* r, float slope, float* G, float* gd, float* colpart) {
    int Hs = H / 2 + H % 2, Ws = W / 2 + W % 2;
    int64_t n = (int64_t)B * Hs * Ws * (Cp / 4);
    if (n == 0) return DINCAE_OK;
    int nb = colpart ? k2_colpart_blocks(ctx, n, Cp) : k2_row_blocks(ctx, (int64_t)B * Hs, n);
    DC_REQUIRE(nb > 0, DINCAE_EINVAL, "k2_upbwd: column partials not supported for %d channels", Cp);
    DC_WORK(ctx, 0.0, 4.0 * Cp * ((double)B * H * W + (double)B * Hs * Ws * ((r ? 1 : 0) + (G ? 1 : 0) + (gd ? 1 : 0))));   // gu in; r in, G and gd out at half resolution
    DC_LAUNCH_PDL(ctx, upbwd_kernel, nb, 256, 0, gu, B, H, W, Cp, Hs, Ws, r, slope, G, gd, (int)ctx->tf32_acts, colpart);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

int32_t k2_poolbwd(dincae_ctx* ctx, const float* xg, const float* a, int B, int H, int W, int Cp, float slope, float* gc, float* colpart) {
    int H2 = H / 2 + H % 2, W2 = W / 2 + W % 2;
    int64_t n = (int64_t)B * H * W * (Cp / 4);
    if (n == 0) return DINCAE_OK;
    int nb = colpart ? k2_colpart_blocks(ctx, n, Cp) : k2_row_blocks(ctx, (int64_t)B * H, n);
    DC_REQUIRE(nb > 0, DINCAE_EINVAL, "k2_poolbwd: column partials not supported for %d channels", Cp);
    DC_WORK(ctx, 0.0, 4.0 * Cp * ((double)B * H * W * 2 + (double)B * H2 * W2));       // a in, gc out, xg in at half resolution
    DC_LAUNCH_PDL(ctx, poolbwd_kernel, nb, 256, 0, xg, a, (const uint32_t*)nullptr, B, H, W, Cp, H2, W2, slope, gc, (int)ctx->tf32_acts, colpart);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

int32_t k2_poolbwd_mask(dincae_ctx* ctx, const float* xg, const uint32_t* mask, int B, int H, int W, int Cp, float slope, float* gc, float* colpart) {
    int H2 = H / 2 + H % 2, W2 = W / 2 + W % 2;
    int64_t n = (int64_t)B * H * W * (Cp / 4);
    if (n == 0) return DINCAE_OK;
    DC_REQUIRE((Cp % 32) == 0, DINCAE_EINVAL, "k2_poolbwd_mask: %d channels are not whole mask words", Cp);
    int nb = colpart ? k2_colpart_blocks(ctx, n, Cp) : k2_row_blocks(ctx, (int64_t)B * H, n);
    DC_REQUIRE(nb > 0, DINCAE_EINVAL, "k2_poolbwd_mask: column partials not supported for %d channels", Cp);
    DC_WORK(ctx, 0.0, 4.0 * Cp * ((double)B * H * W * (1.0 + 1.0 / 32) + (double)B * H2 * W2));     // gc out, sign bits + quarter-resolution xg in
    DC_LAUNCH_PDL(ctx, poolbwd_kernel, nb, 256, 0, xg, (const float*)nullptr, mask, B, H, W, Cp, H2, W2, slope, gc, (int)ctx->tf32_acts, colpart);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

int32_t k2_up_bilinear(dincae_ctx* ctx, const float* src0, const float* src1, int B, int Hs, int Ws, int Cp, int H, int W, float* dst, int split) {
    int64_t n = (int64_t)B * H * W * (Cp / 4);
    if (n == 0) return DINCAE_OK;
    DC_REQUIRE(!split || ctx->tf32_acts, DINCAE_EINVAL, "k2_up_bilinear: split layout needs TF32 mode");
    DC_LAUNCH(ctx, up_bilinear_kernel, (unsigned)dc_cdiv(n, 256), 256, 0, src0, src1, B, Hs, Ws, Cp, H, W, dst, (int)ctx->tf32_acts, split);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

int32_t k2_up_nearest(dincae_ctx* ctx, const float* src0, const float* src1, int B, int Hs, int Ws, int Cp, int H, int W, float* dst, int split) {
    int64_t n = (int64_t)B * H * W * (Cp / 4);
    if (n == 0) return DINCAE_OK;
    DC_REQUIRE(!split || ctx->tf32_acts, DINCAE_EINVAL, "k2_up_nearest: split layout needs TF32 mode");
    DC_WORK(ctx, 0.0, 4.0 * Cp * ((double)B * Hs * Ws * (src1 ? 2 : 1) + (double)B * H * W * (split ? 2 : 1)));
    DC_LAUNCH_PDL(ctx, up_nearest_kernel, (unsigned)k2_row_blocks(ctx, (int64_t)B * Hs, n / 4), 256, 0, src0, src1, B, Hs, Ws, Cp, H, W, dst, (int)ctx->tf32_acts, split);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

int32_t k2_up_bilinear_bwd(dincae_ctx* ctx, const float* gu, int B, int H, int W, int Cp, int Hs, int Ws, const float* r, float slope,
                           float* G, float* gd) {
    int64_t n = (int64_t)B * Hs * Ws * (Cp / 4);
    if (n == 0) return DINCAE_OK;
    DC_LAUNCH(ctx, up_bilinear_bwd_kernel, (unsigned)dc_cdiv(n, 256), 256, 0, gu, B, H, W, Cp, Hs, Ws, r, slope, G, gd, (int)ctx->tf32_acts);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}

int32_t k2_make_wT(dincae_ctx* ctx, const float* w, float* wT, float* wK, float* wTK, const WLayerDesc* d_table, int nlayers, int64_t total) {
    if (total == 0 || nlayers == 0) return DINCAE_OK;
    if (wK && wTK && (total % 1024) == 0) {                 // TF32 mode: every layer is padded to multiples of 32 channels
        DC_LAUNCH(ctx, make_wT_tiled_kernel, (unsigned)(total / 1024), 256, 0, w, wT, wK, wTK, d_table, nlayers);
        DC_CHECK_LAUNCH();
        return DINCAE_OK;
    }
    dim3 grid(48u, (unsigned)nlayers);                      // layers differ in size; 48 x 256 threads stride over each
    DC_LAUNCH(ctx, make_wT_kernel, grid, 256, 0, w, wT, wK, wTK, d_table, nlayers);
    DC_CHECK_LAUNCH();
    return DINCAE_OK;
}
