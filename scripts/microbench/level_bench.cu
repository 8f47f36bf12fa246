// Stand-alone timing of layout variants of the pool-backward level kernel (not part of the library):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o scripts/microbench/level_bench scripts/microbench/level_bench.cu
// gc[b,y,x,c] = 0.25 * xg[b,y/2,x/2,c] * (bit(mask[b,y,x], c) ? 1 : slope); block column sums of gc -> colpart.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)
__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 scale(float4 g, uint32_t bits, float slope) {
    g.x *= 0.25f * ((bits & 1u) ? 1.f : slope); g.y *= 0.25f * ((bits & 2u) ? 1.f : slope);
    g.z *= 0.25f * ((bits & 4u) ? 1.f : slope); g.w *= 0.25f * ((bits & 8u) ? 1.f : slope);
    return g;
}
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
// V0: one block walks output rows (the library's layout): thread = (c4, lane), x = lane, lane + PL, ...
template <int MINB>
__global__ void __launch_bounds__(256, MINB) v0_rows(const float* __restrict__ xg, const uint32_t* __restrict__ mask, int B, int H, int W, int Cp,
                                                     float slope, float* __restrict__ gc, float* __restrict__ colpart) {
    const int C4 = Cp >> 2, CW = C4 < 256 ? C4 : 256, PL = 256 / CW, c4 = threadIdx.x % CW, lane = threadIdx.x / CW;
    const int H2 = H / 2, W2 = W / 2;
    float4 cacc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int row = blockIdx.x; row < B * H; row += gridDim.x) {
        const int b = row / H, y = row - b * H;
        const float* rx = xg + ((size_t)b * H2 + (y >> 1)) * W2 * Cp;
        const size_t ro = (size_t)row * W * Cp;
        for (int x = lane; x < W; x += PL) {
            float4 g = ldg4(rx + (size_t)(x >> 1) * Cp + c4 * 4);
            const uint32_t bits = __ldg(mask + ((size_t)row * W + x) * (Cp >> 5) + (c4 >> 3)) >> ((c4 & 7) * 4);
            g = scale(g, bits, slope);
            *reinterpret_cast<float4*>(gc + ro + (size_t)x * Cp + c4 * 4) = g;
            cacc.x += g.x; cacc.y += g.y; cacc.z += g.z; cacc.w += g.w;
        }
    }
    block_colpart(cacc, C4, Cp, colpart);
}
// V3: flat over SOURCE float4 (one xg load + 4 mask words -> 4 stores), grid-stride; total threads are a multiple of C4
template <int MINB>
__global__ void __launch_bounds__(256, MINB) v3_src(const float* __restrict__ xg, const uint32_t* __restrict__ mask, int B, int H, int W, int Cp,
                                                    float slope, float* __restrict__ gc, float* __restrict__ colpart) {
    const int C4 = Cp >> 2, H2 = H / 2, W2 = W / 2;
    const unsigned n = (unsigned)(B * H2 * W2 * C4), T = gridDim.x * 256u;
    float4 cacc = make_float4(0.f, 0.f, 0.f, 0.f);
    const int mw = Cp >> 5;
    for (unsigned i = blockIdx.x * 256u + threadIdx.x; i < n; i += T) {
        const int c4 = (int)(i % C4);
        unsigned p = i / C4;
        const int x2 = (int)(p % W2); p /= W2;
        const int y2 = (int)(p % H2); const int b = (int)(p / H2);
        const float4 g = ldg4(xg + (size_t)i * 4);
        const size_t pa = ((size_t)b * H + 2 * y2) * W + 2 * x2, pb = pa + W;
        const int sh = (c4 & 7) * 4, wi = c4 >> 3;
        const uint32_t b00 = __ldg(mask + pa * mw + wi) >> sh, b01 = __ldg(mask + (pa + 1) * mw + wi) >> sh;
        const uint32_t b10 = __ldg(mask + pb * mw + wi) >> sh, b11 = __ldg(mask + (pb + 1) * mw + wi) >> sh;
        const float4 o00 = scale(g, b00, slope), o01 = scale(g, b01, slope), o10 = scale(g, b10, slope), o11 = scale(g, b11, slope);
        *reinterpret_cast<float4*>(gc + pa * Cp + c4 * 4) = o00;
        *reinterpret_cast<float4*>(gc + (pa + 1) * Cp + c4 * 4) = o01;
        *reinterpret_cast<float4*>(gc + pb * Cp + c4 * 4) = o10;
        *reinterpret_cast<float4*>(gc + (pb + 1) * Cp + c4 * 4) = o11;
        cacc.x += (o00.x + o01.x) + (o10.x + o11.x); cacc.y += (o00.y + o01.y) + (o10.y + o11.y);
        cacc.z += (o00.z + o01.z) + (o10.z + o11.z); cacc.w += (o00.w + o01.w) + (o10.w + o11.w);
    }
    block_colpart(cacc, C4, Cp, colpart);
}
// V4: flat over OUTPUT float4, grid-stride, U elements per iteration with the loads first
template <int U, int MINB>
__global__ void __launch_bounds__(256, MINB) v4_out(const float* __restrict__ xg, const uint32_t* __restrict__ mask, int B, int H, int W, int Cp,
                                                    float slope, float* __restrict__ gc, float* __restrict__ colpart) {
    const int C4 = Cp >> 2, H2 = H / 2, W2 = W / 2, mw = Cp >> 5;
    const unsigned n = (unsigned)(B * H * W * C4), T = gridDim.x * 256u;
    float4 cacc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (unsigned i0 = blockIdx.x * 256u + threadIdx.x; i0 < n; i0 += T * U) {
        float4 g[U]; uint32_t bits[U];
#pragma unroll
        for (int k = 0; k < U; k++) {
            const unsigned i = i0 + k * T;
            if (i < n) {
                const int c4 = (int)(i % C4);
                unsigned p = i / C4;
                const int x = (int)(p % W); unsigned q = p / W;
                const int y = (int)(q % H); const int b = (int)(q / H);
                g[k] = ldg4(xg + (((size_t)b * H2 + (y >> 1)) * W2 + (x >> 1)) * Cp + c4 * 4);
                bits[k] = __ldg(mask + (size_t)p * mw + (c4 >> 3)) >> ((c4 & 7) * 4);
            }
        }
#pragma unroll
        for (int k = 0; k < U; k++) {
            const unsigned i = i0 + k * T;
            if (i < n) {
                const float4 o = scale(g[k], bits[k], slope);
                *reinterpret_cast<float4*>(gc + (size_t)i * 4) = o;
                cacc.x += o.x; cacc.y += o.y; cacc.z += o.z; cacc.w += o.w;
            }
        }
    }
    block_colpart(cacc, C4, Cp, colpart);
}
// pure streaming write of the same volume (upper bound of what the memory system does for this shape)
__global__ void __launch_bounds__(256) v9_fill(float* __restrict__ gc, int64_t n4) {
    const int64_t T = (int64_t)gridDim.x * 256;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += T) reinterpret_cast<float4*>(gc)[i] = make_float4(1.f, 2.f, 3.f, 4.f);
}

template <class F> static float time_us(F f, int reps = 50) {
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    for (int i = 0; i < 5; i++) f();
    CK(cudaEventRecord(a));
    for (int i = 0; i < reps; i++) f();
    CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    CK(cudaGetLastError());
    return ms * 1000.f / reps;
}
int main() {
    int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int B = 32;
    const int shapes[4][2] = {{112, 32}, {56, 64}, {28, 128}, {14, 256}};
    float *xg, *gc, *colpart; uint32_t* mask; float* ref;
    const size_t nmax = (size_t)B * 112 * 112 * 32;
    CK(cudaMalloc(&xg, nmax)); CK(cudaMalloc(&gc, nmax * 4)); CK(cudaMalloc(&ref, nmax * 4)); CK(cudaMalloc(&mask, nmax / 8)); CK(cudaMalloc(&colpart, 16384 * 256 * 4));
    std::vector<float> h(nmax / 4); for (size_t i = 0; i < h.size(); i++) h[i] = (float)((i * 2654435761u) % 1000) / 500.f - 1.f;
    CK(cudaMemcpy(xg, h.data(), nmax, cudaMemcpyHostToDevice));
    std::vector<uint32_t> hm(nmax / 32); for (size_t i = 0; i < hm.size(); i++) hm[i] = (uint32_t)(i * 2246822519u) ^ (uint32_t)(i >> 3);
    CK(cudaMemcpy(mask, hm.data(), nmax / 8, cudaMemcpyHostToDevice));
    for (int s = 0; s < 4; s++) {
        const int H = shapes[s][0], W = H, Cp = shapes[s][1], C4 = Cp / 4;
        const int64_t n = (int64_t)B * H * W * C4, nsrc = n / 4;
        const double mb = (n * 16.0 + nsrc * 16.0 + n * 16.0 / 32) / 1e6;
        printf("== %dx%d x %d channels: %.1f MB algorithmic, %.2f us at 6540 GB/s\n", H, W, Cp, mb, mb / 6540e3 * 1e6);
        auto check = [&](const char* name) {
            std::vector<float> a(n * 4), b(n * 4);
            CK(cudaMemcpy(a.data(), gc, n * 16, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(b.data(), ref, n * 16, cudaMemcpyDeviceToHost));
            size_t bad = 0; for (size_t i = 0; i < a.size(); i++) bad += a[i] != b[i];
            if (bad) printf("   !! %s differs from v0 in %zu values\n", name, bad);
        };
        const int rows = B * H;
        auto rb = [&](int per_sm) { int cap = per_sm * sms; return rows < cap ? rows : cap; };
        auto fb = [&](int64_t cnt, int per_sm) { int64_t nb = (cnt + 255) / 256, cap = (int64_t)per_sm * sms; return (int)(nb < cap ? nb : cap); };
        float t;
        t = time_us([&] { v0_rows<1><<<rb(8), 256>>>(xg, mask, B, H, W, Cp, 0.f, ref, colpart); }); printf("   v0 rows, grid %5d (as shipped)        %7.2f us\n", rb(8), t);
        t = time_us([&] { v0_rows<8><<<rb(8), 256>>>(xg, mask, B, H, W, Cp, 0.f, gc, colpart); }); check("v0<8>"); printf("   v0 rows, 8 blocks/SM forced            %7.2f us\n", t);
        t = time_us([&] { v0_rows<1><<<rb(6), 256>>>(xg, mask, B, H, W, Cp, 0.f, gc, colpart); }); printf("   v0 rows, grid %5d                    %7.2f us\n", rb(6), t);
        t = time_us([&] { v0_rows<1><<<rows, 256>>>(xg, mask, B, H, W, Cp, 0.f, gc, colpart); }); printf("   v0 rows, one row per block (%5d)     %7.2f us\n", rows, t);
        for (int per : {4, 8, 16}) {
            t = time_us([&] { v3_src<1><<<fb(nsrc, per), 256>>>(xg, mask, B, H, W, Cp, 0.f, gc, colpart); }); check("v3");
            printf("   v3 per source, grid %5d              %7.2f us\n", fb(nsrc, per), t);
        }
        t = time_us([&] { v3_src<1><<<(unsigned)((nsrc + 255) / 256), 256>>>(xg, mask, B, H, W, Cp, 0.f, gc, colpart); }); printf("   v3 per source, one element per thread  %7.2f us\n", t);
        for (int per : {4, 8}) {
            t = time_us([&] { v4_out<1, 1><<<fb(n, per), 256>>>(xg, mask, B, H, W, Cp, 0.f, gc, colpart); }); check("v4<1>");
            printf("   v4 per output U=1, grid %5d          %7.2f us\n", fb(n, per), t);
            t = time_us([&] { v4_out<4, 1><<<fb(n, per), 256>>>(xg, mask, B, H, W, Cp, 0.f, gc, colpart); }); check("v4<4>");
            printf("   v4 per output U=4, grid %5d          %7.2f us\n", fb(n, per), t);
        }
        t = time_us([&] { v9_fill<<<8 * sms, 256>>>(gc, n); }); printf("   plain fill of the output               %7.2f us\n", t);
    }
    return 0;
}
