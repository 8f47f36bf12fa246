// Micro-benchmark: how fast can the SMs pull L2-resident data into shared memory with bulk async copies?
// Each CTA streams `iters` chunks of CHUNK bytes through a ring of STAGES shared-memory buffers (cp.async.bulk + mbarrier),
// reading a window of a buffer that fits L2.  Prints aggregate GB/s for a few (stages, chunk, CTAs/SM, footprint) settings.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    }
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

template <int STAGES>
__global__ void stream_kernel(const uint8_t* __restrict__ buf, size_t footprint, int chunk, int iters, int same, unsigned long long* sink) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t full[STAGES];
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; s++) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const size_t nchunks = footprint / chunk;
        size_t pos = same ? 0 : ((size_t)blockIdx.x * 977) % nchunks;       // same=1: every CTA reads the same chunks in the same order
        for (int i = 0; i < iters + STAGES; i++) {
            const int s = i % STAGES, u = i / STAGES;
            if (i >= STAGES) mbar_wait(&full[s], (uint32_t)((u - 1) & 1));    // previous use of this stage has landed (nobody consumes it)
            if (i < iters) {
                mbar_expect_tx(&full[s], (uint32_t)chunk);
                bulk_load(smem + (size_t)s * chunk, buf + pos * chunk, (uint32_t)chunk, &full[s]);
                pos = (pos + 1) % nchunks;
            }
        }
        sink[blockIdx.x] = smem[0];
    }
}

int main() {
    int dev = 0; cudaSetDevice(dev);
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, dev);
    const int sms = prop.multiProcessorCount;
    const size_t big = (size_t)256 << 20;
    uint8_t* buf; cudaMalloc(&buf, big); cudaMemset(buf, 1, big);
    unsigned long long* sink; cudaMalloc(&sink, 4096 * 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    printf("SMs %d, L2 %d MB\n", sms, prop.l2CacheSize >> 20);
    const int chunks[] = {4096, 16384, 32768};
    const size_t foots[] = {(size_t)2 << 20, (size_t)32 << 20, big};
    for (size_t fp : foots)
        for (int same = 0; same < 2; same++)
            for (int chunk : chunks)
                for (int per_sm = 1; per_sm <= 2; per_sm++) {
                    constexpr int ST = 6;
                    const int smem = ST * chunk;
                    if (per_sm * (smem + 1024) > 220 * 1024) continue;
                    cudaFuncSetAttribute(stream_kernel<ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
                    const int iters = (int)(((size_t)64 << 20) / chunk / per_sm);      // 64 MB per SM
                    const int grid = sms * per_sm;
                    stream_kernel<ST><<<grid, 32, smem>>>(buf, fp, chunk, iters / 8, same, sink);   // warm L2
                    cudaEventRecord(e0);
                    stream_kernel<ST><<<grid, 32, smem>>>(buf, fp, chunk, iters, same, sink);
                    cudaEventRecord(e1); cudaEventSynchronize(e1);
                    float ms; cudaEventElapsedTime(&ms, e0, e1);
                    if (cudaGetLastError() != cudaSuccess) { printf("launch error\n"); return 1; }
                    const double gb = (double)grid * iters * chunk / 1e9;
                    printf("footprint %4zu MB  same=%d  chunk %5d B  stages %d  CTAs/SM %d : %8.1f GB/s  (%.1f B/clk/SM at 1.965 GHz)\n",
                           fp >> 20, same, chunk, ST, per_sm, gb / (ms * 1e-3), gb * 1e9 / (ms * 1e-3) / sms / 1.965e9);
                }
    return 0;
}
