// What one kernel boundary costs on this GPU (not part of the library): a chain of N dependent kernels of a given size, launched
//   (a) plainly in one stream, (b) with programmatic dependent launch, (c) as a CUDA graph of (a), (d) as a CUDA graph of (b).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o scripts/microbench/launch_floor scripts/microbench/launch_floor.cu
#include <cuda_runtime.h>
#include <chrono>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)
template <bool PDL>
__global__ void __launch_bounds__(256) stream_kernel(const float4* __restrict__ in, float4* __restrict__ out, int64_t n4) {
    if (PDL) { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); asm volatile("griddepcontrol.wait;" ::: "memory"); }
    const int64_t T = (int64_t)gridDim.x * 256;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += T) { float4 v = in[i]; v.x += 1.f; out[i] = v; }
}
template <bool PDL> static void launch(cudaStream_t s, int grid, const float4* in, float4* out, int64_t n4) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(256); cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = PDL ? 1 : 0;
    CK(cudaLaunchKernelEx(&cfg, stream_kernel<PDL>, in, out, n4));
}
int main() {
    int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    cudaStream_t s; CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    float4 *a, *b; const int64_t nmax = 64 << 20;        // 64 Mi float4 = 1 GiB each
    CK(cudaMalloc(&a, nmax * 16 / 8)); CK(cudaMalloc(&b, nmax * 16 / 8)); CK(cudaMemset(a, 0, nmax * 2)); CK(cudaMemset(b, 0, nmax * 2));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const int N = 64;
    printf("chain of %d dependent kernels (ping-pong copy), per-kernel time in us; cpu = host time to enqueue one launch\n", N);
    printf("%12s %10s %10s %10s %10s %10s\n", "MB moved", "plain", "pdl", "graph", "graph+pdl", "cpu/launch");
    for (int64_t mb : {0, 1, 4, 16, 32, 64, 128}) {
        const int64_t n4 = mb == 0 ? 256 : mb * 1000000 / 32;          // read + write = 32 bytes per float4
        const int grid = (int)(n4 / 256 < 8 * sms ? (n4 + 255) / 256 : 8 * sms);
        float res[4]; double cpu_us = 0;
        for (int mode = 0; mode < 4; mode++) {
            const bool pdl = mode & 1, graph = mode >= 2;
            auto chain = [&](cudaStream_t st) {
                for (int k = 0; k < N; k++) {
                    const float4* in = (k & 1) ? b : a; float4* out = (k & 1) ? a : b;
                    if (pdl) launch<true>(st, grid, in, out, n4); else launch<false>(st, grid, in, out, n4);
                }
            };
            cudaGraphExec_t ge = nullptr;
            if (graph) {
                cudaGraph_t g;
                CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
                chain(s);
                CK(cudaStreamEndCapture(s, &g));
                CK(cudaGraphInstantiate(&ge, g, 0));
                CK(cudaGraphDestroy(g));
            }
            auto run = [&] { if (graph) CK(cudaGraphLaunch(ge, s)); else chain(s); };
            for (int w = 0; w < 3; w++) run();
            CK(cudaStreamSynchronize(s));
            const int reps = 10;
            auto t0 = std::chrono::steady_clock::now();
            CK(cudaEventRecord(e0, s));
            for (int r = 0; r < reps; r++) run();
            CK(cudaEventRecord(e1, s));
            auto t1 = std::chrono::steady_clock::now();
            CK(cudaEventSynchronize(e1));
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
            res[mode] = ms * 1000.f / (reps * N);
            if (mode == 1) cpu_us = std::chrono::duration<double, std::micro>(t1 - t0).count() / (reps * N);
            if (ge) CK(cudaGraphExecDestroy(ge));
        }
        printf("%12lld %10.2f %10.2f %10.2f %10.2f %10.2f\n", (long long)mb, res[0], res[1], res[2], res[3], cpu_us);
    }
    return 0;
}
