// K2 (tensor-core path, DINCAE_PREC_TF32): 3x3 convolution as an implicit GEMM on the 5th-generation
// tensor cores.  One CTA = 128 output pixels x BN output channels:
//   * 4 producer warps gather the im2col rows (one pixel row = 32 channels = 128 B per K chunk) straight
//     into the canonical K-major SWIZZLE_128B shared-memory layout, fusing the level operators into the
//     producer (nearest upsample + crop + skip add: src/flux.jl:4, src/model.jl:270-279);
//   * one elected thread issues tcgen05.mma.kind::tf32 (operands FP32 in smem read as TF32, FP32
//     accumulation in TMEM), tcgen05.commit recycles the smem stage through an mbarrier ring;
//   * the same 4 warps drain the accumulator with tcgen05.ld (thread t <-> TMEM lane t <-> pixel row t)
//     and fuse bias + (leaky-)ReLU + residual add into the NHWC store.
// Used for forward convs and (with the transposed, tap-flipped K-major weight copy) data gradients.
#include "k2_conv.cuh"

namespace {

constexpr int TC_BM = 128;      // pixels per CTA (= UMMA M)
constexpr int TC_KC = 32;       // channels per K chunk (= 128 B rows, one swizzle atom wide)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    uint32_t ok = 0;
    while (!ok) {
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}" : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    }
}
// round-to-nearest FP32 -> TF32 (the tensor core would otherwise TRUNCATE the low 13 mantissa bits, a biased error)
__device__ __forceinline__ float to_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}
__device__ __forceinline__ float4 to_tf32(float4 v) { return make_float4(to_tf32(v.x), to_tf32(v.y), to_tf32(v.z), to_tf32(v.w)); }
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
}

// K-major SWIZZLE_128B shared-memory matrix descriptor: 8-row groups of 128-byte rows, groups 1024 B apart
__device__ __forceinline__ uint64_t make_desc_k_sw128(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);          // start address, 16-byte units
    d |= (uint64_t)1 << 16;                           // leading byte offset (unused for swizzled K-major; canonical value 1)
    d |= (uint64_t)(1024 >> 4) << 32;                 // stride byte offset between 8-row groups
    d |= (uint64_t)1 << 46;                           // descriptor version (sm_100)
    d |= (uint64_t)2 << 61;                           // layout type SWIZZLE_128B
    return d;
}

__device__ __forceinline__ float act_apply(float v, int act, float slope) { return (act == ACT_RELU && v < 0.f) ? v * slope : v; }

template <int BN, int STAGES>
__global__ void __launch_bounds__(128) conv3x3_tc(ConvArgs a, int Mtot, int nkc) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    constexpr int A_BYTES = TC_BM * 128, B_BYTES = BN * 128, STAGE_BYTES = A_BYTES + B_BYTES;
    __shared__ uint64_t bar_empty[STAGES];
    __shared__ uint64_t bar_done;
    __shared__ uint32_t tmem_base_s;

    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    const uint32_t smem_base = smem_u32(smem);
    const int t = threadIdx.x, warp = t >> 5;
    const int m0 = blockIdx.x * TC_BM, n0 = blockIdx.y * BN;
    const int HW = a.H * a.W, Cin_p = a.in.C;
    const int nchunks = 9 * nkc;

    if (t == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; s++) mbar_init(&bar_empty[s], 1);
        mbar_init(&bar_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"((uint32_t)BN));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base_s;

    // Producer mapping (coalesced): in pass j a thread moves 16 B chunk (t & 7) of tile row j*16 + (t >> 3), so a warp
    // reads 4 full 128-byte rows per request and a quarter-warp writes one full swizzled smem row (conflict-free).
    const int ch = t & 7;
    int rb_[8], ry[8], rx[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const int mr = m0 + j * 16 + (t >> 3);
        if (mr < Mtot) {
            rb_[j] = mr / HW;
            int r = mr - rb_[j] * HW;
            ry[j] = r / a.W;
            rx[j] = r - ry[j] * a.W;
        } else {
            rb_[j] = -1; ry[j] = 0; rx[j] = 0;
        }
    }

    // instruction descriptor: D = F32, A = B = TF32, both K-major, N = BN, M = 128
    constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
    constexpr int BPASS = BN / 16;                   // weight-tile passes (16 rows per pass)

    float4 ra[8], ra1[8];
    float4 rb[BPASS];
    const bool have1 = a.in.src1 != nullptr;

    auto prefetch = [&](int i) {
        const int tap = i / nkc, kc = i - tap * nkc;
        const int ky = tap / 3, kx = tap - ky * 3;
        const int c0 = kc * TC_KC + ch * 4;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            int yy = ry[j] + ky - 1, xx = rx[j] + kx - 1;
            const bool ok = rb_[j] >= 0 && yy >= 0 && yy < a.H && xx >= 0 && xx < a.W;
            if (ok) {
                if (a.in.mode == SRC_UP_NEAREST) { yy >>= 1; xx >>= 1; }
                const size_t off = (((size_t)rb_[j] * a.in.Hs + yy) * a.in.Ws + xx) * Cin_p + c0;
                ra[j] = __ldg(reinterpret_cast<const float4*>(a.in.src0 + off));
                if (have1) ra1[j] = __ldg(reinterpret_cast<const float4*>(a.in.src1 + off));
            } else {
                ra[j] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (have1) ra1[j] = ra[j];
            }
        }
#pragma unroll
        for (int j = 0; j < BPASS; j++) {
            const int n = n0 + j * 16 + (t >> 3);
            rb[j] = (n < a.Cout_p) ? __ldg(reinterpret_cast<const float4*>(a.wgtK + (size_t)n * a.ldk + (size_t)tap * Cin_p + c0))
                                   : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    };

    prefetch(0);
    for (int i = 0; i < nchunks; i++) {
        const int s = i % STAGES, u = i / STAGES;
        if (u > 0) mbar_wait(&bar_empty[s], (uint32_t)((u - 1) & 1));     // the MMAs that read this stage have retired
        uint8_t* sa = smem + (size_t)s * STAGE_BYTES;
        uint8_t* sb = sa + A_BYTES;
#pragma unroll
        for (int j = 0; j < 8; j++) {       // skip add (SumSkip of the level below) and TF32 rounding happen here, after the loads have landed
            float4 v = ra[j];
            if (have1) { v.x += ra1[j].x; v.y += ra1[j].y; v.z += ra1[j].z; v.w += ra1[j].w; }
            const int r = j * 16 + (t >> 3);
            *reinterpret_cast<float4*>(sa + (r >> 3) * 1024 + (r & 7) * 128 + ((ch ^ (r & 7)) << 4)) = to_tf32(v);
        }
#pragma unroll
        for (int j = 0; j < BPASS; j++) {
            const int r = j * 16 + (t >> 3);
            *reinterpret_cast<float4*>(sb + (r >> 3) * 1024 + (r & 7) * 128 + ((ch ^ (r & 7)) << 4)) = rb[j];
        }
        if (i + 1 < nchunks) prefetch(i + 1);                             // next chunk's global loads fly during the MMAs
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy writes -> visible to the tensor core
        __syncthreads();
        if (t == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint64_t da = make_desc_k_sw128(smem_base + (uint32_t)(s * STAGE_BYTES));
            const uint64_t db = make_desc_k_sw128(smem_base + (uint32_t)(s * STAGE_BYTES + A_BYTES));
#pragma unroll
            for (int k = 0; k < TC_KC / 8; k++)                           // UMMA K = 8 for TF32: 32-byte steps inside the 128-byte row
                umma_tf32(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), IDESC, (i | k) != 0 ? 1u : 0u);
            umma_commit(&bar_empty[s]);
            if (i == nchunks - 1) umma_commit(&bar_done);
        }
    }
    // epilogue row of this thread: TMEM lane t = tile row t
    const int m = m0 + t;
    const bool pv = m < Mtot;
    mbar_wait(&bar_done, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    // epilogue: thread t owns pixel row t = TMEM lane t; 16 accumulator columns per tcgen05.ld
    const uint32_t lane_addr = tmem_d + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
    for (int cb = 0; cb < BN; cb += 16) {
        uint32_t r[16];
        tmem_ld16(lane_addr + (uint32_t)cb, r);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        const int n = n0 + cb;
        if (pv && n < a.Cout_p) {
            const size_t off = (size_t)m * a.Cout_p + n;
#pragma unroll
            for (int q = 0; q < 4; q++) {
                float4 o;
                o.x = __uint_as_float(r[4 * q + 0]); o.y = __uint_as_float(r[4 * q + 1]);
                o.z = __uint_as_float(r[4 * q + 2]); o.w = __uint_as_float(r[4 * q + 3]);
                if (a.bias) {
                    float4 bv = __ldg(reinterpret_cast<const float4*>(a.bias + n + 4 * q));
                    o.x += bv.x; o.y += bv.y; o.z += bv.z; o.w += bv.w;
                }
                o.x = act_apply(o.x, a.act, a.slope); o.y = act_apply(o.y, a.act, a.slope);
                o.z = act_apply(o.z, a.act, a.slope); o.w = act_apply(o.w, a.act, a.slope);
                if (a.addend) {
                    float4 av = __ldg(reinterpret_cast<const float4*>(a.addend + off + 4 * q));
                    o.x += av.x; o.y += av.y; o.z += av.z; o.w += av.w;
                }
                *reinterpret_cast<float4*>(a.dst + off + 4 * q) = o;
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"((uint32_t)BN));
}

template <int BN, int STAGES>
constexpr int tc_smem_bytes() { return STAGES * (TC_BM * 128 + BN * 128) + 1024; }

}  // namespace

#define TC_LAUNCH(BN_, ST_)                                                                                               \
    do {                                                                                                                  \
        auto conv3x3_tc_bn##BN_ = conv3x3_tc<BN_, ST_>;                                                                   \
        constexpr int smem = tc_smem_bytes<BN_, ST_>();                                                                   \
        static bool attr_set = false;                                                                                     \
        if (!attr_set) {                                                                                                  \
            DC_CHECK_CUDA(cudaFuncSetAttribute(conv3x3_tc_bn##BN_, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));   \
            attr_set = true;                                                                                              \
        }                                                                                                                 \
        dim3 grid((unsigned)dc_cdiv(Mtot, TC_BM), (unsigned)dc_cdiv(a.Cout_p, BN_));                                      \
        DC_WORK(ctx, a.work_flops > 0 ? a.work_flops : 2.0 * 9 * a.in.C * a.Cout_p * (double)Mtot, a.work_bytes);                  \
        DC_LAUNCH(ctx, conv3x3_tc_bn##BN_, grid, 128, smem, a, Mtot, a.in.C / TC_KC);                                     \
        DC_CHECK_LAUNCH();                                                                                                \
        return DINCAE_OK;                                                                                                 \
    } while (0)

namespace {
}  // namespace

bool k2_conv_tc_eligible(const ConvArgs& a) {
    return a.wgtK != nullptr && (a.in.C % TC_KC) == 0 && (a.Cout_p % 32) == 0 && (a.in.mode == SRC_DIRECT || a.in.mode == SRC_UP_NEAREST);
}

int32_t k2_conv_tc(dincae_ctx* ctx, const ConvArgs& a) {
    int Mtot = a.B * a.H * a.W;
    if (Mtot == 0) return DINCAE_OK;
    if (a.Cout_p % 256 == 0) TC_LAUNCH(256, 2);
    if (a.Cout_p % 128 == 0) TC_LAUNCH(128, 3);
    if (a.Cout_p % 64 == 0) TC_LAUNCH(64, 4);
    TC_LAUNCH(32, 4);
}
