// K2 (TMA-fed tensor-core path): 3x3 convolution of a plain NHWC tensor as an implicit GEMM.
//   * one elected thread streams, per K chunk, a (32 ch x PW x PH x 1) activation box and a (32 x BN) weight box
//     with cp.async.bulk.tensor (TMA) into a SWIZZLE_128B smem ring; the 3x3 taps are just shifted box
//     coordinates and the conv zero-padding is TMA's out-of-bounds zero fill;
//   * one elected thread issues tcgen05.mma.kind::tf32 into a TMEM accumulator (128 pixels x BN channels);
//   * four warps drain TMEM with tcgen05.ld and fuse bias + (leaky-)ReLU + residual add into the NHWC store.
// Operands must already be TF32-representable (the kernels that produce conv inputs round on store in TF32 mode;
// weights come from the rounded K-major copy).  Used by the encoder convs and all data-gradient convs.
#include "k2_conv.cuh"
#include <cuda.h>
#include <stdlib.h>

namespace {

constexpr int TM_BM = 128;
constexpr int TM_KC = 32;

struct TmaConvArgs {
    int B, H, W;
    int PW, PH, ptx, pty;      // pixel patch PW x PH = 128 and patches per image
    int Cin_p, Cout_p;
    const float* bias;
    int act;
    float slope;
    const float* addend;
    float* dst;
    float* planar_out;
    int planar_C, planar_accumulate, planar_fold;
    float* pool_dst;           // persistent kernel, even H and W: also writes MeanPool((2,2)) of the output, [B][H/2][W/2][Cout_p]
    int pool_rnd;              // round the pooled values to TF32 (they feed the next tensor-core conv)
    uint32_t* mask_dst;        // pool-only mode (dst == NULL, pool_dst set): sign bits of the activation, word [pixel][Cout_p / 32]
};

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
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2, int c3) {
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
// true in exactly one lane of a converged warp (the same lane every time)
__device__ __forceinline__ bool elect_one_sync() {
    uint32_t pred = 0;
    asm volatile(
        "{\n\t"
        ".reg .b32 rx;\n\t"
        ".reg .pred px;\n\t"
        "elect.sync rx|px, %1;\n\t"
        "@px mov.s32 %0, 1;\n\t"
        "}" : "+r"(pred) : "r"(0xFFFFFFFFu));
    return pred != 0;
}
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
// accumulate variant without the predicate set-up (D += A*B)
__device__ __forceinline__ void umma_tf32_acc(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.eq.u32 p, 1, 1;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
}
__device__ __forceinline__ uint64_t make_desc_k_sw128(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
__device__ __forceinline__ float act_apply(float v, int act, float slope) { return (act == ACT_RELU && v < 0.f) ? v * slope : v; }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

// NHWC activation tensor (C, W, H, B), box (32, PW, PH, 1), 128-byte swizzle, zero fill out of bounds
int32_t make_tmap_act(CUtensorMap* m, const float* ptr, int C, int W, int H, int B, int PW, int PH) {
    EncodeTiledFn enc = get_encode();
    DC_REQUIRE(enc != nullptr, DINCAE_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t gstr[3] = {(cuuint64_t)C * 4, (cuuint64_t)W * C * 4, (cuuint64_t)H * W * C * 4};
    cuuint32_t box[4] = {32, (cuuint32_t)PW, (cuuint32_t)PH, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (void*)ptr, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DC_REQUIRE(r == CUDA_SUCCESS, DINCAE_ECUDA, "cuTensorMapEncodeTiled(activation C=%d W=%d H=%d B=%d) failed: %d", C, W, H, B, (int)r);
    return DINCAE_OK;
}
// K-major weight matrix [rows][ldk], box (32, BN)
int32_t make_tmap_w(CUtensorMap* m, const float* ptr, int ldk, int rows, int BN) {
    EncodeTiledFn enc = get_encode();
    DC_REQUIRE(enc != nullptr, DINCAE_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t gdim[2] = {(cuuint64_t)ldk, (cuuint64_t)rows};
    cuuint64_t gstr[1] = {(cuuint64_t)ldk * 4};
    cuuint32_t box[2] = {32, (cuuint32_t)BN};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)ptr, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DC_REQUIRE(r == CUDA_SUCCESS, DINCAE_ECUDA, "cuTensorMapEncodeTiled(weights ldk=%d rows=%d) failed: %d", ldk, rows, (int)r);
    return DINCAE_OK;
}

// the same K-major weight matrix seen as [tap][row][ci]: one box = (32 ci, BN rows, `taps` taps) -> `taps` consecutive K-major
// SW128 tiles in shared memory from ONE TMA instruction (the L2 -> SM stream is latency x bytes-in-flight bound: ~2100 cycles
// per bulk load measured, so few large loads beat many small ones)
int32_t make_tmap_w3(CUtensorMap* m, const float* ptr, int Cin_p, int rows, int BN, int taps) {
    EncodeTiledFn enc = get_encode();
    DC_REQUIRE(enc != nullptr, DINCAE_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t gdim[3] = {(cuuint64_t)Cin_p, (cuuint64_t)rows, 9};
    cuuint64_t gstr[2] = {(cuuint64_t)9 * Cin_p * 4, (cuuint64_t)Cin_p * 4};
    cuuint32_t box[3] = {32, (cuuint32_t)BN, (cuuint32_t)taps};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)ptr, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DC_REQUIRE(r == CUDA_SUCCESS, DINCAE_ECUDA, "cuTensorMapEncodeTiled(weights3 Cin_p=%d rows=%d taps=%d) failed: %d", Cin_p, rows, taps, (int)r);
    return DINCAE_OK;
}


}  // namespace

// -------------------------------------------------------------------------------------------------
// Halo variant: per 32-channel K chunk the (8+2) x (16+2) pixel halo of the output patch is loaded ONCE; the nine
// taps are nine descriptor views of that tile (start address shifted by (dy*10+dx) rows of 128 B, 8-pixel row groups
// one halo pitch = 1280 B apart), so the activation operand crosses L2->SM once instead of nine times.  The weight
// tiles (one per chunk and tap) stream through their own ring.
namespace {

constexpr int HL_PW = 8, HL_PH = 16, HL_HW = HL_PW + 2, HL_HH = HL_PH + 2;
constexpr int HL_A_BYTES = HL_HW * HL_HH * 128;                // 23040
constexpr int HL_A_STRIDE = (HL_A_BYTES + 1023) / 1024 * 1024;  // 23552

__device__ __forceinline__ uint64_t make_desc_k_sw128_halo(uint32_t saddr, int use_base_offset) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)((HL_HW * 128) >> 4) << 32;        // stride between 8-pixel row groups = one halo row pitch
    d |= (uint64_t)1 << 46;
    if (use_base_offset) d |= (uint64_t)((saddr >> 7) & 7) << 49;
    d |= (uint64_t)2 << 61;
    return d;
}

int32_t make_tmap_halo(CUtensorMap* m, const float* ptr, int C, int W, int H, int B) {
    EncodeTiledFn enc = get_encode();
    DC_REQUIRE(enc != nullptr, DINCAE_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t gstr[3] = {(cuuint64_t)C * 4, (cuuint64_t)W * C * 4, (cuuint64_t)H * W * C * 4};
    cuuint32_t box[4] = {32, (cuuint32_t)HL_HW, (cuuint32_t)HL_HH, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (void*)ptr, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DC_REQUIRE(r == CUDA_SUCCESS, DINCAE_ECUDA, "cuTensorMapEncodeTiled(halo C=%d W=%d H=%d B=%d) failed: %d", C, W, H, B, (int)r);
    return DINCAE_OK;
}


}  // namespace

// -------------------------------------------------------------------------------------------------
// Persistent, warp-specialised form of the halo kernel: one CTA per SM loops over output tiles.
//   warp 0 (one lane): TMA producer      warp 1 (one lane): tcgen05.mma issuer
//   warps 2 / 3 (one lane each): TMA-store issuer of epilogue group 0 / 1
//   warps 4-7 / 8-11: epilogue group 0 / 1.  With two groups (one CTA per SM) group g drains TMEM accumulator g, i.e.
//   every other tile of this CTA; narrow layers whose shared-memory plan allows two CTAs per SM run one group (256 threads).
// The FP32 accumulator is double-buffered in TMEM (2 x BN columns), so the epilogues of tiles i and i+1 (tcgen05.ld, bias,
// activation, swizzled shared staging, TMA store) overlap each other and the TMA + MMA main loop of tile i+2.  When the whole K-major weight matrix of the
// layer fits in shared memory (narrow layers) it is loaded once per CTA and stays resident; then only one 23 KB
// activation halo per tile and K chunk crosses L2 -> SM.
namespace {

constexpr int PS_MAX_A = 4, PS_MAX_B = 16, PS_MAX_COUT = 512, PS_MAX_O = 2, PS_THREADS = 384;

struct PersistArgs {
    int num_tiles, ntiles_n;
    int a_stages, b_stages;
    int b_resident;
    int o_stages;              // 16 KB output staging slabs per epilogue group (1 or 2)
    int groups;                // epilogue groups (1: 256 threads, 2: 384 threads)
    int tps;                   // streamed weights: taps per weight stage (3 = one kernel row per TMA load, 1 when shared memory is short)
    int ksteps;                // K = 8 steps over the input channels that are not zero padding (= Cin_p / 8 when unknown)
    long long* trace;          // tuning aid (dincae_debug_conv): clock64 stamps of CTA 0, 3 roles x 64 slots; nullptr in production
};

#define PS_TRACE(role, idx) do { if (q.trace && blockIdx.x == 0 && (idx) < 64) q.trace[(role) * 64 + (idx)] = clock64(); } while (0)

// TMA store of one (32 channel x 8 x 16 pixel) slab, shared -> global; pixels outside the image are clipped by the unit
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* map, const void* src, int c0, int c1, int c2, int c3) {
    asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ float rna_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, float4 v) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w));
}
// One 32-column slab of one accumulator row: bias, leaky ReLU as max(v, v * slope) (0 <= slope <= 1), optional addend, then
// eight 16 B shared-memory stores at the 128B-swizzled chunk positions of this row.
template <bool ADD, bool POOL, bool STORE = true>
__device__ __forceinline__ void epi_slab(const uint32_t (&r)[32], const float* bs, float slope, const float* addp, uint32_t st_row, int sw,
                                         float* poolp, int pool_rnd, uint32_t* maskw = nullptr) {
    float4 av[ADD ? 8 : 1];
    if (ADD) {
#pragma unroll
        for (int k = 0; k < 8; k++) av[k] = addp ? __ldg(reinterpret_cast<const float4*>(addp + 4 * k)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float4 o[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {                // all bias reads first: the shared loads overlap instead of serialising with the stores
        const float4 bv = *reinterpret_cast<const float4*>(bs + 4 * k);
        o[k].x = __uint_as_float(r[4 * k + 0]) + bv.x; o[k].y = __uint_as_float(r[4 * k + 1]) + bv.y;
        o[k].z = __uint_as_float(r[4 * k + 2]) + bv.z; o[k].w = __uint_as_float(r[4 * k + 3]) + bv.w;
    }
#pragma unroll
    for (int k = 0; k < 8; k++) {
        o[k].x = fmaxf(o[k].x, o[k].x * slope); o[k].y = fmaxf(o[k].y, o[k].y * slope);
        o[k].z = fmaxf(o[k].z, o[k].z * slope); o[k].w = fmaxf(o[k].w, o[k].w * slope);
        if (ADD) { o[k].x += av[k].x; o[k].y += av[k].y; o[k].z += av[k].z; o[k].w += av[k].w; }
    }
    if (STORE) {
#pragma unroll
        for (int k = 0; k < 8; k++) st_shared_v4(st_row + (uint32_t)((k ^ sw) << 4), o[k]);
    } else {
        // the full-resolution activation is not kept: only the sign of every element (what the backward pass needs of it)
        uint32_t bits = 0;
#pragma unroll
        for (int k = 0; k < 8; k++)
            bits |= (o[k].x > 0.f ? 1u : 0u) << (4 * k) | (o[k].y > 0.f ? 1u : 0u) << (4 * k + 1) | (o[k].z > 0.f ? 1u : 0u) << (4 * k + 2) |
                    (o[k].w > 0.f ? 1u : 0u) << (4 * k + 3);
        if (maskw) *maskw = bits;
    }
    if (POOL) {
        // MeanPool((2,2)): tile row = h * 8 + w, so the 2 x 2 window of an even (h, w) is lanes {l, l^1, l^8, l^9} of this warp
#pragma unroll
        for (int k = 0; k < 8; k++) {
            float4 s = o[k];
            s.x += __shfl_xor_sync(0xffffffffu, s.x, 1); s.y += __shfl_xor_sync(0xffffffffu, s.y, 1);
            s.z += __shfl_xor_sync(0xffffffffu, s.z, 1); s.w += __shfl_xor_sync(0xffffffffu, s.w, 1);
            s.x += __shfl_xor_sync(0xffffffffu, s.x, 8); s.y += __shfl_xor_sync(0xffffffffu, s.y, 8);
            s.z += __shfl_xor_sync(0xffffffffu, s.z, 8); s.w += __shfl_xor_sync(0xffffffffu, s.w, 8);
            if (poolp) {
                s.x *= 0.25f; s.y *= 0.25f; s.z *= 0.25f; s.w *= 0.25f;
                if (pool_rnd) { s.x = rna_tf32(s.x); s.y = rna_tf32(s.y); s.z = rna_tf32(s.z); s.w = rna_tf32(s.w); }
                *reinterpret_cast<float4*>(poolp + 4 * k) = s;
            }
        }
    }
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// Epilogue of one 128-pixel x BN-channel accumulator, run by the 128 threads of one epilogue group (thread = tile row = TMEM lane).
// Planar output (<= 16 channels) is written directly; NHWC output goes slab by slab (32 channels) through the group's
// swizzled staging buffers to its store issuer.  `tempty` gets one arrival per thread once the accumulator has been read.
struct EpiCtx {
    const float* bias_s;       // shared: the layer bias
    uint32_t smemO_u;          // shared address of this group's staging slabs
    uint64_t *ofull, *oempty;  // this group's hand-off barriers [o_stages]
    int o_stages, obuf, ouse;
    int row, lane;
    int act_mode;
    float slope, slope_eff;
};
template <int BN>
__device__ __forceinline__ void epi_tile(const TmaConvArgs& p, EpiCtx& e, uint32_t lane_addr, uint64_t* tempty, int b, int y0, int x0, int n0) {
    const int h = e.row / HL_PW, w = e.row - h * HL_PW;
    const int y = y0 + h, x = x0 + w;
    const bool pv = b < p.B && y < p.H && x < p.W;
    const size_t m = ((size_t)b * p.H + y) * p.W + x;
    if (p.planar_out) {
        uint32_t r[16];
        tmem_ld16(lane_addr, r);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (pv && n0 == 0) {
#pragma unroll
            for (int k = 0; k < 16; k++) {
                if (k < p.planar_C) {
                    float v = __uint_as_float(r[k]);
                    if (p.planar_fold) v += __uint_as_float(r[(k + 8) & 15]);     // + the low-order weight column (3xTF32 head), k < planar_C <= 8
                    v += e.bias_s[k];
                    v = act_apply(v, p.act, p.slope);
                    const size_t po = (((size_t)b * p.planar_C + k) * p.H + y) * p.W + x;
                    p.planar_out[po] = p.planar_accumulate ? p.planar_out[po] + v : v;
                }
            }
        }
    }
    if (!p.dst && p.pool_dst) {
        // pool-only mode (encoder convs): MeanPool of the tile goes to pool_dst, the activation itself leaves only its sign bits
        float* poolp = (pv && (e.lane & 9) == 0)
                           ? p.pool_dst + ((((size_t)b * (p.H >> 1) + (y >> 1)) * (p.W >> 1) + (x >> 1)) * p.Cout_p + n0) : nullptr;
        uint32_t* maskp = (p.mask_dst && pv) ? p.mask_dst + m * (size_t)(p.Cout_p >> 5) + (n0 >> 5) : nullptr;
        uint32_t r[32];
        tmem_ld16(lane_addr, *reinterpret_cast<uint32_t(*)[16]>(&r[0]));
        tmem_ld16(lane_addr + 16u, *reinterpret_cast<uint32_t(*)[16]>(&r[16]));
#pragma unroll 1
        for (int sl = 0; sl < BN / 32; sl++) {
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            const int nl = sl * 32;
            epi_slab<false, true, false>(r, &e.bias_s[n0 + nl], e.slope_eff, nullptr, 0u, 0, poolp ? poolp + nl : nullptr, p.pool_rnd,
                                         maskp ? maskp + sl : nullptr);
            if (sl + 1 < BN / 32) {
                tmem_ld16(lane_addr + (uint32_t)(nl + 32), *reinterpret_cast<uint32_t(*)[16]>(&r[0]));
                tmem_ld16(lane_addr + (uint32_t)(nl + 48), *reinterpret_cast<uint32_t(*)[16]>(&r[16]));
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        mbar_arrive(tempty);
    } else if (p.dst) {
        const float* addp = (p.addend && pv) ? p.addend + m * p.Cout_p + n0 : nullptr;
        float* poolp = (p.pool_dst && pv && (e.lane & 9) == 0)
                           ? p.pool_dst + ((((size_t)b * (p.H >> 1) + (y >> 1)) * (p.W >> 1) + (x >> 1)) * p.Cout_p + n0) : nullptr;
        uint32_t r[32];
        tmem_ld16(lane_addr, *reinterpret_cast<uint32_t(*)[16]>(&r[0]));
        tmem_ld16(lane_addr + 16u, *reinterpret_cast<uint32_t(*)[16]>(&r[16]));
#pragma unroll 1
        for (int sl = 0; sl < BN / 32; sl++) {
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (sl == BN / 32 - 1) {                 // accumulator drained: the MMA warp may reuse it
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                mbar_arrive(tempty);
            }
            if (e.ouse > 0) mbar_wait(&e.oempty[e.obuf], (uint32_t)((e.ouse - 1) & 1));   // the store that last used this slab has been read out
            const int nl = sl * 32;                  // column of this slab inside the tile
            const uint32_t st_row = e.smemO_u + (uint32_t)(e.obuf * 16384 + e.row * 128);
            const float* bs = &e.bias_s[n0 + nl];
            if (e.act_mode != 2) {
                if (p.pool_dst) epi_slab<false, true>(r, bs, e.slope_eff, nullptr, st_row, e.row & 7, poolp ? poolp + nl : nullptr, p.pool_rnd);
                else if (p.addend) epi_slab<true, false>(r, bs, e.slope_eff, addp ? addp + nl : nullptr, st_row, e.row & 7, nullptr, 0);
                else epi_slab<false, false>(r, bs, e.slope_eff, nullptr, st_row, e.row & 7, nullptr, 0);
            } else {                                 // slopes outside [0, 1): the select form
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    float4 o;
                    o.x = act_apply(__uint_as_float(r[4 * k + 0]) + bs[4 * k + 0], ACT_RELU, e.slope);
                    o.y = act_apply(__uint_as_float(r[4 * k + 1]) + bs[4 * k + 1], ACT_RELU, e.slope);
                    o.z = act_apply(__uint_as_float(r[4 * k + 2]) + bs[4 * k + 2], ACT_RELU, e.slope);
                    o.w = act_apply(__uint_as_float(r[4 * k + 3]) + bs[4 * k + 3], ACT_RELU, e.slope);
                    if (addp) {
                        const float4 av = __ldg(reinterpret_cast<const float4*>(addp + nl + 4 * k));
                        o.x += av.x; o.y += av.y; o.z += av.z; o.w += av.w;
                    }
                    st_shared_v4(st_row + (uint32_t)((k ^ (e.row & 7)) << 4), o);
                }
            }
            if (sl + 1 < BN / 32) {                  // next slab's TMEM read overlaps the hand-off below
                tmem_ld16(lane_addr + (uint32_t)(nl + 32), *reinterpret_cast<uint32_t(*)[16]>(&r[0]));
                tmem_ld16(lane_addr + (uint32_t)(nl + 48), *reinterpret_cast<uint32_t(*)[16]>(&r[16]));
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy writes -> visible to the TMA store
            mbar_arrive(&e.ofull[e.obuf]);
            if (++e.obuf == e.o_stages) { e.obuf = 0; e.ouse++; }
        }
    } else {
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        mbar_arrive(tempty);
    }
}

template <int BN>
__global__ void __launch_bounds__(PS_THREADS) conv3x3_persist(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmO,
                                                       TmaConvArgs p, PersistArgs q) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    constexpr int B_BYTES = BN * 128;
    constexpr int TCOLS = 2 * BN < 32 ? 32 : 2 * BN;
    __shared__ uint64_t bar_fullA[PS_MAX_A], bar_emptyA[PS_MAX_A], bar_fullB[PS_MAX_B], bar_emptyB[PS_MAX_B];
    __shared__ uint64_t bar_bres, bar_tfull[2], bar_tempty[2];
    __shared__ uint64_t bar_ofull[2][PS_MAX_O], bar_oempty[2][PS_MAX_O];     // epilogue group -> its store issuer and back
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(16) float bias_s[PS_MAX_COUT];      // the layer's bias (zeros when absent): broadcast reads in the epilogue

    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* smemB = smem + q.a_stages * HL_A_STRIDE;
    uint8_t* smemO = smemB + (q.b_resident ? (p.Cin_p / TM_KC) * 9 : q.b_stages * q.tps) * B_BYTES;     // 1024-aligned: B_BYTES is a multiple of 4 KB
    const uint32_t smemA_u = smem_u32(smem), smemB_u = smem_u32(smemB);
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int patches = p.ptx * p.pty;
    const int nkc = p.Cin_p / TM_KC;
    if (t == 64) PS_TRACE(2, 62);

    if (t == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
        if (p.dst) asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmO)) : "memory");
        for (int s = 0; s < PS_MAX_A; s++) { mbar_init(&bar_fullA[s], 1); mbar_init(&bar_emptyA[s], 1); }
        for (int s = 0; s < PS_MAX_B; s++) { mbar_init(&bar_fullB[s], 1); mbar_init(&bar_emptyB[s], 1); }
        mbar_init(&bar_bres, 1);
        for (int s = 0; s < 2; s++) { mbar_init(&bar_tfull[s], 1); mbar_init(&bar_tempty[s], 128); }
        for (int g = 0; g < 2; g++)
            for (int s = 0; s < PS_MAX_O; s++) { mbar_init(&bar_ofull[g][s], 128); mbar_init(&bar_oempty[g][s], 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    dc_pdl_launch_dependents();
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"((uint32_t)TCOLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    dc_pdl_wait();                 // everything above overlaps the tail of the previous kernel; global memory only from here on
    for (int i = t; i < p.Cout_p; i += (int)blockDim.x) bias_s[i] = p.bias ? __ldg(p.bias + i) : 0.f;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (t == 64) PS_TRACE(2, 63);
    const uint32_t tmem_d = tmem_base_s;
    constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TM_BM >> 4) << 24);

    if (warp == 0) {
        // ===== TMA producer: warp-uniform loop, one elected lane issues (a divergent lane-0 branch costs ~50 cycles per TMA issue) =====
        if (q.b_resident && elect_one_sync()) {
            mbar_expect_tx(&bar_bres, (uint32_t)(nkc * 9 * B_BYTES));
            for (int kc = 0; kc < nkc; kc++)
                for (int tap = 0; tap < 9; tap++)
                    tma_load_2d(smemB + (size_t)(kc * 9 + tap) * B_BYTES, &tmB, &bar_bres, tap * p.Cin_p + kc * TM_KC, 0);
        }
        __syncwarp();
        int ia = 0, ib = 0;
        if (lane == 0) PS_TRACE(0, 0);
        const uint32_t stage_b = (uint32_t)(q.tps * B_BYTES);
        for (int tile = blockIdx.x; tile < q.num_tiles; tile += gridDim.x) {
            const int pi = tile / q.ntiles_n, n0 = (tile - pi * q.ntiles_n) * BN;
            const int b = pi / patches, pr = pi - b * patches;
            const int y0 = (pr / p.ptx) * HL_PH, x0 = (pr - (pr / p.ptx) * p.ptx) * HL_PW;
            for (int kc = 0; kc < nkc; kc++, ia++) {
                const int sa = ia % q.a_stages, ua = ia / q.a_stages;
                if (ua > 0) mbar_wait(&bar_emptyA[sa], (uint32_t)((ua - 1) & 1));
                if (lane == 0) PS_TRACE(0, 1 + ia);
                if (elect_one_sync()) {
                    mbar_expect_tx(&bar_fullA[sa], (uint32_t)HL_A_BYTES);
                    tma_load_4d(smem + (size_t)sa * HL_A_STRIDE, &tmA, &bar_fullA[sa], kc * TM_KC, x0 - 1, y0 - 1, b);
                }
                __syncwarp();
                if (!q.b_resident) {
                    for (int t0 = 0; t0 < 9; t0 += q.tps, ib++) {           // one weight stage = q.tps taps from one 3-D box
                        const int sb = ib % q.b_stages, ub = ib / q.b_stages;
                        if (ub > 0) mbar_wait(&bar_emptyB[sb], (uint32_t)((ub - 1) & 1));
                        if (elect_one_sync()) {
                            mbar_expect_tx(&bar_fullB[sb], stage_b);
                            tma_load_3d(smemB + (size_t)sb * stage_b, &tmB, &bar_fullB[sb], kc * TM_KC, n0, t0);
                        }
                        __syncwarp();
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: the whole warp walks the loop (warp-uniform control flow), one elected lane issues =====
        // (inside a divergent `lane == 0` branch the compiler wraps every tcgen05.mma in its own elect/branch loop, ~50 cycles each)
        if (lane == 0) PS_TRACE(1, 0);
        if (q.b_resident) mbar_wait(&bar_bres, 0);
        if (lane == 0) PS_TRACE(1, 1);
        int ia = 0, ib = 0, it = 0;
        for (int tile = blockIdx.x; tile < q.num_tiles; tile += gridDim.x, it++) {
            const int acc = it & 1, ut = it >> 1;
            if (ut > 0) mbar_wait(&bar_tempty[acc], (uint32_t)((ut - 1) & 1));      // the epilogue has drained this accumulator
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t dcol = tmem_d + (uint32_t)(acc * BN);
            for (int kc = 0; kc < nkc; kc++, ia++) {
                const int sa = ia % q.a_stages, ua = ia / q.a_stages;
                mbar_wait(&bar_fullA[sa], (uint32_t)(ua & 1));
                if (lane == 0) PS_TRACE(1, 2 + 2 * ia);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                // descriptor bases once per chunk; the 36 (tap, k) offsets below are compile-time constants
                const uint64_t da0 = make_desc_k_sw128_halo(smemA_u + (uint32_t)(sa * HL_A_STRIDE), 0);
                if (q.b_resident) {
                    const uint64_t db0 = make_desc_k_sw128(smemB_u + (uint32_t)(kc * 9 * B_BYTES));
                    const int ks = q.ksteps - kc * (TM_KC / 8);           // >= 1: steps of this chunk that are not zero padding
                    if (elect_one_sync()) {
#pragma unroll
                        for (int tap = 0; tap < 9; tap++) {
#pragma unroll
                            for (int k = 0; k < TM_KC / 8; k++) {
                                const uint64_t da = da0 + (uint64_t)((((tap / 3) * HL_HW + (tap % 3)) * 128 >> 4) + 2 * k);
                                const uint64_t db = db0 + (uint64_t)((tap * B_BYTES >> 4) + 2 * k);
                                if (tap == 0 && k == 0) umma_tf32(dcol, da, db, IDESC, kc != 0 ? 1u : 0u);
                                else if (k < ks) umma_tf32_acc(dcol, da, db, IDESC);
                            }
                        }
                        umma_commit(&bar_emptyA[sa]);
                        if (kc == nkc - 1) umma_commit(&bar_tfull[acc]);
                    }
                } else {
#pragma unroll 1
                    for (int t0 = 0; t0 < 9; t0 += q.tps, ib++) {
                        const int sb = ib % q.b_stages, ub = ib / q.b_stages;
                        mbar_wait(&bar_fullB[sb], (uint32_t)(ub & 1));
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        const uint64_t dbs = make_desc_k_sw128(smemB_u + (uint32_t)(sb * q.tps * B_BYTES));
                        if (elect_one_sync()) {
                            for (int tl = 0; tl < q.tps; tl++) {
                                const int tap = t0 + tl;
                                const uint64_t db0 = dbs + (uint64_t)((tl * B_BYTES) >> 4);
                                const uint64_t dat = da0 + (uint64_t)(((tap / 3) * HL_HW + (tap % 3)) * 128 >> 4);
                                if (tap == 0) umma_tf32(dcol, dat, db0, IDESC, kc != 0 ? 1u : 0u);
                                else umma_tf32_acc(dcol, dat, db0, IDESC);
#pragma unroll
                                for (int k = 1; k < TM_KC / 8; k++) umma_tf32_acc(dcol, dat + (uint64_t)(2 * k), db0 + (uint64_t)(2 * k), IDESC);
                            }
                            umma_commit(&bar_emptyB[sb]);
                            if (t0 + q.tps >= 9) {
                                umma_commit(&bar_emptyA[sa]);
                                if (kc == nkc - 1) umma_commit(&bar_tfull[acc]);
                            }
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) PS_TRACE(1, 3 + 2 * ia);
            }
        }
    } else if (warp >= 4) {
        // ===== epilogue warps 4..11: group g; TMEM lane quarter = warp % 4 =====
        // NHWC output: each 32-channel slab of the tile is staged in shared memory in the 128B-swizzled box layout
        // (conflict-free 16 B writes, one 128 B row per thread) and handed to the group's store issuer, whose single TMA
        // store also clips the pixels of edge patches that fall outside the image.  Planar output (<= 16 channels) is
        // written directly.
        const int g = (warp - 4) >> 2;
        EpiCtx e;
        e.bias_s = bias_s;
        e.smemO_u = smem_u32(smemO) + (uint32_t)(g * q.o_stages * 16384);
        e.ofull = bar_ofull[g]; e.oempty = bar_oempty[g];
        e.o_stages = q.o_stages; e.obuf = 0; e.ouse = 0;
        e.row = (warp & 3) * 32 + lane; e.lane = lane;           // tile row = TMEM lane
        e.slope = p.slope;
        e.act_mode = p.act != ACT_RELU ? 0 : ((p.slope >= 0.f && p.slope < 1.f) ? 1 : 2);
        e.slope_eff = e.act_mode == 0 ? 1.f : p.slope;           // max(v, v * 1) = v: no activation
        const bool tr0 = (t == 128);
        for (int tile = blockIdx.x + g * gridDim.x, it = g; tile < q.num_tiles; tile += q.groups * gridDim.x, it += q.groups) {
            const int acc = it & 1, ut = it >> 1;
            const int pi = tile / q.ntiles_n, n0 = (tile - pi * q.ntiles_n) * BN;
            const int b = pi / patches, pr = pi - b * patches;
            const int y0 = (pr / p.ptx) * HL_PH, x0 = (pr - (pr / p.ptx) * p.ptx) * HL_PW;
            if (tr0) PS_TRACE(2, 2 * (it / q.groups));
            mbar_wait(&bar_tfull[acc], (uint32_t)(ut & 1));
            if (tr0) PS_TRACE(2, 1 + 2 * (it / q.groups));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t lane_addr = tmem_d + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(acc * BN);
            epi_tile<BN>(p, e, lane_addr, &bar_tempty[acc], b, y0, x0, n0);
        }
    } else if (lane == 0 && p.dst && warp - 2 < q.groups) {
        // ===== store issuer of group g: one TMA store per staged slab, then hand the slab back once it has been read =====
        const int g = warp - 2;
        int obuf = 0, ouse = 0;
        for (int tile = blockIdx.x + g * gridDim.x; tile < q.num_tiles; tile += q.groups * gridDim.x) {
            const int pi = tile / q.ntiles_n, n0 = (tile - pi * q.ntiles_n) * BN;
            const int b = pi / patches, pr = pi - b * patches;
            const int y0 = (pr / p.ptx) * HL_PH, x0 = (pr - (pr / p.ptx) * p.ptx) * HL_PW;
            for (int sl = 0; sl < BN / 32; sl++) {
                mbar_wait(&bar_ofull[g][obuf], (uint32_t)(ouse & 1));
                tma_store_4d(&tmO, smemO + (size_t)(g * q.o_stages + obuf) * 16384, n0 + sl * 32, x0, y0, b);
                tma_store_commit();
                tma_store_wait_read<0>();
                mbar_arrive(&bar_oempty[g][obuf]);
                if (++obuf == q.o_stages) { obuf = 0; ouse++; }
            }
        }
    }
    __syncwarp();
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (t == 0) PS_TRACE(0, 63);
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"((uint32_t)TCOLS));
}


// Paired-tile form for layers whose weights do not stay resident (deep levels: few pixels, wide channels).  There the
// L2 -> SM weight stream is the limit: every 128-pixel tile re-reads the whole 9 * Cin * BN weight slab.  This kernel gives one
// CTA TWO pixel tiles of the same channel block at a time: each weight stage that lands in shared memory feeds the MMAs of both
// tiles (two accumulators), which halves the weight traffic per output.  Epilogue group g drains tile g of every pair, the pair
// of accumulators is double-buffered (4 x BN TMEM columns), everything else is as in conv3x3_persist.
template <int BN>
__global__ void __launch_bounds__(PS_THREADS) conv3x3_pair(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmO,
                                                    TmaConvArgs p, PersistArgs q) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    constexpr int B_BYTES = BN * 128;
    constexpr int TCOLS = 4 * BN;
    static_assert(TCOLS <= 512, "two double-buffered accumulator pairs must fit TMEM");
    __shared__ uint64_t bar_fullA[PS_MAX_A], bar_emptyA[PS_MAX_A], bar_fullB[PS_MAX_B], bar_emptyB[PS_MAX_B];
    __shared__ uint64_t bar_tfull[2], bar_tempty[2];
    __shared__ uint64_t bar_ofull[2][PS_MAX_O], bar_oempty[2][PS_MAX_O];
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(16) float bias_s[PS_MAX_COUT];

    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    constexpr int A_PAIR = 2 * HL_A_STRIDE;                                  // one activation stage = the halos of both tiles
    uint8_t* smemB = smem + q.a_stages * A_PAIR;
    uint8_t* smemO = smemB + q.b_stages * q.tps * B_BYTES;
    const uint32_t smemA_u = smem_u32(smem), smemB_u = smem_u32(smemB);
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int patches = p.ptx * p.pty;
    const int nkc = p.Cin_p / TM_KC;
    const int num_units = q.num_tiles;                                       // (pixel-tile pair, channel block)

    if (t == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
        if (p.dst) asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmO)) : "memory");
        for (int s = 0; s < PS_MAX_A; s++) { mbar_init(&bar_fullA[s], 1); mbar_init(&bar_emptyA[s], 1); }
        for (int s = 0; s < PS_MAX_B; s++) { mbar_init(&bar_fullB[s], 1); mbar_init(&bar_emptyB[s], 1); }
        for (int s = 0; s < 2; s++) { mbar_init(&bar_tfull[s], 1); mbar_init(&bar_tempty[s], 256); }
        for (int g = 0; g < 2; g++)
            for (int s = 0; s < PS_MAX_O; s++) { mbar_init(&bar_ofull[g][s], 128); mbar_init(&bar_oempty[g][s], 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    dc_pdl_launch_dependents();
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"((uint32_t)TCOLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    dc_pdl_wait();                 // everything above overlaps the tail of the previous kernel; global memory only from here on
    for (int i = t; i < p.Cout_p; i += (int)blockDim.x) bias_s[i] = p.bias ? __ldg(p.bias + i) : 0.f;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base_s;
    constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TM_BM >> 4) << 24);

    // unit -> pixel tiles 2*pp and 2*pp+1 (the second may lie past the last image: TMA then reads zeros and stores nothing)
    auto tile_coords = [&](int pi, int& b, int& y0, int& x0) {
        b = pi / patches;
        const int pr = pi - b * patches;
        y0 = (pr / p.ptx) * HL_PH; x0 = (pr - (pr / p.ptx) * p.ptx) * HL_PW;
    };

    if (warp == 0) {
        // ===== TMA producer (warp-uniform loop, one elected lane issues) =====
        int ia = 0, ib = 0;
        const uint32_t stage_b = (uint32_t)(q.tps * B_BYTES);
        for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
            const int pp = unit / q.ntiles_n, n0 = (unit - pp * q.ntiles_n) * BN;
            int b0, y00, x00, b1, y01, x01;
            tile_coords(2 * pp, b0, y00, x00);
            tile_coords(2 * pp + 1, b1, y01, x01);
            for (int kc = 0; kc < nkc; kc++, ia++) {
                const int sa = ia % q.a_stages, ua = ia / q.a_stages;
                if (ua > 0) mbar_wait(&bar_emptyA[sa], (uint32_t)((ua - 1) & 1));
                if (elect_one_sync()) {
                    mbar_expect_tx(&bar_fullA[sa], (uint32_t)(2 * HL_A_BYTES));
                    tma_load_4d(smem + (size_t)sa * A_PAIR, &tmA, &bar_fullA[sa], kc * TM_KC, x00 - 1, y00 - 1, b0);
                    tma_load_4d(smem + (size_t)sa * A_PAIR + HL_A_STRIDE, &tmA, &bar_fullA[sa], kc * TM_KC, x01 - 1, y01 - 1, b1);
                }
                __syncwarp();
                for (int t0 = 0; t0 < 9; t0 += q.tps, ib++) {               // one weight stage = q.tps taps from one 3-D box
                    const int sb = ib % q.b_stages, ub = ib / q.b_stages;
                    if (ub > 0) mbar_wait(&bar_emptyB[sb], (uint32_t)((ub - 1) & 1));
                    if (elect_one_sync()) {
                        mbar_expect_tx(&bar_fullB[sb], stage_b);
                        tma_load_3d(smemB + (size_t)sb * stage_b, &tmB, &bar_fullB[sb], kc * TM_KC, n0, t0);
                    }
                    __syncwarp();
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (warp-uniform loop, one elected lane issues) =====
        int ia = 0, ib = 0, iu = 0;
        for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x, iu++) {
            const int ubuf = iu & 1, ut = iu >> 1;
            if (ut > 0) mbar_wait(&bar_tempty[ubuf], (uint32_t)((ut - 1) & 1));     // both epilogue groups have drained this accumulator pair
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t dcol0 = tmem_d + (uint32_t)(ubuf * 2 * BN), dcol1 = dcol0 + (uint32_t)BN;
            for (int kc = 0; kc < nkc; kc++, ia++) {
                const int sa = ia % q.a_stages, ua = ia / q.a_stages;
                mbar_wait(&bar_fullA[sa], (uint32_t)(ua & 1));
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint64_t da00 = make_desc_k_sw128_halo(smemA_u + (uint32_t)(sa * A_PAIR), 0);
                const uint64_t da10 = make_desc_k_sw128_halo(smemA_u + (uint32_t)(sa * A_PAIR + HL_A_STRIDE), 0);
#pragma unroll 1
                for (int t0 = 0; t0 < 9; t0 += q.tps, ib++) {
                    const int sb = ib % q.b_stages, ub = ib / q.b_stages;
                    mbar_wait(&bar_fullB[sb], (uint32_t)(ub & 1));
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint64_t dbs = make_desc_k_sw128(smemB_u + (uint32_t)(sb * q.tps * B_BYTES));
                    if (elect_one_sync()) {
                        for (int tl = 0; tl < q.tps; tl++) {
                            const int tap = t0 + tl;
                            const uint64_t db0 = dbs + (uint64_t)((tl * B_BYTES) >> 4);
                            const uint64_t toff = (uint64_t)(((tap / 3) * HL_HW + (tap % 3)) * 128 >> 4);
                            if (tap == 0) { umma_tf32(dcol0, da00 + toff, db0, IDESC, kc != 0 ? 1u : 0u); umma_tf32(dcol1, da10 + toff, db0, IDESC, kc != 0 ? 1u : 0u); }
                            else { umma_tf32_acc(dcol0, da00 + toff, db0, IDESC); umma_tf32_acc(dcol1, da10 + toff, db0, IDESC); }
#pragma unroll
                            for (int k = 1; k < TM_KC / 8; k++) {
                                umma_tf32_acc(dcol0, da00 + toff + (uint64_t)(2 * k), db0 + (uint64_t)(2 * k), IDESC);
                                umma_tf32_acc(dcol1, da10 + toff + (uint64_t)(2 * k), db0 + (uint64_t)(2 * k), IDESC);
                            }
                        }
                        umma_commit(&bar_emptyB[sb]);
                        if (t0 + q.tps >= 9) {
                            umma_commit(&bar_emptyA[sa]);
                            if (kc == nkc - 1) umma_commit(&bar_tfull[ubuf]);
                        }
                    }
                    __syncwarp();
                }
            }
        }
    } else if (warp >= 4) {
        // ===== epilogue: group g drains tile g of every pair =====
        const int g = (warp - 4) >> 2;
        EpiCtx e;
        e.bias_s = bias_s;
        e.smemO_u = smem_u32(smemO) + (uint32_t)(g * q.o_stages * 16384);
        e.ofull = bar_ofull[g]; e.oempty = bar_oempty[g];
        e.o_stages = q.o_stages; e.obuf = 0; e.ouse = 0;
        e.row = (warp & 3) * 32 + lane; e.lane = lane;
        e.slope = p.slope;
        e.act_mode = p.act != ACT_RELU ? 0 : ((p.slope >= 0.f && p.slope < 1.f) ? 1 : 2);
        e.slope_eff = e.act_mode == 0 ? 1.f : p.slope;
        int iu = 0;
        for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x, iu++) {
            const int ubuf = iu & 1, ut = iu >> 1;
            const int pp = unit / q.ntiles_n, n0 = (unit - pp * q.ntiles_n) * BN;
            int b, y0, x0;
            tile_coords(2 * pp + g, b, y0, x0);
            mbar_wait(&bar_tfull[ubuf], (uint32_t)(ut & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t lane_addr = tmem_d + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((ubuf * 2 + g) * BN);
            epi_tile<BN>(p, e, lane_addr, &bar_tempty[ubuf], b, y0, x0, n0);
        }
    } else if (lane == 0 && p.dst) {
        // ===== store issuer of group g =====
        const int g = warp - 2;
        int obuf = 0, ouse = 0;
        for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
            const int pp = unit / q.ntiles_n, n0 = (unit - pp * q.ntiles_n) * BN;
            int b, y0, x0;
            tile_coords(2 * pp + g, b, y0, x0);
            for (int sl = 0; sl < BN / 32; sl++) {
                mbar_wait(&bar_ofull[g][obuf], (uint32_t)(ouse & 1));
                tma_store_4d(&tmO, smemO + (size_t)(g * q.o_stages + obuf) * 16384, n0 + sl * 32, x0, y0, b);
                tma_store_commit();
                tma_store_wait_read<0>();
                mbar_arrive(&bar_oempty[g][obuf]);
                if (++obuf == q.o_stages) { obuf = 0; ouse++; }
            }
        }
    }
    __syncwarp();
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"((uint32_t)TCOLS));
}

}  // namespace

static long long* g_persist_trace = nullptr;     // set by dincae_debug_conv_trace only

#define PERSIST_LAUNCH(BN_)                                                                                                \
    do {                                                                                                                   \
        auto conv3x3_persist_bn##BN_ = conv3x3_persist<BN_>;                                                               \
        static int attr_smem = 0;                                                                                          \
        if (smem > attr_smem) {                                                                                            \
            DC_CHECK_CUDA(cudaFuncSetAttribute(conv3x3_persist_bn##BN_, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); \
            attr_smem = smem;                                                                                              \
        }                                                                                                                  \
        if (q.b_resident) DC_TRY(make_tmap_w(&tmB, a.wgtK, a.ldk, a.Cout_p, BN_));                                         \
        else DC_TRY(make_tmap_w3(&tmB, a.wgtK, a.in.C, a.Cout_p, BN_, q.tps));                                             \
        DC_WORK(ctx, a.work_flops > 0 ? a.work_flops : 2.0 * 9 * a.in.C * a.Cout_p * (double)a.B * a.H * a.W, a.work_bytes); \
        DC_LAUNCH_PDL(ctx, conv3x3_persist_bn##BN_, grid, q.groups == 2 ? PS_THREADS : 256, smem, tmA, tmB, tmO, p, q);                                         \
        DC_CHECK_LAUNCH();                                                                                                 \
        return DINCAE_OK;                                                                                                  \
    } while (0)

#define PAIR_LAUNCH(BN_)                                                                                                   \
    do {                                                                                                                   \
        auto conv3x3_pair_bn##BN_ = conv3x3_pair<BN_>;                                                                     \
        static int attr_smem = 0;                                                                                          \
        if (smem > attr_smem) {                                                                                            \
            DC_CHECK_CUDA(cudaFuncSetAttribute(conv3x3_pair_bn##BN_, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));  \
            attr_smem = smem;                                                                                              \
        }                                                                                                                  \
        if (q.b_resident) DC_TRY(make_tmap_w(&tmB, a.wgtK, a.ldk, a.Cout_p, BN_));                                         \
        else DC_TRY(make_tmap_w3(&tmB, a.wgtK, a.in.C, a.Cout_p, BN_, q.tps));                                             \
        DC_WORK(ctx, a.work_flops > 0 ? a.work_flops : 2.0 * 9 * a.in.C * a.Cout_p * (double)a.B * a.H * a.W, a.work_bytes); \
        DC_LAUNCH_PDL(ctx, conv3x3_pair_bn##BN_, grid, PS_THREADS, smem, tmA, tmB, tmO, p, q);                                 \
        DC_CHECK_LAUNCH();                                                                                                 \
        return DINCAE_OK;                                                                                                  \
    } while (0)

static int32_t k2_conv_persist(dincae_ctx* ctx, const ConvArgs& a) {
    TmaConvArgs p;
    p.B = a.B; p.H = a.H; p.W = a.W;
    p.PW = HL_PW; p.PH = HL_PH;
    p.ptx = (int)dc_cdiv(a.W, p.PW); p.pty = (int)dc_cdiv(a.H, p.PH);
    p.Cin_p = a.in.C; p.Cout_p = a.Cout_p;
    p.bias = a.bias; p.act = a.act; p.slope = a.slope; p.addend = a.addend; p.dst = a.dst;
    p.planar_out = a.planar_out; p.planar_C = a.planar_C; p.planar_accumulate = a.planar_accumulate; p.planar_fold = a.planar_fold;
    DC_REQUIRE(!a.planar_fold || (a.planar_out && a.planar_C <= 8 && a.Cout_p >= 16), DINCAE_EINVAL, "conv3x3_persist: planar_fold needs a planar output of at most 8 channels");
    p.pool_dst = a.pool_dst; p.pool_rnd = ctx->tf32_acts ? 1 : 0; p.mask_dst = a.mask_dst;
    DC_REQUIRE(!a.pool_dst || k2_conv_fuses_pool(a), DINCAE_EINVAL, "conv3x3_persist: pool fusion needs even H, W, a leaky slope in [0,1) and no addend");
    DC_REQUIRE(!a.mask_dst || (a.pool_dst && !a.dst && a.act == ACT_RELU), DINCAE_EINVAL, "conv3x3_persist: the sign mask belongs to the pool-only mode");
    int bn = a.Cout_p % 256 == 0 ? 256 : (a.Cout_p % 128 == 0 ? 128 : (a.Cout_p % 64 == 0 ? 64 : 32));
    const int m_tiles = p.ptx * p.pty * a.B;
    while (bn > 64 && m_tiles * (a.Cout_p / bn) < ctx->num_sms) bn /= 2;     // few pixel tiles (deep levels): split N so every SM has work
    const int nkc = a.in.C / TM_KC;
    PersistArgs q;
    q.ntiles_n = a.Cout_p / bn;
    q.num_tiles = m_tiles * q.ntiles_n;
    q.trace = g_persist_trace;
    q.ksteps = a.in.C / 8;
    if (a.in_real_C > 0 && a.in_real_C < a.in.C) q.ksteps = (int)dc_cdiv(a.in_real_C, 8);
    const int bbytes = bn * 128;
    const int resident_bytes = nkc * 9 * bbytes;
    // shared-memory plan.  Narrow resident layers: two CTAs per SM, each with one epilogue group, two activation stages and
    // one staging slab (the second CTA hides the bubbles between tiles).  Everything else: one CTA per SM with two epilogue
    // groups; resident weights if they fit beside two activation stages, else a ring of weight stages; whatever is left
    // deepens the activation ring (TMA latency is 2-3k cycles).
    q.tps = 1;
    const int static_smem = 5120;                     // barriers + bias (~3.2 KB) and the 1 KB alignment slack, rounded up
    const int small_plan = 2 * HL_A_STRIDE + resident_bytes + (a.dst ? 16384 : 0) + 1024;
    const bool two_per_sm = q.ntiles_n == 1 && 2 * (small_plan + static_smem) <= 227 * 1024 && 4 * bn <= 512;
    int smem, per_sm;
    if (two_per_sm) {
        q.groups = 1; q.b_resident = 1; q.b_stages = 0; q.a_stages = 2; q.o_stages = a.dst ? 1 : 0;
        DC_REQUIRE(q.ntiles_n == 1 && small_plan + static_smem <= 227 * 1024, DINCAE_EUNSUPPORTED, "conv3x3_persist: one-group plan does not fit");
        smem = small_plan;
        per_sm = two_per_sm ? 2 : 1;
    } else {
        q.groups = 2;
        const int budget = 227 * 1024 - static_smem - 1024;
        q.o_stages = a.dst ? 2 : 0;
        int left = budget - 2 * q.o_stages * 16384;
        q.b_resident = (q.ntiles_n == 1 && resident_bytes + 2 * HL_A_STRIDE <= left) ? 1 : 0;
        if (q.b_resident) {
            q.b_stages = 0;
            left -= resident_bytes;
        } else {
            q.tps = 3;                                           // one kernel row of taps per weight stage ...
            int bs = (left - 2 * HL_A_STRIDE) / (q.tps * bbytes);
            if (bs < 2) { q.tps = 1; bs = (left - 2 * HL_A_STRIDE) / bbytes; }      // ... unless two such stages do not fit
            if (bs > PS_MAX_B) bs = PS_MAX_B;
            DC_REQUIRE(bs >= 2, DINCAE_EUNSUPPORTED, "conv3x3_persist: no room for two weight stages (BN=%d)", bn);
            q.b_stages = bs;
            left -= bs * q.tps * bbytes;
        }
        q.a_stages = left / HL_A_STRIDE;
        if (q.a_stages > PS_MAX_A) q.a_stages = PS_MAX_A;
        smem = q.a_stages * HL_A_STRIDE + (q.b_resident ? resident_bytes : q.b_stages * q.tps * bbytes) + 2 * q.o_stages * 16384 + 1024;
        per_sm = 1;
    }
    // paired tiles: streamed weights (deep levels), two pixel tiles per CTA share every weight stage
    if (q.groups == 2 && !q.b_resident && bn <= 128 && m_tiles >= 2 && ((int)dc_cdiv(m_tiles, 2) * q.ntiles_n >= (ctx->num_sms * 3) / 4)) {
        const int budget = 227 * 1024 - static_smem - 1024;
        q.o_stages = a.dst ? 1 : 0;
        int left = budget - 2 * q.o_stages * 16384 - 2 * (2 * HL_A_STRIDE);
        q.tps = 3;
        int bs = left / (q.tps * bbytes);
        if (bs < 2) { q.tps = 1; bs = left / bbytes; }
        if (bs > PS_MAX_B) bs = PS_MAX_B;
        if (bs >= 2) {
            q.b_stages = bs; q.a_stages = 2;
            left -= bs * q.tps * bbytes;
            if (left >= 2 * HL_A_STRIDE) q.a_stages = 3;
            smem = q.a_stages * 2 * HL_A_STRIDE + q.b_stages * q.tps * bbytes + 2 * q.o_stages * 16384 + 1024;
            q.num_tiles = (int)dc_cdiv(m_tiles, 2) * q.ntiles_n;          // units
            int ctas = ctx->num_sms < q.num_tiles ? ctx->num_sms : q.num_tiles;
            dim3 grid((unsigned)ctas);
            alignas(64) CUtensorMap tmA, tmB, tmO;
            DC_TRY(make_tmap_halo(&tmA, a.in.src0, a.in.C, a.W, a.H, a.B));
            if (a.dst) DC_TRY(make_tmap_act(&tmO, a.dst, a.Cout_p, a.W, a.H, a.B, HL_PW, HL_PH));
            else tmO = tmA;
            if (bn == 128) PAIR_LAUNCH(128);
            if (bn == 64) PAIR_LAUNCH(64);
            PAIR_LAUNCH(32);
        }
    }
    int ctas = ctx->num_sms * per_sm;                 // persistent
    if (ctas > q.num_tiles) ctas = q.num_tiles;
    dim3 grid((unsigned)ctas);
    alignas(64) CUtensorMap tmA, tmB, tmO;
    DC_TRY(make_tmap_halo(&tmA, a.in.src0, a.in.C, a.W, a.H, a.B));
    if (a.dst) DC_TRY(make_tmap_act(&tmO, a.dst, a.Cout_p, a.W, a.H, a.B, HL_PW, HL_PH));
    else tmO = tmA;
    if (bn == 256) PERSIST_LAUNCH(256);
    if (bn == 128) PERSIST_LAUNCH(128);
    if (bn == 64) PERSIST_LAUNCH(64);
    PERSIST_LAUNCH(32);
}

// true when k2_conv() on these arguments can also emit MeanPool((2,2)) of its output into a.pool_dst (even H and W so that no window
// straddles the padding, activation in the max(v, v*slope) form, no residual add, NHWC or no full-resolution output)
bool k2_conv_fuses_pool(const ConvArgs& a) {
    static const int off = dc_env_int("DINCAE_NO_POOL_FUSION", 0);
    return !off && k2_conv_tma_eligible(a) && a.planar_out == nullptr && a.addend == nullptr && (a.H % 2) == 0 && (a.W % 2) == 0 &&
           a.B * a.H * a.W > 0 && (a.act != ACT_RELU || (a.slope >= 0.f && a.slope < 1.f));
}
// the persistent tcgen05 kernel takes it; everything else (more than 512 output channels, a wide planar output, a fused
// SIMT producer, no tensor-map encoder in the driver) goes to the FP32 CUDA-core kernel
bool k2_conv_tma_eligible(const ConvArgs& a) {
    return a.wgtK != nullptr && a.in.mode == SRC_DIRECT && a.in.src1 == nullptr && (a.in.C % TM_KC) == 0 && (a.Cout_p % 32) == 0 &&
           a.in_is_tf32 && a.Cout_p <= PS_MAX_COUT && (a.planar_out == nullptr || a.planar_C <= 16) && get_encode() != nullptr;
}

int32_t k2_conv_tma(dincae_ctx* ctx, const ConvArgs& a) {
    if (a.B * a.H * a.W == 0) return DINCAE_OK;
    return k2_conv_persist(ctx, a);
}

// =================================================================================================
// Weight gradient on the tensor cores:  dW[tap][ci][co] = sum over pixels of in[pixel + tap][ci] * g[pixel][co].
// Both operands are pixel-major NHWC tensors, i.e. MN-major for the MMA (channels contiguous, the reduction index
// = pixel is the slow one).  Per stage one 8x4 pixel patch is streamed by TMA as 32-channel boxes (4 KB each, 32 pixel
// rows x 128 B, 128B swizzle with 32B atoms = the canonical MN-major TF32 layout: 4-row atoms 512 B apart, 32-channel blocks LBO apart);
// the tap is a shift of the `in` box, padding is TMA zero fill.  D (128 x N, FP32) lives in TMEM; the reduction over
// pixels is split over CTAs and finished by a deterministic second-stage sum (no float atomics).
namespace {


struct WgradTmaArgs {
    int B, H, W;
    int ptx, pty;            // 8x4 patches per image
    int Cin_p, Cout_p;
    int in_is_m;             // 1: M = ci (in), N = co (g);  0: M = co (g), N = ci (in)
    int m_ch, n_ch;          // channels of the M / N operand
    int64_t npatches, patches_per_split;
    float* partial;          // [splits][9*Cin_p*Cout_p]
};

// MN-major TF32 operands must use the 128-byte swizzle with 32-byte atomicity (layout type SWIZZLE_128B_BASE32B,
// Swizzle<2,5,2>: the 32-byte chunk index is XORed with the pixel-row index mod 4); its K atom is 4 rows (512 B), so one
// K = 8 MMA spans two atoms SBO apart.  TMA writes the same pattern with CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B.

__global__ void __launch_bounds__(256) reduce_splits_kernel(const float* __restrict__ partial, int splits, int64_t n, float* __restrict__ dst) {
    __shared__ float sh[8][32];
    const int c = threadIdx.x & 31, zg = threadIdx.x >> 5;
    const int64_t i = (int64_t)blockIdx.x * 32 + c;
    float s = 0.f;
    if (i < n) {
        int z = zg;
        for (; z + 24 < splits; z += 32) {
            float v0 = __ldg(partial + (size_t)z * n + i), v1 = __ldg(partial + (size_t)(z + 8) * n + i);
            float v2 = __ldg(partial + (size_t)(z + 16) * n + i), v3 = __ldg(partial + (size_t)(z + 24) * n + i);
            s += (v0 + v1) + (v2 + v3);
        }
        for (; z < splits; z += 8) s += __ldg(partial + (size_t)z * n + i);
    }
    sh[zg][c] = s;
    __syncthreads();
    if (zg == 0 && i < n) {
        float tot = 0.f;
#pragma unroll
        for (int k = 0; k < 8; k++) tot += sh[k][c];
        dst[i] = tot;
    }
}

// One block = `cols` float4 column groups x (256/cols) split groups.  Wide entries (dW) use 32 x 8, narrow entries with many
// partial rows (bias column partials: n = C, splits ~ 1000) use 8 x 32 so that enough threads share the long sums.
// Every entry has n % 4 == 0 and 16-byte aligned rows (checked on the host), so each thread streams float4s; four rows are
// in flight per thread.  The summation order is fixed: rows zg, zg + G, ... per thread, then groups 0..G-1.
__global__ void __launch_bounds__(256) reduce_batched_kernel(const __grid_constant__ RedList L) {
    __shared__ float4 sh[256];
    int e = 0;
    while (e + 1 < L.count && (int)blockIdx.x >= L.blk_start[e + 1]) e++;
    const RedEntry en = L.e[e];
    const int cols = en.n <= 512 ? 8 : 32, ngroups = 256 / cols;
    const int c = threadIdx.x % cols, zg = threadIdx.x / cols;
    const int64_t n4 = en.n >> 2;
    const int64_t i = (int64_t)((int)blockIdx.x - L.blk_start[e]) * cols + c;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < n4) {
        const float4* src = reinterpret_cast<const float4*>(en.part) + i;
        int z = zg;
        for (; z + 3 * ngroups < en.splits; z += 4 * ngroups) {
            const float4 v0 = __ldg(src + (size_t)z * n4), v1 = __ldg(src + (size_t)(z + ngroups) * n4);
            const float4 v2 = __ldg(src + (size_t)(z + 2 * ngroups) * n4), v3 = __ldg(src + (size_t)(z + 3 * ngroups) * n4);
            s.x += (v0.x + v1.x) + (v2.x + v3.x); s.y += (v0.y + v1.y) + (v2.y + v3.y);
            s.z += (v0.z + v1.z) + (v2.z + v3.z); s.w += (v0.w + v1.w) + (v2.w + v3.w);
        }
        for (; z < en.splits; z += ngroups) {
            const float4 v = __ldg(src + (size_t)z * n4);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
    }
    sh[threadIdx.x] = s;
    __syncthreads();
    if (zg == 0 && i < n4) {
        float4 tot = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int k = 0; k < ngroups; k++) {
            const float4 v = sh[k * cols + c];
            tot.x += v.x; tot.y += v.y; tot.z += v.z; tot.w += v.w;
        }
        reinterpret_cast<float4*>(en.dst)[i] = tot;
    }
}

// bias gradient: column sums of g [M][C] (C % 4 == 0) in two deterministic stages.
// stage 1: block b sums rows b, b+gridDim, ... ; thread = (row lane, float4 column group); fixed-order smem combine.
__global__ void __launch_bounds__(256) colsum_partial_kernel(const float* __restrict__ g, int64_t M, int C, float* __restrict__ part) {
    const int C4 = C >> 2;                       // float4 column groups (<= 64 per pass)
    __shared__ float4 sh[256];
    for (int cg0 = 0; cg0 < C4; cg0 += 64) {
        const int ncg = C4 - cg0 < 64 ? C4 - cg0 : 64;
        const int lanes = 256 / ncg;             // row lanes
        const int cg = threadIdx.x % ncg, rl = threadIdx.x / ncg;
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        if (rl < lanes) {
            const int64_t stride = (int64_t)gridDim.x * lanes;
            int64_t r = (int64_t)blockIdx.x * lanes + rl;
            for (; r + 3 * stride < M; r += 4 * stride) {
                float4 v0 = __ldg(reinterpret_cast<const float4*>(g + r * C) + cg0 + cg);
                float4 v1 = __ldg(reinterpret_cast<const float4*>(g + (r + stride) * C) + cg0 + cg);
                float4 v2 = __ldg(reinterpret_cast<const float4*>(g + (r + 2 * stride) * C) + cg0 + cg);
                float4 v3 = __ldg(reinterpret_cast<const float4*>(g + (r + 3 * stride) * C) + cg0 + cg);
                s.x += (v0.x + v1.x) + (v2.x + v3.x); s.y += (v0.y + v1.y) + (v2.y + v3.y);
                s.z += (v0.z + v1.z) + (v2.z + v3.z); s.w += (v0.w + v1.w) + (v2.w + v3.w);
            }
            for (; r < M; r += stride) {
                float4 v = __ldg(reinterpret_cast<const float4*>(g + r * C) + cg0 + cg);
                s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
            }
        }
        __syncthreads();
        sh[threadIdx.x] = s;
        __syncthreads();
        if (threadIdx.x < ncg) {
            float4 tot = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int k = 0; k < lanes; k++) { float4 v = sh[k * ncg + threadIdx.x]; tot.x += v.x; tot.y += v.y; tot.z += v.z; tot.w += v.w; }
            reinterpret_cast<float4*>(part + (size_t)blockIdx.x * C)[cg0 + threadIdx.x] = tot;
        }
    }
}
// stage 2: one warp per column; lane l sums partials l, l+32, ... then a fixed-order shuffle tree

// Three taps per CTA (dx = -1, 0, +1 of one kernel row ky): the `in` operand is loaded once per patch as a (32 ch x 10 x 4)
// box with a one-pixel halo in x; the three taps are three descriptor views of it (start shifted by dx pixel rows of 128 B,
// the four 8-pixel k-steps one box row = 1280 B apart) feeding three TMEM accumulators, while the g box is shared.
__device__ __forceinline__ uint64_t make_desc_mn_sw128_lbo(uint32_t saddr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)(lbo_bytes >> 4) << 16;
    d |= (uint64_t)(512 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)1 << 61;
    return d;
}

constexpr int W9_PR = 8;
constexpr int W9_MBOX = 10 * W9_PR * 128;          // (10 x PR) box of the 32-channel operand
constexpr int W9_MREGION = W9_MBOX + 1024;         // + zeroed overrun read by the (discarded) padding block
constexpr int W9_NBOX = 8 * (W9_PR + 2) * 128;     // (8 x (PR+2)) box per 32-channel block of the other operand (multiple of 1024)

template <int NB, int STAGES>
__global__ void __launch_bounds__(128) wgrad9_kernel(const __grid_constant__ CUtensorMap tmM, const __grid_constant__ CUtensorMap tmN, WgradTmaArgs p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    constexpr int NCB = NB / 32;                   // 32-channel blocks of the N operand
    constexpr int STAGE = W9_MREGION + NCB * W9_NBOX;
    constexpr uint32_t TCOLS = NCB == 1 ? 128u : 256u;
    __shared__ uint64_t bar_full[STAGES];
    __shared__ uint64_t bar_empty[STAGES];
    __shared__ uint64_t bar_done;
    __shared__ uint32_t tmem_base_s;
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    const uint32_t smem_base = smem_u32(smem);
    const int t = threadIdx.x, warp = t >> 5;
    const int split = blockIdx.x;
    const int64_t p_lo = (int64_t)split * p.patches_per_split;
    int64_t p_hi = p_lo + p.patches_per_split;
    if (p_hi > p.npatches) p_hi = p.npatches;
    const int nst = (int)(p_hi - p_lo);

    // the overrun of the (discarded) 4th M block must stay finite
    for (int s = 0; s < STAGES; s++)
        for (int i = t; i < 1024 / 16; i += 128)
            reinterpret_cast<float4*>(smem + (size_t)s * STAGE + W9_MBOX)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (t == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmM)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmN)) : "memory");
#pragma unroll
        for (int s = 0; s < STAGES; s++) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], 1); }
        mbar_init(&bar_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(TCOLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base_s;
    constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(96 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);

    if (t == 0) {
        const int per_img = p.ptx * p.pty;
        for (int i = 0; i < nst; i++) {
            const int s = i % STAGES, u = i / STAGES;
            if (u > 0) mbar_wait(&bar_empty[s], (uint32_t)((u - 1) & 1));
            const int64_t pp = p_lo + i;
            const int b = (int)(pp / per_img);
            const int r = (int)(pp - (int64_t)b * per_img);
            const int y0 = (r / p.ptx) * W9_PR, x0 = (r - (r / p.ptx) * p.ptx) * 8;
            uint8_t* st = smem + (size_t)s * STAGE;
            mbar_expect_tx(&bar_full[s], (uint32_t)(W9_MBOX + NCB * W9_NBOX));
            tma_load_4d(st, &tmM, &bar_full[s], 0, x0 - 1, y0, b);                                // (32, 10, PR, 1)
#pragma unroll
            for (int cb = 0; cb < NCB; cb++)
                tma_load_4d(st + W9_MREGION + cb * W9_NBOX, &tmN, &bar_full[s], cb * 32, x0, y0 - 1, b);   // (32, 8, PR+2, 1)
        }
    } else if (warp == 1) {
        // MMA issuer: warp-uniform loop, one elected lane issues
        for (int i = 0; i < nst; i++) {
            const int s = i % STAGES, u = i / STAGES;
            mbar_wait(&bar_full[s], (uint32_t)(u & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t abase = smem_base + (uint32_t)(s * STAGE);
            const uint64_t da0 = make_desc_mn_sw128_lbo(abase, 128u);                       // M block i = the box shifted by i pixels
            const uint64_t db0 = make_desc_mn_sw128_lbo(abase + (uint32_t)W9_MREGION, 1024u);  // N block j = the box shifted by j rows
            if (elect_one_sync()) {
#pragma unroll
                for (int h = 0; h < W9_PR; h++) {                                           // one k-step = the 8 pixels of box row h
#pragma unroll
                    for (int cb = 0; cb < NCB; cb++) {
                        const uint64_t da = da0 + (uint64_t)((h * 10 * 128) >> 4), db = db0 + (uint64_t)((cb * W9_NBOX + h * 1024) >> 4);
                        if (h == 0) umma_tf32(tmem_d + (uint32_t)(cb * 96), da, db, IDESC, i != 0 ? 1u : 0u);
                        else umma_tf32_acc(tmem_d + (uint32_t)(cb * 96), da, db, IDESC);
                    }
                }
                umma_commit(&bar_empty[s]);
                if (i == nst - 1) umma_commit(&bar_done);
            }
            __syncwarp();
        }
    }
    __syncwarp();
    float* out = p.partial + (size_t)split * 9 * p.Cin_p * p.Cout_p;
    if (nst > 0) {
        mbar_wait(&bar_done, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    const uint32_t lane_addr = tmem_d + ((uint32_t)(warp * 32) << 16);
    const int bi = t >> 5, cm = t & 31;            // M block (warp) and channel of the 32-channel operand
#pragma unroll 1
    for (int cb = 0; cb < NCB; cb++) {
#pragma unroll 1
        for (int j = 0; j < 3; j++) {
#pragma unroll 1
            for (int c0 = 0; c0 < 32; c0 += 16) {
                uint32_t r[16];
                if (nst > 0) {
                    tmem_ld16(lane_addr + (uint32_t)(cb * 96 + j * 32 + c0), r);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                } else {
#pragma unroll
                    for (int k = 0; k < 16; k++) r[k] = 0u;
                }
                if (bi < 3) {
                    if (p.in_is_m) {               // thread = ci, 16 consecutive co
                        const int tap = (2 - j) * 3 + bi;
                        float* o = out + ((size_t)tap * p.Cin_p + cm) * p.Cout_p + cb * 32 + c0;
#pragma unroll
                        for (int k = 0; k < 4; k++)
                            *reinterpret_cast<float4*>(o + 4 * k) = make_float4(__uint_as_float(r[4 * k]), __uint_as_float(r[4 * k + 1]), __uint_as_float(r[4 * k + 2]), __uint_as_float(r[4 * k + 3]));
                    } else {                       // thread = co (lanes coalesce), column = ci
                        const int tap = j * 3 + (2 - bi);
#pragma unroll
                        for (int k = 0; k < 16; k++)
                            out[((size_t)tap * p.Cin_p + cb * 32 + c0 + k) * p.Cout_p + cm] = __uint_as_float(r[k]);
                    }
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(TCOLS));
}

// The same idea for wide layers: M = 128 channels of the wider operand (four 32-channel boxes of the (10 x 4) patch halo), N = 32
// channels of the other operand with the three dy shifts stacked (one (8 x 6) box), and the three dx shifts as three start
// offsets of the M descriptor into three accumulators (3 x 96 TMEM columns).  One CTA = one (128-channel, 32-channel) block of
// all nine taps; per 32 pixels it loads 26 KB where the one-tap-per-CTA kernel loads 9 x (16 + 4*BN/32) KB -- the deep levels
// are bound by exactly this L2 -> SM stream.
//   in_is_m = 1:  M = in  (offset dx),      N = g  (dy = 2 - block j)
//   in_is_m = 0:  M = g   (offset 2 - dx),  N = in (dy = block j)
constexpr int WM_MREGION = 4 * W9_MBOX, WM_STAGE = WM_MREGION + W9_NBOX;      // 50 KB

template <int STAGES>
__global__ void __launch_bounds__(128) wgradm_kernel(const __grid_constant__ CUtensorMap tmM, const __grid_constant__ CUtensorMap tmN, WgradTmaArgs p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ uint64_t bar_full[STAGES];
    __shared__ uint64_t bar_empty[STAGES];
    __shared__ uint64_t bar_done;
    __shared__ uint32_t tmem_base_s;
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    const uint32_t smem_base = smem_u32(smem);
    const int t = threadIdx.x, warp = t >> 5;
    const int split = blockIdx.x, m0 = blockIdx.y * 128, n0 = blockIdx.z * 32;
    const int64_t p_lo = (int64_t)split * p.patches_per_split;
    int64_t p_hi = p_lo + p.patches_per_split;
    if (p_hi > p.npatches) p_hi = p.npatches;
    const int nst = (int)(p_hi - p_lo);

    if (t == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmM)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmN)) : "memory");
#pragma unroll
        for (int s = 0; s < STAGES; s++) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], 1); }
        mbar_init(&bar_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512u));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base_s;
    constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(96 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);

    if (t == 0) {
        const int per_img = p.ptx * p.pty;
        for (int i = 0; i < nst; i++) {
            const int s = i % STAGES, u = i / STAGES;
            if (u > 0) mbar_wait(&bar_empty[s], (uint32_t)((u - 1) & 1));
            const int64_t pp = p_lo + i;
            const int b = (int)(pp / per_img);
            const int r = (int)(pp - (int64_t)b * per_img);
            const int y0 = (r / p.ptx) * W9_PR, x0 = (r - (r / p.ptx) * p.ptx) * 8;
            uint8_t* st = smem + (size_t)s * WM_STAGE;
            mbar_expect_tx(&bar_full[s], (uint32_t)WM_STAGE);
#pragma unroll
            for (int j = 0; j < 4; j++) tma_load_4d(st + j * W9_MBOX, &tmM, &bar_full[s], m0 + j * 32, x0 - 1, y0, b);   // channels past the tensor read zeros
            tma_load_4d(st + WM_MREGION, &tmN, &bar_full[s], n0, x0, y0 - 1, b);
        }
    } else if (warp == 1) {
        for (int i = 0; i < nst; i++) {
            const int s = i % STAGES, u = i / STAGES;
            mbar_wait(&bar_full[s], (uint32_t)(u & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t abase = smem_base + (uint32_t)(s * WM_STAGE);
            const uint64_t da0 = make_desc_mn_sw128_lbo(abase, (uint32_t)W9_MBOX);                // M block = next 32-channel box
            const uint64_t db0 = make_desc_mn_sw128_lbo(abase + (uint32_t)WM_MREGION, 1024u);   // N block j = the box shifted by j rows
            if (elect_one_sync()) {
#pragma unroll
                for (int h = 0; h < W9_PR; h++) {
#pragma unroll
                    for (int dx = 0; dx < 3; dx++) {
                        const uint64_t da = da0 + (uint64_t)((h * 1280 + dx * 128) >> 4), db = db0 + (uint64_t)((h * 1024) >> 4);
                        if (h == 0) umma_tf32(tmem_d + (uint32_t)(dx * 96), da, db, IDESC, i != 0 ? 1u : 0u);
                        else umma_tf32_acc(tmem_d + (uint32_t)(dx * 96), da, db, IDESC);
                    }
                }
                umma_commit(&bar_empty[s]);
                if (i == nst - 1) umma_commit(&bar_done);
            }
            __syncwarp();
        }
    }
    __syncwarp();
    float* out = p.partial + (size_t)split * 9 * p.Cin_p * p.Cout_p;
    if (nst > 0) {
        mbar_wait(&bar_done, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    const uint32_t lane_addr = tmem_d + ((uint32_t)(warp * 32) << 16);
    const int mch = m0 + t;
#pragma unroll 1
    for (int o = 0; o < 3; o++) {                  // descriptor offset o: dx = o (in_is_m) or 2 - o
#pragma unroll 1
        for (int j = 0; j < 3; j++) {
#pragma unroll 1
            for (int c0 = 0; c0 < 32; c0 += 16) {
                uint32_t r[16];
                if (nst > 0) {
                    tmem_ld16(lane_addr + (uint32_t)(o * 96 + j * 32 + c0), r);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                } else {
#pragma unroll
                    for (int k = 0; k < 16; k++) r[k] = 0u;
                }
                if (mch < p.m_ch) {
                    if (p.in_is_m) {               // thread = ci, 16 consecutive co
                        const int tap = (2 - j) * 3 + o;
                        float* dst = out + ((size_t)tap * p.Cin_p + mch) * p.Cout_p + n0 + c0;
#pragma unroll
                        for (int k = 0; k < 4; k++)
                            *reinterpret_cast<float4*>(dst + 4 * k) = make_float4(__uint_as_float(r[4 * k]), __uint_as_float(r[4 * k + 1]), __uint_as_float(r[4 * k + 2]), __uint_as_float(r[4 * k + 3]));
                    } else {                       // thread = co (lanes coalesce), column = ci
                        const int tap = j * 3 + (2 - o);
#pragma unroll
                        for (int k = 0; k < 16; k++)
                            out[((size_t)tap * p.Cin_p + n0 + c0 + k) * p.Cout_p + mch] = __uint_as_float(r[k]);
                    }
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(512u));
}

int32_t make_tmap_box(CUtensorMap* m, const float* ptr, int C, int W, int H, int B, int bw, int bh) {
    EncodeTiledFn enc = get_encode();
    DC_REQUIRE(enc != nullptr, DINCAE_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t gstr[3] = {(cuuint64_t)C * 4, (cuuint64_t)W * C * 4, (cuuint64_t)H * W * C * 4};
    cuuint32_t box[4] = {32, (cuuint32_t)bw, (cuuint32_t)bh, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (void*)ptr, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DC_REQUIRE(r == CUDA_SUCCESS, DINCAE_ECUDA, "cuTensorMapEncodeTiled(box %dx%d C=%d W=%d H=%d B=%d) failed: %d", bw, bh, C, W, H, B, (int)r);
    return DINCAE_OK;
}




// wgrad9_kernel: a 32-channel side against 32 or 64 channels; wgradm_kernel: every other shape
static bool wgrad9_selected(int Cin_p, int Cout_p) {
    return (Cin_p == 32 && (Cout_p == 32 || Cout_p == 64)) || (Cin_p == 64 && Cout_p == 32);
}
constexpr int WGRAD_PATCH_ROWS = 8;

void wgrad_tma_plan(int num_sms, const WgradArgs& a, int& bn, int& splits, int64_t& npatches, int64_t& pps, int& in_is_m) {
    const int Cin_p = a.in.C, Cout_p = a.Cout_p;
    in_is_m = Cin_p >= Cout_p;
    const int m_ch = in_is_m ? Cin_p : Cout_p, n_ch = in_is_m ? Cout_p : Cin_p;
    bn = n_ch % 256 == 0 ? 256 : (n_ch % 128 == 0 ? 128 : (n_ch % 64 == 0 ? 64 : 32));
    const int mtiles = (m_ch + 127) / 128;
    npatches = (int64_t)a.B * dc_cdiv(a.H, WGRAD_PATCH_ROWS) * dc_cdiv(a.W, 8);
    if (wgrad9_selected(Cin_p, Cout_p)) {                  // wgrad9_kernel: all nine taps in every MMA; M = the 32-channel side
        in_is_m = Cin_p == 32;
        splits = ((Cin_p == 32 && Cout_p == 32) ? 3 : 2) * num_sms;       // CTAs per SM the kernel fits (smem, TMEM)
    } else {                                               // wgradm_kernel: one CTA per SM, (128-channel, 32-channel) blocks of all nine taps
        const int tiles = mtiles * (n_ch / 32);
        splits = (num_sms + tiles - 1) / tiles;
    }
    int64_t maxs = dc_cdiv(npatches, 2);
    if (splits > maxs) splits = (int)maxs;
    // the split partials are written and re-read once: keep them below ~16 MB per layer
    static const int cap_mb = dc_env_int("DINCAE_WGRAD_CAP_MB", 16);
    int64_t cap = ((int64_t)cap_mb << 20) / ((int64_t)9 * Cin_p * Cout_p * 4);
    if (cap < 1) cap = 1;
    if (splits > cap) splits = (int)cap;
    if (splits < 1) splits = 1;
    pps = dc_cdiv(npatches, splits);
    splits = (int)dc_cdiv(npatches, pps);
}

}  // namespace

bool k2_wgrad_tma_eligible(const WgradArgs& a) {
    return a.tf32 && a.in.mode == SRC_DIRECT && a.in.src1 == nullptr && (a.in.C % 32) == 0 && (a.Cout_p % 32) == 0 && get_encode() != nullptr;
}

size_t k2_wgrad_tma_scratch_bytes(int num_sms, int B, int H, int W, int Cin_p, int Cout_p) {
    if (Cin_p % 32 || Cout_p % 32) return 0;
    WgradArgs a{};
    a.in.C = Cin_p; a.Cout_p = Cout_p; a.B = B; a.H = H; a.W = W;
    int bn, splits, in_is_m;
    int64_t np, pps;
    wgrad_tma_plan(num_sms, a, bn, splits, np, pps, in_is_m);
    return ((size_t)splits * 9 * Cin_p * Cout_p + 1024 * (size_t)Cout_p) * sizeof(float);
}

int32_t k2_wgrad_tma(dincae_ctx* ctx, const WgradArgs& a, void* scratch, size_t scratch_bytes) {
    const int Cin_p = a.in.C, Cout_p = a.Cout_p;
    int bn, splits, in_is_m;
    int64_t np, pps;
    wgrad_tma_plan(ctx->num_sms, a, bn, splits, np, pps, in_is_m);
    const int64_t nW = (int64_t)9 * Cin_p * Cout_p;
    const int64_t M = (int64_t)a.B * a.H * a.W;
    const int cs_blocks = (int)(M < (int64_t)4 * ctx->num_sms * 32 ? dc_cdiv(M, 32) : 4 * ctx->num_sms);
    DC_REQUIRE(scratch_bytes >= ((size_t)splits * nW + (size_t)cs_blocks * Cout_p) * sizeof(float), DINCAE_ESIZE, "k2_wgrad_tma: scratch too small");
    WgradTmaArgs p;
    p.B = a.B; p.H = a.H; p.W = a.W; p.ptx = (int)dc_cdiv(a.W, 8); p.pty = (int)dc_cdiv(a.H, WGRAD_PATCH_ROWS);
    p.Cin_p = Cin_p; p.Cout_p = Cout_p; p.in_is_m = in_is_m;
    p.m_ch = in_is_m ? Cin_p : Cout_p; p.n_ch = in_is_m ? Cout_p : Cin_p;
    p.npatches = np; p.patches_per_split = pps; p.partial = (float*)scratch;
    if (wgrad9_selected(Cin_p, Cout_p)) {
        alignas(64) CUtensorMap tmM, tmN;
        DC_TRY(make_tmap_box(&tmM, in_is_m ? a.in.src0 : a.gout, 32, a.W, a.H, a.B, 10, W9_PR));
        DC_TRY(make_tmap_box(&tmN, in_is_m ? a.gout : a.in.src0, p.n_ch, a.W, a.H, a.B, 8, W9_PR + 2));
        DC_WORK(ctx, a.work_flops > 0 ? a.work_flops : 2.0 * 9 * Cin_p * Cout_p * (double)a.B * a.H * a.W, a.work_bytes);
        if (p.n_ch == 32) {
            auto wgrad9_n32 = wgrad9_kernel<32, 3>;
            constexpr int smem = 3 * (W9_MREGION + W9_NBOX) + 1024;
            static bool attr_set = false;
            if (!attr_set) { DC_CHECK_CUDA(cudaFuncSetAttribute(wgrad9_n32, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); attr_set = true; }
            DC_LAUNCH(ctx, wgrad9_n32, (unsigned)splits, 128, smem, tmM, tmN, p);
        } else {
            auto wgrad9_n64 = wgrad9_kernel<64, 3>;
            constexpr int smem = 3 * (W9_MREGION + 2 * W9_NBOX) + 1024;
            static bool attr_set = false;
            if (!attr_set) { DC_CHECK_CUDA(cudaFuncSetAttribute(wgrad9_n64, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); attr_set = true; }
            DC_LAUNCH(ctx, wgrad9_n64, (unsigned)splits, 128, smem, tmM, tmN, p);
        }
        DC_CHECK_LAUNCH();
    } else {
        alignas(64) CUtensorMap tmM, tmN;
        DC_TRY(make_tmap_box(&tmM, in_is_m ? a.in.src0 : a.gout, p.m_ch, a.W, a.H, a.B, 10, W9_PR));
        DC_TRY(make_tmap_box(&tmN, in_is_m ? a.gout : a.in.src0, p.n_ch, a.W, a.H, a.B, 8, W9_PR + 2));
        DC_WORK(ctx, a.work_flops > 0 ? a.work_flops : 2.0 * 9 * Cin_p * Cout_p * (double)a.B * a.H * a.W, a.work_bytes);
        auto wgradm = wgradm_kernel<4>;
        constexpr int smem = 4 * WM_STAGE + 1024;
        static bool attr_set = false;
        if (!attr_set) { DC_CHECK_CUDA(cudaFuncSetAttribute(wgradm, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); attr_set = true; }
        dim3 grid((unsigned)splits, (unsigned)((p.m_ch + 127) / 128), (unsigned)(p.n_ch / 32));
        DC_LAUNCH(ctx, wgradm, grid, 128, smem, tmM, tmN, p);
        DC_CHECK_LAUNCH();
    }
    // second stages: split-K sum of dW and column-partial sum of the bias gradient (queued when a.defer is set)
    const float* cpart = a.colpart;
    int cblocks = a.colpart_blocks;
    if (!cpart) {
        float* cp = p.partial + (size_t)splits * nW;
        DC_LAUNCH(ctx, colsum_partial_kernel, cs_blocks, 256, 0, a.gout, M, Cout_p, cp);
        DC_CHECK_LAUNCH();
        cpart = cp;
        cblocks = cs_blocks;
    }
    if (a.defer) {
        DC_REQUIRE(a.defer->count + 2 <= RED_MAX, DINCAE_ESIZE, "k2_wgrad_tma: too many deferred reductions");
        a.defer->e[a.defer->count++] = RedEntry{p.partial, a.dW, nW, splits};
        a.defer->e[a.defer->count++] = RedEntry{cpart, a.db, (int64_t)Cout_p, cblocks};
    } else {
        DC_LAUNCH(ctx, reduce_splits_kernel, (unsigned)dc_cdiv(nW, 32), 256, 0, p.partial, splits, nW, a.dW);
        DC_CHECK_LAUNCH();
        DC_LAUNCH(ctx, reduce_splits_kernel, (unsigned)dc_cdiv(Cout_p, 32), 256, 0, cpart, cblocks, (int64_t)Cout_p, a.db);
        DC_CHECK_LAUNCH();
    }
    return DINCAE_OK;
}

// ---- test hook (not part of the public header): run one weight gradient through either kernel family ----------
extern "C" int32_t dincae_debug_wgrad(dincae_ctx* ctx, const float* d_in, const float* d_g, int32_t B, int32_t H, int32_t W,
                                      int32_t Cin_p, int32_t Cout_p, float* d_dW, float* d_db, int32_t use_tma) {
    DC_REQUIRE(ctx && d_in && d_g && d_dW && d_db, DINCAE_EINVAL, "dincae_debug_wgrad: NULL argument");
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    WgradArgs a{};
    a.in = ConvSrc{d_in, nullptr, SRC_DIRECT, H, W, Cin_p};
    a.B = B; a.H = H; a.W = W; a.gout = d_g; a.Cout_p = Cout_p; a.dW = d_dW; a.db = d_db; a.tf32 = use_tma;
    size_t need = k2_wgrad_scratch_bytes(ctx->num_sms, B, H, W, Cin_p, Cout_p);
    size_t need2 = k2_wgrad_tma_scratch_bytes(ctx->num_sms, B, H, W, Cin_p, Cout_p);
    if (need2 > need) need = need2;
    void* sc = nullptr;
    DC_TRY(dc_scratch(ctx, need, &sc));
    return k2_wgrad(ctx, a, sc, ctx->scratch_bytes);
}

// ---- test hook (not part of the public header): one 3x3 convolution through a chosen kernel family ----------------
//   variant 0: FP32 CUDA-core kernel   4: persistent warp-specialised tcgen05 kernel (plain or paired tiles, as k2_conv picks)
// d_w: [round_up(9*Cin_p,16)][Cout_p] forward layout; d_wK: K-major [Cout_p][9*Cin_p] copy (TF32-representable values).
extern "C" int32_t dincae_debug_conv(dincae_ctx* ctx, const float* d_in, const float* d_w, const float* d_wK, const float* d_bias,
                                     int32_t B, int32_t H, int32_t W, int32_t Cin_p, int32_t Cout_p, int32_t relu, float* d_out, int32_t variant) {
    DC_REQUIRE(ctx && d_in && d_w && d_wK && d_out, DINCAE_EINVAL, "dincae_debug_conv: NULL argument");
    DC_REQUIRE(variant == 0 || variant == 4, DINCAE_EUNSUPPORTED, "dincae_debug_conv: variant %d (0 = FP32 CUDA cores, 4 = persistent tcgen05)", variant);
    DC_CHECK_CUDA(cudaSetDevice(ctx->device));
    ConvArgs a{};
    a.in = ConvSrc{d_in, nullptr, SRC_DIRECT, H, W, Cin_p};
    a.B = B; a.H = H; a.W = W; a.wgt = d_w; a.wgtK = variant == 0 ? nullptr : d_wK; a.ldk = 9 * Cin_p; a.in_is_tf32 = variant != 0;
    a.bias = d_bias; a.Cout_p = Cout_p; a.act = relu ? ACT_RELU : ACT_NONE; a.slope = 0.f; a.dst = d_out;
    if (variant == 0) return k2_conv(ctx, a);
    DC_REQUIRE(k2_conv_tma_eligible(a), DINCAE_EUNSUPPORTED, "dincae_debug_conv: shape not eligible for the tcgen05 kernel");
    return k2_conv_persist(ctx, a);
}

extern "C" int32_t dincae_debug_conv_trace(long long* d_trace) { g_persist_trace = d_trace; return DINCAE_OK; }

int32_t k2_reduce_batched(dincae_ctx* ctx, RedList& list) {
    if (list.count == 0) return DINCAE_OK;
    int blocks = 0;
    for (int e = 0; e < list.count; e++) {
        list.blk_start[e] = blocks;
        DC_REQUIRE(list.e[e].n % 4 == 0 && ((uintptr_t)list.e[e].part & 15) == 0 && ((uintptr_t)list.e[e].dst & 15) == 0, DINCAE_EINVAL,
                   "k2_reduce_batched: entry %d is not float4-aligned (n=%lld)", e, (long long)list.e[e].n);
        blocks += (int)dc_cdiv(list.e[e].n / 4, list.e[e].n <= 512 ? 8 : 32);
    }
    list.blk_start[list.count] = blocks;
    double rbytes = 0.0;
    for (int e = 0; e < list.count; e++) rbytes += 4.0 * list.e[e].n * (list.e[e].splits + 1);      // every partial row read once, the sum written
    DC_WORK(ctx, 0.0, rbytes);
    DC_LAUNCH(ctx, reduce_batched_kernel, blocks, 256, 0, list);
    DC_CHECK_LAUNCH();
    list.count = 0;
    return DINCAE_OK;
}
