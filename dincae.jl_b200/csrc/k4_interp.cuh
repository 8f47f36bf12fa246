// K4: bilinear gather and deterministic (sorted, segmented) adjoint scatter-add.
// Replaces interpnd!/interp_adjn! (src/points.jl:6-77) and the CUDA.jl kernels
// interpnd_d!/interp_adjn_d! (ext/CUDAExt.jl:8-81; the latter: one 256-thread block + float atomics).
#pragma once
#include "common.cuh"

// Batched problem description: point p lies on plane plane_of[p] (NULL = plane 0).
// Field f of plane q starts at base + (q / inner)*outer_stride + (q % inner)*plane_stride + f*field_stride
// (inner <= 0 means a single level: base + q*plane_stride), each field a (W,H) column-major grid.
struct K4Batch {
    const float* pos;          // (P,2) interleaved 1-based (x,y)
    const int32_t* plane_of;   // (P) or NULL
    const uint8_t* keep;       // (P) or NULL: 0 drops the point (track dropout, src/points.jl:284-300)
    int64_t P;
    int32_t W, H;
    int32_t nplanes;
    int64_t plane_stride, field_stride;
    int32_t nfields;           // 1 or 2
    int32_t inner = 0;         // two-level plane addressing (see above)
    int64_t outer_stride = 0;
    int32_t zero_out = 1;      // 0: the caller has already zeroed the target grids
};

size_t k4_scatter_scratch_bytes(int64_t P);
// vals[f]: (P) values of field f.  out: grids as described by K4Batch, zeroed here first.
int32_t k4_scatter_sorted(dincae_ctx* ctx, const K4Batch& b, const float* vals0, const float* vals1, float* out,
                          void* scratch, size_t scratch_bytes);
// out[f*P + p] = bilinear gather of field f at point p
int32_t k4_gather(dincae_ctx* ctx, const K4Batch& b, const float* grids, float* out);
