// K1: device-resident gridded data cube and masked-input batch assembly
// (NCData ctor src/data.jl:250-369, getxy! :391-517, DataBatches.iterate :831-868).
#pragma once
#include "common.cuh"
#include <vector>

struct dincae_gridded {
    dincae_ctx* ctx = nullptr;
    int W = 0, H = 0, V = 0, T = 0, nout = 0;
    int64_t N = 0;
    float* d_data = nullptr;        // [N][V][H][W]: data - mean, NaN = missing (ONE float per value)
    float* d_err2 = nullptr;        // optional [N][V][H][W]: per-pixel observation error VARIANCE (errvarname, src/data.jl:66-67, :323-326)
    float* d_mean = nullptr;        // [V][H][W]
    float* d_mean_out = nullptr;    // [nout][H][W]: meandata[:,:,findall(isoutput)]
    float* d_lon = nullptr;         // [W] scaled to [-1,1]
    float* d_lat = nullptr;         // [H]
    float* d_tcos = nullptr;        // [N]
    float* d_tsin = nullptr;        // [N]
    uint8_t* d_landmask = nullptr;  // [H][W] or NULL
    float* d_params = nullptr;      // [V] obs_err_std^2, [V] jitter_std
    int* d_outvars = nullptr;       // [nout]
    int64_t* d_idx = nullptr;       // staging: time_idx[maxB], imask[maxB]
    int idx_cap = 0;
    std::vector<int> outvars;
    std::vector<std::vector<int64_t>> valid_counts;   // [nout][N]
};

// CUDA-graph support: the (single) assembly node of a captured step and the per-step update of its sample indices
int32_t k1_check_indices(const dincae_gridded* ds, const int64_t* h_time_idx, const int64_t* h_imask, int B);
int k1_assemble_max_single_launch();
cudaGraphNode_t k1_find_assemble_node(cudaGraph_t graph);
int32_t k1_update_assemble_node(cudaGraphExec_t exec, cudaGraphNode_t node, dincae_gridded* ds, const int64_t* h_time_idx, const int64_t* h_imask,
                                uint64_t seed, uint64_t ctr0, int B, float* d_xin, float* d_xtrue);
// device-pointer core used by dincae_gridded_assemble and dincae_reconstruct_range
int32_t k1_assemble(dincae_gridded* ds, const int64_t* h_time_idx, const int64_t* h_imask, uint64_t seed, uint64_t ctr0,
                    int B, float* d_xin, float* d_xtrue);
