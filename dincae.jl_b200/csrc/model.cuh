// Model object of libdincae_b200: topology (genmodel/recmodel4, src/model.jl:242-363), flat parameter /
// gradient / Adam buffers and the per-level activation workspace.
#pragma once
#include "common.cuh"
#include "k2_conv.cuh"
#include "k3_loss.cuh"
#include <vector>

struct LayerInfo {
    int Cin, Cout, Cin_p, Cout_p;
    int Kpad;            // round_up(9*Cin_p,16): rows of the forward weight matrix
    int KpadT;           // round_up(9*Cout_p,16): rows of the transposed matrix used by the data gradient
    int64_t w_off, b_off;    // offsets in the internal flat parameter buffer
    int64_t wT_off;          // offset in the transposed-weight buffer
    int64_t wk_off;          // offset in the K-major TF32 forward copy wK (tensor-core path; expanded layout for 3xTF32 layers)
    int64_t wtk_off;         // offset in the K-major TF32 data-gradient copy wTK
    int Cin_pe;              // channels of the conv's INPUT TENSOR on the tensor-core path: Cin_p, or the expanded [hi|lo|hi] /
                             // [hi|lo] operand layout of a 3xTF32 layer (split_mode 1 = first conv of a chain, 2 = head conv)
    int split_mode;
    int64_t gexp_off;        // 3xTF32 layers: offset of the expanded weight gradient in g_exp
    int64_t flux_w_off, flux_b_off;   // offsets in the Flux.trainables order
};

struct ChainBufs {
    float* X[DINCAE_MAX_LEVELS + 1];   // X[l]: input of level l (X[0] = packed input), NHWC
    float* A[DINCAE_MAX_LEVELS];       // relu(conv_enc_l); NULL where the conv epilogue pools and keeps only the sign bits MK[l]
    uint32_t* MK[DINCAE_MAX_LEVELS];   // sign bits of relu(conv_enc_l), word [pixel][Cout_p / 32] (levels with the fused pool)
    float* R[DINCAE_MAX_LEVELS];       // relu(conv_dec_l), l >= 1
    float* U[DINCAE_MAX_LEVELS];       // bilinear only: materialised upsampled input of conv_dec_l
    float* RAW;                        // planar (W,H,2*nout,B): raw chain output
    float* GRAW;                       // planar gradient w.r.t. RAW
};

// One captured gradient step (forward + loss + backward of every chain, both streams): the ~80 launches of a step are replayed as
// one CUDA graph, which takes the per-kernel boundary from 2-4 us (stream launches, with or without programmatic dependent launch)
// to about 1.2 us (measured: scripts/microbench/launch_floor.cu).  Keyed on what the launches were recorded with.
struct StepGraph {
    const float* dx = nullptr;       // device input / truth the step reads
    const float* dy = nullptr;
    int B = 0;
    bool count_here = false;         // the step counts the valid targets itself (no global count handed in)
    bool apply = false;              // the optimiser update is part of the step (dincae_train_step / _slot)
    const void* ds = nullptr;        // the step starts with the batch assembly from this dataset (dincae_train_step_gridded)
    cudaGraphNode_t asm_node = nullptr;
    cudaGraph_t graph = nullptr;     // kept for the optimiser node, whose arguments are updated before every replay
    cudaGraphNode_t adam_node = nullptr;
    cudaGraphExec_t exec = nullptr;  // nullptr: the key has been seen once (that step ran launch by launch, which also did every lazy setup)
    int64_t launches = 0;            // kernels one replay launches
    uint64_t last_use = 0;
};
constexpr int DC_STEP_GRAPHS = 4;
// the forward pass of reconstruction (staging buffer in, (m, sigma2) buffer out) for one batch size
struct FwdGraph {
    int B = 0;                       // 0 = free
    cudaGraphExec_t exec = nullptr;  // nullptr: seen once
    int64_t launches = 0;
};

struct dincae_model {
    StepGraph step_graph[DC_STEP_GRAPHS];
    FwdGraph fwd_graph[2];           // full batches and the last partial one
    uint64_t graph_tick = 0;
    bool graph_off = false;      // capture failed once: launch by launch from then on
    bool side_pending = false;   // weight gradients in flight on the context's side stream
    // peer-memory all-reduce of the gradient buffer (p2p.cu): the other ranks' workspaces mapped through CUDA IPC
    int p2p_state = 0;                 // 0 = not tried, 1 = ready, -1 = unavailable (NCCL is used)
    void* p2p_peer_base[16] = {};      // peer rank -> its all_mem mapped here (own rank: all_mem)
    uint32_t* p2p_flags = nullptr;     // this rank's flag words (inside all_mem, zero at creation)
    float* p2p_stage = nullptr;        // this rank's staging area: the peers push their parts of this rank's slice here
    uint32_t p2p_epoch = 0;
    bool dp_overlap = false;     // dincae_train_step_dp: all-reduce every chain's gradient bucket as soon as its backward pass ends
    int64_t chain_off[DINCAE_MAX_CHAINS + 1];     // parameter range of chain c in the flat buffers: [chain_off[c], chain_off[c+1])
    SplitTable split_chain[DINCAE_MAX_CHAINS];    // the split-operand layers of chain c

    dincae_ctx* ctx = nullptr;
    dincae_config cfg;
    int L = 0, nchains = 0, O = 0, nco = 4, maxB = 0;
    int Hl[DINCAE_MAX_LEVELS + 1], Wl[DINCAE_MAX_LEVELS + 1];
    std::vector<std::vector<LayerInfo>> enc, dec;     // [chain][level]
    int64_t n_flux = 0, n_int = 0, n_wT = 0, n_wT_copy = 0;
    float *w = nullptr, *g = nullptr, *m = nullptr, *v = nullptr, *wT = nullptr;
    float *wK = nullptr, *wTK = nullptr;     // K-major TF32-rounded copies (NULL in FP32 mode)
    int64_t n_wk = 0, n_wtk = 0;
    SplitTable split;                        // the 3xTF32 layers (DINCAE_PREC_TF32): first conv and head conv of every chain
    float* g_exp = nullptr;                  // their weight gradients in the expanded layout, folded into g after the reductions
    int64_t n_gexp = 0;
    WLayerDesc* d_wtab = nullptr;
    int n_wlayers = 0;
    bool wT_dirty = true;
    ChainBufs cb[DINCAE_MAX_CHAINS];
    // gradient workspace shared by all chains
    float* GD0 = nullptr;
    float *GU[DINCAE_MAX_LEVELS], *G[DINCAE_MAX_LEVELS + 1], *GD[DINCAE_MAX_LEVELS], *GC[DINCAE_MAX_LEVELS], *XG[DINCAE_MAX_LEVELS + 1];
    float* MS = nullptr;               // planar (m, sigma2)
    float* GMS = nullptr;              // planar gradient w.r.t. (m, sigma2) (point loss)
    double* d_loss = nullptr;          // [0] cost terms, [1] regularisation
    float* d_inv_n = nullptr;          // [nout]
    unsigned long long* d_nvalid = nullptr;
    void* red_scratch = nullptr;
    size_t red_scratch_bytes = 0;
    void* wg_scratch = nullptr;        // arena: every weight gradient of a backward pass gets its own slice (deferred reductions)
    size_t wg_scratch_bytes = 0, wg_off = 0;
    RedList red;
    float *in_xin = nullptr, *in_xtrue = nullptr;     // device staging of host mini-batches
    // prefetch slots (dincae_batch_upload): the next mini-batch crosses PCIe on a copy stream while this one trains
    float* up_buf[2] = {nullptr, nullptr};
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t up_done[2] = {nullptr, nullptr}, slot_free[2] = {nullptr, nullptr};
    bool slot_used[2] = {false, false};
    // point-loss workspace (grown on demand)
    float* pt_buf = nullptr;
    size_t pt_cap = 0;                 // in points
    void* all_mem = nullptr;
    int64_t t_step = 0;
    double bt1 = 1.0, bt2 = 1.0;       // beta^t
};

int32_t model_forward(dincae_model* M, const float* d_xin, int B, bool train);
int32_t model_head_last(dincae_model* M, int B, float* d_out);
// SUM of M->g[0, n) over all ranks, in place, on the context's stream: two-shot all-reduce over NVLink peer memory when the ranks'
// workspaces could be mapped (one process per GPU, CUDA IPC), else ncclAllReduce
int32_t model_allreduce_grads(dincae_model* M);
void model_p2p_release(dincae_model* M);
int32_t model_p2p_check(dincae_model* M);    // (m, sigma2) of the last chain into d_out (planar)
