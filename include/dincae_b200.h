/* libdincae_b200 -- C ABI of the B200-native DINCAE training + reconstruction hot path.
 *
 * Every entry point replaces one piece of the reference's (gher-uliege/DINCAE.jl v2.0.6)
 * Atype-backend surface; the replaced interface is cited as file:line relative to the
 * reference tree.  Conventions:
 *   - all functions return int32_t status (0 = DINCAE_OK); no C++ exception crosses the ABI;
 *     dincae_last_error() returns the message of the last failure on the calling thread.
 *   - external tensors keep the reference's memory order, Julia column-major WHCN
 *     (xin[i,j,c,b], i = lon fastest: src/data.jl:411-416), Float32; positions are
 *     interleaved (x,y) Float32 1-based fractional indices (src/points.jl:108-121).
 *   - `loc` says where caller buffers live.  DINCAE_HOST pointers are borrowed and copied on the context's
 *     stream: a call that returns a host scalar or host array is complete on return; a call that returns
 *     nothing to the host (dincae_grad_step, dincae_apply_step with loss = NULL, dincae_batch_upload, ...)
 *     may leave the copy in flight, so the buffer must stay untouched until dincae_sync (or the matching
 *     *_slot call) returns.  DINCAE_DEVICE pointers must be valid on the context's device and the call is
 *     asynchronous on the context's stream unless a host scalar is returned.
 *   - no CPU fallback exists anywhere behind this header.
 */
#ifndef DINCAE_B200_H
#define DINCAE_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum { DINCAE_OK = 0, DINCAE_EINVAL = 1, DINCAE_EUNSUPPORTED = 2, DINCAE_ECUDA = 3, DINCAE_ENOMEM = 4, DINCAE_ESIZE = 5 };
enum { DINCAE_HOST = 0, DINCAE_DEVICE = 1,
       DINCAE_DEVICE_XIN = 2 /* *_points entry points only: xin on the device (e.g. written by dincae_points_assemble), the point truth on the host */ };
enum { DINCAE_UP_NEAREST = 0, DINCAE_UP_BILINEAR = 1 };
/* conv arithmetic; accumulation is always FP32.
 *   FP32   : CUDA-core kernels, exact FP32 operands.
 *   TF32   : tcgen05 tensor cores (default).  Operands rounded to TF32, except the first conv and the 2*noutput-channel head conv
 *            of every chain (the thin full-resolution layers whose rounding dominates the output error), which run as 3xTF32
 *            (hi/lo operand split, FP32-grade products): (m, sigma) stay inside 1e-3 of the FP32 reference.
 *   TF32X1 : single-pass TF32 on every conv (what cuDNN's default math mode does); slightly faster, error ~2.5x larger. */
enum { DINCAE_PREC_FP32 = 0, DINCAE_PREC_TF32 = 1, DINCAE_PREC_TF32X1 = 2 };

typedef struct dincae_ctx dincae_ctx;
typedef struct dincae_model dincae_model;
typedef struct dincae_gridded dincae_gridded;
typedef struct dincae_points dincae_points;

#define DINCAE_MAX_LEVELS 8
#define DINCAE_MAX_CHAINS 4

/* Flattened keyword surface of DINCAE.reconstruct / genmodel (src/model.jl:286-299, :446-472). */
typedef struct dincae_config {
    int32_t W, H;                    /* grid size (lon, lat) = sz[1:2] */
    int32_t ninput;                  /* sz[end-1] of the first mini-batch (src/model.jl:546) */
    int32_t noutput;                 /* number of output variables */
    int32_t nlevels;                 /* length(enc_nfilter_internal) */
    int32_t nfilter[DINCAE_MAX_LEVELS]; /* enc_nfilter_internal */
    uint32_t skip_mask;              /* bit (l-1) set <=> level l in skipconnections (1-based, src/model.jl:277) */
    int32_t upsampling;              /* upsampling_method: DINCAE_UP_NEAREST | DINCAE_UP_BILINEAR */
    int32_t nchains;                 /* length(loss_weights_refine): 1, or 2 with the refinement stage (src/model.jl:322-340);
                                        3 and 4 repeat the refinement wiring once more each (an extension, not in the reference) */
    float loss_weights[DINCAE_MAX_CHAINS]; /* loss_weights_refine */
    float gamma;                     /* log(min_std_err^-2), src/model.jl:318 */
    int32_t truth_uncertain;
    float l1_beta, l2_beta;          /* regularization_L1_beta / _L2_beta (all trainables incl. biases, src/flux.jl:42) */
    float laplacian_penalty, laplacian_error_penalty;
    float clip_grad;                 /* Optimisers.ClipGrad: element-wise clamp; <= 0 disables (src/flux.jl:10-14) */
    float adam_beta1, adam_beta2, adam_eps;
    float leaky_slope;               /* 0 = ReLU (the reference, src/model.jl:251-256) */
    int32_t precision;               /* DINCAE_PREC_* */
    int32_t max_batch;               /* largest mini-batch that will be passed */
} dincae_config;

/* ---- context (replaces _to_device, src/flux.jl:47-49, ext/CUDAExt.jl:83-85) -------------------- */
int32_t dincae_ctx_create(int32_t device, dincae_ctx** out);
int32_t dincae_ctx_destroy(dincae_ctx* ctx);
int32_t dincae_ctx_set_stream(dincae_ctx* ctx, void* cuda_stream);   /* use an external cudaStream_t */
int32_t dincae_sync(dincae_ctx* ctx);                                /* CUDA.@sync, ext/CUDAExt.jl:32 */
const char* dincae_last_error(void);
int32_t dincae_version(void);
/* number of kernels this library has launched on ctx since creation (bench.py's gpu_launches) */
int64_t dincae_launch_count(dincae_ctx* ctx);
/* measurement aid (no counterpart in the reference, which only wraps loops in @time, src/model.jl:588,625):
 * between _begin and _end every kernel launch is bracketed by CUDA events on the context's stream; _end writes a
 * JSON array [{kernel, launches, ms, flops, bytes}] (flops/bytes = algorithmic work declared by the launcher). */
int32_t dincae_profile_begin(dincae_ctx* ctx);
int32_t dincae_profile_end(dincae_ctx* ctx, char* json, int64_t cap);

/* ---- K4: point gather / adjoint scatter (src/points.jl:6-77; ext/CUDAExt.jl:8-81) ---------------
 * interp:  out[p] = sum over the 4 corners of A[floor(pos)+off] * w      (interpnd!, overwrites out)
 * adjoint: out (W,H) zeroed, then out[floor(pos)+off] += values[p]*w     (interp_adjn!)
 *          deterministic: points are sorted by cell and reduced by segment, no float atomics.
 * Precondition as the reference: 1 <= pos < (W,H) (src/points.jl:45-50); violated => DINCAE_EINVAL
 * is NOT raised (device data); out-of-domain points are skipped. */
int32_t dincae_interp(dincae_ctx* ctx, const float* pos, int64_t P, const float* A, int32_t W, int32_t H,
                      float* out, int32_t loc);
int32_t dincae_interp_adjoint(dincae_ctx* ctx, const float* pos, const float* values, int64_t P,
                              int32_t W, int32_t H, float* out, int32_t loc);
/* dincae_interp_adjoint without any device allocation (no counterpart in the reference: its interp_adjn! gets
 * its temporaries from Julia's own GPU pool).  For hosts whose device memory is held by another allocator --
 * e.g. CUDA.jl's caching pool around a ccall, where a foreign cudaMalloc can fail or make the pool trim in the
 * middle of a training step -- or that give a plug-in no memory budget at all.
 *   - pos, values and out are DEVICE pointers the caller owns; out (W,H) is cleared, then accumulated into.
 *   - the sort scratch is a fixed array inside the library image (about 4.1 MiB on every GPU that loads the
 *     library, held until exit); at most 65536 points per call, more returns DINCAE_ESIZE.
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress); make that
 *     first call outside of, and before, any stream capture.
 *   - all calls on a device share the one array: calls that may overlap on the GPU (different streams or
 *     contexts) race on it and must be serialised by the caller.
 *   - asynchronous on the context's stream, same bit-exact result as dincae_interp_adjoint. */
int32_t dincae_interp_adjoint_noalloc(dincae_ctx* ctx, const float* pos, const float* values, int64_t P,
                                      int32_t W, int32_t H, float* out);
/* integer part only: cell[p] = (floor(x), floor(y)) 1-based int32 pairs (bit-exact check) */
int32_t dincae_interp_cells(dincae_ctx* ctx, const float* pos, int64_t P, int32_t* cells, int32_t loc);

/* ---- model (genmodel src/model.jl:286-363, StepModel :151-189, train_init src/flux.jl:7-21) ------ */
int32_t dincae_model_create(dincae_ctx* ctx, const dincae_config* cfg, dincae_model** out);
int32_t dincae_model_destroy(dincae_model* m);
int32_t dincae_model_num_params(dincae_model* m, int64_t* n);     /* sum(length, Flux.trainables), src/flux.jl:19 */
/* Flux.trainables order, each W as (kw,kh,cin,cout) column-major (Flux.state, src/model.jl:603) */
int32_t dincae_model_get_params(dincae_model* m, float* host, int64_t n);
int32_t dincae_model_set_params(dincae_model* m, const float* host, int64_t n);   /* Flux.loadmodel!, src/model.jl:395 */
int32_t dincae_model_init_glorot(dincae_model* m, uint64_t seed);   /* Flux glorot_uniform weights, zero biases */
/* Adam state (m, v in Flux order; t = steps taken) -- the reference does not save it (src/model.jl:599-620) */
int32_t dincae_model_get_opt_state(dincae_model* m, float* host_m, float* host_v, int64_t n, int64_t* t);
int32_t dincae_model_set_opt_state(dincae_model* m, const float* host_m, const float* host_v, int64_t n, int64_t t);

/* model(xin): (W,H,ninput,B) -> (W,H,2*noutput,B) = (mean, sigma^2) of the LAST chain (src/model.jl:182-189) */
int32_t dincae_forward(dincae_model* m, const float* xin, int32_t B, float* out, int32_t loc);
/* loss_function(model, xin, xtrue) (src/model.jl:192-222); xtrue (W,H,2*noutput,B) */
int32_t dincae_loss(dincae_model* m, const float* xin, const float* xtrue, int32_t B, int32_t loc, double* loss);
/* one mini-batch of train_epoch! (src/flux.jl:28-36): withgradient + ClipGrad + Adam; lr per call (src/flux.jl:25) */
int32_t dincae_train_step(dincae_model* m, const float* xin, const float* xtrue, int32_t B, float lr,
                          int32_t loc, float* loss);
/* The gridded gradient step (dincae_train_step / _slot / _dp / dincae_grad_step) replays its ~80 kernel launches as one CUDA graph
 * once the same (device input, device truth, B) has been seen twice: first step launch by launch, second step captured, then replay
 * (up to four keys are kept, least recently used replaced; DINCAE_GRAPH=0 in the environment disables it for the process).  Same
 * kernels, same order, same results; no counterpart in the reference (Flux launches kernel by kernel, src/flux.jl:28-36).
 * enable = 0 / 1 switches it for this model (0 also drops the captured graphs), -1 only queries; captured (may be NULL) <- number
 * of instantiated graphs. */
int32_t dincae_model_step_graph(dincae_model* m, int32_t enable, int32_t* captured);
/* one-batch-ahead prefetch (PrefetchDataIter, src/data.jl:872-904): upload mini-batch i+1 into the other slot (0/1) on a copy
 * stream while mini-batch i trains; host buffers (pinned for a truly asynchronous copy) must stay valid until the matching
 * dincae_train_step_slot returns. */
int32_t dincae_batch_upload(dincae_model* m, int32_t slot, const float* xin_host, const float* xtrue_host, int32_t B);
int32_t dincae_train_step_slot(dincae_model* m, int32_t slot, int32_t B, float lr, float* loss);
/* data-parallel form of the same: gradient sums of the uploaded shard, normalised by the GLOBAL valid counts (see dincae_grad_step);
 * the caller all-reduces dincae_grad_buffer and calls dincae_apply_step */
int32_t dincae_grad_step_slot(dincae_model* m, int32_t slot, int32_t B, const int64_t* n_valid_global);
/* point-cloud truth (src/points.jl:354-426): pos (sumP,2), x (sumP,2) = (value, 1/err var), offsets (B+1) int64 */
int32_t dincae_loss_points(dincae_model* m, const float* xin, const float* pos, const float* x,
                           const int64_t* offsets, int32_t B, int32_t loc, double* loss);
int32_t dincae_train_step_points(dincae_model* m, const float* xin, const float* pos, const float* x,
                                 const int64_t* offsets, int32_t B, float lr, int32_t loc, float* loss);

/* vector2_costfun(xrec, xtrue, truth_uncertain, directionobs) (src/vector2.jl:53-100, the cost behind output_ndims = 2): a 2-D vector
 * field observed as radial components at nsites sites.  xrec (W,H,5,B) = (u', v', L11, L21, L22) is the raw model output (the
 * vector model has no final transform, src/model.jl:352-362), xtrue (W,H,2*nsites,B) = (m / sigma2, 1 / sigma2) per site,
 * directionobs (W,H,nsites) or, with dir_has_batch != 0, (W,H,nsites,B): bearings in degrees.  loss <- the cost (host double);
 * grad (may be NULL; same shape and place as xrec) <- d cost / d xrec.  A NaN cost is an error, as in the reference (:95-97).
 * loc = DINCAE_HOST or DINCAE_DEVICE for all four arrays. */
int32_t dincae_vector2_loss(dincae_ctx* ctx, const float* xrec, const float* xtrue, const float* directionobs, int32_t W, int32_t H,
                            int32_t nsites, int32_t B, int32_t dir_has_batch, int32_t truth_uncertain, int32_t loc, double* loss,
                            float* grad);

/* Data-parallel split of train_step (new capability; the reference is single-device):
 *   grad_step   : forward + backward of the local shard; gradients are normalised by n_valid_global
 *                 (per output variable; NULL = count this shard's own) so a SUM over ranks gives the
 *                 full mini-batch gradient; fills the flat gradient buffer.
 *   grad_buffer : device pointer + length of that buffer; element [n-1] carries the local loss term, so one
 *                 all-reduce moves gradients and loss together.
 *   apply_step  : L1/L2 + ClipGrad + Adam on the (all-reduced) buffer; returns the mini-batch loss. */
int32_t dincae_grad_step(dincae_model* m, const float* xin, const float* xtrue, int32_t B,
                         const int64_t* n_valid_global, int32_t loc);
int32_t dincae_grad_step_points(dincae_model* m, const float* xin, const float* pos, const float* x,
                                const int64_t* offsets, int32_t B, int64_t n_valid_global, int32_t loc);
int32_t dincae_grad_buffer(dincae_model* m, float** dev_ptr, int64_t* n);
int32_t dincae_apply_step(dincae_model* m, float lr, float* loss);
/* ---- data-parallel gradient exchange (new capability: the reference trains on one device, src/flux.jl:23-40) -------------
 * One process per GPU, one context per process.  The library owns an NCCL communicator per context (NCCL is bound with
 * dlopen at the first call: an NCCL already loaded by the host process is shared, else libnccl.so.2, else $DINCAE_NCCL_LIB).
 *   comm_unique_id : rank 0 creates the 128-byte rendezvous id (ncclGetUniqueId) and hands it to the other ranks by the
 *                    host's own means (MPI.bcast, a file, torch.distributed ...).
 *   comm_init_rank : collective over all ranks; binds the communicator to ctx's device.
 *   allreduce      : in-place SUM of a device buffer over all ranks, asynchronous on the context's stream.
 *   allreduce_grads: the same for the model's flat gradient buffer + loss slot (between grad_step* and apply_step).
 *   train_step_dp  : grad_step + all-reduce + apply_step in one call.  n_valid_global as in grad_step (must not be NULL here).
 *                    From 4 ranks on the all-reduce is the library's own kernel over NVLink peer memory (the ranks'
 *                    workspaces are mapped into each other with CUDA IPC at the first step: one process per GPU; reduce-
 *                    scatter and all-gather as posted stores, two flag rounds; DINCAE_P2P=0/1 forces NCCL / this kernel);
 *                    it falls back to ncclAllReduce when the mapping is not possible.  Every element is summed by one rank
 *                    in rank order, so the replicas stay bit-identical.  While the mapping exists dincae_model_destroy is
 *                    collective (every rank must call it).  DINCAE_DP_OVERLAP=1 selects per-chain NCCL buckets behind the
 *                    backward pass instead (measured slower: the persistent kernels own every SM).
 *   train_step_dp_slot: the same on a mini-batch shard uploaded with dincae_batch_upload. */
#define DINCAE_COMM_ID_BYTES 128
int32_t dincae_comm_unique_id(void* id128);
int32_t dincae_comm_init_rank(dincae_ctx* ctx, int32_t nranks, int32_t rank, const void* id128);
int32_t dincae_comm_destroy(dincae_ctx* ctx);
int32_t dincae_comm_info(dincae_ctx* ctx, int32_t* nranks, int32_t* rank, int32_t* nccl_version);
int32_t dincae_allreduce(dincae_ctx* ctx, float* dev_buf, int64_t n);
int32_t dincae_allreduce_grads(dincae_model* m);
int32_t dincae_train_step_dp(dincae_model* m, const float* xin, const float* xtrue, int32_t B, const int64_t* n_valid_global,
                             float lr, int32_t loc, float* loss);
int32_t dincae_train_step_dp_slot(dincae_model* m, int32_t slot, int32_t B, const int64_t* n_valid_global, float lr, float* loss);
/* max over all parameters of |w_rank - w_rank0| (replica drift check; 0 when the replicas are bit-identical); collective */
int32_t dincae_replica_divergence(dincae_model* m, float* max_abs_diff);

/* number of valid (xtrue[:,:,2v,:] != 0) targets per output variable in a batch (n_noncloud, src/model.jl:61) */
int32_t dincae_count_valid(dincae_model* m, const float* xtrue, int32_t B, int32_t loc, int64_t* n_valid);

/* ---- K1: gridded dataset + batch assembly (NCData src/data.jl:250-369, getxy! :391-517,
 *      DataBatches.iterate :831-868) ------------------------------------------------------------
 * data_full (W,H,V,N) Float32 with NaN = missing; obs_err_std, jitter_std, isoutput per variable;
 * mean_data (W,H,V) or NULL (computed like src/data.jl:266-270 when remove_mean != 0);
 * landmask (W,H) uint8, 1 = valid, or NULL; lon_scaled (W), lat_scaled (H), time_cos/sin (N) are host
 * precomputed.  The cube lives on the device as ONE Float32 per value. */
int32_t dincae_gridded_create(dincae_ctx* ctx, const float* data_full, int32_t W, int32_t H, int32_t V, int64_t N,
                              const float* obs_err_std, const float* jitter_std, const uint8_t* isoutput,
                              int32_t remove_mean, const float* mean_data, const uint8_t* landmask,
                              const float* lon_scaled, const float* lat_scaled,
                              const float* time_cos, const float* time_sin, int32_t ntime_win,
                              int32_t loc, dincae_gridded** out);
int32_t dincae_gridded_destroy(dincae_gridded* ds);
/* per-pixel observation error field (errvarname: src/data.jl:34,66-67 and the 4-D obs_err_std branch :323-326):
 * obs_err_std (W,H,V,N) replaces the per-variable value, +Inf where the error field is missing (both channels become 0). */
int32_t dincae_gridded_set_error(dincae_gridded* ds, const float* obs_err_std, int32_t loc);
int32_t dincae_gridded_get_mean(dincae_gridded* ds, float* host_mean);   /* meandata (W,H,V) */
/* time_idx[B] 0-based sample indices (perm[j]); imask[B] 0-based time index whose missing mask is
 * transplanted onto the centre step, NULL = no perturbation (train=false); jitter noise is
 * jitter_std[v] * Philox4x32-10 normal keyed by (seed, sample_ctr0 + b). Outputs are device or host per loc. */
int32_t dincae_gridded_assemble(dincae_gridded* ds, const int64_t* time_idx, const int64_t* imask,
                                uint64_t seed, uint64_t sample_ctr0, int32_t B,
                                float* xin, float* xtrue, int32_t loc);
/* One mini-batch of train_epoch! over DataBatches straight from the device-resident dataset (iterate, src/data.jl:831-868, then
 * src/flux.jl:28-36): getxy! of the samples time_idx / imask (arguments as in dincae_gridded_assemble) into the model's own batch
 * buffers, then dincae_train_step on them -- one captured CUDA graph whose sample indices, learning rate and Adam step are patched per
 * call, so consecutive steps run back to back on the GPU.  Same results as dincae_gridded_assemble + dincae_train_step.
 * The data-parallel form takes this rank's shard and the GLOBAL valid counts (dincae_train_step_dp). */
int32_t dincae_train_step_gridded(dincae_model* m, dincae_gridded* ds, const int64_t* time_idx, const int64_t* imask, uint64_t seed,
                                  uint64_t sample_ctr0, int32_t B, float lr, float* loss);
int32_t dincae_train_step_dp_gridded(dincae_model* m, dincae_gridded* ds, const int64_t* time_idx, const int64_t* imask, uint64_t seed,
                                     uint64_t sample_ctr0, int32_t B, const int64_t* n_valid_global, float lr, float* loss);
/* valid-target count of every time step for output variable v (static; lets data-parallel ranks
 * know n_noncloud of the global batch without a collective) */
int32_t dincae_gridded_valid_counts(dincae_gridded* ds, int32_t v, int64_t* host_counts);

/* full-record reconstruction of steps [t0,t1) (save loop src/model.jl:624-644 + savesample arithmetic
 * src/data.jl:615-619,:653-675): assemble (no perturbation) -> model -> rec = m + mean, sigma = sqrt(sigma2),
 * NaN outside the land mask, running mean ((count-1)*old + new)/count with NaN -> fill_value.
 * m_out / sigma_out: (W,H,t1-t0) per output variable v-major, i.e. (W,H,t1-t0,noutput). */
int32_t dincae_reconstruct_range(dincae_model* m, dincae_gridded* ds, int64_t t0, int64_t t1,
                                 int32_t count, float fill_value, float* m_out, float* sigma_out, int32_t loc);

/* ---- points dataset + gridding (PointCloud src/points.jl:137-227, getxy! :237-313) --------------------
 * pos (sumP,2) interleaved and x (sumP,2) column-major (x[0:sumP] = value, x[sumP:2sumP] = 1/err var) are the
 * ragged per-time-step vectors AFTER posindices! filtering, concatenated; offsets (N+1) host int64;
 * aux (W,H,naux,N) or NULL.  All pointers are HOST pointers (one-time set-up). */
int32_t dincae_points_create(dincae_ctx* ctx, int32_t W, int32_t H, int64_t N, const float* pos, const float* x,
                             const int64_t* offsets, const float* aux, int32_t naux,
                             const float* lon_scaled, const float* lat_scaled,
                             const float* doy_cos, const float* doy_sin, int32_t ntime_win,
                             float jitter_std_x, float jitter_std_y, dincae_points** out);
int32_t dincae_points_destroy(dincae_points* ds);
/* lon/lat -> fractional index + inside flag (posindices!, src/points.jl:98-134), bit-exact integer work */
/* grid given by its first/last coordinate and length (only those enter the formula); evaluated in Float64 as the
 * reference does for Float64 grid ranges, pos stored as Float32 */
int32_t dincae_posindices(dincae_ctx* ctx, const float* lon, const float* lat, int64_t P,
                          double gridx_first, double gridx_last, int32_t W,
                          double gridy_first, double gridy_last, int32_t H,
                          float* pos, uint8_t* inside, int32_t loc);
/* time_idx[B]; perturb != 0 => position jitter (pertpos :229-235, Philox); keep: per sample a uint8 flag for every
 * point of its centre step (track dropout drawn by the caller, :284-300), concatenated, keep_offsets (B+1); NULL = all.
 * xin (W,H,4+naux+2*ntime_win,B). */
int32_t dincae_points_assemble(dincae_points* ds, const int64_t* time_idx, int32_t B, int32_t perturb,
                               uint64_t seed, uint64_t sample_ctr0, const uint8_t* keep, const int64_t* keep_offsets,
                               float* xin, int32_t loc);

#ifdef __cplusplus
}
#endif
#endif
