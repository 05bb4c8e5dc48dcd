/*
 * posteriflow_b200.h -- C-ABI of the B200-native (sm_100a) posterior-flow hot path.
 *
 * This is the drop-in boundary for ONE path of bibinthomas123/PosteriFlow: the conditional
 * masked-autoregressive rational-quadratic Neural Spline Flow behind LeanNPE sampling and NLL.
 * The reference has no FFI layer of its own (it is pure Python/PyTorch over the third-party
 * `nflows` package); each entry point below names the reference interface it replaces.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes, no torch / C++ types in any signature.
 *   - every pointer is a DEVICE pointer on the device given at pf_flow_create, unless it is
 *     marked [host]; all tensors are dense row-major float32 unless stated.
 *   - the caller allocates every input, output and workspace buffer; the library owns only the
 *     opaque packed-weights handle.  No call synchronises the stream except pf_flow_create/destroy,
 *     pf_flow_check, and pf_flow_pack when mask buffers are handed in (validation) or a source
 *     pointer changed since the previous pack (its job table is re-uploaded).
 *   - all work is enqueued on the given cudaStream_t (passed as void*; 0 = legacy default).
 *   - return value: PF_OK (0) or a negative pf_status; never throws, never exits.
 *   - one in-flight call per (handle, stream); handles are independent of each other.
 *   - there is no CPU fallback: without a CUDA device every entry point returns PF_ERR_CUDA.
 *
 * Context addressing.  The flow is conditioned on a context row per flow row.  The reference
 * materialises a repeated [R, C] tensor (lean_npe.py:328).  Here rows address contexts through
 *     row2ctx != NULL : ctx row of flow row r is row2ctx[r]           (int64, device)
 *     row2ctx == NULL : ctx row of flow row r is r / n_per_ctx         (n_per_ctx >= 1)
 * so n_per_ctx = 1 is the reference's dense [R, C] form and n_per_ctx = N is "N posterior
 * samples per event" with the context projected once per event.
 */
#ifndef POSTERIFLOW_B200_H
#define POSTERIFLOW_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

#define PF_ABI_VERSION 1

typedef int32_t pf_status;
enum {
  PF_OK = 0,
  PF_ERR_INVALID = -1,     /* bad argument / unsupported shape                         */
  PF_ERR_CUDA = -2,        /* CUDA runtime error (see pf_last_cuda_error)              */
  PF_ERR_NOT_PACKED = -3,  /* compute call before pf_flow_pack                          */
  PF_ERR_WORKSPACE = -4,   /* workspace too small                                       */
  PF_ERR_MASK = -5,        /* mask buffers differ from nflows' deterministic MADE masks */
  PF_ERR_UNSUPPORTED = -6  /* valid request that this build has no kernel for           */
};

/* arithmetic tier of the conditioner GEMMs */
enum {
  PF_TIER_FP32 = 0, /* FP32 FMA everywhere: parity tier, |dtheta|,|dlogq| <= 1e-4 rel      */
  PF_TIER_BF16 = 1, /* bf16 operands on tcgen05 tensor cores, fp32 accumulate + spline     */
  /* Split-bf16 operands on tcgen05 (each factor as a bf16 hi + lo pair, three MMAs per product, fp32 accumulate) and the
   * exact fp32 spline: the tensor-core tier that meets the FP32 tolerance.  Serves the DENSITY direction (forward /
   * log-prob / NLL, and the fused training step); the sampling direction of such a handle runs on the FP32-FMA
   * kernels -- ten split activation arrays of a 128-row tile do not fit in one SM's shared + tensor memory. */
  PF_TIER_BF16X3 = 2
};

/* flags for pf_flow_inverse / pf_flow_forward / pf_sample_fused */
enum {
  PF_FLAG_SANITIZE_CTX = 1u << 0, /* nan->0, +inf->1e-3, -inf->-1e-3 on ctx (flows.py:626-632)      */
  PF_FLAG_CLAMP = 1u << 1,        /* nan_to_num(0,+1,-1) then clamp +-clamp_bound (flows.py:650-654) */
  PF_FLAG_WRAP = 1u << 2,         /* ParamScaler.wrap (lean_npe.py:100-104)                          */
  PF_FLAG_DENORMALIZE = 1u << 3,  /* ParamScaler.denormalize (lean_npe.py:111-114)                   */
  PF_FLAG_CTX_IS_PROJ = 1u << 4,  /* `ctx` argument is already pf_flow_ctxproj output                */
  PF_FLAG_SORT_MASSES = 1u << 5   /* m1 >= m2 on dims 0,1 of the output (pipeline.py:184-186)        */
};

/* workspace selectors for pf_flow_workspace_bytes */
enum { PF_OP_CTXPROJ = 0, PF_OP_INVERSE = 1, PF_OP_FORWARD = 2, PF_OP_BACKWARD = 3, PF_OP_SAMPLE = 4 };

typedef struct pf_flow pf_flow; /* opaque packed-weights handle */

/* Shape of one NSFPosteriorFlow.  Replaces the constructor arguments at
 * /root/reference/src/ahsd/models/flows.py:379-383 as used by lean_npe.py:285-295. */
typedef struct pf_flow_desc {
  int32_t features;         /* D: 11                                                 */
  int32_t context_features; /* C: 288 = 256 encoder + 32 rank embedding              */
  int32_t hidden_features;  /* H: 256                                                */
  int32_t num_layers;       /* L: 10  ([ReversePermutation, MAF-RQS] pairs)          */
  int32_t num_bins;         /* K: 16                                                 */
  int32_t num_blocks;       /* residual blocks per MADE; only 2 is supported         */
  float tail_bound;         /* 5.0  (spline domain, linear tails outside)            */
  float clamp_bound;        /* 3.0  FLOW_NORM_BOUND, parameter_scalers.py:27         */
  int32_t tier;             /* PF_TIER_*                                             */
  int32_t device;           /* CUDA device ordinal                                   */
} pf_flow_desc;

/* Raw weights of ONE MADE conditioner in nflows' own layout and state-dict naming
 * (`transform._transforms.{2l+1}.autoregressive_net.*`).  Masks may be
 * NULL (nflows' deterministic masks are then assumed); if given they are checked against them. */
typedef struct pf_layer_ptrs {
  const float* initial_w;       /* initial_layer.weight  [H, D]  */
  const float* initial_b;       /* initial_layer.bias    [H]     */
  const float* initial_mask;    /* initial_layer.mask    [H, D]  */
  const float* ctx_w;           /* context_layer.weight  [H, C]  */
  const float* ctx_b;           /* context_layer.bias    [H]     */
  const float* blk_ctx_w[2];    /* blocks.k.context_layer.weight [H, C] */
  const float* blk_ctx_b[2];    /* blocks.k.context_layer.bias   [H]    */
  const float* blk_lin_w[2][2]; /* blocks.k.linear_layers.j.weight [H, H] */
  const float* blk_lin_b[2][2]; /* blocks.k.linear_layers.j.bias   [H]    */
  const float* blk_lin_mask[2][2];
  const float* final_w;         /* final_layer.weight [D*(3K-1), H] */
  const float* final_b;         /* final_layer.bias   [D*(3K-1)]    */
  const float* final_mask;
} pf_layer_ptrs;

/* Gradient destinations, same layout as pf_layer_ptrs (accumulated INTO: caller zeroes). */
typedef struct pf_layer_grads {
  float* initial_w;
  float* initial_b;
  float* ctx_w;
  float* ctx_b;
  float* blk_ctx_w[2];
  float* blk_ctx_b[2];
  float* blk_lin_w[2][2];
  float* blk_lin_b[2][2];
  float* final_w;
  float* final_b;
} pf_layer_grads;

/* ParamScaler tensors (lean_npe.py:87-98), device pointers, D entries each. */
typedef struct pf_scaler {
  const float* lo;
  const float* hi;
  const uint8_t* log_mask;
  const uint8_t* circ_mask;
} pf_scaler;

/* ---- lifetime ------------------------------------------------------------------------------ */

/* Replaces NSFPosteriorFlow.__init__ (flows.py:379-548): allocates the packed-weight arena. */
pf_status pf_flow_create(const pf_flow_desc* desc, pf_flow** out);
void pf_flow_destroy(pf_flow* flow);

/* Replaces the per-call `self.weight * self.mask` of nflows' MaskedLinear.forward and the
 * implicit weight layout: masks are pre-multiplied, hidden units are sorted by autoregressive
 * degree (block-triangular masks), transposed copies are made for the backward GEMMs and the
 * sampler's per-degree weight blocks are laid out in streaming order.  Call again after every
 * optimizer step / load_state_dict: a re-pack from unchanged pointers is three kernel launches, does not
 * synchronise and may be captured in a CUDA graph.  `layers` is a [host] array of L structs. */
pf_status pf_flow_pack(pf_flow* flow, const pf_layer_ptrs* layers, void* stream);

/* Barrier-timeout flag of the handle's tensor-core kernels.  A tcgen05 kernel whose internal producer/consumer
 * protocol times out (a bug, or a wedged SM) raises a flag in pinned host memory, writes NaN into every output row
 * of the tiles it abandons, and exits; the call that launched it has long returned PF_OK by then.  The flag is
 * reported -- PF_ERR_CUDA, text in pf_last_cuda_error() -- by the NEXT compute call on the handle (without any
 * synchronisation) and by pf_flow_check, which synchronises `stream` first.  Reading clears it.
 * (No reference counterpart: PyTorch kernels cannot fail this way.) */
pf_status pf_flow_check(pf_flow* flow, void* stream);

/* ---- context projection -------------------------------------------------------------------- */

/* Replaces the three context Linear layers of every MADE (nflows MADE.context_layer and
 * MaskedResidualBlock.context_layer, built at flows.py:511-524), evaluated ONCE per distinct
 * context row instead of once per sample and per autoregressive pass:
 *   proj[e, l, 0, :] = relu(Wc0_l ctx_e + b),  proj[e, l, 1|2, :] = sigmoid(Wc{1|2}_l ctx_e + b)
 * proj is [E, L*3*H] float32, hidden units in the handle's packed (degree-sorted) order. */
pf_status pf_flow_ctxproj(pf_flow* flow, const float* ctx, int64_t E, uint32_t flags, float* proj, void* stream);

/* ---- sampling direction (z -> theta) -------------------------------------------------------- */

/* Replaces NSFPosteriorFlow.inverse (flows.py:620-655) = CompositeTransform.inverse over
 * nflows' 11-pass AutoregressiveTransform.inverse, computed incrementally (one conditioner
 * pass of work per layer).  x [R, D]; logdet [R] or NULL (= sum over layers of -logabsdet,
 * as the reference returns it).  flags: SANITIZE_CTX | CLAMP reproduce the reference method;
 * WRAP | DENORMALIZE additionally fuse ParamScaler (scaler must then be non-NULL). */
pf_status pf_flow_inverse(pf_flow* flow, const float* z, const float* ctx, const int64_t* row2ctx,
                          int64_t n_per_ctx, int64_t R, int64_t E, float* x, float* logdet,
                          uint32_t flags, const pf_scaler* scaler, void* workspace,
                          size_t workspace_bytes, void* stream);

/* Replaces LeanNPE.sample_posterior after the encoder (lean_npe.py:326-332): context broadcast,
 * base noise, inverse, wrap, denormalize -> theta [E, n, D] in physical units, event-major.
 * z == NULL draws the base noise in-kernel (Philox4x32-10 keyed by seed, counter = row index);
 * z != NULL ([E*n, D]) reproduces the reference exactly for a given noise tensor.
 * railed [E*n] (uint8) or NULL: pipeline.py:177 rail mask. */
pf_status pf_sample_fused(pf_flow* flow, const float* ctx, const pf_scaler* scaler, const float* z,
                          uint64_t seed, int64_t E, int64_t n, float* theta, float* logdet,
                          uint8_t* railed, uint32_t flags, void* workspace, size_t workspace_bytes,
                          void* stream);

/* pf_sample_fused for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO counterpart in the reference.
 * Same arguments and results as pf_sample_fused minus the workspace: the sampler's scratch (the per-event context
 * projection) lives in a file-scope __device__ array of 32 MiB inside the library (csrc/pf_static.cu), held by every
 * GPU that loads the library until the process exits.  The call itself allocates nothing (no cudaMalloc / cudaFree /
 * allocating library routine).  Limit: pf_flow_workspace_bytes-equivalent scratch <= 32 MiB, i.e. E <= 1075 events per
 * call at the LeanNPE shape on the tensor-core tier (any n); beyond it PF_ERR_WORKSPACE, text in pf_last_cuda_error().
 * All calls on a device share the array (it keeps the last call's contents; nothing in it needs to start at zero):
 * calls that overlap on the GPU race -- serialise them.  The first call on a device resolves the array's device
 * address (cudaGetSymbolAddress): make it outside graph capture. */
pf_status pf_sample_fused_static(pf_flow* flow, const float* ctx, const pf_scaler* scaler, const float* z,
                                 uint64_t seed, int64_t E, int64_t n, float* theta, float* logdet,
                                 uint8_t* railed, uint32_t flags, void* stream);

/* ---- density direction (theta -> z) and its backward ---------------------------------------- */

/* Replaces NSFPosteriorFlow.forward (flows.py:610-618): z [R, D], logdet [R].
 * nll != NULL additionally fuses compute_psd_aware_nll (flows.py:727-779) with the
 * PSDScaledNormal base (flows.py:73-85); log_sigma [R, D] or NULL (= zeros, as LeanNPE passes).
 * saved != NULL keeps the activations pf_flow_backward needs (pf_flow_saved_bytes(R, E)). */
pf_status pf_flow_forward(pf_flow* flow, const float* x, const float* ctx, const int64_t* row2ctx,
                          int64_t n_per_ctx, int64_t R, int64_t E, float* z, float* logdet,
                          float* nll, const float* log_sigma, uint32_t flags, void* saved,
                          void* workspace, size_t workspace_bytes, void* stream);

/* Replaces autograd through nflows for the density direction (loss.backward() at
 * experiments/train_lean_npe.py:365).  Given gz [R, D] and glogdet [R] (either may be NULL = 0)
 * accumulates weight grads into `grads` ([host] array of L structs, nflows layout, masked
 * entries get zero gradient exactly as d(W*mask)/dW does) and writes gx [R, D] (or NULL) and
 * gctx [E, C] (or NULL). */
pf_status pf_flow_backward(pf_flow* flow, const void* saved, const float* ctx, const int64_t* row2ctx,
                           int64_t n_per_ctx, int64_t R, int64_t E, const float* gz,
                           const float* glogdet, float* gx, float* gctx, const pf_layer_grads* grads,
                           void* workspace, size_t workspace_bytes, void* stream);

/* ---- glue of the fused training step ---------------------------------------------------------- */

/* Replaces the tensor glue of `batch_nll` (experiments/train_lean_npe.py:108-127) around the two flow calls, in the form
 * that runs EVERY (event, rank) slot as a flow row with weight [rank < nsig] / count (posteriflow_b200/train.py):
 *   x[b*S + r, :]     = ParamScaler.normalize(params[b, r, :])  (lean_npe.py:100-109)   or 0 for an absent slot
 *   cfull[b*S + r, :] = context[b, :] | rank_emb[r, :]           (LeanNPE._full_context, lean_npe.py:296-304)
 *   w[b*S + r]        = [r < nsig[b]] * inv,   inv = *inv_count if use_inv && *use_inv > 0 (the global count of a
 *                       data-parallel job) else 1 / max(1, sum_b min(nsig[b], S))       (train_lean_npe.py:127)
 * params [B, S, D] physical, nsig [B] int64, context [B, context_dim], rank_emb [S, C - context_dim]; x [B*S, D],
 * cfull [B*S, C], w [B*S].  One kernel instead of ~20 elementwise launches. */
pf_status pf_slot_prepare(const float* params, const int64_t* nsig, const float* context, const float* rank_emb,
                          const pf_scaler* scaler, const float* inv_count, const float* use_inv, int32_t B, int32_t S,
                          int32_t D, int32_t context_dim, int32_t C, float* x, float* cfull, float* w, int32_t device,
                          void* stream);

/* Replaces `loss = (nll * w).sum()` of the same function with nll = compute_psd_aware_nll(log_sigma = 0)
 * (flows.py:727-779) and the first step of its backward: loss[0] = sum_r w_r (0.5 (|z_r|^2 + D log 2 pi) - logdet_r),
 * gz[r, :] = w_r z[r, :], glogdet[r] = -w_r.  One cluster of eight CTAs, partial sums combined
 * through distributed shared memory in a fixed order (deterministic, no scratch, no atomics). */
pf_status pf_slot_loss(const float* z, const float* logdet, const float* w, int64_t rows, int32_t D, float* loss, float* gz,
                       float* glogdet, int32_t device, void* stream);

/* Replaces autograd's backward of the context gather and of the rank embedding (lean_npe.py:296-304):
 * gcontext[b, :] = sum_r gcfull[b*S + r, :context_dim],  grank_emb[r, :] = sum_b gcfull[b*S + r, context_dim:]. */
pf_status pf_slot_ctx_grad(const float* gcfull, int32_t B, int32_t S, int32_t context_dim, int32_t C, float* gcontext,
                           float* grank_emb, int32_t device, void* stream);

/* The compact form of the same glue: only the PRESENT (event, rank) pairs become flow rows (event-major), for batches whose
 * slots would not fit one wave of tiles.  pf_rows_prepare writes offs[b] = sum_{b' < b} min(nsig[b'], S) (offs[B] = R, the
 * number of rows; int32 [B + 1]) and rows [0, cap): row offs[b] + k = (event b, rank k) as in pf_slot_prepare with
 * w = inv, inv = *inv_count if use_inv && *use_inv > 0 else 1 / max(1, R); rows [R, cap) get x = 0, w = 0.
 * pf_rows_ctx_grad sums d loss / d full-context rows back per event and per rank.  Replaces signal_rows' nonzero + the
 * index_select / cat / normalize / index_add kernels of batch_nll and its autograd (train_lean_npe.py:108-127). */
pf_status pf_rows_prepare(const float* params, const int64_t* nsig, const float* context, const float* rank_emb,
                          const pf_scaler* scaler, const float* inv_count, const float* use_inv, int32_t B, int32_t S,
                          int32_t D, int32_t context_dim, int32_t C, int32_t cap, int32_t* offs, float* x, float* cfull,
                          float* w, int32_t device, void* stream);
pf_status pf_rows_ctx_grad(const float* gcfull, const int32_t* offs, int32_t B, int32_t S, int32_t context_dim, int32_t C,
                           float* gcontext, float* grank_emb, int32_t device, void* stream);

/* ---- serving epilogue ------------------------------------------------------------------------ */

/* Replaces the tail of one iteration of infer()'s sampling loop (inference/pipeline.py:180, 57-76, 184-186):
 * given wrapped normalised samples y [R, D] and their flow NLL nll [R] (pf_flow_forward(nll != NULL) on y),
 *   theta[r, :] = ParamScaler.denormalize(y[r, :])                       (lean_npe.py:111-114)
 *   logq[r]     = -nll[r] + sum_j (log 2 - log(hi_j - lo_j)) - sum_{log dims} log(max(theta_rj, 1e-6))
 * and, with PF_FLAG_SORT_MASSES, dims 0/1 of theta swapped so that m1 >= m2 (logq is not affected).
 * No handle: the arithmetic involves the scaler only.  theta / logq may be NULL. */
pf_status pf_physical_logq(const float* y, const float* nll, const pf_scaler* scaler, int32_t D, int64_t R,
                           float* theta, float* logq, uint32_t flags, int32_t device, void* stream);

/* ---- on-device posterior summaries ----------------------------------------------------------- */

/* Replaces the host-side numpy summaries the reference computes after pulling every sample off the device
 * (scripts/overlap_benchmark.py:67-89 calib_block, scripts/validate_checkpoint.py:102-128,
 * experiments/train_lean_npe.py:160-167), per event e and dimension j over the n samples theta[e, :, j]:
 *   quant[e, i, j]      = np.quantile(theta[e, :, j], q[i])          ('linear' method = torch.quantile), i < Q <= 8
 *   lower_median[e, j]  = torch.median(theta[e, :, j]).values        (lower of the two middle elements)
 *   below[e, j]         = (theta[e, :, j] < truth[e, j]).mean()      (the SBC rank statistic; needs truth [E, D])
 *   rail_frac[e]        = railed[e*n : (e+1)*n].mean()               (pipeline.py:188, validate_checkpoint.py:126)
 * theta [E, n, D] float32 as pf_sample_fused writes it; q is a HOST array; every output may be NULL.
 * Order statistics are exact (radix select on the float keys), the interpolation is numpy's _lerp in double. */
pf_status pf_posterior_summary(const float* theta, int64_t E, int64_t n, int32_t D, const double* q /* [host] Q */,
                               int32_t Q, const float* truth, const uint8_t* railed, float* quant,
                               float* lower_median, float* below, float* rail_frac, int32_t device, void* stream);

/* ---- sizes / diagnostics -------------------------------------------------------------------- */

size_t pf_flow_workspace_bytes(const pf_flow* flow, int32_t op, int64_t R, int64_t E);
size_t pf_flow_saved_bytes(const pf_flow* flow, int64_t R, int64_t E);
/* 1 if pf_flow_forward(saved) / pf_flow_backward of this (packed) handle run as the fused persistent kernels (tensor-core
 * tier, LeanNPE-class shape, dense context form), 0 if they run GEMM per launch: callers size their batches by it */
int32_t pf_flow_fused_training(const pf_flow* flow);
/* number of kernels the last call on this handle launched (bench.py's gpu_launches counter) */
int64_t pf_flow_last_launch_count(const pf_flow* flow);
/* hidden-unit permutation of the packed layout: perm[s] = nflows unit index at sorted slot s */
pf_status pf_flow_hidden_perm(const pf_flow* flow, int32_t* perm_host /* [host] H entries */);
const char* pf_status_str(pf_status s);
const char* pf_last_cuda_error(void);
int32_t pf_abi_version(void);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* POSTERIFLOW_B200_H */
