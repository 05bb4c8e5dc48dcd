// Internal definitions shared by the posteriflow_b200 CUDA translation units (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/posteriflow_b200.h"

#define PF_MAX_D 16      // autoregressive features supported by the kernels
#define PF_MAX_BINS 32   // spline bins supported
#define PF_ROW_CHUNK 16384  // rows processed per pass of the FP32 tier (keeps activations L2-resident)

namespace pf {

extern thread_local char g_cuda_err[256];

inline pf_status cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  snprintf(g_cuda_err, sizeof(g_cuda_err), "%s: %s (%s:%d)", what, cudaGetErrorString(e), file, line);
  return PF_ERR_CUDA;
}
#define PF_CUDA(expr)                                                        \
  do {                                                                       \
    cudaError_t _e = (expr);                                                 \
    if (_e != cudaSuccess) return pf::cuda_fail(_e, #expr, __FILE__, __LINE__); \
  } while (0)
#define PF_LAUNCH_CHECK(flow)                                                \
  do {                                                                       \
    (flow)->launches++;                                                      \
    cudaError_t _e = cudaGetLastError();                                     \
    if (_e != cudaSuccess) return pf::cuda_fail(_e, "kernel launch", __FILE__, __LINE__); \
  } while (0)
#define PF_TRY(expr)                 \
  do {                               \
    pf_status _s = (expr);           \
    if (_s != PF_OK) return _s;      \
  } while (0)

static inline int round_up(int x, int m) { return (x + m - 1) / m * m; }
static inline int64_t round_up64(int64_t x, int64_t m) { return (x + m - 1) / m * m; }
static inline int cdiv(int a, int b) { return (a + b - 1) / b; }

// nflows' deterministic degrees (MaskedLinear._get_mask_and_degrees, random_mask=False).
__host__ __device__ __forceinline__ int hidden_degree(int u, int D) {
  int mx = D - 1 > 1 ? D - 1 : 1;
  int mn = D - 1 < 1 ? D - 1 : 1;
  return u % mx + mn;
}
enum MaskRule { RULE_NONE = 0, RULE_INITIAL = 1, RULE_HIDDEN = 2, RULE_FINAL = 3 };
// mask[r][c] of a MaskedLinear in nflows' own (unsorted) index space
__host__ __device__ __forceinline__ float rule_mask(int rule, int r, int c, int D, int MULT) {
  switch (rule) {
    case RULE_INITIAL: return hidden_degree(r, D) >= c + 1 ? 1.f : 0.f;
    case RULE_HIDDEN: return hidden_degree(r, D) >= hidden_degree(c, D) ? 1.f : 0.f;
    case RULE_FINAL: return (r / MULT + 1) > hidden_degree(c, D) ? 1.f : 0.f;
    default: return 1.f;
  }
}

// Raise the handle's error flag from a kernel (pinned mapped host memory: plain store + system fence; the first
// code to land wins often enough -- the flag only has to be non-zero).
__device__ __forceinline__ void raise_flag(int* err, int code) {
  *reinterpret_cast<volatile int*>(err) = code;
  __threadfence_system();
}

// Row -> context-row addressing shared by every kernel (see header comment in the public .h).
struct CtxMap {
  const int64_t* row2ctx;
  int64_t n_per_ctx;
  int64_t row0;  // global index of local row 0 (calls process rows in chunks)
  __device__ __forceinline__ int64_t operator()(int64_t r) const {
    r += row0;
    return row2ctx ? row2ctx[r] : (n_per_ctx == 1 ? r : r / n_per_ctx);
  }
};

}  // namespace pf

// Packed-weight handle.  All per-layer arrays are stored [L][...] contiguously in one device arena,
// hidden units in degree-sorted order (slot s holds nflows unit perm[s]); masks are pre-multiplied.
struct pf_flow {
  pf_flow_desc desc;
  int D, C, H, L, NB, MULT, P, PLD, DP, G;  // P = D*MULT params/row, PLD = padded ld, DP = padded D
  int off[PF_MAX_D + 1];                    // off[g] = #hidden units with degree <= g  (off[0] = 0)
  int max_group;                            // largest degree group
  int* perm_host;                           // [H]
  bool packed;
  int64_t launches;
  int sm_count;

  float* arena;
  size_t arena_floats;
  int* d_perm;    // [H] device copy of perm
  int* d_flag;    // mask-validation flag

  // dense packed matrices, "N x K" = [out][in] row-major (sorted hidden order on both sides)
  float *W0, *b0;        // [L][H][DP], [L][H]
  float *Wctx, *bctx;    // [L][3H][C], [L][3H]      sections: c0 | gate1 | gate2
  float *Wa, *ba;        // [L][2][H][H], [L][2][H]  blocks.k.linear_layers.0
  float *Wb, *bb;        // [L][2][H][H], [L][2][H]  blocks.k.linear_layers.1
  float *Wf, *bf;        // [L][PLD][H], [L][PLD]    final layer (rows = dim-major params)
  // transposed copies ([in][out]) for the data-gradient GEMMs
  float *WaT, *WbT;      // [L][2][H][H]
  float *WfT;            // [L][H][PLD]
  float *WctxT;          // [C][L*3H]   (all layers side by side: one GEMM for d(ctx))
  float* W0T;            // [L][DP][H]  (d(x) through the initial layer)
  // bf16 tensor-core tier (pf_tc.cu): opaque state (weight image + padded side arrays), built by pf_flow_pack
  void* tc;              // NULL when tier != PF_TIER_BF16
  int UF;                // padded per-dim parameter count (MULT rounded up to 4)
  void* dense;           // dense density-direction kernel state (pf_tc_dense.cu), NULL when the shape is not covered
  void* train;           // fused training-step state (pf_tc_train.cu): backward weight images
  // training path of the tensor-core tier (pf_tc_gemm.cuh): split-bf16 tcgen05 GEMMs instead of the FP32-FMA ones
  bool train_tc;         // tier == PF_TIER_BF16 and PF_TRAIN_TC != 0
  int train_tc_min_rows; // calls with fewer flow rows stay on the FP32-FMA GEMMs (PF_TRAIN_TC_MIN_ROWS, default 512)
  bool train_debug;      // PF_TC_DEBUG=1: synchronise and check the error flag after every tensor-core call
  // Error flag of every tcgen05 kernel of this handle (barrier timeout = protocol bug or a wedged SM).  It lives in
  // pinned, mapped HOST memory: a kernel raises it with a system-scope store, the host reads it WITHOUT a stream
  // synchronisation at the start of the next call on the handle and in pf_flow_check().  `tg_err` is the device
  // alias of `h_err`.
  int* tg_err;
  volatile int* h_err;
  void* pack_state;      // host-side packer state (pf_pack.cu: cached job table)
  bool arena_zeroed;     // the packed arena's pad entries are zero-filled once (first pf_flow_pack)
  // backward: side stream for the gradients that are off the layer-to-layer chain (fork / join with events)
  static constexpr int NSIDE = 4;   // final layer | residual block 1 | residual block 0 | initial layer + context path
  cudaStream_t side[NSIDE];
  cudaEvent_t ev_main[2], ev_side[2][NSIDE];
  bool train_streams;    // PF_TRAIN_STREAMS != 0
  bool train_pdl;        // PF_TRAIN_PDL != 0: programmatic dependent launch of the tcgen05 training GEMMs
  bool train_fused;      // PF_TRAIN_FUSED != 0: forward-with-saves / backward / weight gradients as three persistent kernels
};

// Host side of the error flag: non-blocking read (+ clear).  PF_ERR_CUDA with a message when a kernel of an EARLIER,
// already finished call on this handle hit a barrier timeout.
pf_status pf_check_flag(pf_flow* f, const char* where);

// bf16 tensor-core tier entry points (pf_tc.cu)
pf_status pf_tc_pack(pf_flow* f, cudaStream_t st);
pf_status pf_tc_inverse(pf_flow* f, cudaStream_t st, const float* z, uint64_t seed, const float* ctx,
                        const int64_t* row2ctx, int64_t n_per_ctx, int64_t R, int64_t E, float* x, float* logdet,
                        uint8_t* railed, uint32_t flags, const pf_scaler* scaler, void* workspace,
                        size_t workspace_bytes);
pf_status pf_tc_forward(pf_flow* f, cudaStream_t st, const float* x, const float* ctx, const int64_t* row2ctx,
                        int64_t n_per_ctx, int64_t R, int64_t E, float* z, float* logdet, uint32_t flags,
                        void* workspace, size_t workspace_bytes);
pf_status pf_tc_ctxproj(pf_flow* f, cudaStream_t st, const float* ctx, int64_t E, uint32_t flags, float* proj);
size_t pf_tc_workspace_bytes(const pf_flow* f, int64_t R, int64_t E);
void pf_tc_destroy(pf_flow* f);
// dense tcgen05 density direction (pf_tc_dense.cu): H = 256, 16 bins, D <= 11
bool pf_dn_supported(const pf_flow* f);
pf_status pf_dn_pack(pf_flow* f, cudaStream_t st);
void pf_dn_destroy(pf_flow* f);
pf_status pf_dn_forward(pf_flow* f, cudaStream_t st, const float* x, const float* ctx, const int64_t* row2ctx,
                        int64_t n_per_ctx, int64_t R, int64_t E, float* z, float* logdet, uint32_t flags,
                        void* workspace, size_t workspace_bytes, int* err, int split);
pf_status pf_dn_forward_save(pf_flow* f, cudaStream_t st, const float* x, const float* proj, int64_t R, float* z,
                             float* logdet, uint8_t* img, uint8_t* xt, float* P, float* XS);
// fused training step, backward side (pf_tc_train.cu)
pf_status pf_tt_pack(pf_flow* f, cudaStream_t st);
void pf_tt_destroy(pf_flow* f);
