// Internal definitions shared by the posteriflow_b200 CUDA translation units (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/posteriflow_b200.h"

#define PF_MAX_D 16      // autoregressive features supported by the kernels
#define PF_MAX_BINS 32   // spline bins supported
#define PF_KC 32         // k-rows per streamed weight chunk of the sampler

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
  // sampler stream: per layer, in consumption order, chunks of PF_KC k-rows
  float* stream;         // [L][stream_layer_floats]
  size_t stream_layer_floats;
  int UH, UF;            // padded unit counts of a hidden-group block / a final (per-dim) block
};
