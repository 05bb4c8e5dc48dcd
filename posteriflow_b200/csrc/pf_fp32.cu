// FP32 parity tier (PF_TIER_FP32) of the posterior-flow hot path: every conditioner matrix product is an
// FP32-FMA GEMM with the bias / context / GLU / residual arithmetic fused into its epilogue, the spline
// work is one fused kernel per layer (density direction) or per autoregressive dimension (sampling
// direction).  Sampling is INCREMENTAL: hidden units are stored sorted by autoregressive degree, so the
// masked matrices are block lower-triangular and dimension d's 3K-1 spline parameters need only the
// hidden groups of degree < d -- one conditioner pass of work per layer instead of nflows' D passes
// (SURVEY.md 0.2, Appendix A.2-A.4).
//
// Reference interfaces replaced (file:line under /root/reference/src/ahsd/models):
//   flows.py:610-618  NSFPosteriorFlow.forward          -> pf_flow_forward
//   flows.py:620-655  NSFPosteriorFlow.inverse          -> pf_flow_inverse
//   flows.py:727-779  compute_psd_aware_nll             -> pf_flow_forward(nll != NULL)
//   lean_npe.py:318-332 LeanNPE.sample_posterior (post-encoder) -> pf_sample_fused
//   loss.backward() through nflows (experiments/train_lean_npe.py:365) -> pf_flow_backward
#include <stdlib.h>

#include "pf_common.cuh"
#include "pf_gemm.cuh"
#include "pf_tc_gemm.cuh"
#include "pf_spline.cuh"
#include "pf_tail.cuh"
#include "pf_train.cuh"

namespace pf {

// Launch of a kernel on the training dependency chain: with programmatic dependent launch (tensor-core tier) the grid
// may start while its predecessor in the stream drains; such kernels execute griddepcontrol.wait before their first
// global read (a no-op under a normal launch).
template <class... KArgs, class... Args>
static pf_status launch_chain(pf_flow* f, void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, Args... args) {
  if (f->train_tc && f->train_pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    PF_CUDA(cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...));
    f->launches++;
    return PF_OK;
  }
  kernel<<<grid, block, 0, st>>>(KArgs(args)...);
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}
__device__ __forceinline__ void chain_kernel_entry() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// ------------------------------------------------------------------------------------------------
// workspace carving
// ------------------------------------------------------------------------------------------------
struct Carver {
  char* base;
  size_t cap, used;
  bool ok;
  Carver(void* p, size_t n) : base((char*)p), cap(n), used(0), ok(true) {}
  float* f32(size_t n) {
    size_t b = (n * sizeof(float) + 255) / 256 * 256;
    if (!base || used + b > cap) { ok = false; used += b; return nullptr; }
    float* r = (float*)(base + used);
    used += b;
    return r;
  }
};
static inline size_t pad256(size_t floats) { return (floats * sizeof(float) + 255) / 256 * 256; }

// ------------------------------------------------------------------------------------------------
// GEMM epilogues
// ------------------------------------------------------------------------------------------------
// proj[e, n] = relu | sigmoid (acc + b[n]) : section (n / H) % 3 == 0 is MADE.context_layer (relu),
// 1|2 are the residual blocks' GLU gates (sigmoid).
struct EpiCtxProj {
  float* out; int ld; const float* bias; int H;
  __device__ __forceinline__ void operator()(int m, int n, float acc) const {
    float v = acc + bias[n];
    int sec = (n / H) % 3;
    out[(int64_t)m * ld + n] = sec == 0 ? fmaxf(v, 0.f) : 1.f / (1.f + expf(-v));
  }
};
// The same values in the tile-transposed layout of the fused training step (pf_train.cuh): out[tile][n][row in tile].
// kLanesAlongRows: the tensor-core GEMM's epilogue then runs its lanes along m, so these stores are coalesced.
struct EpiCtxProjT {
  static constexpr bool kLanesAlongRows = true;
  float* out; int ld; const float* bias; int H;
  __device__ __forceinline__ void operator()(int m, int n, float acc) const {
    float v = acc + bias[n];
    int sec = (n / H) % 3;
    out[((int64_t)(m / TT_ROWS) * ld + n) * TT_ROWS + (m % TT_ROWS)] = sec == 0 ? fmaxf(v, 0.f) : 1.f / (1.f + expf(-v));
  }
};
// Epilogues may split into fetch(m, n) -> operands read from global memory, and apply(m, n, acc, fetched): a caller
// that handles several elements per thread (pf_tc_gemm.cuh) issues all fetches of a batch before the first apply, so
// the loads overlap instead of costing one memory latency per element.  operator() stays the single-call form.
// out[m, n0 + n] = acc + bias[n0 + n]
struct EpiBias {
  float* out; int ld; const float* bias; int n0;
  __device__ __forceinline__ void operator()(int m, int n, float acc) const {
    out[(int64_t)m * ld + n0 + n] = acc + bias[n0 + n];
  }
  __device__ __forceinline__ float fetch(int m, int n) const { return bias[n0 + n]; }
  __device__ __forceinline__ void apply(int m, int n, float acc, float b) const { out[(int64_t)m * ld + n0 + n] = acc + b; }
};
// hout = hin + gate * (acc + bias)   (GLU-gated residual, nflows MaskedResidualBlock); optional v store
struct EpiGluResidual {
  const float* hin; float* hout; float* vout; int ld; const float* bias; int n0;
  const float* gate; int64_t ldg; CtxMap cm;
  __device__ __forceinline__ void operator()(int m, int n, float acc) const {
    float v = acc + bias[n0 + n];
    float g = gate[cm(m) * ldg + n0 + n];
    int64_t i = (int64_t)m * ld + n0 + n;
    if (vout) vout[i] = v;
    hout[i] = hin[i] + g * v;
  }
  __device__ __forceinline__ float3 fetch(int m, int n) const {
    return make_float3(bias[n0 + n], gate[cm(m) * ldg + n0 + n], hin[(int64_t)m * ld + n0 + n]);
  }
  __device__ __forceinline__ void apply(int m, int n, float acc, float3 t) const {
    const float v = acc + t.x;
    const int64_t i = (int64_t)m * ld + n0 + n;
    if (vout) vout[i] = v;
    hout[i] = t.z + t.y * v;
  }
};

// ------------------------------------------------------------------------------------------------
// elementwise kernels
// ------------------------------------------------------------------------------------------------
// MADE initial layer for units [u0, u0+nu): h0 = W0[u, :kmax] . x + b0 + c0(ctx)   (K <= D: not a GEMM)
__global__ void initial_kernel(const float* __restrict__ X, int ldx, int reversed, int D, const float* __restrict__ W0,
                               int DP, const float* __restrict__ b0, const float* __restrict__ c0, int64_t ldc,
                               CtxMap cm, int R, int u0, int nu, int kmax, float* __restrict__ Hout, int ldh) {
  chain_kernel_entry();
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)R * nu) return;
  int r = (int)(idx / nu), u = u0 + (int)(idx % nu);
  float acc = b0[u];
  const float* xr = X + (int64_t)r * ldx;
  const float* w = W0 + (int64_t)u * DP;
  for (int k = 0; k < kmax; ++k) acc = fmaf(w[k], xr[reversed ? D - 1 - k : k], acc);
  Hout[(int64_t)r * ldh + u] = acc + c0[cm(r) * ldc + u];
}

struct RegParams {
  const float* q;
  __device__ __forceinline__ float operator()(int i) const { return q[i]; }
};

// Density direction, one layer: thread (dim = threadIdx.x < D, row = blockIdx.x*16 + threadIdx.y).
// xin = Xin[r, D-1-j] (the ReversePermutation in front of every MAF), Y[r, j] = spline_j(xin),
// logdet[r] (+)= sum_j lad.
template <int NB>
__global__ void __launch_bounds__(256)
spline_fwd_kernel(const float* __restrict__ Xin, const float* __restrict__ P, int PLD, int D, float B, int R,
                  float* __restrict__ Y, float* __restrict__ logdet, int accumulate) {
  chain_kernel_entry();
  constexpr int MULT = 3 * NB - 1;
  const int j = threadIdx.x;
  const int r = blockIdx.x * 16 + threadIdx.y;
  float lad = 0.f;
  if (r < R && j < D) {
    float q[MULT];
    const float* p = P + (int64_t)r * PLD + j * MULT;
#pragma unroll
    for (int i = 0; i < MULT; ++i) q[i] = p[i];
    float x = Xin[(int64_t)r * D + (D - 1 - j)];
    Y[(int64_t)r * D + j] = rqs_apply<NB, false>(RegParams{q}, x, B, lad);
  }
#pragma unroll
  for (int o = 8; o >= 1; o >>= 1) lad += __shfl_xor_sync(0xffffffffu, lad, o, 16);
  if (r < R && j == 0) logdet[r] = accumulate ? logdet[r] + lad : lad;
}

// Sampling direction, one (layer, dim j): thread per row; parameters of this dim in Pd [R, UF].
template <int NB>
__global__ void __launch_bounds__(128)
spline_inv_kernel(const float* __restrict__ Yin, int reversed, const float* __restrict__ Pd, int UF, int D, int j,
                  float B, int R, float* __restrict__ U, float* __restrict__ logdet, int accumulate) {
  constexpr int MULT = 3 * NB - 1;
  constexpr int NV = (MULT + 3) / 4;
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R) return;
  float q[NV * 4];
  const float4* p = reinterpret_cast<const float4*>(Pd + (int64_t)r * UF);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    float4 v = p[i];
    q[4 * i] = v.x; q[4 * i + 1] = v.y; q[4 * i + 2] = v.z; q[4 * i + 3] = v.w;
  }
  float y = Yin[(int64_t)r * D + (reversed ? D - 1 - j : j)];
  float lad;
  U[(int64_t)r * D + j] = rqs_apply<NB, true>(RegParams{q}, y, B, lad);
  logdet[r] = accumulate ? logdet[r] + lad : lad;
}

// PSDScaledNormal.log_prob + compute_psd_aware_nll (flows.py:73-85, 767-770), thread per row
__global__ void nll_kernel(const float* __restrict__ Z, const float* __restrict__ logdet,
                           const float* __restrict__ log_sigma, int D, int R, float* __restrict__ nll) {
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R) return;
  float quad = 0.f, var = 0.f;
  for (int j = 0; j < D; ++j) {
    float z = Z[(int64_t)r * D + j];
    if (log_sigma) {
      float ls = log_sigma[(int64_t)r * D + j];
      z = z / expf(ls);
      var += ls;
    }
    quad = fmaf(z, z, quad);
  }
  float logp = -0.5f * (quad + 2.f * var + (float)D * 1.8378770664093453f);
  nll[r] = -(logp + logdet[r]);
}

__global__ void copy_rows_kernel(const float* __restrict__ src, float* __restrict__ dst, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[i];
}

__global__ void philox_normal_kernel(float* __restrict__ Z, int D, int R, int64_t row0, uint64_t seed) {
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R) return;
  float n[PF_MAX_D];
  philox_normal_row((uint64_t)(row0 + r), seed, D, n);
  for (int j = 0; j < D; ++j) Z[(int64_t)r * D + j] = n[j];
}

// Tail of NSFPosteriorFlow.inverse (flows.py:648-654) + ParamScaler.wrap / denormalize
// (lean_npe.py:100-114) + rail mask (pipeline.py:177) + m1>=m2 ordering (pipeline.py:184-186).
// U holds the last MAF's inverse output; the trailing ReversePermutation is folded into the read.
__global__ void finish_inverse_kernel(const float* __restrict__ U, int D, int R, uint32_t flags, float clamp_bound,
                                      ScalerDev sc, float* __restrict__ X, uint8_t* __restrict__ railed) {
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R) return;
  float v[PF_MAX_D];
#pragma unroll
  for (int j = 0; j < PF_MAX_D; ++j) v[j] = j < D ? U[(int64_t)r * D + (D - 1 - j)] : 0.f;
  bool rail = finish_row(v, D, flags, clamp_bound, sc);
#pragma unroll
  for (int j = 0; j < PF_MAX_D; ++j) {
    if (j >= D) break;
    X[(int64_t)r * D + j] = v[j];
  }
  if (railed) railed[r] = rail ? 1 : 0;
}

}  // namespace pf
using namespace pf;

#define PF_NB_DISPATCH(NBV, CALL)                          \
  switch (NBV) {                                           \
    case 8: { constexpr int NB_ = 8; CALL; } break;        \
    case 16: { constexpr int NB_ = 16; CALL; } break;      \
    case 24: { constexpr int NB_ = 24; CALL; } break;      \
    case 32: { constexpr int NB_ = 32; CALL; } break;      \
    default: return PF_ERR_UNSUPPORTED;                    \
  }

// ------------------------------------------------------------------------------------------------
// context projection (A3)
// ------------------------------------------------------------------------------------------------
static pf_status ctxproj_tiled(pf_flow* f, cudaStream_t st, const float* ctx, int64_t E, uint32_t flags, float* proj) {
  const int N = f->L * 3 * f->H;
  if (E % TT_ROWS) {   // rows past E of the last tile are read (never used) by the chain kernels: keep them finite
    const int64_t t = E / TT_ROWS;
    PF_CUDA(cudaMemsetAsync(proj + t * N * TT_ROWS, 0, (size_t)N * TT_ROWS * sizeof(float), st));
  }
  EpiCtxProjT epi{proj, N, f->bctx, f->H};
  if (flags & PF_FLAG_SANITIZE_CTX)
    return gemm_nt<EpiCtxProjT, A_SANITIZE, 64>(f, st, ctx, f->C, f->Wctx, f->C, (int)E, N, f->C, epi);
  return gemm_nt<EpiCtxProjT, A_PLAIN, 64>(f, st, ctx, f->C, f->Wctx, f->C, (int)E, N, f->C, epi);
}

static pf_status ctxproj_impl(pf_flow* f, cudaStream_t st, const float* ctx, int64_t E, uint32_t flags, float* proj) {
  const int N = f->L * 3 * f->H;
  EpiCtxProj epi{proj, N, f->bctx, f->H};
  if (flags & PF_FLAG_SANITIZE_CTX)
    return gemm_nt<EpiCtxProj, A_SANITIZE, 64>(f, st, ctx, f->C, f->Wctx, f->C, (int)E, N, f->C, epi);
  return gemm_nt<EpiCtxProj, A_PLAIN, 64>(f, st, ctx, f->C, f->Wctx, f->C, (int)E, N, f->C, epi);
}

// PF_TC_DEBUG=1: synchronise and surface a barrier timeout of the split-bf16 tcgen05 GEMMs (pf_tc_gemm.cuh)
static pf_status train_tc_check(pf_flow* f, cudaStream_t st, const char* where) {
  if (!f->train_tc || !f->train_debug) return PF_OK;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cap) != cudaSuccess || cap != cudaStreamCaptureStatusNone) return PF_OK;  // no sync inside a capture
  PF_CUDA(cudaStreamSynchronize(st));
  return pf_check_flag(f, where);
}

static pf_status check_call(pf_flow* f, int64_t R, int64_t E, const int64_t* row2ctx, int64_t n_per_ctx) {
  if (!f) return PF_ERR_INVALID;
  if (!f->packed) return PF_ERR_NOT_PACKED;
  if (R < 0 || E < 0 || R > (int64_t)1 << 40) return PF_ERR_INVALID;
  if (!row2ctx) {
    if (n_per_ctx < 1) return PF_ERR_INVALID;
    if (R > 0 && (R - 1) / n_per_ctx >= E) return PF_ERR_INVALID;  // not enough context rows
  }
  if (E > 8000000) return PF_ERR_INVALID;
  // a barrier timeout of an earlier, finished call on this handle surfaces here (no stream synchronisation)
  return pf_check_flag(f, "previous call");
}

extern "C" pf_status pf_flow_ctxproj(pf_flow* f, const float* ctx, int64_t E, uint32_t flags, float* proj,
                                     void* stream) {
  if (!f || !ctx || !proj || E < 0 || E > 8000000) return PF_ERR_INVALID;
  if (!f->packed) return PF_ERR_NOT_PACKED;
  PF_CUDA(cudaSetDevice(f->desc.device));
  f->launches = 0;
  if (f->desc.tier == PF_TIER_BF16) return pf_tc_ctxproj(f, (cudaStream_t)stream, ctx, E, flags, proj);
  return ctxproj_impl(f, (cudaStream_t)stream, ctx, E, flags, proj);
}

// ------------------------------------------------------------------------------------------------
// saved-activation layout shared by forward (writer) and backward (reader)
// ------------------------------------------------------------------------------------------------
struct SavedLayout {
  float* proj;   // [E][L*3H]
  float* XS;     // [L][R][D]   XS[l] = input of layer l (before its ReversePermutation)
  float* act;    // [L][7][R][H]  H0 U1 V1 H1 U2 V2 H2
  float* P;      // [L][R][PLD]
  size_t bytes;
  SavedLayout(const pf_flow* f, void* base, int64_t R, int64_t E) {
    Carver c(base ? base : (void*)256, (size_t)-1 / 2);  // counting mode when base == NULL (pointers unused)
    proj = c.f32((size_t)E * f->L * 3 * f->H);
    XS = c.f32((size_t)f->L * R * f->D);
    act = c.f32((size_t)f->L * 7 * R * f->H);
    P = c.f32((size_t)f->L * R * f->PLD);
    bytes = c.used;
  }
  float* a(const pf_flow* f, int64_t R, int l, int which) const { return act + ((size_t)l * 7 + which) * R * f->H; }
};

// The fused training step (pf_train.cuh) serves the dense-context form (one context row per flow row: what batch_nll /
// LeanNPE.nll pass) of handles whose shape the dense tensor-core kernels cover.
static inline bool fused_train(const pf_flow* f, const int64_t* row2ctx, int64_t n_per_ctx) {
  return f->train_fused && f->train && f->dense && !row2ctx && n_per_ctx == 1 && f->C <= TT_CIMG_COLS;
}

extern "C" int32_t pf_flow_fused_training(const pf_flow* f) {
  return (f && f->train_fused && f->train && f->dense && f->C <= TT_CIMG_COLS) ? 1 : 0;
}

extern "C" size_t pf_flow_saved_bytes(const pf_flow* f, int64_t R, int64_t E) {
  if (!f || R < 0 || E < 0) return 0;
  size_t b = SavedLayout(f, nullptr, R, E).bytes;
  if (f->train_fused && f->train && f->dense) {
    const size_t t = TtSaved(f, nullptr, R).bytes;
    if (t > b) b = t;
  }
  return b;
}

// ------------------------------------------------------------------------------------------------
// density direction (A4-A7, A9, A10)
// ------------------------------------------------------------------------------------------------
struct FwdWs {
  float *proj, *h, *u, *p, *xa, *xb, *zbuf, *ldbuf;
  size_t bytes; bool ok;
  FwdWs(const pf_flow* f, void* base, size_t cap, int64_t chunk, int64_t proj_rows, bool with_act) {
    Carver c(base, base ? cap : (size_t)-1 / 2);
    proj = c.f32((size_t)proj_rows * f->L * 3 * f->H);
    if (with_act) {
      h = c.f32((size_t)chunk * f->H); u = c.f32((size_t)chunk * f->H); p = c.f32((size_t)chunk * f->PLD);
      xa = c.f32((size_t)chunk * f->D); xb = c.f32((size_t)chunk * f->D);
    } else { h = u = p = xa = xb = nullptr; }
    zbuf = c.f32((size_t)chunk * f->D); ldbuf = c.f32((size_t)chunk);
    bytes = c.used; ok = c.ok;
  }
};

static inline int64_t proj_rows_needed(bool dense, bool is_proj, int64_t chunk, int64_t E) {
  return is_proj ? 0 : (dense ? chunk : E);
}

template <int NB>
static pf_status launch_spline_fwd(pf_flow* f, cudaStream_t st, const float* Xin, const float* P, int R, float* Y,
                                   float* logdet, int accumulate) {
  dim3 block(16, 16);
  return launch_chain(f, spline_fwd_kernel<NB>, dim3(cdiv(R, 16)), block, st, Xin, P, f->PLD, f->D, f->desc.tail_bound, R,
                      Y, logdet, accumulate);
}

extern "C" pf_status pf_flow_forward(pf_flow* f, const float* x, const float* ctx, const int64_t* row2ctx,
                                     int64_t n_per_ctx, int64_t R, int64_t E, float* z, float* logdet, float* nll,
                                     const float* log_sigma, uint32_t flags, void* saved, void* workspace,
                                     size_t workspace_bytes, void* stream) {
  PF_TRY(check_call(f, R, E, row2ctx, n_per_ctx));
  f->launches = 0;
  if (R == 0) return PF_OK;
  if (!x || !ctx) return PF_ERR_INVALID;
  PF_CUDA(cudaSetDevice(f->desc.device));
  cudaStream_t st = (cudaStream_t)stream;
  const int D = f->D, H = f->H, L = f->L, PLD = f->PLD, P = f->P;
  const bool x3_dense = f->desc.tier == PF_TIER_BF16X3 && f->dense && !(flags & PF_FLAG_CTX_IS_PROJ);
  if ((f->desc.tier == PF_TIER_BF16 || x3_dense) && !saved) {
    // inference form (no activations kept): the density direction runs on the tensor-core kernels too -- bf16 operands
    // (PF_TIER_BF16) or split operands + exact spline (PF_TIER_BF16X3: the dense kernel at the FP32 tolerance)
    const size_t tcb = (pf_tc_workspace_bytes(f, R, E) + 255) / 256 * 256;
    const size_t need = tcb + pad256((size_t)R * D) + pad256((size_t)R);
    if (!workspace || workspace_bytes < need) return PF_ERR_WORKSPACE;
    float* zb = z ? z : reinterpret_cast<float*>((char*)workspace + tcb);
    float* lb = logdet ? logdet : reinterpret_cast<float*>((char*)workspace + tcb + pad256((size_t)R * D));
    if (x3_dense) {
      PF_TRY(pf_dn_forward(f, st, x, ctx, row2ctx, n_per_ctx, R, E, zb, lb, flags & PF_FLAG_SANITIZE_CTX, workspace, tcb,
                           f->tg_err, 3));
      if (f->train_debug) {
        PF_CUDA(cudaStreamSynchronize(st));
        PF_TRY(pf_check_flag(f, "split dense density kernel"));
      }
    } else {
      PF_TRY(pf_tc_forward(f, st, x, ctx, row2ctx, n_per_ctx, R, E, zb, lb, flags, workspace, tcb));
    }
    if (nll) {
      if (R > 0x7fffffff) return PF_ERR_INVALID;
      nll_kernel<<<(unsigned)((R + 255) / 256), 256, 0, st>>>(zb, lb, log_sigma, D, (int)R, nll);
      PF_LAUNCH_CHECK(f);
    }
    return PF_OK;
  }
  if (saved && fused_train(f, row2ctx, n_per_ctx)) {
    // ---- fused training forward: context projection GEMM + ONE persistent kernel over all layers (pf_tc_dense.cu) ----
    if (R > 8000000) return PF_ERR_INVALID;
    TtSaved sv(f, saved, R);
    const int64_t ldp = (int64_t)L * 3 * H;
    (void)ldp;
    if (flags & PF_FLAG_CTX_IS_PROJ) return PF_ERR_UNSUPPORTED;   // a pre-projected context is row-major: GEMM-per-launch form only
    {
      const char* ct = getenv("PF_CTX_TC");
      if (f->C <= TT_CIMG_COLS && !(flags & PF_FLAG_SANITIZE_CTX) && !(ct && ct[0] == '0'))
        PF_TRY(pf_tt_ctxproj(f, st, ctx, R, sv.proj, sv.c16));       // tcgen05 projection kernel (pf_tc_train.cu)
      else
        PF_TRY(ctxproj_tiled(f, st, ctx, R, flags, sv.proj));        // generic tensor-core GEMM, fp32 operands split on the fly
    }
    const size_t need = pad256((size_t)R * D) + pad256((size_t)R);
    if ((!z || !logdet) && (!workspace || workspace_bytes < need)) return PF_ERR_WORKSPACE;
    float* zb = z ? z : reinterpret_cast<float*>(workspace);
    float* lb = logdet ? logdet : reinterpret_cast<float*>((char*)workspace + pad256((size_t)R * D));
    PF_TRY(pf_dn_forward_save(f, st, x, sv.proj, R, zb, lb, sv.img, sv.xt, sv.P, sv.XS));
    if (nll) {
      nll_kernel<<<(unsigned)((R + 255) / 256), 256, 0, st>>>(zb, lb, log_sigma, D, (int)R, nll);
      PF_LAUNCH_CHECK(f);
    }
    return train_tc_check(f, st, "pf_flow_forward (fused)");
  }
  const bool is_proj = flags & PF_FLAG_CTX_IS_PROJ;
  const bool dense = !row2ctx && n_per_ctx == 1;
  const int64_t chunk = saved ? R : (R < PF_ROW_CHUNK ? R : PF_ROW_CHUNK);
  if (chunk > 8000000) return PF_ERR_INVALID;
  const int64_t ldproj = (int64_t)L * 3 * H;
  SavedLayout sv(f, saved, R, E);
  FwdWs ws(f, workspace, workspace_bytes, chunk, saved ? 0 : proj_rows_needed(dense, is_proj, chunk, E), !saved);
  if (!ws.ok) return PF_ERR_WORKSPACE;

  const float* proj_all = nullptr;  // projection addressed by global ctx row
  if (saved) {
    if (is_proj) {
      int64_t n = E * ldproj;
      copy_rows_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(ctx, sv.proj, n);
      PF_LAUNCH_CHECK(f);
    } else {
      PF_TRY(ctxproj_impl(f, st, ctx, E, flags, sv.proj));
    }
    proj_all = sv.proj;
  } else if (is_proj) {
    proj_all = ctx;
  } else if (!dense) {
    PF_TRY(ctxproj_impl(f, st, ctx, E, flags, ws.proj));
    proj_all = ws.proj;
  }

  for (int64_t r0 = 0; r0 < R; r0 += chunk) {
    const int rows = (int)((R - r0) < chunk ? (R - r0) : chunk);
    const float* proj = proj_all;
    CtxMap cm{row2ctx, n_per_ctx, r0};
    if (!proj_all) {  // dense [R, C] context: project this chunk's rows only
      PF_TRY(ctxproj_impl(f, st, ctx + r0 * f->C, rows, flags, ws.proj));
      proj = ws.proj;
      cm = CtxMap{nullptr, 1, 0};
    }
    float* zc = z ? z + r0 * D : ws.zbuf;
    float* ldc = logdet ? logdet + r0 : ws.ldbuf;
    if (saved) {
      int64_t n = (int64_t)rows * D;
      copy_rows_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(x, sv.XS, n);
      PF_LAUNCH_CHECK(f);
    }
    for (int l = 0; l < L; ++l) {
      const float* xin = saved ? sv.XS + (size_t)l * R * D : (l == 0 ? x + r0 * D : ((l & 1) ? ws.xa : ws.xb));
      float* yout = (l == L - 1) ? zc : (saved ? sv.XS + (size_t)(l + 1) * R * D : ((l & 1) ? ws.xb : ws.xa));
      float* H0 = saved ? sv.a(f, R, l, 0) : ws.h;
      float* U1 = saved ? sv.a(f, R, l, 1) : ws.u;
      float* V1 = saved ? sv.a(f, R, l, 2) : nullptr;
      float* H1 = saved ? sv.a(f, R, l, 3) : ws.h;
      float* U2 = saved ? sv.a(f, R, l, 4) : ws.u;
      float* V2 = saved ? sv.a(f, R, l, 5) : nullptr;
      float* H2 = saved ? sv.a(f, R, l, 6) : ws.h;
      float* Pl = saved ? sv.P + (size_t)l * R * PLD : ws.p;
      const float* pl = proj + (size_t)l * 3 * H;
      {
        int64_t n = (int64_t)rows * H;
        PF_TRY(launch_chain(f, initial_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), st, xin, D, 1, D,
                            f->W0 + (size_t)l * H * f->DP, f->DP, f->b0 + (size_t)l * H, pl, ldproj, cm, rows, 0, H, D, H0, H));
      }
      const float* Wa1 = f->Wa + ((size_t)l * 2 + 0) * H * H; const float* ba1 = f->ba + ((size_t)l * 2 + 0) * H;
      const float* Wb1 = f->Wb + ((size_t)l * 2 + 0) * H * H; const float* bb1 = f->bb + ((size_t)l * 2 + 0) * H;
      const float* Wa2 = f->Wa + ((size_t)l * 2 + 1) * H * H; const float* ba2 = f->ba + ((size_t)l * 2 + 1) * H;
      const float* Wb2 = f->Wb + ((size_t)l * 2 + 1) * H * H; const float* bb2 = f->bb + ((size_t)l * 2 + 1) * H;
      PF_TRY((gemm_nt<EpiBias, A_RELU, 64>(f, st, H0, H, Wa1, H, rows, H, H, EpiBias{U1, H, ba1, 0})));
      PF_TRY((gemm_nt<EpiGluResidual, A_RELU, 64>(f, st, U1, H, Wb1, H, rows, H, H,
                                                   EpiGluResidual{H0, H1, V1, H, bb1, 0, pl + H, ldproj, cm})));
      PF_TRY((gemm_nt<EpiBias, A_RELU, 64>(f, st, H1, H, Wa2, H, rows, H, H, EpiBias{U2, H, ba2, 0})));
      PF_TRY((gemm_nt<EpiGluResidual, A_RELU, 64>(f, st, U2, H, Wb2, H, rows, H, H,
                                                   EpiGluResidual{H1, H2, V2, H, bb2, 0, pl + 2 * H, ldproj, cm})));
      PF_TRY((gemm_nt<EpiBias, A_PLAIN, 64>(f, st, H2, H, f->Wf + (size_t)l * PLD * H, H, rows, P, H,
                                             EpiBias{Pl, PLD, f->bf + (size_t)l * PLD, 0})));
      PF_NB_DISPATCH(f->NB, PF_TRY(launch_spline_fwd<NB_>(f, st, xin, Pl, rows, yout, ldc, l > 0)));
    }
    if (nll) {
      nll_kernel<<<cdiv(rows, 256), 256, 0, st>>>(zc, ldc, log_sigma ? log_sigma + r0 * D : nullptr, D, rows,
                                                  nll + r0);
      PF_LAUNCH_CHECK(f);
    }
  }
  return train_tc_check(f, st, "pf_flow_forward");
}

// ------------------------------------------------------------------------------------------------
// sampling direction (A2, A4-A6, A8-A11): incremental autoregressive inverse
// ------------------------------------------------------------------------------------------------
struct InvWs {
  float *proj, *a[5], *pd, *u0, *u1, *zbuf, *ldbuf;
  size_t bytes; bool ok;
  InvWs(const pf_flow* f, void* base, size_t cap, int64_t chunk, int64_t proj_rows) {
    Carver c(base, base ? cap : (size_t)-1 / 2);
    proj = c.f32((size_t)proj_rows * f->L * 3 * f->H);
    for (int i = 0; i < 5; ++i) a[i] = c.f32((size_t)chunk * f->H);
    pd = c.f32((size_t)chunk * f->UF);
    u0 = c.f32((size_t)chunk * f->D); u1 = c.f32((size_t)chunk * f->D);
    zbuf = c.f32((size_t)chunk * f->D); ldbuf = c.f32((size_t)chunk);
    bytes = c.used; ok = c.ok;
  }
};

template <int NB>
static pf_status launch_spline_inv(pf_flow* f, cudaStream_t st, const float* Yin, int reversed, const float* Pd, int j,
                                   int R, float* U, float* logdet, int accumulate) {
  spline_inv_kernel<NB><<<cdiv(R, 128), 128, 0, st>>>(Yin, reversed, Pd, f->UF, f->D, j, f->desc.tail_bound, R, U,
                                                     logdet, accumulate);
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}

static pf_status inverse_impl(pf_flow* f, cudaStream_t st, const float* z, uint64_t seed, const float* ctx,
                              const int64_t* row2ctx, int64_t n_per_ctx, int64_t R, int64_t E, float* x,
                              float* logdet, uint8_t* railed, uint32_t flags, const pf_scaler* scaler, void* workspace,
                              size_t workspace_bytes) {
  const int D = f->D, H = f->H, L = f->L, G = f->G, MULT = f->MULT, UF = f->UF;
  const bool is_proj = flags & PF_FLAG_CTX_IS_PROJ;
  const bool dense = !row2ctx && n_per_ctx == 1;
  const int64_t chunk = R < PF_ROW_CHUNK ? R : PF_ROW_CHUNK;
  const int64_t ldproj = (int64_t)L * 3 * H;
  InvWs ws(f, workspace, workspace_bytes, chunk, proj_rows_needed(dense, is_proj, chunk, E));
  if (!ws.ok) return PF_ERR_WORKSPACE;
  ScalerDev sc{nullptr, nullptr, nullptr, nullptr};
  if (flags & (PF_FLAG_WRAP | PF_FLAG_DENORMALIZE)) {
    if (!scaler || !scaler->lo || !scaler->hi || !scaler->log_mask || !scaler->circ_mask) return PF_ERR_INVALID;
    sc = ScalerDev{scaler->lo, scaler->hi, scaler->log_mask, scaler->circ_mask};
  }
  const float* proj_all = nullptr;
  if (is_proj) proj_all = ctx;
  else if (!dense) {
    PF_TRY(ctxproj_impl(f, st, ctx, E, flags, ws.proj));
    proj_all = ws.proj;
  }
  for (int64_t r0 = 0; r0 < R; r0 += chunk) {
    const int rows = (int)((R - r0) < chunk ? (R - r0) : chunk);
    const float* proj = proj_all;
    CtxMap cm{row2ctx, n_per_ctx, r0};
    if (!proj_all) {
      PF_TRY(ctxproj_impl(f, st, ctx + r0 * f->C, rows, flags, ws.proj));
      proj = ws.proj;
      cm = CtxMap{nullptr, 1, 0};
    }
    const float* zc;
    if (z) zc = z + r0 * D;
    else {
      philox_normal_kernel<<<cdiv(rows, 128), 128, 0, st>>>(ws.zbuf, D, rows, r0, seed);
      PF_LAUNCH_CHECK(f);
      zc = ws.zbuf;
    }
    float* ldc = logdet ? logdet + r0 : ws.ldbuf;
    const float* yin = zc;
    float* ucur = ws.u0;
    for (int l = L - 1; l >= 0; --l) {
      const int reversed = (l != L - 1);
      const float* pl = proj + (size_t)l * 3 * H;
      const float* Wf = f->Wf + (size_t)l * f->PLD * H;
      const float* bfl = f->bf + (size_t)l * f->PLD;
      for (int d = 1; d <= D; ++d) {
        // spline parameters of dim d-1 from hidden groups of degree < d
        PF_TRY((sgemm_nt<EpiBias, A_PLAIN, 64>(f, st, ws.a[4], H, Wf + (size_t)(d - 1) * MULT * H, H, rows, MULT,
                                               f->off[d - 1], EpiBias{ws.pd, UF, bfl + (d - 1) * MULT, 0})));
        PF_NB_DISPATCH(f->NB, PF_TRY(launch_spline_inv<NB_>(f, st, yin, reversed, ws.pd, d - 1, rows, ucur, ldc,
                                                            !(l == L - 1 && d == 1))));
        if (d > G) continue;
        // hidden group of degree d through all five stages (needs x_1..x_d and groups <= d)
        const int u0 = f->off[d - 1], nu = f->off[d] - f->off[d - 1], K = f->off[d];
        if (nu == 0) continue;
        {
          int64_t n = (int64_t)rows * nu;
          initial_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(ucur, D, 0, D, f->W0 + (size_t)l * H * f->DP,
                                                                     f->DP, f->b0 + (size_t)l * H, pl, ldproj, cm, rows,
                                                                     u0, nu, d, ws.a[0], H);
          PF_LAUNCH_CHECK(f);
        }
        for (int k = 0; k < 2; ++k) {
          const float* Wa = f->Wa + ((size_t)l * 2 + k) * H * H + (size_t)u0 * H;
          const float* Wb = f->Wb + ((size_t)l * 2 + k) * H * H + (size_t)u0 * H;
          const float* ba = f->ba + ((size_t)l * 2 + k) * H;
          const float* bb = f->bb + ((size_t)l * 2 + k) * H;
          float* hin = ws.a[2 * k];          // raw h entering the block (a0 or a2)
          float* umid = ws.a[2 * k + 1];     // raw u (a1 or a3)
          float* hout = ws.a[2 * k + 2];     // raw h leaving the block (a2 or a4)
          PF_TRY((sgemm_nt<EpiBias, A_RELU, 32>(f, st, hin, H, Wa, H, rows, nu, K, EpiBias{umid, H, ba, u0})));
          PF_TRY((sgemm_nt<EpiGluResidual, A_RELU, 32>(
              f, st, umid, H, Wb, H, rows, nu, K,
              EpiGluResidual{hin, hout, nullptr, H, bb, u0, pl + (size_t)(k + 1) * H, ldproj, cm})));
        }
      }
      yin = ucur;
      ucur = (ucur == ws.u0) ? ws.u1 : ws.u0;
    }
    // yin now holds the last MAF's output; the trailing ReversePermutation is folded into the finish read
    finish_inverse_kernel<<<cdiv(rows, 128), 128, 0, st>>>(yin, D, rows, flags, f->desc.clamp_bound, sc, x + r0 * D,
                                                          railed ? railed + r0 : nullptr);
    PF_LAUNCH_CHECK(f);
  }
  return PF_OK;
}

extern "C" pf_status pf_flow_inverse(pf_flow* f, const float* z, const float* ctx, const int64_t* row2ctx,
                                     int64_t n_per_ctx, int64_t R, int64_t E, float* x, float* logdet, uint32_t flags,
                                     const pf_scaler* scaler, void* workspace, size_t workspace_bytes, void* stream) {
  PF_TRY(check_call(f, R, E, row2ctx, n_per_ctx));
  f->launches = 0;
  if (R == 0) return PF_OK;
  if (!z || !ctx || !x) return PF_ERR_INVALID;
  PF_CUDA(cudaSetDevice(f->desc.device));
  if (f->desc.tier == PF_TIER_BF16)
    return pf_tc_inverse(f, (cudaStream_t)stream, z, 0, ctx, row2ctx, n_per_ctx, R, E, x, logdet, nullptr, flags, scaler,
                         workspace, workspace_bytes);
  return inverse_impl(f, (cudaStream_t)stream, z, 0, ctx, row2ctx, n_per_ctx, R, E, x, logdet, nullptr, flags, scaler,
                      workspace, workspace_bytes);
}

extern "C" pf_status pf_sample_fused(pf_flow* f, const float* ctx, const pf_scaler* scaler, const float* z,
                                     uint64_t seed, int64_t E, int64_t n, float* theta, float* logdet, uint8_t* railed,
                                     uint32_t flags, void* workspace, size_t workspace_bytes, void* stream) {
  if (!f || n < 1 || E < 0) return PF_ERR_INVALID;
  const int64_t R = E * n;
  PF_TRY(check_call(f, R, E, nullptr, n));
  f->launches = 0;
  if (R == 0) return PF_OK;
  if (!ctx || !theta) return PF_ERR_INVALID;
  PF_CUDA(cudaSetDevice(f->desc.device));
  if (f->desc.tier == PF_TIER_BF16)
    return pf_tc_inverse(f, (cudaStream_t)stream, z, seed, ctx, nullptr, n, R, E, theta, logdet, railed, flags, scaler,
                         workspace, workspace_bytes);
  return inverse_impl(f, (cudaStream_t)stream, z, seed, ctx, nullptr, n, R, E, theta, logdet, railed, flags, scaler,
                      workspace, workspace_bytes);
}

// ------------------------------------------------------------------------------------------------
// backward of the density direction (A12)
// ------------------------------------------------------------------------------------------------
namespace pf {

struct GradStore {
  float* g;
  __device__ __forceinline__ void operator()(int i, float v) const { g[i] = v; }
};

// thread (dim j, row): d(params) of this (row, dim) -> GP[r, j*MULT + i]; direct d(x) -> GXd[r, j]
template <int NB>
__global__ void __launch_bounds__(256)
spline_bwd_kernel(const float* __restrict__ Xin, const float* __restrict__ P, int PLD, int D, float B, int R,
                  const float* __restrict__ GY, const float* __restrict__ GL, float* __restrict__ GP,
                  float* __restrict__ GXd) {
  chain_kernel_entry();
  constexpr int MULT = 3 * NB - 1;
  const int j = threadIdx.x;
  const int r = blockIdx.x * 16 + threadIdx.y;
  if (r >= R || j >= D) return;
  float q[MULT];
  const float* p = P + (int64_t)r * PLD + j * MULT;
#pragma unroll
  for (int i = 0; i < MULT; ++i) q[i] = p[i];
  float x = Xin[(int64_t)r * D + (D - 1 - j)];
  float gy = GY ? GY[(int64_t)r * D + j] : 0.f;
  float gl = GL ? GL[r] : 0.f;
  GXd[(int64_t)r * D + j] =
      rqs_forward_backward<NB>(RegParams{q}, x, B, gy, gl, GradStore{GP + (int64_t)r * PLD + j * MULT});
}

// dst[map(c)] += sum_r src[r, c]   (bias gradients); grid (ceil(ncols/32), slabs), block (32, 8)
struct ColMap {
  const int* perm; int H; int sections;  // sections > 0: c -> (c / H) selects dst pointer, perm on c % H
  float* dst[3];
  __device__ __forceinline__ float* operator()(int c) const {
    if (sections) {
      float* d = dst[c / H];
      return d ? d + perm[c % H] : nullptr;
    }
    return dst[0] ? dst[0] + (perm ? perm[c] : c) : nullptr;
  }
};
__global__ void colsum_kernel(const float* __restrict__ src, int ld, int R, int ncols, int rows_per_slab, ColMap map) {
  __shared__ float red[8][33];
  int c = blockIdx.x * 32 + threadIdx.x;
  int rbeg = blockIdx.y * rows_per_slab, rend = min(R, rbeg + rows_per_slab);
  float s = 0.f;
  if (c < ncols)
    for (int r = rbeg + threadIdx.y; r < rend; r += 8) s += src[(int64_t)r * ld + c];
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < ncols) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][threadIdx.x];
    float* d = map(c);
    if (d) atomicAdd(d, t);
  }
}

// data-gradient epilogues
struct EpiStore {  // out = acc
  float* out; int ld;
  __device__ __forceinline__ void operator()(int m, int n, float acc) const { out[(int64_t)m * ld + n] = acc; }
};
struct EpiReluMaskStore {  // out = pre > 0 ? acc : 0
  float* out; const float* pre; int ld;
  __device__ __forceinline__ void operator()(int m, int n, float acc) const {
    int64_t i = (int64_t)m * ld + n;
    out[i] = pre[i] > 0.f ? acc : 0.f;
  }
  __device__ __forceinline__ float fetch(int m, int n) const { return pre[(int64_t)m * ld + n]; }
  __device__ __forceinline__ void apply(int m, int n, float acc, float p) const {
    out[(int64_t)m * ld + n] = p > 0.f ? acc : 0.f;
  }
};
struct EpiReluMaskAdd {  // out += pre > 0 ? acc : 0
  float* out; const float* pre; int ld;
  __device__ __forceinline__ void operator()(int m, int n, float acc) const {
    int64_t i = (int64_t)m * ld + n;
    if (pre[i] > 0.f) out[i] += acc;
  }
  __device__ __forceinline__ float2 fetch(int m, int n) const {
    const int64_t i = (int64_t)m * ld + n;
    return make_float2(pre[i], out[i]);
  }
  __device__ __forceinline__ void apply(int m, int n, float acc, float2 t) const {
    if (t.x > 0.f) out[(int64_t)m * ld + n] = t.y + acc;
  }
};
// Data-gradient epilogue fused with the elementwise kernel that used to follow it on the dependency chain:
//   g = acc (pre == NULL: final-layer data gradient)  or  gh + [pre > 0] acc (second Linear of a residual block);  gh = g;
//   then, with t = tvec[ctx(m), n]:  V != NULL: GLU backward of the block below (h_out = h_in + t v, t = sigmoid
//   gate): gv = g t, gproj[m, sec H + n] = g V t (1 - t);   V == NULL: relu(context_layer) backward, t = relu'd
//   context term c0: gproj[m, n] = [t > 0] g.
struct EpiGradFuse {
  float* gh; const float* pre; int ld;
  const float* V; const float* tvec; int64_t ldt; CtxMap cm; float* gv; float* gproj; int ldgp; int sec;
  __device__ __forceinline__ float4 fetch(int m, int n) const {
    const int64_t i = (int64_t)m * ld + n;
    return make_float4(pre ? pre[i] : 1.f, pre ? gh[i] : 0.f, tvec[cm(m) * ldt + n], V ? V[i] : 0.f);
  }
  __device__ __forceinline__ void apply(int m, int n, float acc, float4 f) const {
    const int64_t i = (int64_t)m * ld + n;
    const float g = pre ? (f.x > 0.f ? f.y + acc : f.y) : acc;
    if (!pre || f.x > 0.f) gh[i] = g;
    if (V) {
      gv[i] = g * f.z;
      gproj[(int64_t)m * ldgp + sec * ld + n] = g * f.w * f.z * (1.f - f.z);
    } else {
      gproj[(int64_t)m * ldgp + n] = f.z > 0.f ? g : 0.f;
    }
  }
  __device__ __forceinline__ void operator()(int m, int n, float acc) const { apply(m, n, acc, fetch(m, n)); }
};
struct EpiGxOut {  // Gnext[m, D-1-n] = acc + GXd[m, n]   (undo the layer's ReversePermutation)
  float* out; const float* gxd; int D;
  __device__ __forceinline__ void operator()(int m, int n, float acc) const {
    out[(int64_t)m * D + (D - 1 - n)] = acc + gxd[(int64_t)m * D + n];
  }
  __device__ __forceinline__ float fetch(int m, int n) const { return gxd[(int64_t)m * D + n]; }
  __device__ __forceinline__ void apply(int m, int n, float acc, float g) const { out[(int64_t)m * D + (D - 1 - n)] = acc + g; }
};
struct EpiCtxAtomic {  // gctx[ctx(m), n] += acc
  float* out; int C; CtxMap cm;
  __device__ __forceinline__ void operator()(int m, int n, float acc) const { atomicAdd(out + cm(m) * C + n, acc); }
};
// weight-gradient epilogue: scatter into nflows' own layout, masked entries receive nothing
struct EpiWgrad {
  float* dst; int ld; const int* rperm; const int* cperm; int rule; int D; int MULT;
  __device__ __forceinline__ void operator()(int n1, int n2, float acc) const {
    int r = rperm ? rperm[n1] : n1, c = cperm ? cperm[n2] : n2;
    if (rule_mask(rule, r, c, D, MULT) != 0.f) atomicAdd(dst + (int64_t)r * ld + c, acc);
  }
};
struct EpiWgradCtx {  // rows = [3][H] sections of the three context Linear layers
  float* dst[3]; int C; int H; const int* perm;
  __device__ __forceinline__ void operator()(int n1, int n2, float acc) const {
    float* d = dst[n1 / H];
    if (d) atomicAdd(d + (int64_t)perm[n1 % H] * C + n2, acc);
  }
};
struct LoadRev {  // X[m, D-1-n]
  const float* p; int D;
  __device__ __forceinline__ float operator()(int m, int n) const { return p[(int64_t)m * D + (D - 1 - n)]; }
};
struct LoadCtx {
  const float* ctx; int C; CtxMap cm;
  __device__ __forceinline__ float operator()(int m, int n) const { return ctx[cm(m) * C + n]; }
};

// Gradient activations of one layer.  Two sets (layer parity): the weight / bias / context gradients of layer l
// (none of which is on the layer-to-layer dependency chain) run on the handle's side stream while the main
// stream already works on layer l-1, so the set of layer l must survive until layer l-2 starts.
struct BwdSet { float *gp, *gh, *gu[2], *gv[2], *gproj; };
struct BwdWs {
  BwdSet set[2];
  float *gxd, *ga, *gb;
  size_t bytes; bool ok;
  BwdWs(const pf_flow* f, void* base, size_t cap, int64_t R) {
    Carver c(base, base ? cap : (size_t)-1 / 2);
    for (int i = 0; i < 2; ++i) {
      BwdSet& b = set[i];
      b.gp = c.f32((size_t)R * f->PLD); b.gh = c.f32((size_t)R * f->H);
      for (int k = 0; k < 2; ++k) { b.gu[k] = c.f32((size_t)R * f->H); b.gv[k] = c.f32((size_t)R * f->H); }
      b.gproj = c.f32((size_t)R * 3 * f->H);
    }
    gxd = c.f32((size_t)R * f->D);
    ga = c.f32((size_t)R * f->D); gb = c.f32((size_t)R * f->D);
    bytes = c.used; ok = c.ok;
  }
};

static pf_status colsum(pf_flow* f, cudaStream_t st, const float* src, int ld, int R, int ncols, const ColMap& map) {
  int slabs = cdiv(R, 64);
  if (slabs > 4 * f->sm_count) slabs = 4 * f->sm_count;
  int rps = cdiv(R, slabs);
  dim3 grid(cdiv(ncols, 32), cdiv(R, rps)), block(32, 8);
  colsum_kernel<<<grid, block, 0, st>>>(src, ld, R, ncols, rps, map);
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}

template <int NB>
static pf_status launch_spline_bwd(pf_flow* f, cudaStream_t st, const float* Xin, const float* P, int R,
                                   const float* GY, const float* GL, float* GP, float* GXd) {
  dim3 block(16, 16);
  return launch_chain(f, spline_bwd_kernel<NB>, dim3(cdiv(R, 16)), block, st, Xin, P, f->PLD, f->D, f->desc.tail_bound, R,
                      GY, GL, GP, GXd);
}

}  // namespace pf

extern "C" pf_status pf_flow_backward(pf_flow* f, const void* saved, const float* ctx, const int64_t* row2ctx,
                                      int64_t n_per_ctx, int64_t R64, int64_t E, const float* gz, const float* glogdet,
                                      float* gx, float* gctx, const pf_layer_grads* grads, void* workspace,
                                      size_t workspace_bytes, void* stream) {
  PF_TRY(check_call(f, R64, E, row2ctx, n_per_ctx));
  if (!saved || R64 > 8000000) return PF_ERR_INVALID;
  if ((gctx || (grads && grads[0].ctx_w)) && !ctx) return PF_ERR_INVALID;
  PF_CUDA(cudaSetDevice(f->desc.device));
  cudaStream_t st = (cudaStream_t)stream;
  f->launches = 0;
  if (R64 == 0) return PF_OK;
  const int R = (int)R64, D = f->D, H = f->H, L = f->L, PLD = f->PLD, P = f->P, C = f->C, MULT = f->MULT;
  const int64_t ldproj = (int64_t)L * 3 * H;
  if (fused_train(f, row2ctx, n_per_ctx)) {
    // ---- fused training backward (pf_tc_train.cu): data-gradient chain kernel, weight-gradient kernel, and the context
    // path as three calls over ALL layers at once (the gate / context gradients of every layer sit side by side in gproj)
    TtSaved sv(f, const_cast<void*>(saved), R);
    TtWork ws(f, workspace, R);
    if (!workspace || ws.bytes > workspace_bytes) return PF_ERR_WORKSPACE;
    PF_TRY(pf_tt_backward(f, st, sv, ws, R, gz, glogdet, gx, grads));
    PF_TRY(pf_tt_wgrad(f, st, sv, ws, R, ctx, grads));
    if (gctx) PF_TRY(pf_tt_dctx(f, st, ws, R, gctx));   // d ctx [R, C] = (gate / context gradients) . Wctx, all layers
    return train_tc_check(f, st, "pf_flow_backward (fused)");
  }
  SavedLayout sv(f, const_cast<void*>(saved), R, E);
  BwdWs ws(f, workspace, workspace_bytes, R);
  if (!ws.ok) return PF_ERR_WORKSPACE;
  CtxMap cm{row2ctx, n_per_ctx, 0};
  const int* perm = f->d_perm;

  // `st` carries the dependency chain (spline backward and the data gradients, 7 kernels per layer); the 13 kernels per
  // layer whose results nobody downstream waits for go to the handle's four side streams, forked from and joined back
  // into `st` with events (capturable in a CUDA graph): s0 final-layer weight / bias gradient, s1 / s2 the two residual
  // blocks', s3 initial layer + context path.  All of these kernels are latency-bound at a sharded batch's row count,
  // so running several at a time is what fills the GPU (948 rows, one graph replay: 1 stream 3.95 ms, 2 streams 3.27,
  // 3 streams 2.39 with the tcgen05 GEMMs).
  const bool two = f->side[0] != nullptr && f->train_streams;
  cudaStream_t sx[pf_flow::NSIDE];
  for (int j = 0; j < pf_flow::NSIDE; ++j) sx[j] = two ? f->side[j] : st;
  const float* gy = gz;      // gradient wrt the current layer's output
  float* gnext = ws.ga;
  for (int l = L - 1; l >= 0; --l) {
    const pf_layer_grads* g = grads ? &grads[l] : nullptr;
    const int par = (L - 1 - l) & 1;
    const BwdSet& b = ws.set[par];
    const float* xin = sv.XS + (size_t)l * R * D;
    const float* H0 = sv.a(f, R, l, 0); const float* U1 = sv.a(f, R, l, 1); const float* V1 = sv.a(f, R, l, 2);
    const float* H1 = sv.a(f, R, l, 3); const float* U2 = sv.a(f, R, l, 4); const float* V2 = sv.a(f, R, l, 5);
    const float* H2 = sv.a(f, R, l, 6);
    const float* Pl = sv.P + (size_t)l * R * PLD;
    const float* pl = sv.proj + (size_t)l * 3 * H;
    // this parity's buffers were last read by the side work of layer l+2
    if (two && l + 2 <= L - 1)
      for (int j = 0; j < pf_flow::NSIDE; ++j) PF_CUDA(cudaStreamWaitEvent(st, f->ev_side[par][j], 0));
    // ---- chain ----
    // 1. spline: d(params), direct d(x)
    PF_NB_DISPATCH(f->NB, PF_TRY(launch_spline_bwd<NB_>(f, st, xin, Pl, R, gy, glogdet, b.gp, ws.gxd)));
    // 2. final layer (+ GLU backward of block 1 in its epilogue)
    PF_TRY((gemm_nt<EpiGradFuse, A_PLAIN, 64>(
        f, st, b.gp, PLD, f->WfT + (size_t)l * H * PLD, PLD, R, H, P,
        EpiGradFuse{b.gh, nullptr, H, V2, pl + (size_t)2 * H, ldproj, cm, b.gv[1], b.gproj, 3 * H, 2})));
    // 3. residual blocks, last first: d relu(u) = gv Wb, then gh += [h > 0] (gu Wa) with, in the same epilogue, the
    //    GLU backward of block 0 (after block 1) or the relu(context_layer) backward (after block 0)
    for (int k = 1; k >= 0; --k) {
      const float* Hin = k ? H1 : H0;
      const float* U = k ? U2 : U1;
      PF_TRY((gemm_nt<EpiReluMaskStore, A_PLAIN, 64>(f, st, b.gv[k], H, f->WbT + ((size_t)l * 2 + k) * H * H, H, R, H, H,
                                                     EpiReluMaskStore{b.gu[k], U, H})));
      const EpiGradFuse epi = k ? EpiGradFuse{b.gh, Hin, H, V1, pl + (size_t)H, ldproj, cm, b.gv[0], b.gproj, 3 * H, 1}
                                : EpiGradFuse{b.gh, Hin, H, nullptr, pl, ldproj, cm, nullptr, b.gproj, 3 * H, 0};
      PF_TRY((gemm_nt<EpiGradFuse, A_PLAIN, 64>(f, st, b.gu[k], H, f->WaT + ((size_t)l * 2 + k) * H * H, H, R, H, H, epi)));
    }
    float* gout = (l == 0) ? gx : gnext;
    if (gout)
      PF_TRY((gemm_nt<EpiGxOut, A_PLAIN, 32>(f, st, b.gh, H, f->W0T + (size_t)l * f->DP * H, H, R, D, H,
                                             EpiGxOut{gout, ws.gxd, D})));
    if (two) {
      PF_CUDA(cudaEventRecord(f->ev_main[par], st));
      for (int j = 0; j < pf_flow::NSIDE; ++j) PF_CUDA(cudaStreamWaitEvent(sx[j], f->ev_main[par], 0));
    }
    // ---- off the chain: weight, bias and context gradients of this layer ----
    if (g && g->final_w)
      PF_TRY((gemm_tn(f, sx[0], LoadPlain{b.gp, PLD}, LoadPlain{H2, H}, R, P, H,
                      EpiWgrad{g->final_w, H, nullptr, perm, RULE_FINAL, D, MULT})));
    if (g && g->final_b) PF_TRY(colsum(f, sx[0], b.gp, PLD, R, P, ColMap{nullptr, H, 0, {g->final_b, nullptr, nullptr}}));
    for (int k = 1; k >= 0; --k) {
      const float* Hin = k ? H1 : H0;
      const float* U = k ? U2 : U1;
      if (g && g->blk_lin_w[k][1])
        PF_TRY((gemm_tn(f, sx[2 - k], LoadPlain{b.gv[k], H}, LoadRelu{U, H}, R, H, H,
                        EpiWgrad{g->blk_lin_w[k][1], H, perm, perm, RULE_HIDDEN, D, MULT})));
      if (g && g->blk_lin_b[k][1])
        PF_TRY(colsum(f, sx[2 - k], b.gv[k], H, R, H, ColMap{perm, H, 0, {g->blk_lin_b[k][1], nullptr, nullptr}}));
      if (g && g->blk_lin_w[k][0])
        PF_TRY((gemm_tn(f, sx[2 - k], LoadPlain{b.gu[k], H}, LoadRelu{Hin, H}, R, H, H,
                        EpiWgrad{g->blk_lin_w[k][0], H, perm, perm, RULE_HIDDEN, D, MULT})));
      if (g && g->blk_lin_b[k][0])
        PF_TRY(colsum(f, sx[2 - k], b.gu[k], H, R, H, ColMap{perm, H, 0, {g->blk_lin_b[k][0], nullptr, nullptr}}));
    }
    if (g && g->initial_w)
      PF_TRY((gemm_tn(f, sx[3], LoadPlain{b.gh, H}, LoadRev{xin, D}, R, H, D,
                      EpiWgrad{g->initial_w, D, perm, nullptr, RULE_INITIAL, D, MULT})));
    if (g && g->initial_b) PF_TRY(colsum(f, sx[3], b.gh, H, R, H, ColMap{perm, H, 0, {g->initial_b, nullptr, nullptr}}));
    // context path: d(ctx) and the three context Linear layers
    if (gctx)
      PF_TRY((gemm_nt<EpiCtxAtomic, A_PLAIN, 64>(f, sx[3], b.gproj, 3 * H, f->WctxT + (size_t)l * 3 * H, (int)ldproj, R,
                                                 C, 3 * H, EpiCtxAtomic{gctx, C, cm})));
    if (g && (g->ctx_w || g->blk_ctx_w[0] || g->blk_ctx_w[1]))
      PF_TRY((gemm_tn(f, sx[3], LoadPlain{b.gproj, 3 * H}, LoadCtx{ctx, C, cm}, R, 3 * H, C,
                      EpiWgradCtx{{g->ctx_w, g->blk_ctx_w[0], g->blk_ctx_w[1]}, C, H, perm})));
    if (g && (g->ctx_b || g->blk_ctx_b[0] || g->blk_ctx_b[1]))
      PF_TRY(colsum(f, sx[3], b.gproj, 3 * H, R, 3 * H, ColMap{perm, H, 3, {g->ctx_b, g->blk_ctx_b[0], g->blk_ctx_b[1]}}));
    if (two)
      for (int j = 0; j < pf_flow::NSIDE; ++j) PF_CUDA(cudaEventRecord(f->ev_side[par][j], sx[j]));
    gy = gnext;
    gnext = (gnext == ws.ga) ? ws.gb : ws.ga;
  }
  if (two) {   // join: everything the side stream produced is ordered before what follows on `st`
    for (int i = 0; i < (L > 1 ? 2 : 1); ++i)
      for (int j = 0; j < pf_flow::NSIDE; ++j) PF_CUDA(cudaStreamWaitEvent(st, f->ev_side[i][j], 0));
  }
  return train_tc_check(f, st, "pf_flow_backward");
}

// ------------------------------------------------------------------------------------------------
// sizes
// ------------------------------------------------------------------------------------------------
extern "C" size_t pf_flow_workspace_bytes(const pf_flow* f, int32_t op, int64_t R, int64_t E) {
  if (!f || R < 0 || E < 0) return 0;
  const int64_t chunk = R < PF_ROW_CHUNK ? R : PF_ROW_CHUNK;
  // worst case over addressing modes: dense needs `chunk` projection rows, broadcast needs E
  const int64_t prow = E > chunk ? E : chunk;
  switch (op) {
    case PF_OP_CTXPROJ: return 0;
    case PF_OP_FORWARD: {  // + full-length z / logdet scratch for the saved (training) form
      size_t b = FwdWs(f, nullptr, 0, chunk, prow, true).bytes + pad256((size_t)R * f->D) + pad256((size_t)R) + 256;
      if (f->desc.tier == PF_TIER_BF16 || f->desc.tier == PF_TIER_BF16X3) {
        size_t t = (pf_tc_workspace_bytes(f, R, E) + 255) / 256 * 256 + pad256((size_t)R * f->D) + pad256((size_t)R) + 256;
        if (t > b) b = t;
      }
      return b;
    }
    case PF_OP_INVERSE:
    case PF_OP_SAMPLE:
      if (f->desc.tier == PF_TIER_BF16) return pf_tc_workspace_bytes(f, R, E);
      return InvWs(f, nullptr, 0, chunk, prow).bytes + 256;
    case PF_OP_BACKWARD: {
      size_t b = BwdWs(f, nullptr, 0, R).bytes + 256;
      if (f->train_fused && f->train && f->dense) {
        const size_t t = TtWork(f, nullptr, R).bytes + 256;
        if (t > b) b = t;
      }
      return b;
    }
  }
  return 0;
}

// ------------------------------------------------------------------------------------------------
// serving epilogue: physical-units samples + log-density (pipeline.py:57-76, 180, 184-186)
// ------------------------------------------------------------------------------------------------
namespace pf {
__global__ void physical_logq_kernel(const float* __restrict__ Y, const float* __restrict__ nll, ScalerDev sc, int D,
                                     int64_t R, float* __restrict__ theta, float* __restrict__ logq, uint32_t flags) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R) return;
  float v[PF_MAX_D];
  float jac = 0.f, slog = 0.f;
#pragma unroll
  for (int j = 0; j < PF_MAX_D; ++j) {
    if (j >= D) break;
    const float lo = sc.lo[j], hi = sc.hi[j];
    const float y = fminf(fmaxf(Y[r * D + j], -1.f), 1.f);
    float x = (y + 1.f) / 2.f * (hi - lo) + lo;
    if (sc.log_mask[j]) {
      x = expf(x);
      slog += logf(fmaxf(x, 1e-6f));
    }
    jac += 0.6931471805599453f - logf(hi - lo);
    v[j] = x;
  }
  if (logq) logq[r] = -nll[r] + jac - slog;
  if (theta) {
    if ((flags & PF_FLAG_SORT_MASSES) && D >= 2 && v[1] > v[0]) { float t = v[0]; v[0] = v[1]; v[1] = t; }
#pragma unroll
    for (int j = 0; j < PF_MAX_D; ++j) {
      if (j >= D) break;
      theta[r * D + j] = v[j];
    }
  }
}
}  // namespace pf

extern "C" pf_status pf_physical_logq(const float* y, const float* nll, const pf_scaler* scaler, int32_t D, int64_t R,
                                      float* theta, float* logq, uint32_t flags, int32_t device, void* stream) {
  if (!y || !scaler || !scaler->lo || !scaler->hi || !scaler->log_mask || D < 1 || D > PF_MAX_D || R < 0)
    return PF_ERR_INVALID;
  if (logq && !nll) return PF_ERR_INVALID;
  if (R == 0) return PF_OK;
  PF_CUDA(cudaSetDevice(device));
  ScalerDev sc{scaler->lo, scaler->hi, scaler->log_mask, scaler->circ_mask};
  physical_logq_kernel<<<(unsigned)((R + 255) / 256), 256, 0, (cudaStream_t)stream>>>(y, nll, sc, D, R, theta, logq, flags);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return pf::cuda_fail(e, "physical_logq launch", __FILE__, __LINE__);
  return PF_OK;
}
