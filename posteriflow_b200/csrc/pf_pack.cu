// Handle lifetime + weight packing (replaces nflows' per-call `weight * mask`).
#include <vector>

#include <stdlib.h>

#include "pf_common.cuh"

namespace pf {
thread_local char g_cuda_err[256] = "";

// dst[(transpose ? c : r) * dst_ld + (transpose ? r : c)] = src[rp, cp] * mask(rp, cp)
// with rp = rperm ? rperm[r0 + r] : r0 + r, cp = cperm ? cperm[c] : c   (r < rows, c < cols)
__global__ void pack_matrix_kernel(float* __restrict__ dst, int dst_ld, const float* __restrict__ src, int src_ld,
                                   const float* __restrict__ mask, int rule, int rows, int cols, int r0,
                                   const int* __restrict__ rperm, const int* __restrict__ cperm, int transpose,
                                   int D, int MULT) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)rows * cols) return;
  int r = (int)(i / cols), c = (int)(i % cols);
  int rp = rperm ? rperm[r0 + r] : r0 + r;
  int cp = cperm ? cperm[c] : c;
  float m = mask ? mask[(int64_t)rp * src_ld + cp] : rule_mask(rule, rp, cp, D, MULT);
  float v = src[(int64_t)rp * src_ld + cp] * m;
  if (transpose) dst[(int64_t)c * dst_ld + r] = v;
  else dst[(int64_t)r * dst_ld + c] = v;
}

__global__ void pack_vector_kernel(float* __restrict__ dst, const float* __restrict__ src, int n,
                                   const int* __restrict__ perm) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[perm ? perm[i] : i];
}

__global__ void check_mask_kernel(const float* __restrict__ mask, int rule, int rows, int cols, int D, int MULT,
                                  int* __restrict__ flag) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)rows * cols) return;
  int r = (int)(i / cols), c = (int)(i % cols);
  if (mask[i] != rule_mask(rule, r, c, D, MULT)) atomicOr(flag, 1);
}

static pf_status pack_matrix(pf_flow* f, cudaStream_t st, float* dst, int dst_ld, const float* src, int src_ld,
                             const float* mask, int rule, int rows, int cols, int r0, const int* rperm,
                             const int* cperm, int transpose) {
  if (rows <= 0 || cols <= 0) return PF_OK;
  int64_t n = (int64_t)rows * cols;
  pack_matrix_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(dst, dst_ld, src, src_ld, mask, rule, rows, cols,
                                                                  r0, rperm, cperm, transpose, f->D, f->MULT);
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}

static pf_status pack_vector(pf_flow* f, cudaStream_t st, float* dst, const float* src, int n, const int* perm) {
  pack_vector_kernel<<<(n + 255) / 256, 256, 0, st>>>(dst, src, n, perm);
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}

static pf_status check_mask(pf_flow* f, cudaStream_t st, const float* mask, int rule, int rows, int cols) {
  if (!mask) return PF_OK;
  int64_t n = (int64_t)rows * cols;
  check_mask_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(mask, rule, rows, cols, f->D, f->MULT, f->d_flag);
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}

}  // namespace pf

using namespace pf;

extern "C" pf_status pf_flow_create(const pf_flow_desc* d, pf_flow** out) {
  if (!d || !out) return PF_ERR_INVALID;
  *out = nullptr;
  if (d->features < 2 || d->features > PF_MAX_D || d->num_bins < 2 || d->num_bins > PF_MAX_BINS ||
      d->hidden_features < d->features - 1 || d->hidden_features % 4 != 0 || d->hidden_features > 1024 ||
      d->context_features < 1 || d->num_layers < 1 || d->num_blocks != 2 || !(d->tail_bound > 0.f))
    return PF_ERR_INVALID;
  if (d->tier != PF_TIER_FP32 && d->tier != PF_TIER_BF16) return PF_ERR_INVALID;
  // nflows raises when min_bin_width * num_bins > 1
  if (1e-3 * d->num_bins > 1.0) return PF_ERR_INVALID;
  int ndev = 0;
  PF_CUDA(cudaGetDeviceCount(&ndev));
  if (d->device < 0 || d->device >= ndev) return PF_ERR_INVALID;
  PF_CUDA(cudaSetDevice(d->device));

  pf_flow* f = new pf_flow();
  memset(f, 0, sizeof(*f));
  f->desc = *d;
  f->D = d->features; f->C = d->context_features; f->H = d->hidden_features; f->L = d->num_layers;
  f->NB = d->num_bins; f->MULT = 3 * f->NB - 1; f->P = f->D * f->MULT; f->PLD = round_up(f->P, 4);
  f->DP = round_up(f->D, 4); f->G = f->D - 1;
  cudaDeviceGetAttribute(&f->sm_count, cudaDevAttrMultiProcessorCount, d->device);
  {
    const char* tt = getenv("PF_TRAIN_TC");
    const char* dbg = getenv("PF_TC_DEBUG");
    f->train_tc = d->tier == PF_TIER_BF16 && !(tt && tt[0] == '0');
    f->train_debug = dbg && dbg[0] == '1';
    const char* mr = getenv("PF_TRAIN_TC_MIN_ROWS");
    f->train_tc_min_rows = mr ? atoi(mr) : 512;
    const char* ts = getenv("PF_TRAIN_STREAMS");
    f->train_streams = !(ts && ts[0] == '0');
    const char* pd = getenv("PF_TRAIN_PDL");
    f->train_pdl = !(pd && pd[0] == '0');
  }

  // degree-sorted hidden permutation (stable): slot s <- nflows unit perm[s]
  f->perm_host = new int[f->H];
  int s = 0;
  f->off[0] = 0;
  f->max_group = 0;
  for (int g = 1; g <= f->G; ++g) {
    int n0 = s;
    for (int u = 0; u < f->H; ++u)
      if (hidden_degree(u, f->D) == g) f->perm_host[s++] = u;
    f->off[g] = s;
    if (s - n0 > f->max_group) f->max_group = s - n0;
  }
  for (int g = f->G + 1; g <= PF_MAX_D; ++g) f->off[g] = s;
  if (s != f->H) { delete[] f->perm_host; delete f; return PF_ERR_INVALID; }
  f->UF = round_up(f->MULT, 4);

  const size_t L = f->L, H = f->H, C = f->C, PLD = f->PLD, DP = f->DP;
  size_t n = 0;
  auto take = [&](size_t cnt) { size_t o = n; n += (cnt + 3) / 4 * 4; return o; };
  size_t oW0 = take(L * H * DP), ob0 = take(L * H), oWctx = take(L * 3 * H * C), obctx = take(L * 3 * H);
  size_t oWa = take(L * 2 * H * H), oba = take(L * 2 * H), oWb = take(L * 2 * H * H), obb = take(L * 2 * H);
  size_t oWf = take(L * PLD * H), obf = take(L * PLD);
  size_t oWaT = take(L * 2 * H * H), oWbT = take(L * 2 * H * H), oWfT = take(L * H * PLD);
  size_t oWctxT = take(C * L * 3 * H), oW0T = take(L * DP * H);
  f->arena_floats = n;
  cudaError_t e = cudaMalloc(&f->arena, n * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(&f->d_perm, H * sizeof(int));
  if (e == cudaSuccess) e = cudaMalloc(&f->d_flag, sizeof(int));
  if (e == cudaSuccess) e = cudaMalloc(&f->tg_err, sizeof(int));
  if (e == cudaSuccess) e = cudaMemset(f->tg_err, 0, sizeof(int));
  for (int j = 0; j < pf_flow::NSIDE && e == cudaSuccess; ++j) e = cudaStreamCreateWithFlags(&f->side[j], cudaStreamNonBlocking);
  for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
    e = cudaEventCreateWithFlags(&f->ev_main[i], cudaEventDisableTiming);
    for (int j = 0; j < pf_flow::NSIDE && e == cudaSuccess; ++j)
      e = cudaEventCreateWithFlags(&f->ev_side[i][j], cudaEventDisableTiming);
  }
  if (e == cudaSuccess) e = cudaMemcpy(f->d_perm, f->perm_host, H * sizeof(int), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    pf_status st = cuda_fail(e, "arena allocation", __FILE__, __LINE__);
    pf_flow_destroy(f);
    return st;
  }
  float* a = f->arena;
  f->W0 = a + oW0; f->b0 = a + ob0; f->Wctx = a + oWctx; f->bctx = a + obctx;
  f->Wa = a + oWa; f->ba = a + oba; f->Wb = a + oWb; f->bb = a + obb; f->Wf = a + oWf; f->bf = a + obf;
  f->WaT = a + oWaT; f->WbT = a + oWbT; f->WfT = a + oWfT; f->WctxT = a + oWctxT; f->W0T = a + oW0T;
  *out = f;
  return PF_OK;
}

extern "C" void pf_flow_destroy(pf_flow* f) {
  if (!f) return;
  pf_tc_destroy(f);
  pf_dn_destroy(f);
  if (f->arena) cudaFree(f->arena);
  if (f->d_perm) cudaFree(f->d_perm);
  if (f->d_flag) cudaFree(f->d_flag);
  if (f->tg_err) cudaFree(f->tg_err);
  for (int i = 0; i < 2; ++i) {
    if (f->ev_main[i]) cudaEventDestroy(f->ev_main[i]);
    for (int j = 0; j < pf_flow::NSIDE; ++j)
      if (f->ev_side[i][j]) cudaEventDestroy(f->ev_side[i][j]);
  }
  for (int j = 0; j < pf_flow::NSIDE; ++j)
    if (f->side[j]) cudaStreamDestroy(f->side[j]);
  delete[] f->perm_host;
  delete f;
}

extern "C" pf_status pf_flow_pack(pf_flow* f, const pf_layer_ptrs* layers, void* stream) {
  if (!f || !layers) return PF_ERR_INVALID;
  cudaStream_t st = (cudaStream_t)stream;
  PF_CUDA(cudaSetDevice(f->desc.device));
  f->launches = 0;
  f->packed = false;
  const int L = f->L, H = f->H, C = f->C, D = f->D, P = f->P, PLD = f->PLD, DP = f->DP;
  PF_CUDA(cudaMemsetAsync(f->arena, 0, f->arena_floats * sizeof(float), st));
  PF_CUDA(cudaMemsetAsync(f->d_flag, 0, sizeof(int), st));
  const int* perm = f->d_perm;
  for (int l = 0; l < L; ++l) {
    const pf_layer_ptrs& w = layers[l];
    if (!w.initial_w || !w.initial_b || !w.ctx_w || !w.ctx_b || !w.final_w || !w.final_b) return PF_ERR_INVALID;
    PF_TRY(check_mask(f, st, w.initial_mask, RULE_INITIAL, H, D));
    PF_TRY(check_mask(f, st, w.final_mask, RULE_FINAL, P, H));
    // initial layer [H][DP] sorted rows
    PF_TRY(pack_matrix(f, st, f->W0 + (size_t)l * H * DP, DP, w.initial_w, D, nullptr, RULE_INITIAL, H, D, 0, perm,
                       nullptr, 0));
    PF_TRY(pack_matrix(f, st, f->W0T + (size_t)l * DP * H, H, w.initial_w, D, nullptr, RULE_INITIAL, H, D, 0, perm,
                       nullptr, 1));
    PF_TRY(pack_vector(f, st, f->b0 + (size_t)l * H, w.initial_b, H, perm));
    // context layers: section 0 = MADE.context_layer, 1|2 = blocks.k.context_layer
    const float* cw[3] = {w.ctx_w, w.blk_ctx_w[0], w.blk_ctx_w[1]};
    const float* cb[3] = {w.ctx_b, w.blk_ctx_b[0], w.blk_ctx_b[1]};
    for (int s = 0; s < 3; ++s) {
      if (!cw[s] || !cb[s]) return PF_ERR_INVALID;
      PF_TRY(pack_matrix(f, st, f->Wctx + ((size_t)l * 3 + s) * H * C, C, cw[s], C, nullptr, RULE_NONE, H, C, 0,
                         perm, nullptr, 0));
      PF_TRY(pack_vector(f, st, f->bctx + ((size_t)l * 3 + s) * H, cb[s], H, perm));
      // WctxT [C][L*3H]
      PF_TRY(pack_matrix(f, st, f->WctxT + ((size_t)l * 3 + s) * H, L * 3 * H, cw[s], C, nullptr, RULE_NONE, H, C, 0,
                         perm, nullptr, 1));
    }
    for (int k = 0; k < 2; ++k) {
      for (int j = 0; j < 2; ++j) {
        const float* lw = w.blk_lin_w[k][j];
        const float* lb = w.blk_lin_b[k][j];
        if (!lw || !lb) return PF_ERR_INVALID;
        PF_TRY(check_mask(f, st, w.blk_lin_mask[k][j], RULE_HIDDEN, H, H));
        float* dstW = (j == 0 ? f->Wa : f->Wb) + ((size_t)l * 2 + k) * H * H;
        float* dstT = (j == 0 ? f->WaT : f->WbT) + ((size_t)l * 2 + k) * H * H;
        float* dstb = (j == 0 ? f->ba : f->bb) + ((size_t)l * 2 + k) * H;
        PF_TRY(pack_matrix(f, st, dstW, H, lw, H, nullptr, RULE_HIDDEN, H, H, 0, perm, perm, 0));
        PF_TRY(pack_matrix(f, st, dstT, H, lw, H, nullptr, RULE_HIDDEN, H, H, 0, perm, perm, 1));
        PF_TRY(pack_vector(f, st, dstb, lb, H, perm));
      }
    }
    PF_TRY(pack_matrix(f, st, f->Wf + (size_t)l * PLD * H, H, w.final_w, H, nullptr, RULE_FINAL, P, H, 0, nullptr,
                       perm, 0));
    PF_TRY(pack_matrix(f, st, f->WfT + (size_t)l * H * PLD, PLD, w.final_w, H, nullptr, RULE_FINAL, P, H, 0, nullptr,
                       perm, 1));
    PF_TRY(pack_vector(f, st, f->bf + (size_t)l * PLD, w.final_b, P, nullptr));

  }
  PF_TRY(pf_tc_pack(f, st));
  PF_TRY(pf_dn_pack(f, st));
  int flag = 0;
  PF_CUDA(cudaMemcpyAsync(&flag, f->d_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
  PF_CUDA(cudaStreamSynchronize(st));
  if (flag) return PF_ERR_MASK;
  f->packed = true;
  return PF_OK;
}

extern "C" pf_status pf_flow_hidden_perm(const pf_flow* f, int32_t* perm_host) {
  if (!f || !perm_host) return PF_ERR_INVALID;
  for (int i = 0; i < f->H; ++i) perm_host[i] = f->perm_host[i];
  return PF_OK;
}

extern "C" int64_t pf_flow_last_launch_count(const pf_flow* f) { return f ? f->launches : 0; }

extern "C" const char* pf_status_str(pf_status s) {
  switch (s) {
    case PF_OK: return "ok";
    case PF_ERR_INVALID: return "invalid argument or unsupported shape";
    case PF_ERR_CUDA: return "CUDA runtime error";
    case PF_ERR_NOT_PACKED: return "pf_flow_pack has not been called";
    case PF_ERR_WORKSPACE: return "workspace too small";
    case PF_ERR_MASK: return "mask buffers differ from nflows' deterministic MADE masks";
    case PF_ERR_UNSUPPORTED: return "no kernel for this request in this build";
    default: return "unknown status";
  }
}

extern "C" const char* pf_last_cuda_error(void) { return pf::g_cuda_err; }
extern "C" int32_t pf_abi_version(void) { return PF_ABI_VERSION; }
