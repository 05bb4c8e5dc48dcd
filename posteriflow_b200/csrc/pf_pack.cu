// Handle lifetime + weight packing (replaces nflows' per-call `weight * mask`).
#include <vector>

#include <stdlib.h>

#include "pf_common.cuh"

namespace pf {
thread_local char g_cuda_err[256] = "";

__device__ __forceinline__ float src_masked(float v, int rule, int r, int c, int D, int MULT) {
  return rule == RULE_NONE ? v : v * rule_mask(rule, r, c, D, MULT);
}

// One packing job: dst[(transpose ? c : r) * dst_ld + (transpose ? r : c)] = src[rp, cp] * mask_rule(rp, cp)
// with rp = rperm ? perm[r] : r, cp = cperm ? perm[c] : c  (r < rows, c < cols).  A vector is a 1 x n matrix.
// ALL jobs of a flow (27 per layer) run in ONE launch: blockIdx.y = job, grid-stride over its elements.
struct PackJob {
  float* dst;
  const float* src;
  int dst_ld, src_ld, rows, cols;
  uint8_t rule, rperm, cperm, transpose;
  uint32_t zero;   // explicit tail padding: tables are compared with memcmp
};
constexpr int PACK_BLOCKS_PER_JOB = 24;

__global__ void __launch_bounds__(256)
pack_jobs_kernel(const PackJob* __restrict__ jobs, const int* __restrict__ perm, int D, int MULT) {
  const PackJob j = jobs[blockIdx.y];
  const int n = j.rows * j.cols;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    // walk the DESTINATION in its own row-major order when transposing, so the stores coalesce
    int r, c;
    if (j.transpose) { c = i / j.rows; r = i % j.rows; } else { r = i / j.cols; c = i % j.cols; }
    const int rp = j.rperm ? perm[r] : r;
    const int cp = j.cperm ? perm[c] : c;
    const float v = src_masked(j.src[(int64_t)rp * j.src_ld + cp], j.rule, rp, cp, D, MULT);
    if (j.transpose) j.dst[(int64_t)c * j.dst_ld + r] = v;
    else j.dst[(int64_t)r * j.dst_ld + c] = v;
  }
}

__global__ void check_mask_kernel(const float* __restrict__ mask, int rule, int rows, int cols, int D, int MULT,
                                  int* __restrict__ flag) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)rows * cols) return;
  int r = (int)(i / cols), c = (int)(i % cols);
  if (mask[i] != rule_mask(rule, r, c, D, MULT)) atomicOr(flag, 1);
}

// host-side state of the packer: the job table of the last pack (re-uploaded only when a pointer changed)
struct PackState {
  std::vector<PackJob> jobs;
  PackJob* d_jobs;
  size_t d_cap;
};

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
  if (d->tier != PF_TIER_FP32 && d->tier != PF_TIER_BF16 && d->tier != PF_TIER_BF16X3) return PF_ERR_INVALID;
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
    f->train_tc = (d->tier == PF_TIER_BF16 || d->tier == PF_TIER_BF16X3) && !(tt && tt[0] == '0');
    f->train_debug = dbg && dbg[0] == '1';
    const char* mr = getenv("PF_TRAIN_TC_MIN_ROWS");
    f->train_tc_min_rows = mr ? atoi(mr) : 512;
    const char* ts = getenv("PF_TRAIN_STREAMS");
    f->train_streams = !(ts && ts[0] == '0');
    const char* pd = getenv("PF_TRAIN_PDL");
    f->train_pdl = !(pd && pd[0] == '0');
    const char* fu = getenv("PF_TRAIN_FUSED");
    f->train_fused = !(fu && fu[0] == '0');
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
  if (e == cudaSuccess) e = cudaHostAlloc((void**)&f->h_err, sizeof(int), cudaHostAllocMapped);
  if (e == cudaSuccess) {
    *f->h_err = 0;
    e = cudaHostGetDevicePointer((void**)&f->tg_err, (void*)f->h_err, 0);
  }
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
  if (f->h_err) cudaFreeHost((void*)f->h_err);
  for (int i = 0; i < 2; ++i) {
    if (f->ev_main[i]) cudaEventDestroy(f->ev_main[i]);
    for (int j = 0; j < pf_flow::NSIDE; ++j)
      if (f->ev_side[i][j]) cudaEventDestroy(f->ev_side[i][j]);
  }
  for (int j = 0; j < pf_flow::NSIDE; ++j)
    if (f->side[j]) cudaStreamDestroy(f->side[j]);
  if (f->pack_state) {
    PackState* ps = reinterpret_cast<PackState*>(f->pack_state);
    if (ps->d_jobs) cudaFree(ps->d_jobs);
    delete ps;
  }
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
  if (!f->arena_zeroed) {   // pad entries (DP > D, PLD > P) are never written by a job: zero them once
    PF_CUDA(cudaMemsetAsync(f->arena, 0, f->arena_floats * sizeof(float), st));
    f->arena_zeroed = true;
  }
  PackState* ps = reinterpret_cast<PackState*>(f->pack_state);
  if (!ps) {
    ps = new PackState();
    ps->d_jobs = nullptr;
    ps->d_cap = 0;
    f->pack_state = ps;
  }
  std::vector<PackJob> jobs;
  jobs.reserve((size_t)L * 27);
  auto mat = [&](float* dst, int dst_ld, const float* src, int src_ld, int rule, int rows, int cols, bool rperm,
                 bool cperm, bool transpose) {
    jobs.push_back(PackJob{dst, src, dst_ld, src_ld, rows, cols, (uint8_t)rule, (uint8_t)rperm, (uint8_t)cperm,
                           (uint8_t)transpose, 0u});
  };
  auto vec = [&](float* dst, const float* src, int n, bool perm) { mat(dst, n, src, n, RULE_NONE, 1, n, false, perm, false); };
  bool any_mask = false;
  for (int l = 0; l < L; ++l) {
    const pf_layer_ptrs& w = layers[l];
    if (!w.initial_w || !w.initial_b || !w.ctx_w || !w.ctx_b || !w.final_w || !w.final_b) return PF_ERR_INVALID;
    any_mask = any_mask || w.initial_mask || w.final_mask;
    // initial layer [H][DP] sorted rows (+ transposed copy for d(x))
    mat(f->W0 + (size_t)l * H * DP, DP, w.initial_w, D, RULE_INITIAL, H, D, true, false, false);
    mat(f->W0T + (size_t)l * DP * H, H, w.initial_w, D, RULE_INITIAL, H, D, true, false, true);
    vec(f->b0 + (size_t)l * H, w.initial_b, H, true);
    // context layers: section 0 = MADE.context_layer, 1|2 = blocks.k.context_layer
    const float* cw[3] = {w.ctx_w, w.blk_ctx_w[0], w.blk_ctx_w[1]};
    const float* cb[3] = {w.ctx_b, w.blk_ctx_b[0], w.blk_ctx_b[1]};
    for (int s = 0; s < 3; ++s) {
      if (!cw[s] || !cb[s]) return PF_ERR_INVALID;
      mat(f->Wctx + ((size_t)l * 3 + s) * H * C, C, cw[s], C, RULE_NONE, H, C, true, false, false);
      vec(f->bctx + ((size_t)l * 3 + s) * H, cb[s], H, true);
      mat(f->WctxT + ((size_t)l * 3 + s) * H, L * 3 * H, cw[s], C, RULE_NONE, H, C, true, false, true);   // [C][L*3H]
    }
    for (int k = 0; k < 2; ++k) {
      for (int j = 0; j < 2; ++j) {
        const float* lw = w.blk_lin_w[k][j];
        const float* lb = w.blk_lin_b[k][j];
        if (!lw || !lb) return PF_ERR_INVALID;
        any_mask = any_mask || w.blk_lin_mask[k][j];
        float* dstW = (j == 0 ? f->Wa : f->Wb) + ((size_t)l * 2 + k) * H * H;
        float* dstT = (j == 0 ? f->WaT : f->WbT) + ((size_t)l * 2 + k) * H * H;
        float* dstb = (j == 0 ? f->ba : f->bb) + ((size_t)l * 2 + k) * H;
        mat(dstW, H, lw, H, RULE_HIDDEN, H, H, true, true, false);
        mat(dstT, H, lw, H, RULE_HIDDEN, H, H, true, true, true);
        vec(dstb, lb, H, true);
      }
    }
    mat(f->Wf + (size_t)l * PLD * H, H, w.final_w, H, RULE_FINAL, P, H, false, true, false);
    mat(f->WfT + (size_t)l * H * PLD, PLD, w.final_w, H, RULE_FINAL, P, H, false, true, true);
    vec(f->bf + (size_t)l * PLD, w.final_b, P, false);
  }
  const bool same = jobs.size() == ps->jobs.size() &&
                    memcmp(jobs.data(), ps->jobs.data(), jobs.size() * sizeof(PackJob)) == 0;
  if (!same) {
    // the table moves to the device only when a source pointer changed (first pack, .to(device), load_state_dict
    // into new storage): an optimizer step updates the weights in place and re-packs with the cached table, which
    // is what lets the re-pack sit inside a captured CUDA graph
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cap) == cudaSuccess && cap != cudaStreamCaptureStatusNone) return PF_ERR_INVALID;
    if (jobs.size() > ps->d_cap) {
      if (ps->d_jobs) cudaFree(ps->d_jobs);
      ps->d_jobs = nullptr;
      ps->d_cap = 0;
      PF_CUDA(cudaMalloc(&ps->d_jobs, jobs.size() * sizeof(PackJob)));
      ps->d_cap = jobs.size();
    }
    PF_CUDA(cudaStreamSynchronize(st));   // an earlier pack on this stream may still read the old table
    PF_CUDA(cudaMemcpy(ps->d_jobs, jobs.data(), jobs.size() * sizeof(PackJob), cudaMemcpyHostToDevice));
    ps->jobs = jobs;
  }
  if (any_mask) {
    PF_CUDA(cudaMemsetAsync(f->d_flag, 0, sizeof(int), st));
    for (int l = 0; l < L; ++l) {
      const pf_layer_ptrs& w = layers[l];
      PF_TRY(check_mask(f, st, w.initial_mask, RULE_INITIAL, H, D));
      PF_TRY(check_mask(f, st, w.final_mask, RULE_FINAL, P, H));
      for (int k = 0; k < 2; ++k)
        for (int j = 0; j < 2; ++j) PF_TRY(check_mask(f, st, w.blk_lin_mask[k][j], RULE_HIDDEN, H, H));
    }
  }
  {
    dim3 grid(PACK_BLOCKS_PER_JOB, (unsigned)jobs.size());
    pack_jobs_kernel<<<grid, 256, 0, st>>>(ps->d_jobs, f->d_perm, D, f->MULT);
    PF_LAUNCH_CHECK(f);
  }
  PF_TRY(pf_tc_pack(f, st));
  PF_TRY(pf_dn_pack(f, st));
  if (any_mask) {   // only a caller that hands in mask buffers pays the round trip that validates them
    int flag = 0;
    PF_CUDA(cudaMemcpyAsync(&flag, f->d_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
    PF_CUDA(cudaStreamSynchronize(st));
    if (flag) return PF_ERR_MASK;
  }
  f->packed = true;
  return PF_OK;
}

pf_status pf_check_flag(pf_flow* f, const char* where) {
  if (!f || !f->h_err) return PF_OK;
  const int code = *f->h_err;
  if (!code) return PF_OK;
  *f->h_err = 0;
  snprintf(pf::g_cuda_err, sizeof(pf::g_cuda_err),
           "%s: a tensor-core kernel of this handle hit a barrier timeout (code %d); rows of the affected tiles were "
           "written as NaN", where, code);
  return PF_ERR_CUDA;
}

extern "C" pf_status pf_flow_check(pf_flow* f, void* stream) {
  if (!f) return PF_ERR_INVALID;
  PF_CUDA(cudaSetDevice(f->desc.device));
  PF_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return pf_check_flag(f, "pf_flow_check");
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
