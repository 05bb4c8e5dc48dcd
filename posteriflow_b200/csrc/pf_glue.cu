// Glue of the fused `batch_nll` step (N1 of SURVEY.md 8f; experiments/train_lean_npe.py:108-127) around the two flow calls,
// in the "slot" form of posteriflow_b200/train.py: every (event, rank) slot is a flow row with weight [rank < nsig] / count.
// Three small kernels replace ~30 ATen launches (arange / compare / where / normalize's six elementwise ops / broadcast
// copies / sum / clamp / reciprocal / mul / neg / two reductions), which were 0.16 ms of a 1.10 ms step:
//   pf_slot_prepare   params, nsig, context, rank embedding -> normalised x rows, full context rows, slot weights
//   pf_slot_loss      z, logdet, weights -> loss = sum w nll ; d loss / d z = w z ; d loss / d logdet = -w
//   pf_slot_ctx_grad  d loss / d full-context rows -> d / d context (sum over ranks), d / d rank embedding (sum over events)
// Deterministic (no atomics): fixed reduction trees.
#include "pf_common.cuh"

namespace pf {

// One warp per slot row (b, r): x[row, :] = ParamScaler.normalize(params[b, r, :]) (lean_npe.py:100-109) or 0 for an absent
// slot; cfull[row, :] = context[b, :] | emb[r, :] (LeanNPE._full_context, lean_npe.py:296-304); w[row] = [r < nsig[b]] * inv
// with inv = *inv_count if *use_inv > 0 (global signal count under data parallelism) else 1 / max(1, sum_b min(nsig[b], S)).
__global__ void __launch_bounds__(256) slot_prepare_kernel(const float* __restrict__ params, const int64_t* __restrict__ nsig,
                                                           const float* __restrict__ context, const float* __restrict__ emb,
                                                           const float* __restrict__ lo, const float* __restrict__ hi,
                                                           const uint8_t* __restrict__ log_mask, const float* __restrict__ inv_count,
                                                           const float* __restrict__ use_inv, int B, int S, int D, int cd, int C,
                                                           float* __restrict__ x, float* __restrict__ cfull, float* __restrict__ w) {
  __shared__ int s_part[8];
  __shared__ float s_inv;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // the signal count, recomputed by every block (B integers: cheaper than a second launch)
  int c = 0;
  for (int b = tid; b < B; b += 256) {
    const int64_t n = nsig[b];
    c += (int)(n < 0 ? 0 : (n > S ? S : n));
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (lane == 0) s_part[warp] = c;
  __syncthreads();
  if (tid == 0) {
    int tot = 0;
    for (int i = 0; i < 8; ++i) tot += s_part[i];
    const bool ext = use_inv && inv_count && *use_inv > 0.f;
    s_inv = ext ? *inv_count : 1.f / (float)(tot > 1 ? tot : 1);
  }
  __syncthreads();
  const float inv = s_inv;
  const int R = B * S;
  for (int row = blockIdx.x * 8 + warp; row < R; row += gridDim.x * 8) {
    const int b = row / S, r = row % S;
    const bool present = (int64_t)r < nsig[b];
    if (lane < D) {
      float v = 0.f;
      if (present) {
        float p = params[((size_t)b * S + r) * D + lane];
        if (log_mask[lane]) p = logf(fmaxf(p, 1e-6f));
        v = 2.f * (p - lo[lane]) / (hi[lane] - lo[lane]) - 1.f;
        v = fminf(fmaxf(v, -1.f), 1.f);
      }
      x[(size_t)row * D + lane] = v;
    }
    if (lane == 0) w[row] = present ? inv : 0.f;
    float* dst = cfull + (size_t)row * C;
    const float* cs = context + (size_t)b * cd;
    for (int i = lane; i < cd; i += 32) dst[i] = cs[i];
    const float* es = emb + (size_t)r * (C - cd);
    for (int i = lane; i < C - cd; i += 32) dst[cd + i] = es[i];
  }
}

// One block.  nll[row] = 0.5 (|z|^2 + D log 2 pi) - logdet (flows.py:727-779 with log_sigma = 0); loss = sum_row w nll.
__global__ void __launch_bounds__(1024) slot_loss_kernel(const float* __restrict__ z, const float* __restrict__ ld,
                                                         const float* __restrict__ w, int rows, int D, float* __restrict__ loss,
                                                         float* __restrict__ gz, float* __restrict__ gl) {
  __shared__ float s_part[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float cst = (float)D * 1.8378770664093453f;      // log(2 pi)
  float acc = 0.f;
  for (int row = tid; row < rows; row += 1024) {
    const float wr = w[row];
    float s = 0.f;
    for (int j = 0; j < D; ++j) {
      const float v = z[(size_t)row * D + j];
      s = fmaf(v, v, s);
      gz[(size_t)row * D + j] = v * wr;
    }
    gl[row] = -wr;
    const float nll = 0.5f * (s + cst) - ld[row];
    acc += wr != 0.f ? nll * wr : 0.f;                   // an absent slot contributes exactly nothing, whatever its row holds
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) s_part[warp] = acc;
  __syncthreads();
  if (warp == 0) {
    float v = s_part[lane];
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) *loss = v;
  }
}

// blocks [0, B): gcontext[b, :] = sum_r g[b * S + r, :cd]      (backward of the row gather)
// blocks [B, B + S): gemb[r, :] = sum_b g[b * S + r, cd:]      (backward of the rank embedding), fixed tree over b
__global__ void __launch_bounds__(256) slot_ctx_grad_kernel(const float* __restrict__ g, int B, int S, int cd, int C,
                                                            float* __restrict__ gcontext, float* __restrict__ gemb) {
  const int tid = threadIdx.x;
  if ((int)blockIdx.x < B) {
    const int b = blockIdx.x;
    for (int c = tid; c < cd; c += 256) {
      float s = 0.f;
      for (int r = 0; r < S; ++r) s += g[((size_t)b * S + r) * C + c];
      gcontext[(size_t)b * cd + c] = s;
    }
    return;
  }
  __shared__ float s_part[8][33];
  const int r = (int)blockIdx.x - B, ce = C - cd;
  const int col = tid & 31, grp = tid >> 5;
  for (int c0 = 0; c0 < ce; c0 += 32) {
    float s = 0.f;
    if (c0 + col < ce)
      for (int b = grp; b < B; b += 8) s += g[((size_t)b * S + r) * C + cd + c0 + col];
    s_part[grp][col] = s;
    __syncthreads();
    if (grp == 0 && c0 + col < ce) {
      float t = 0.f;
      for (int i = 0; i < 8; ++i) t += s_part[i][col];
      gemb[(size_t)r * ce + c0 + col] = t;
    }
    __syncthreads();
  }
}

}  // namespace pf

extern "C" pf_status pf_slot_prepare(const float* params, const int64_t* nsig, const float* context, const float* rank_emb,
                                     const pf_scaler* scaler, const float* inv_count, const float* use_inv, int32_t B,
                                     int32_t S, int32_t D, int32_t context_dim, int32_t C, float* x, float* cfull, float* w,
                                     int32_t device, void* stream) {
  if (!params || !nsig || !context || !rank_emb || !scaler || !scaler->lo || !scaler->hi || !scaler->log_mask || !x ||
      !cfull || !w)
    return PF_ERR_INVALID;
  if (B < 0 || S < 1 || D < 1 || D > 32 || context_dim < 0 || C <= context_dim) return PF_ERR_INVALID;
  if (B == 0) return PF_OK;
  PF_CUDA(cudaSetDevice(device));
  const int rows = B * S;
  int grid = (rows + 7) / 8;
  if (grid > 1184) grid = 1184;
  pf::slot_prepare_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(params, nsig, context, rank_emb, scaler->lo, scaler->hi,
                                                                 scaler->log_mask, inv_count, use_inv, B, S, D, context_dim, C,
                                                                 x, cfull, w);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return pf::cuda_fail(e, "slot_prepare launch", __FILE__, __LINE__);
  return PF_OK;
}

extern "C" pf_status pf_slot_loss(const float* z, const float* logdet, const float* w, int64_t rows, int32_t D, float* loss,
                                  float* gz, float* glogdet, int32_t device, void* stream) {
  if (!z || !logdet || !w || !loss || !gz || !glogdet || rows < 0 || rows >= (1ll << 31) || D < 1) return PF_ERR_INVALID;
  PF_CUDA(cudaSetDevice(device));
  pf::slot_loss_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(z, logdet, w, (int)rows, D, loss, gz, glogdet);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return pf::cuda_fail(e, "slot_loss launch", __FILE__, __LINE__);
  return PF_OK;
}

extern "C" pf_status pf_slot_ctx_grad(const float* gcfull, int32_t B, int32_t S, int32_t context_dim, int32_t C,
                                      float* gcontext, float* grank_emb, int32_t device, void* stream) {
  if (!gcfull || !gcontext || !grank_emb || B < 0 || S < 1 || context_dim < 0 || C <= context_dim) return PF_ERR_INVALID;
  if (B == 0) return PF_OK;
  PF_CUDA(cudaSetDevice(device));
  pf::slot_ctx_grad_kernel<<<B + S, 256, 0, (cudaStream_t)stream>>>(gcfull, B, S, context_dim, C, gcontext, grank_emb);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return pf::cuda_fail(e, "slot_ctx_grad launch", __FILE__, __LINE__);
  return PF_OK;
}
