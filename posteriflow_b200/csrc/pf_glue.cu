// Glue of the fused `batch_nll` step (N1 of SURVEY.md 8f; experiments/train_lean_npe.py:108-127) around the two flow calls,
// in the "slot" form of posteriflow_b200/train.py: every (event, rank) slot is a flow row with weight [rank < nsig] / count.
// Three small kernels replace ~30 ATen launches (arange / compare / where / normalize's six elementwise ops / broadcast
// copies / sum / clamp / reciprocal / mul / neg / two reductions), which were 0.16 ms of a 1.10 ms step:
//   pf_slot_prepare   params, nsig, context, rank embedding -> normalised x rows, full context rows, slot weights
//   pf_slot_loss      z, logdet, weights -> loss = sum w nll ; d loss / d z = w z ; d loss / d logdet = -w
//   pf_slot_ctx_grad  d loss / d full-context rows -> d / d context (sum over ranks), d / d rank embedding (sum over events)
//   pf_rows_prepare / pf_rows_ctx_grad   the same for the COMPACT form (only the present pairs become rows, event-major; the
//                     (event, rank) of a row comes from a prefix sum of nsig on the device): full batches beyond one wave
// Deterministic (no atomics): fixed reduction trees; the loss and the embedding gradient are reduced across a cluster of
// 8 CTAs through distributed shared memory (one CTA took 47 us for 7.6k rows, the embedding blocks 56 us).
#include "pf_common.cuh"
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

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

// One cluster of 8 CTAs.  nll[row] = 0.5 (|z|^2 + D log 2 pi) - logdet (flows.py:727-779 with log_sigma = 0); loss = sum_row w nll,
// reduced warp -> CTA -> cluster (partials written into CTA 0's shared memory) in a fixed order: deterministic, no scratch.
template <int DC>      // DC > 0: D known at compile time (the row's loads go out together); 0: runtime D
__device__ __forceinline__ float loss_row(const float* __restrict__ z, float* __restrict__ gz, size_t row, int D, float wr) {
  float s = 0.f;
  if (DC > 0) {
    float v[DC > 0 ? DC : 1];
#pragma unroll
    for (int j = 0; j < DC; ++j) v[j] = z[row * DC + j];
#pragma unroll
    for (int j = 0; j < DC; ++j) {
      s = fmaf(v[j], v[j], s);
      gz[row * DC + j] = v[j] * wr;
    }
  } else {
    for (int j = 0; j < D; ++j) {
      const float v = z[row * D + j];
      s = fmaf(v, v, s);
      gz[row * D + j] = v * wr;
    }
  }
  return s;
}

constexpr int GL_CLUSTER = 8;

__global__ void __cluster_dims__(GL_CLUSTER, 1, 1) __launch_bounds__(1024)
slot_loss_kernel(const float* __restrict__ z, const float* __restrict__ ld, const float* __restrict__ w, int rows, int D,
                 float* __restrict__ loss, float* __restrict__ gz, float* __restrict__ gl) {
  __shared__ float s_part[32];
  __shared__ float s_cl[GL_CLUSTER];
  cg::cluster_group cl = cg::this_cluster();
  const int crank = (int)cl.block_rank();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float cst = (float)D * 1.8378770664093453f;      // log(2 pi)
  float acc = 0.f;
  for (int row = crank * 1024 + tid; row < rows; row += GL_CLUSTER * 1024) {
    const float wr = w[row];
    const float ldr = ld[row];
    const float s = D == 11 ? loss_row<11>(z, gz, (size_t)row, D, wr) : loss_row<0>(z, gz, (size_t)row, D, wr);
    gl[row] = -wr;
    const float nll = 0.5f * (s + cst) - ldr;
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
    if (lane == 0) cl.map_shared_rank(s_cl, 0)[crank] = v;
  }
  cl.sync();
  if (crank == 0 && tid == 0) {
    float t = 0.f;
    for (int i = 0; i < GL_CLUSTER; ++i) t += s_cl[i];
    *loss = t;
  }
}

// d loss / d full-context rows -> d / d context (sum over an event's rows) and d / d rank embedding (sum over events).
// offs == nullptr: slot layout (event b owns rows [b S, b S + S); absent slots hold exact zeros); else the compact layout
// (event b owns rows [offs[b], offs[b + 1]), row offs[b] + k is rank k).
// Blocks [0, nctx) (nctx a multiple of 8): four events each, one thread per context column.
// Blocks [nctx, nctx + 8 S): one cluster of 8 CTAs per rank k; warp gw of the cluster's 256 takes events gw, gw + 256, ...
// (lane = embedding column, loads batched four events deep), then warp -> CTA -> cluster sums in a fixed order.
__global__ void __cluster_dims__(GL_CLUSTER, 1, 1) __launch_bounds__(1024)
ctx_grad_kernel(const float* __restrict__ g, const int* __restrict__ offs, int B, int S, int cd, int C, int nctx,
                float* __restrict__ gcontext, float* __restrict__ gemb) {
  const int tid = threadIdx.x;
  if ((int)blockIdx.x < nctx) {
    const int b = blockIdx.x * 4 + (tid >> 8);
    if (b >= B) return;
    const int o = offs ? offs[b] : b * S, n = offs ? offs[b + 1] - o : S;
    for (int c = tid & 255; c < cd; c += 256) {
      float s = 0.f;
      for (int k = 0; k < n; ++k) s += g[((size_t)o + k) * C + c];
      gcontext[(size_t)b * cd + c] = s;
    }
    return;
  }
  __shared__ float s_part[32][33];
  __shared__ float s_cl[GL_CLUSTER][32];
  cg::cluster_group cl = cg::this_cluster();
  const int crank = (int)cl.block_rank();
  const int k = ((int)blockIdx.x - nctx) / GL_CLUSTER, ce = C - cd;
  const int lane = tid & 31, warp = tid >> 5;
  const int gw = crank * 32 + warp;
  constexpr int NW = GL_CLUSTER * 32;
  for (int c0 = 0; c0 < ce; c0 += 32) {
    const bool colok = c0 + lane < ce;
    float s = 0.f;
    for (int b = gw; b < B; b += 4 * NW) {
      int o[4], n[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int bb = b + u * NW;
        n[u] = 0;
        o[u] = 0;
        if (bb < B) {
          o[u] = offs ? offs[bb] : bb * S;
          n[u] = offs ? offs[bb + 1] - o[u] : S;
        }
      }
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = (n[u] > k && colok) ? g[((size_t)o[u] + k) * C + cd + c0 + lane] : 0.f;
#pragma unroll
      for (int u = 0; u < 4; ++u) s += v[u];
    }
    s_part[warp][lane] = s;
    __syncthreads();
    if (warp == 0) {
      float t = 0.f;
      for (int i = 0; i < 32; ++i) t += s_part[i][lane];
      cl.map_shared_rank(&s_cl[0][0], 0)[crank * 32 + lane] = t;
    }
    cl.sync();
    if (crank == 0 && warp == 0 && colok) {
      float t = 0.f;
      for (int i = 0; i < GL_CLUSTER; ++i) t += s_cl[i][lane];
      gemb[(size_t)k * ce + c0 + lane] = t;
    }
    cl.sync();      // s_cl / s_part are rewritten by the next column chunk
  }
}

// ---- compact ("rows") form: only the present (event, rank) pairs become flow rows, event-major ----------------------
// offs[b] = sum_{b' < b} min(nsig[b'], S), offs[B] = number of rows.  One block; B up to 2^20.
__global__ void __launch_bounds__(1024) rows_scan_kernel(const int64_t* __restrict__ nsig, int B, int S, int* __restrict__ offs) {
  __shared__ int s_sum[1024];
  const int tid = threadIdx.x;
  const int per = (B + 1023) / 1024;
  const int b0 = tid * per, b1 = min(B, b0 + per);
  int loc = 0;
  for (int b = b0; b < b1; ++b) {
    const int64_t n = nsig[b];
    loc += (int)(n < 0 ? 0 : (n > S ? S : n));
  }
  s_sum[tid] = loc;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {           // inclusive Hillis-Steele scan of the per-thread sums
    const int v = tid >= o ? s_sum[tid - o] : 0;
    __syncthreads();
    s_sum[tid] += v;
    __syncthreads();
  }
  int run = s_sum[tid] - loc;                      // exclusive prefix of this thread's chunk
  for (int b = b0; b < b1; ++b) {
    offs[b] = run;
    const int64_t n = nsig[b];
    run += (int)(n < 0 ? 0 : (n > S ? S : n));
  }
  if (tid == 1023) offs[B] = s_sum[1023];
}

// One warp per flow row r < cap: r < offs[B] -> (event b, rank k) by binary search in offs; rows beyond get x = 0, w = 0.
__global__ void __launch_bounds__(256) rows_prepare_kernel(const float* __restrict__ params, const int* __restrict__ offs,
                                                           const float* __restrict__ context, const float* __restrict__ emb,
                                                           const float* __restrict__ lo, const float* __restrict__ hi,
                                                           const uint8_t* __restrict__ log_mask, const float* __restrict__ inv_count,
                                                           const float* __restrict__ use_inv, int B, int S, int D, int cd, int C,
                                                           int cap, float* __restrict__ x, float* __restrict__ cfull,
                                                           float* __restrict__ w) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int R = offs[B];
  const bool ext = use_inv && inv_count && *use_inv > 0.f;
  const float inv = ext ? *inv_count : 1.f / (float)(R > 1 ? R : 1);
  for (int row = blockIdx.x * 8 + warp; row < cap; row += gridDim.x * 8) {
    if (row >= R) {
      if (lane < D) x[(size_t)row * D + lane] = 0.f;
      if (lane == 0) w[row] = 0.f;
      continue;
    }
    int a = 0, z = B;                              // largest b with offs[b] <= row (offs is non-decreasing, offs[B] = R > row)
    while (z - a > 1) {
      const int m = (a + z) >> 1;
      if (offs[m] <= row) a = m; else z = m;
    }
    const int b = a, r = row - offs[a];
    if (lane < D) {
      float p = params[((size_t)b * S + r) * D + lane];
      if (log_mask[lane]) p = logf(fmaxf(p, 1e-6f));
      float v = 2.f * (p - lo[lane]) / (hi[lane] - lo[lane]) - 1.f;
      x[(size_t)row * D + lane] = fminf(fmaxf(v, -1.f), 1.f);
    }
    if (lane == 0) w[row] = inv;
    float* dst = cfull + (size_t)row * C;
    const float* cs = context + (size_t)b * cd;
    for (int i = lane; i < cd; i += 32) dst[i] = cs[i];
    const float* es = emb + (size_t)r * (C - cd);
    for (int i = lane; i < C - cd; i += 32) dst[cd + i] = es[i];
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
  pf::slot_loss_kernel<<<pf::GL_CLUSTER, 1024, 0, (cudaStream_t)stream>>>(z, logdet, w, (int)rows, D, loss, gz, glogdet);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return pf::cuda_fail(e, "slot_loss launch", __FILE__, __LINE__);
  return PF_OK;
}

extern "C" pf_status pf_slot_ctx_grad(const float* gcfull, int32_t B, int32_t S, int32_t context_dim, int32_t C,
                                      float* gcontext, float* grank_emb, int32_t device, void* stream) {
  if (!gcfull || !gcontext || !grank_emb || B < 0 || S < 1 || context_dim < 0 || C <= context_dim) return PF_ERR_INVALID;
  if (B == 0) return PF_OK;
  PF_CUDA(cudaSetDevice(device));
  const int nctx = ((B + 3) / 4 + pf::GL_CLUSTER - 1) / pf::GL_CLUSTER * pf::GL_CLUSTER;
  pf::ctx_grad_kernel<<<nctx + pf::GL_CLUSTER * S, 1024, 0, (cudaStream_t)stream>>>(gcfull, nullptr, B, S, context_dim, C, nctx,
                                                                                    gcontext, grank_emb);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return pf::cuda_fail(e, "slot_ctx_grad launch", __FILE__, __LINE__);
  return PF_OK;
}

extern "C" pf_status pf_rows_prepare(const float* params, const int64_t* nsig, const float* context, const float* rank_emb,
                                     const pf_scaler* scaler, const float* inv_count, const float* use_inv, int32_t B,
                                     int32_t S, int32_t D, int32_t context_dim, int32_t C, int32_t cap, int32_t* offs, float* x,
                                     float* cfull, float* w, int32_t device, void* stream) {
  if (!params || !nsig || !context || !rank_emb || !scaler || !scaler->lo || !scaler->hi || !scaler->log_mask || !offs ||
      !x || !cfull || !w)
    return PF_ERR_INVALID;
  if (B < 1 || B > (1 << 20) || S < 1 || D < 1 || D > 32 || context_dim < 0 || C <= context_dim || cap < 1) return PF_ERR_INVALID;
  PF_CUDA(cudaSetDevice(device));
  pf::rows_scan_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(nsig, B, S, offs);
  int grid = (cap + 7) / 8;
  if (grid > 1184) grid = 1184;
  pf::rows_prepare_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(params, offs, context, rank_emb, scaler->lo, scaler->hi,
                                                                 scaler->log_mask, inv_count, use_inv, B, S, D, context_dim, C,
                                                                 cap, x, cfull, w);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return pf::cuda_fail(e, "rows_prepare launch", __FILE__, __LINE__);
  return PF_OK;
}

extern "C" pf_status pf_rows_ctx_grad(const float* gcfull, const int32_t* offs, int32_t B, int32_t S, int32_t context_dim,
                                      int32_t C, float* gcontext, float* grank_emb, int32_t device, void* stream) {
  if (!gcfull || !offs || !gcontext || !grank_emb || B < 1 || S < 1 || context_dim < 0 || C <= context_dim) return PF_ERR_INVALID;
  PF_CUDA(cudaSetDevice(device));
  const int nctx = ((B + 3) / 4 + pf::GL_CLUSTER - 1) / pf::GL_CLUSTER * pf::GL_CLUSTER;
  pf::ctx_grad_kernel<<<nctx + pf::GL_CLUSTER * S, 1024, 0, (cudaStream_t)stream>>>(gcfull, offs, B, S, context_dim, C, nctx,
                                                                                    gcontext, grank_emb);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return pf::cuda_fail(e, "rows_ctx_grad launch", __FILE__, __LINE__);
  return PF_OK;
}
