// N3 of SURVEY.md 8(f): on-device posterior summaries.  The reference pulls every sample to the host and runs
// numpy there (scripts/overlap_benchmark.py:67-89 calib_block, scripts/validate_checkpoint.py:102-128,
// experiments/train_lean_npe.py:160-167): np.quantile / np.median over the sample axis, SBC ranks
// (samples < truth).mean(axis=1), railing fraction.  At the sampler's rate the raw samples are 1.7 GB/s per GPU of
// output; the summaries are a few hundred bytes per event.
//
// One CTA per (event, dimension).  Exact order statistics by a 4-pass most-significant-digit radix SELECT on the
// order-preserving 32-bit key of the float (no sort): pass p histograms bits [31-8p, 24-8p] of the elements
// that still match a target's key prefix and narrows every target rank to one bin.  All targets (two order
// statistics per requested quantile + the lower median) are resolved in the same 4 sweeps; targets that share a
// prefix share a histogram.  Keys are cached in shared memory after the first sweep when the event fits
// (n <= 50 k), so global memory is read once: HBM-bound work, 4 B per (sample, dim).
#include "pf_common.cuh"

namespace pf {

constexpr int SM_THREADS = 512;
constexpr int SM_MAXQ = 8;
constexpr int SM_MAXT = 2 * SM_MAXQ + 1;

struct SummaryPlan {
  int Q, T;
  int k[SM_MAXT];          // 0-based order-statistic index of each target
  double gamma[SM_MAXQ];   // interpolation weight of quantile i between targets 2i and 2i+1
};

__device__ __forceinline__ uint32_t f2key(float v) {
  uint32_t b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key2f(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

__global__ void __launch_bounds__(SM_THREADS)
posterior_summary_kernel(const float* __restrict__ theta, int64_t n, int D, SummaryPlan plan,
                         const float* __restrict__ truth, const uint8_t* __restrict__ railed,
                         float* __restrict__ quant, float* __restrict__ lower_median, float* __restrict__ below,
                         float* __restrict__ rail_frac, int cache_cap) {
  extern __shared__ uint32_t sm_dyn[];
  __shared__ uint32_t hist[SM_MAXT][256];
  __shared__ uint32_t prefix[SM_MAXT];
  __shared__ uint32_t krem[SM_MAXT];
  __shared__ int rep[SM_MAXT];       // representative target with the same prefix (histogram owner)
  __shared__ int uniq[SM_MAXT];      // list of representatives
  __shared__ int n_uniq;
  __shared__ unsigned long long cnt_below, cnt_rail;

  const int j = (int)(blockIdx.x % (unsigned)D);     // dimension fastest: the D CTAs of an event share its lines in L2
  const int64_t e = blockIdx.x / (unsigned)D;
  const int tid = threadIdx.x, lane = tid & 31;
  const int T = plan.T;
  const float* src = theta + e * n * D + j;
  const bool cached = n <= (int64_t)cache_cap;
  uint32_t* keys = sm_dyn;

  if (tid < T) { prefix[tid] = 0u; krem[tid] = (uint32_t)plan.k[tid]; rep[tid] = 0; }
  if (tid == 0) { n_uniq = 1; uniq[0] = 0; cnt_below = 0ull; cnt_rail = 0ull; }
  const float tr = truth ? truth[e * D + j] : 0.f;
  const int64_t n_round = (n + 31) / 32 * 32;

  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 24 - 8 * pass;
    const uint32_t himask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
    __syncthreads();
    const int nu = n_uniq;
    for (int i = tid; i < nu * 256; i += SM_THREADS) hist[uniq[i >> 8]][i & 255] = 0u;
    __syncthreads();
    unsigned my_below = 0;
    for (int64_t i = tid; i < n_round; i += SM_THREADS) {
      const bool valid = i < n;
      uint32_t key = 0u;
      if (valid) {
        if (pass == 0 || !cached) {
          const float v = src[i * D];
          key = f2key(v);
          if (pass == 0) {
            if (cached) keys[i] = key;
            if (truth && v < tr) ++my_below;
          }
        } else {
          key = keys[i];
        }
      }
      const uint32_t bin = (key >> shift) & 255u;
      if (nu == 1) {
        // one shared histogram (always the case in the first sweep, and in the second for narrow distributions):
        // aggregate equal bins inside the warp so that a peaked histogram does not serialise its atomics
        const bool hit = valid && (key & himask) == prefix[uniq[0]];
        const uint32_t tag = hit ? bin : 0xffffffffu;
        const unsigned peers = __match_any_sync(0xffffffffu, tag);
        if (hit && lane == __ffs(peers) - 1) atomicAdd(&hist[uniq[0]][bin], (uint32_t)__popc(peers));
      } else if (valid) {
        const uint32_t hi = key & himask;
        for (int u = 0; u < nu; ++u) {
          const int t = uniq[u];
          if (hi == prefix[t]) atomicAdd(&hist[t][bin], 1u);
        }
      }
    }
    if (pass == 0 && truth) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) my_below += __shfl_xor_sync(0xffffffffu, my_below, o);
      if (lane == 0 && my_below) atomicAdd(&cnt_below, (unsigned long long)my_below);
    }
    __syncthreads();
    // one warp per target: find the bin that holds rank krem[t] among the elements matching prefix[t]
    for (int t = tid >> 5; t < T; t += SM_THREADS / 32) {
      const uint32_t* h = hist[rep[t]];
      const uint32_t kk = krem[t];
      uint32_t c[8], s = 0;
#pragma unroll
      for (int i = 0; i < 8; ++i) { c[i] = h[lane * 8 + i]; s += c[i]; }
      uint32_t incl = s;                       // inclusive scan of the per-lane sums
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      uint32_t run = incl - s;                 // elements in bins before this lane's 8
      int found = -1;
      uint32_t before = 0;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (found < 0 && kk >= run && kk < run + c[i]) { found = lane * 8 + i; before = run; }
        run += c[i];
      }
      const unsigned who = __ballot_sync(0xffffffffu, found >= 0);
      if (who) {
        const int srcl = __ffs(who) - 1;
        const int b = __shfl_sync(0xffffffffu, found, srcl);
        const uint32_t bf = __shfl_sync(0xffffffffu, before, srcl);
        if (lane == 0) { prefix[t] |= (uint32_t)b << shift; krem[t] = kk - bf; }
      }
    }
    __syncthreads();
    if (tid == 0) {   // regroup targets by prefix for the next sweep
      int nu2 = 0;
      for (int t = 0; t < T; ++t) {
        int r = -1;
        for (int u = 0; u < nu2; ++u)
          if (prefix[uniq[u]] == prefix[t]) { r = uniq[u]; break; }
        if (r < 0) { r = t; uniq[nu2++] = t; }
        rep[t] = r;
      }
      n_uniq = nu2;
    }
  }
  __syncthreads();
  if (tid < plan.Q && quant) {
    const double a = (double)key2f(prefix[2 * tid]), b = (double)key2f(prefix[2 * tid + 1]);
    const double g = plan.gamma[tid], diff = b - a;
    double r = a + diff * g;                    // numpy's _lerp (function_base.py), also torch.quantile 'linear'
    if (g >= 0.5) r = b - diff * (1.0 - g);
    quant[(e * plan.Q + tid) * D + j] = (float)r;
  }
  if (tid == 0 && lower_median) lower_median[e * D + j] = key2f(prefix[T - 1]);
  if (tid == 0 && below) below[e * D + j] = (float)((double)cnt_below / (double)n);
  if (j == 0 && railed && rail_frac) {
    unsigned my = 0;
    for (int64_t i = tid; i < n; i += SM_THREADS) my += railed[e * n + i] ? 1u : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) my += __shfl_xor_sync(0xffffffffu, my, o);
    if (lane == 0 && my) atomicAdd(&cnt_rail, (unsigned long long)my);
    __syncthreads();
    if (tid == 0) rail_frac[e] = (float)((double)cnt_rail / (double)n);
  }
}

}  // namespace pf

extern "C" pf_status pf_posterior_summary(const float* theta, int64_t E, int64_t n, int32_t D, const double* q,
                                          int32_t Q, const float* truth, const uint8_t* railed, float* quant,
                                          float* lower_median, float* below, float* rail_frac, int32_t device,
                                          void* stream) {
  using namespace pf;
  if (!theta || E < 0 || n < 1 || D < 1 || Q < 0 || Q > SM_MAXQ || (Q > 0 && (!q || !quant))) return PF_ERR_INVALID;
  if (n >= (1ll << 31) || E * D >= (1ll << 31)) return PF_ERR_UNSUPPORTED;
  if (below && !truth) return PF_ERR_INVALID;
  if (E == 0) return PF_OK;
  SummaryPlan plan;
  memset(&plan, 0, sizeof(plan));
  plan.Q = Q;
  plan.T = 2 * Q + 1;
  for (int i = 0; i < Q; ++i) {
    if (!(q[i] >= 0.0 && q[i] <= 1.0)) return PF_ERR_INVALID;
    const double pos = q[i] * (double)(n - 1);   // numpy 'linear': virtual index (n - 1) q
    int64_t lo = (int64_t)pos;
    if (lo > n - 1) lo = n - 1;
    const int64_t hi = lo + 1 < n ? lo + 1 : n - 1;
    plan.k[2 * i] = (int)lo;
    plan.k[2 * i + 1] = (int)hi;
    plan.gamma[i] = pos - (double)lo;
  }
  plan.k[2 * Q] = (int)((n - 1) / 2);   // torch.median: the lower of the two middle elements
  PF_CUDA(cudaSetDevice(device));
  static int cap_cache[64] = {};
  int cap = 0;
  if (device >= 0 && device < 64 && cap_cache[device]) cap = cap_cache[device];
  else {
    int optin = 0;
    PF_CUDA(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    cudaFuncAttributes fa;
    PF_CUDA(cudaFuncGetAttributes(&fa, posterior_summary_kernel));
    cap = (optin - (int)fa.sharedSizeBytes - 1024) / 4;
    if (cap < 0) cap = 0;
    PF_CUDA(cudaFuncSetAttribute(posterior_summary_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, cap * 4));
    if (device >= 0 && device < 64) cap_cache[device] = cap;
  }
  const size_t dyn = n <= cap ? (size_t)n * 4 : 0;
  dim3 grid((unsigned)(E * D));
  posterior_summary_kernel<<<grid, SM_THREADS, dyn, (cudaStream_t)stream>>>(theta, n, D, plan, truth, railed, quant,
                                                                         lower_median, below, rail_frac, cap);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return pf::cuda_fail(e, "posterior_summary launch", __FILE__, __LINE__);
  return PF_OK;
}
