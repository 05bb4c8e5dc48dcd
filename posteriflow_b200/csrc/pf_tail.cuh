// Row-level pieces shared by both arithmetic tiers: the in-kernel base-noise generator and the tail of
// NSFPosteriorFlow.inverse + ParamScaler (so the two tiers cannot drift apart).
#pragma once
#include "pf_common.cuh"

namespace pf {

// Philox4x32-10 (Salmon et al. 2011) keyed by the call's seed; counter = (row, quad index).
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t out[4]) {
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float& n0, float& n1) {
  float u1 = ((float)a + 1.f) * 2.3283064365386963e-10f;  // (0, 1]
  float u2 = (float)b * 2.3283064365386963e-10f;
  float rad = sqrtf(-2.f * logf(u1));
  float s, c;
  sincospif(2.f * u2, &s, &c);
  n0 = rad * c; n1 = rad * s;
}
// z[0..D) ~ N(0, I) of global flow row `gr` (replaces torch.randn at lean_npe.py:329 in throughput mode)
__device__ __forceinline__ void philox_normal_row(uint64_t gr, uint64_t seed, int D, float* z) {
#pragma unroll
  for (int q = 0; q < PF_MAX_D / 4; ++q) {
    if (q * 4 >= D) break;
    uint32_t o[4];
    philox4x32_10((uint32_t)gr, (uint32_t)(gr >> 32), (uint32_t)q, 0u, (uint32_t)seed, (uint32_t)(seed >> 32), o);
    box_muller(o[0], o[1], z[4 * q], z[4 * q + 1]);
    box_muller(o[2], o[3], z[4 * q + 2], z[4 * q + 3]);
  }
}

struct ScalerDev { const float* lo; const float* hi; const uint8_t* log_mask; const uint8_t* circ_mask; };

// v[0..D): output of the last inverse layer in the caller's dimension order.  Applies, per flags:
// nan_to_num(0,+1,-1) + clamp (flows.py:650-654), ParamScaler.wrap (lean_npe.py:100-104),
// denormalize (lean_npe.py:111-114), m1 >= m2 (pipeline.py:184-186).  Returns the rail mask
// (pipeline.py:177: any non-circular |y| > 0.999 after wrap).
__device__ __forceinline__ bool finish_row(float* v, int D, uint32_t flags, float clamp_bound, const ScalerDev& sc) {
  bool rail = false;
#pragma unroll
  for (int j = 0; j < PF_MAX_D; ++j) {
    if (j >= D) break;
    float x = v[j];
    if (flags & PF_FLAG_CLAMP) {
      if (x != x) x = 0.f;
      else if (isinf(x)) x = x > 0.f ? 1.f : -1.f;
      x = fminf(fmaxf(x, -clamp_bound), clamp_bound);
    }
    if (flags & PF_FLAG_WRAP) {
      if (sc.circ_mask[j]) {
        float m = fmodf(x + 1.f, 2.f);  // torch.remainder: sign follows the divisor
        if (m != 0.f && m < 0.f) m += 2.f;
        x = m - 1.f;
      } else {
        x = fminf(fmaxf(x, -1.f), 1.f);
        rail = rail || fabsf(x) > 0.999f;
      }
    }
    if (flags & PF_FLAG_DENORMALIZE) {
      float y = fminf(fmaxf(x, -1.f), 1.f);
      float lo = sc.lo[j], hi = sc.hi[j];
      x = (y + 1.f) / 2.f * (hi - lo) + lo;
      if (sc.log_mask[j]) x = expf(x);
    }
    v[j] = x;
  }
  if ((flags & PF_FLAG_SORT_MASSES) && D >= 2 && v[1] > v[0]) { float t = v[0]; v[0] = v[1]; v[1] = t; }
  return rail;
}

}  // namespace pf
