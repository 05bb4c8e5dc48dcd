// Rational-quadratic spline arithmetic (Durkan et al. 2019) exactly as nflows evaluates it
// (SURVEY.md Appendix A.5-A.7): width/height softmax with 1e-3 floors, knot cumsum scaled to
// [-B, B] with pinned end knots, derivative softplus with unit boundary derivatives, bin search
// with +1e-6 on the last knot, forward / inverse map and log|det|, linear tails outside [-B, B].
// Everything stays in registers: one (row, dim) per thread, parameters read through a functor.
#pragma once
#include <math_constants.h>

#include "pf_common.cuh"

namespace pf {

constexpr float kMinBin = 1e-3f;         // DEFAULT_MIN_BIN_WIDTH / HEIGHT
constexpr float kMinDeriv = 1e-3f;       // DEFAULT_MIN_DERIVATIVE
// np.log(np.exp(1 - 1e-3) - 1): the unnormalised derivative nflows pads at both ends
constexpr float kEdgeLogit = 0.5397424172369522f;

__device__ __forceinline__ float softplus_f(float x) {  // torch F.softplus, beta=1, threshold=20
  return x > 20.f ? x : log1pf(expf(x));
}

struct RqsBin {
  float xk, wk, yk, hk, dk, dk1;  // left knot, bin width, bottom knot, bin height, knot derivatives
  int k;
};

// Locate the bin of `v` (x for the forward map, y for the inverse) and gather its six numbers.
// P(i) returns unnormalised parameter i of this (row, dim): [0,NB) widths, [NB,2NB) heights,
// [2NB, 3NB-1) interior derivatives.
template <int NB, bool INVERSE, class P>
__device__ __forceinline__ RqsBin rqs_locate(const P& prm, float v, float B) {
  float ew[NB], eh[NB];
  float mw = -CUDART_INF_F, mh = -CUDART_INF_F;
#pragma unroll
  for (int i = 0; i < NB; ++i) {
    ew[i] = prm(i);
    eh[i] = prm(NB + i);
    mw = fmaxf(mw, ew[i]);
    mh = fmaxf(mh, eh[i]);
  }
  float sw = 0.f, sh = 0.f;
#pragma unroll
  for (int i = 0; i < NB; ++i) {
    ew[i] = expf(ew[i] - mw);
    eh[i] = expf(eh[i] - mh);
    sw += ew[i];
    sh += eh[i];
  }
  const float scale = 1.f - kMinBin * NB;
  const float rw = 1.f / sw, rh = 1.f / sh;
  RqsBin b;
  b.k = 0;
  // pass 1: search along the axis of v
  {
    float cum = 0.f, prev = -B;
    float lo = -B, sz = 0.f;
#pragma unroll
    for (int i = 0; i < NB; ++i) {
      float e = INVERSE ? eh[i] * rh : ew[i] * rw;
      cum += kMinBin + scale * e;
      float next = (i == NB - 1) ? B : (2.f * B) * cum + (-B);
      if (i == 0 || v >= prev) { b.k = i; lo = prev; sz = next - prev; }
      prev = next;
    }
    if (INVERSE) { b.yk = lo; b.hk = sz; } else { b.xk = lo; b.wk = sz; }
  }
  // pass 2: the other axis at bin k
  {
    float cum = 0.f, prev = -B;
    float lo = -B, sz = 0.f;
#pragma unroll
    for (int i = 0; i < NB; ++i) {
      float e = INVERSE ? ew[i] * rw : eh[i] * rh;
      cum += kMinBin + scale * e;
      float next = (i == NB - 1) ? B : (2.f * B) * cum + (-B);
      if (i == b.k) { lo = prev; sz = next - prev; }
      prev = next;
    }
    if (INVERSE) { b.xk = lo; b.wk = sz; } else { b.yk = lo; b.hk = sz; }
  }
  // derivatives at knots k and k+1 (knots 0 and NB carry the padded edge logit => derivative 1)
  float u0 = kEdgeLogit, u1 = kEdgeLogit;
#pragma unroll
  for (int i = 0; i < NB - 1; ++i) {
    float u = prm(2 * NB + i);   // interior knot i+1
    if (i + 1 == b.k) u0 = u;
    if (i == b.k) u1 = u;
  }
  b.dk = kMinDeriv + softplus_f(u0);
  b.dk1 = kMinDeriv + softplus_f(u1);
  return b;
}

// forward map inside [-B, B]: returns y, writes log|dy/dx|
__device__ __forceinline__ float rqs_forward(const RqsBin& b, float x, float& lad) {
  float s = b.hk / b.wk;
  float th = (x - b.xk) / b.wk;
  float tt = th * (1.f - th);
  float num = b.hk * (s * th * th + b.dk * tt);
  float den = s + (b.dk + b.dk1 - 2.f * s) * tt;
  float omt = 1.f - th;
  float dnum = s * s * (b.dk1 * th * th + 2.f * s * tt + b.dk * omt * omt);
  lad = logf(dnum) - 2.f * logf(den);
  return b.yk + num / den;
}

// inverse map inside [-B, B]: returns x, writes -log|dy/dx| (the sign nflows returns)
__device__ __forceinline__ float rqs_inverse(const RqsBin& b, float y, float& lad) {
  float s = b.hk / b.wk;
  float dy = y - b.yk;
  float t = b.dk + b.dk1 - 2.f * s;
  float qa = dy * t + b.hk * (s - b.dk);
  float qb = b.hk * b.dk - dy * t;
  float qc = -s * dy;
  // nflows asserts disc >= 0 and the reference then falls back for the WHOLE batch (flows.py:638-642);
  // per element we clamp instead -- unreachable for finite inputs (documented in DESIGN.md, H6).
  float disc = fmaxf(qb * qb - 4.f * qa * qc, 0.f);
  float root = (2.f * qc) / (-qb - sqrtf(disc));
  float tt = root * (1.f - root);
  float den = s + t * tt;
  float omr = 1.f - root;
  float dnum = s * s * (b.dk1 * root * root + 2.f * s * tt + b.dk * omr * omr);
  lad = -(logf(dnum) - 2.f * logf(den));
  return root * b.wk + b.xk;
}

// One (row, dim) of MaskedPiecewiseRationalQuadraticAutoregressiveTransform._elementwise with
// tails='linear'.
template <int NB, bool INVERSE, class P>
__device__ __forceinline__ float rqs_apply(const P& prm, float v, float B, float& lad) {
  if (!(v >= -B && v <= B)) {  // outside (or NaN): identity, logdet 0
    lad = 0.f;
    return v;
  }
  RqsBin b = rqs_locate<NB, INVERSE>(prm, v, B);
  return INVERSE ? rqs_inverse(b, v, lad) : rqs_forward(b, v, lad);
}

}  // namespace pf
