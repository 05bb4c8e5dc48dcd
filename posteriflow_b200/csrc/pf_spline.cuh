// Rational-quadratic spline arithmetic (Durkan et al. 2019) exactly as nflows evaluates it
// (SURVEY.md Appendix A.5-A.7): width/height softmax with 1e-3 floors, knot cumsum scaled to
// [-B, B] with pinned end knots, derivative softplus with unit boundary derivatives, bin search
// with +1e-6 on the last knot, forward / inverse map and log|det|, linear tails outside [-B, B].
// Everything stays in registers: one (row, dim) per thread, parameters read through a functor.
#pragma once
#include <math.h>

#include "pf_common.cuh"

#ifndef PF_HD
#define PF_HD __host__ __device__ __forceinline__
#endif

namespace pf {

constexpr float kMinBin = 1e-3f;         // DEFAULT_MIN_BIN_WIDTH / HEIGHT
constexpr float kMinDeriv = 1e-3f;       // DEFAULT_MIN_DERIVATIVE
// np.log(np.exp(1 - 1e-3) - 1): the unnormalised derivative nflows pads at both ends
constexpr float kEdgeLogit = 0.5397424172369522f;

PF_HD float softplus_f(float x) {  // torch F.softplus, beta=1, threshold=20
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
PF_HD RqsBin rqs_locate(const P& prm, float v, float B) {
  float ew[NB], eh[NB];
  float mw = -INFINITY, mh = -INFINITY;
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
PF_HD float rqs_forward(const RqsBin& b, float x, float& lad) {
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
PF_HD float rqs_inverse(const RqsBin& b, float y, float& lad) {
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
PF_HD float rqs_apply(const P& prm, float v, float B, float& lad) {
  if (!(v >= -B && v <= B)) {  // outside (or NaN): identity, logdet 0
    lad = 0.f;
    return v;
  }
  RqsBin b = rqs_locate<NB, INVERSE>(prm, v, B);
  return INVERSE ? rqs_inverse(b, v, lad) : rqs_forward(b, v, lad);
}


// Reverse mode of the forward map (theta -> z direction) for ONE (row, dim): given the upstream
// gradients gy (wrt the transformed value) and gl (wrt this dim's log|det|), writes the gradient
// wrt each of the 3*NB-1 unnormalised parameters through G(i, value) and returns d/dx.
// Follows exactly what autograd does through nflows' rational_quadratic_spline (Appendix A.7):
// pinned end knots and the two padded edge derivatives carry no gradient; outside the tails the
// map is the identity (gx = gy, no parameter gradient).
template <int NB, class P, class GOut>
PF_HD float rqs_forward_backward(const P& prm, float x, float B, float gy, float gl, const GOut& G) {
  if (!(x >= -B && x <= B)) {
#pragma unroll
    for (int i = 0; i < 3 * NB - 1; ++i) G(i, 0.f);
    return gy;
  }
  float pw[NB], ph[NB];
  float mw = -INFINITY, mh = -INFINITY;
#pragma unroll
  for (int i = 0; i < NB; ++i) {
    pw[i] = prm(i);
    ph[i] = prm(NB + i);
    mw = fmaxf(mw, pw[i]);
    mh = fmaxf(mh, ph[i]);
  }
  float sw = 0.f, sh = 0.f;
#pragma unroll
  for (int i = 0; i < NB; ++i) {
    pw[i] = expf(pw[i] - mw);
    ph[i] = expf(ph[i] - mh);
    sw += pw[i];
    sh += ph[i];
  }
  const float scale = 1.f - kMinBin * NB;
  const float rw = 1.f / sw, rh = 1.f / sh;
  // knots along x (search axis) and the bin
  int k = 0;
  float xk = -B, wk = 0.f, yk = -B, hk = 0.f;
  {
    float cum = 0.f, prev = -B;
#pragma unroll
    for (int i = 0; i < NB; ++i) {
      pw[i] *= rw;  // softmax probabilities from here on
      cum += kMinBin + scale * pw[i];
      float next = (i == NB - 1) ? B : (2.f * B) * cum + (-B);
      if (i == 0 || x >= prev) { k = i; xk = prev; wk = next - prev; }
      prev = next;
    }
    cum = 0.f; prev = -B;
#pragma unroll
    for (int i = 0; i < NB; ++i) {
      ph[i] *= rh;
      cum += kMinBin + scale * ph[i];
      float next = (i == NB - 1) ? B : (2.f * B) * cum + (-B);
      if (i == k) { yk = prev; hk = next - prev; }
      prev = next;
    }
  }
  float u0 = kEdgeLogit, u1 = kEdgeLogit;
#pragma unroll
  for (int i = 0; i < NB - 1; ++i) {
    float u = prm(2 * NB + i);
    if (i + 1 == k) u0 = u;
    if (i == k) u1 = u;
  }
  const float dk = kMinDeriv + softplus_f(u0), dk1 = kMinDeriv + softplus_f(u1);

  // forward intermediates
  const float s = hk / wk;
  const float th = (x - xk) / wk;
  const float omt = 1.f - th;
  const float tt = th * omt;
  const float inner = s * th * th + dk * tt;
  const float num = hk * inner;
  const float cf = dk + dk1 - 2.f * s;
  const float den = s + cf * tt;
  const float Q = dk1 * th * th + 2.f * s * tt + dk * omt * omt;
  const float dnum = s * s * Q;
  // reverse
  const float g_num = gy / den;
  const float g_den = -gy * num / (den * den) - 2.f * gl / den;
  const float g_dnum = gl / dnum;
  const float g_Q = g_dnum * s * s;
  float g_s = g_dnum * 2.f * s * Q + g_Q * 2.f * tt + g_den * (1.f - 2.f * tt) + g_num * hk * th * th;
  float g_dk = g_Q * omt * omt + g_den * tt + g_num * hk * tt;
  float g_dk1 = g_Q * th * th + g_den * tt;
  float g_tt = g_Q * 2.f * s + g_den * cf + g_num * hk * dk;
  float g_th = g_Q * (2.f * dk1 * th - 2.f * dk * omt) + g_num * hk * s * 2.f * th + g_tt * (1.f - 2.f * th);
  float g_h = g_num * inner + g_s / wk;
  float g_w = -g_th * th / wk - g_s * s / wk;
  const float gx = g_th / wk;
  const float g_xk = -g_th / wk;
  const float g_yk = gy;
  // knots -> cumulative sums.  cw_j = 2B * sum_{i<j} w_i - B for 0 < j < NB (ends pinned).
  // bin k uses knots k (left) and k+1 (right): grad wrt knot k: g_xk - g_w ; knot k+1: g_w.
  const float Gk = (k > 0) ? (g_xk - g_w) : 0.f;
  const float Gk1 = (k + 1 < NB) ? g_w : 0.f;
  const float Hk = (k > 0) ? (g_yk - g_h) : 0.f;
  const float Hk1 = (k + 1 < NB) ? g_h : 0.f;
  // g wrt normalised width i: 2B * (Gk [i < k] + Gk1 [i < k+1]); softmax backward
  float dotw = 0.f, doth = 0.f;
#pragma unroll
  for (int i = 0; i < NB; ++i) {
    float gw_i = (2.f * B) * ((i < k ? Gk : 0.f) + (i < k + 1 ? Gk1 : 0.f));
    float gh_i = (2.f * B) * ((i < k ? Hk : 0.f) + (i < k + 1 ? Hk1 : 0.f));
    dotw += gw_i * pw[i];
    doth += gh_i * ph[i];
  }
#pragma unroll
  for (int i = 0; i < NB; ++i) {
    float gw_i = (2.f * B) * ((i < k ? Gk : 0.f) + (i < k + 1 ? Gk1 : 0.f));
    float gh_i = (2.f * B) * ((i < k ? Hk : 0.f) + (i < k + 1 ? Hk1 : 0.f));
    G(i, scale * pw[i] * (gw_i - dotw));
    G(NB + i, scale * ph[i] * (gh_i - doth));
  }
  // softplus'(u) = sigmoid(u) (torch: 1 above the threshold of 20)
  const float sg0 = u0 > 20.f ? 1.f : 1.f / (1.f + expf(-u0));
  const float sg1 = u1 > 20.f ? 1.f : 1.f / (1.f + expf(-u1));
#pragma unroll
  for (int i = 0; i < NB - 1; ++i) {
    float g = 0.f;
    if (i + 1 == k) g += g_dk * sg0;
    if (i == k) g += g_dk1 * sg1;
    G(2 * NB + i, g);
  }
  return gx;
}

// ------------------------------------------------------------------------------------------------
// Tensor-core tier variant of the inverse map (same formulas, Appendix A.5-A.7) built for a lone warp per
// scheduler: hardware ex2 / lg2 / rcp approximations (about 2^-21 relative, far below that tier's bf16
// operand rounding), one logarithm for the log-determinant, and the two softmax/cumsum chains written
// side by side so their dependent instruction chains interleave.  q[0,NB) widths, [NB,2NB) heights,
// [2NB,3NB-1) interior knot derivatives (unnormalised).
#ifdef __CUDACC__
// single-instruction MUFU forms (flush-to-zero, no range fix-up code around them)
__device__ __forceinline__ float fx_ex2(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float fx_lg2(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float fx_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float fx_sqrt(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float fx_exp(float x) { return fx_ex2(x * 1.4426950408889634f); }
__device__ __forceinline__ float fx_log(float x) { return fx_lg2(x) * 0.6931471805599453f; }

template <int NB>
__device__ __forceinline__ float rqs_inverse_fast(const float* q, float y, float B, float& lad) {
  if (!(y >= -B && y <= B)) {
    lad = 0.f;
    return y;
  }
  float mw = q[0], mh = q[NB];
#pragma unroll
  for (int i = 1; i < NB; ++i) {
    mw = fmaxf(mw, q[i]);
    mh = fmaxf(mh, q[NB + i]);
  }
  float ew[NB], eh[NB];
  float sw0 = 0.f, sw1 = 0.f, sh0 = 0.f, sh1 = 0.f;
#pragma unroll
  for (int i = 0; i < NB; i += 2) {
    ew[i] = fx_ex2((q[i] - mw) * 1.4426950408889634f);
    ew[i + 1] = fx_exp(q[i + 1] - mw);
    eh[i] = fx_exp(q[NB + i] - mh);
    eh[i + 1] = fx_exp(q[NB + i + 1] - mh);
    sw0 += ew[i]; sw1 += ew[i + 1];
    sh0 += eh[i]; sh1 += eh[i + 1];
  }
  const float scale = 1.f - kMinBin * NB;
  const float cw = (2.f * B) * scale * fx_rcp(sw0 + sw1);   // knot spacing contributed by e_i
  const float ch = (2.f * B) * scale * fx_rcp(sh0 + sh1);
  const float base = (2.f * B) * kMinBin;
  // walk both knot sequences; the bin is found on the height axis, the width axis is gathered at it
  int k = 0;
  float yk = -B, hk = 0.f, xk = -B, wk = 0.f;
  float py = -B, px = -B, cy = 0.f, cx = 0.f;
#pragma unroll
  for (int i = 0; i < NB; ++i) {
    cy += fmaf(ch, eh[i], base);
    cx += fmaf(cw, ew[i], base);
    const float ny = (i == NB - 1) ? B : cy - B;
    const float nx = (i == NB - 1) ? B : cx - B;
    if (i == 0 || y >= py) { k = i; yk = py; hk = ny - py; xk = px; wk = nx - px; }
    py = ny; px = nx;
  }
  float u0 = kEdgeLogit, u1 = kEdgeLogit;
#pragma unroll
  for (int i = 0; i < NB - 1; ++i) {
    const float u = q[2 * NB + i];
    if (i + 1 == k) u0 = u;
    if (i == k) u1 = u;
  }
  const float dk = kMinDeriv + (u0 > 20.f ? u0 : fx_log(1.f + fx_exp(u0)));
  const float dk1 = kMinDeriv + (u1 > 20.f ? u1 : fx_log(1.f + fx_exp(u1)));
  const float s = hk * fx_rcp(wk);
  const float dy = y - yk;
  const float t = dk + dk1 - 2.f * s;
  const float qa = fmaf(dy, t, hk * (s - dk));
  const float qb = fmaf(hk, dk, -dy * t);
  const float qc = -s * dy;
  const float disc = fmaxf(fmaf(qb, qb, -4.f * qa * qc), 0.f);
  const float root = 2.f * qc * fx_rcp(-qb - fx_sqrt(disc));
  const float tt = root * (1.f - root);
  const float den = fmaf(t, tt, s);
  const float omr = 1.f - root;
  const float dnum = s * s * (dk1 * root * root + 2.f * s * tt + dk * omr * omr);
  lad = -fx_log(dnum * fx_rcp(den * den));
  return fmaf(root, wk, xk);
}

// Reverse mode of the forward map with the hardware ex2 / lg2 / rcp approximations (~2^-22 relative): the variant the
// fused backward kernel of the tensor-core tiers runs (its operands are split-bf16, ~1e-5; gradient tolerance 2e-3).
// Same formulas, same argument meaning as rqs_forward_backward above.
template <int NB, class P, class GOut>
__device__ __forceinline__ float rqs_forward_backward_fast(const P& prm, float x, float B, float gy, float gl, const GOut& G) {
  if (!(x >= -B && x <= B)) {
#pragma unroll
    for (int i = 0; i < 3 * NB - 1; ++i) G(i, 0.f);
    return gy;
  }
  float pw[NB], ph[NB];
  float mw = -INFINITY, mh = -INFINITY;
#pragma unroll
  for (int i = 0; i < NB; ++i) {
    pw[i] = prm(i);
    ph[i] = prm(NB + i);
    mw = fmaxf(mw, pw[i]);
    mh = fmaxf(mh, ph[i]);
  }
  float sw = 0.f, sh = 0.f;
#pragma unroll
  for (int i = 0; i < NB; ++i) {
    pw[i] = fx_exp(pw[i] - mw);
    ph[i] = fx_exp(ph[i] - mh);
    sw += pw[i];
    sh += ph[i];
  }
  const float scale = 1.f - kMinBin * NB;
  const float rw = fx_rcp(sw), rh = fx_rcp(sh);
  // knots along x (search axis) and the bin
  int k = 0;
  float xk = -B, wk = 0.f, yk = -B, hk = 0.f;
  {
    float cum = 0.f, prev = -B;
#pragma unroll
    for (int i = 0; i < NB; ++i) {
      pw[i] *= rw;  // softmax probabilities from here on
      cum += kMinBin + scale * pw[i];
      float next = (i == NB - 1) ? B : (2.f * B) * cum + (-B);
      if (i == 0 || x >= prev) { k = i; xk = prev; wk = next - prev; }
      prev = next;
    }
    cum = 0.f; prev = -B;
#pragma unroll
    for (int i = 0; i < NB; ++i) {
      ph[i] *= rh;
      cum += kMinBin + scale * ph[i];
      float next = (i == NB - 1) ? B : (2.f * B) * cum + (-B);
      if (i == k) { yk = prev; hk = next - prev; }
      prev = next;
    }
  }
  float u0 = kEdgeLogit, u1 = kEdgeLogit;
#pragma unroll
  for (int i = 0; i < NB - 1; ++i) {
    float u = prm(2 * NB + i);
    if (i + 1 == k) u0 = u;
    if (i == k) u1 = u;
  }
  const float dk = kMinDeriv + (u0 > 20.f ? u0 : fx_log(1.f + fx_exp(u0))), dk1 = kMinDeriv + (u1 > 20.f ? u1 : fx_log(1.f + fx_exp(u1)));

  // forward intermediates
  const float rwk = fx_rcp(wk);
  const float s = hk * rwk;
  const float th = (x - xk) * rwk;
  const float omt = 1.f - th;
  const float tt = th * omt;
  const float inner = s * th * th + dk * tt;
  const float num = hk * inner;
  const float cf = dk + dk1 - 2.f * s;
  const float den = s + cf * tt;
  const float Q = dk1 * th * th + 2.f * s * tt + dk * omt * omt;
  const float dnum = s * s * Q;
  // reverse
  const float rden = fx_rcp(den);
  const float g_num = gy * rden;
  const float g_den = -gy * num * rden * rden - 2.f * gl * rden;
  const float g_dnum = gl * fx_rcp(dnum);
  const float g_Q = g_dnum * s * s;
  float g_s = g_dnum * 2.f * s * Q + g_Q * 2.f * tt + g_den * (1.f - 2.f * tt) + g_num * hk * th * th;
  float g_dk = g_Q * omt * omt + g_den * tt + g_num * hk * tt;
  float g_dk1 = g_Q * th * th + g_den * tt;
  float g_tt = g_Q * 2.f * s + g_den * cf + g_num * hk * dk;
  float g_th = g_Q * (2.f * dk1 * th - 2.f * dk * omt) + g_num * hk * s * 2.f * th + g_tt * (1.f - 2.f * th);
  float g_h = g_num * inner + g_s * rwk;
  float g_w = -g_th * th * rwk - g_s * s * rwk;
  const float gx = g_th * rwk;
  const float g_xk = -g_th * rwk;
  const float g_yk = gy;
  // knots -> cumulative sums.  cw_j = 2B * sum_{i<j} w_i - B for 0 < j < NB (ends pinned).
  // bin k uses knots k (left) and k+1 (right): grad wrt knot k: g_xk - g_w ; knot k+1: g_w.
  const float Gk = (k > 0) ? (g_xk - g_w) : 0.f;
  const float Gk1 = (k + 1 < NB) ? g_w : 0.f;
  const float Hk = (k > 0) ? (g_yk - g_h) : 0.f;
  const float Hk1 = (k + 1 < NB) ? g_h : 0.f;
  // g wrt normalised width i: 2B * (Gk [i < k] + Gk1 [i < k+1]); softmax backward
  float dotw = 0.f, doth = 0.f;
#pragma unroll
  for (int i = 0; i < NB; ++i) {
    float gw_i = (2.f * B) * ((i < k ? Gk : 0.f) + (i < k + 1 ? Gk1 : 0.f));
    float gh_i = (2.f * B) * ((i < k ? Hk : 0.f) + (i < k + 1 ? Hk1 : 0.f));
    dotw += gw_i * pw[i];
    doth += gh_i * ph[i];
  }
#pragma unroll
  for (int i = 0; i < NB; ++i) {
    float gw_i = (2.f * B) * ((i < k ? Gk : 0.f) + (i < k + 1 ? Gk1 : 0.f));
    float gh_i = (2.f * B) * ((i < k ? Hk : 0.f) + (i < k + 1 ? Hk1 : 0.f));
    G(i, scale * pw[i] * (gw_i - dotw));
    G(NB + i, scale * ph[i] * (gh_i - doth));
  }
  // softplus'(u) = sigmoid(u) (torch: 1 above the threshold of 20)
  const float sg0 = u0 > 20.f ? 1.f : fx_rcp(1.f + fx_exp(-u0));
  const float sg1 = u1 > 20.f ? 1.f : fx_rcp(1.f + fx_exp(-u1));
#pragma unroll
  for (int i = 0; i < NB - 1; ++i) {
    float g = 0.f;
    if (i + 1 == k) g += g_dk * sg0;
    if (i == k) g += g_dk1 * sg1;
    G(2 * NB + i, g);
  }
  return gx;
}

#endif

}  // namespace pf
