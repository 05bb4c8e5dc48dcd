// Tensor-core tier, density direction (x -> z, NSFPosteriorFlow.forward, flows.py:610-618; the flow part of
// compute_psd_aware_nll, flows.py:727-779) as DENSE per-layer GEMMs: with x known, a MADE layer is one pass
//   h0 = W0 x + b0 + relu(c0);  2 x [ u = Wa relu(h) + ba ; h += gate * (Wb relu(u) + bb) ];  p = Wf h + bf
// so each 128-row tile needs 6 MMA stages of N = 256 / 240 per layer (60 per tile) instead of the 660 skinny
// N = 32 stages of the stepwise schedule the sampling direction is forced into (pf_tc.cu).
//
// One persistent CTA per SM, tile = 128 flow rows = TMEM lanes = UMMA M.  320 threads:
//   warps 0-7  row warps, two threads per flow row (thread pair = column halves / dimension halves)
//   warp 8     MMA issuer (converged warp, one elected lane), owns the TMEM allocation
//   warp 9     weight loader: streams the layer's bf16 blobs (UMMA canonical K-major layout, 786 KB per layer,
//              L2-resident) through a 6 x 16 KB shared-memory ring with cp.async.bulk + mbarrier complete_tx
// Tensor memory: cols [0,256) stage accumulator, [256,512) the fp32 residual stream h (tcgen05.ld/st by the row
// warps); for the final layer the accumulator is [0,240) params of dims 1..5 | [240,480) dims 6..10 (dim 0 has no
// hidden input under the autoregressive mask: bias only).  Shared memory: the bf16 A operand [128 x 256] in UMMA
// layout (written by the epilogues), the x tile (bf16 hi + lo halves, K = 32), the ring, per-layer constants.
#include <cuda_fp16.h>
#include <stdlib.h>

#include "pf_common.cuh"
#include "pf_sm100.cuh"
#include "pf_spline.cuh"
#include "pf_tail.cuh"

namespace pf {
using namespace pf::sm100;

constexpr int DN_ROWS = 128;
constexpr int DN_THREADS = 320;
constexpr int DN_SLOT = 16384;
constexpr int DN_H = 256;
constexpr int DN_NB = 16;
constexpr int DN_MULT = 3 * DN_NB - 1;   // 47
constexpr int DN_HACC = 0, DN_HRES = 256;
constexpr float DN_WSCALE = 16.f;   // split tier: weights are stored as fp16 pairs of 16 W

struct DnShape {
  int D, L, PLD, DP;
  int ga, gb;        // dims 1..ga in final-layer group A, ga+1..ga+gb in group B (dim 0: bias only)
  int NA, NBn;       // UMMA N of the two groups (multiples of 16, <= 256)
  uint32_t layer_bytes;    // bf16 image (SPLIT = 1)
  uint32_t layer_bytes3;   // split image (SPLIT = 3): every K = 16 chunk as a (hi, lo) pair
};

struct DnState {
  DnShape s;
  uint8_t* image;    // [L][layer_bytes]   bf16(W)
  uint8_t* image3;   // [L][layer_bytes3]  bf16 hi | lo of W, chunk by chunk (split tier and training forward)
  uint8_t* imagep;   // [L][2][DP_LAYER_BYTES]  pair form of image3 (pf_tc_pair.cuh), NULL when the pair kernel is off
  uint8_t* image2;   // [L][2][D2_LAYER_BYTES]  pair form of image (pf_tc_pair2.cuh: inference, two tiles in flight), or NULL
  long long* prof;
};

// Activations the forward keeps for the fused backward / weight-gradient kernels (pf_tc_train.cu), written by the
// epilogues that produce them.  Every [128 x 256] activation tile is stored as a 64 KB bf16 IMAGE in the layout an
// operand tile has in shared memory (UMMA K-major no-swizzle: 8-row x 16-byte core matrices, 2048 bytes per 8 columns).
// The backward reads relu masks and the pre-gate values from these images with one 16-byte load per 8 columns, and -- a
// core matrix of "8 rows x 8 K-elements, K contiguous" being byte-identical to one of "8 K-rows x 8 MN-elements, MN
// contiguous" -- the weight-gradient kernel bulk-copies the very same bytes into shared memory and feeds them to
// tcgen05.mma as an MN-major operand (contraction over the tile's rows).
enum { DN_IMG_H0 = 0, DN_IMG_U1 = 1, DN_IMG_H1 = 2, DN_IMG_U2 = 3, DN_IMG_H2 = 4,     // bf16(relu(h0)) .. bf16(h2)
       DN_IMG_V1 = 5, DN_IMG_V1L = 6, DN_IMG_V2 = 7, DN_IMG_V2L = 8, DN_NIMG = 9 };       // pre-gate v: bf16 hi | lo
struct DnSave {
  uint8_t* img;      // [L][DN_NIMG][tiles][65536]
  uint8_t* xt;       // [L][tiles][4096]  bf16 x tile of the layer, [128 x 16] in the same layout (columns = reversed dims)
  float* P;          // [L][tiles][PLD][128]  spline parameters (final-layer output incl. bias), tile-transposed: row minor
  float* XS;         // [L][tiles][D][128]    input of layer l (before its ReversePermutation), tile-transposed
  int64_t R;         // rows of the whole call
  int tiles;         // tiles of the whole call
};
__host__ __device__ inline size_t dn_img_off(int l, int which, int tile, int tiles) {
  return (((size_t)l * DN_NIMG + which) * (size_t)tiles + tile) * 65536;
}

struct DnArgs {
  DnShape s;
  const uint8_t* image;
  const float *b0, *ba, *bb, *bf;   // fp32 packed biases of the handle ([L][H], [L][2][H], [L][2][H], [L][PLD])
  const float* proj;                // [ctx rows][L*3*H]  relu(c0) | sigmoid gate 1 | sigmoid gate 2
  int64_t ldproj;
  CtxMap cm;
  const float* x;                   // [R, D]
  int64_t R;
  int n_tiles;
  float* z; float* logdet;
  float tail;
  int* err;
  long long* prof;                  // NULL, or 8 cycle counters filled by CTA 0 (PF_TC_PROF=1)
  DnSave sv;                        // SAVE instantiation only
};

// bf16 weight image of one layer, chunk after chunk (one chunk = [N x 16] in UMMA K-major no-swizzle layout):
// SPLIT 1:  I (2 chunks, N=256: W0 for the hi half of x, W0 again for the lo half) | A1 B1 A2 B2 (16 chunks each,
//           N=256) | FA (16 chunks, N=NA) | FB (16 chunks, N=NB)
// SPLIT 3:  every chunk as a pair fp16(16 W) | fp16(16 W - hi):  I (1 pair) | A1 B1 A2 B2 (16 pairs each) | FA | FB
__global__ void dn_pack_kernel(DnShape s, uint8_t* __restrict__ image, uint8_t* __restrict__ image3,
                               const float* __restrict__ W0, const float* __restrict__ Wa, const float* __restrict__ Wb,
                               const float* __restrict__ Wf) {
  const int l = blockIdx.y;
  const bool split = blockIdx.x >= 98;
  const int c = split ? blockIdx.x - 98 : blockIdx.x;   // chunk index within the layer
  int stage, cc, N;
  uint32_t off;
  if (!split) {        // 0..1 I, 2..65 hidden, 66..81 FA, 82..97 FB
    if (c < 2) { stage = 0; cc = c; N = 256; off = (uint32_t)c * 8192u; }
    else if (c < 66) { stage = 1 + (c - 2) / 16; cc = (c - 2) % 16; N = 256; off = (uint32_t)c * 8192u; }
    else if (c < 82) { stage = 5; cc = c - 66; N = s.NA; off = 66u * 8192u + (uint32_t)cc * (uint32_t)s.NA * 32u; }
    else { stage = 6; cc = c - 82; N = s.NBn; off = 66u * 8192u + 16u * (uint32_t)s.NA * 32u + (uint32_t)cc * (uint32_t)s.NBn * 32u; }
  } else {             // 0 I, 1..64 hidden, 65..80 FA, 81..96 FB
    if (c < 1) { stage = 0; cc = 0; N = 256; off = 0; }
    else if (c < 65) { stage = 1 + (c - 1) / 16; cc = (c - 1) % 16; N = 256; off = (uint32_t)c * 16384u; }
    else if (c < 81) { stage = 5; cc = c - 65; N = s.NA; off = 65u * 16384u + (uint32_t)cc * (uint32_t)s.NA * 64u; }
    else if (c < 97) { stage = 6; cc = c - 81; N = s.NBn; off = 65u * 16384u + 16u * (uint32_t)s.NA * 64u + (uint32_t)cc * (uint32_t)s.NBn * 64u; }
    else return;
  }
  __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>((split ? image3 + (size_t)l * s.layer_bytes3 : image + (size_t)l * s.layer_bytes) + off);
  const int H = DN_H;
  for (int e = threadIdx.x; e < N * 16; e += blockDim.x) {
    const int n = e / 16, kk = e % 16;
    float v = 0.f;
    if (stage == 0) {
      if (kk < s.D) v = W0[((size_t)l * H + n) * s.DP + kk];
    } else if (stage <= 4) {
      const int blk = (stage - 1) / 2;
      const float* W = ((stage - 1) & 1) ? Wb : Wa;
      v = W[(((size_t)l * 2 + blk) * H + n) * H + cc * 16 + kk];
    } else {
      const int first = stage == 5 ? 1 : 1 + s.ga, cnt = stage == 5 ? s.ga : s.gb;
      if (n < cnt * DN_MULT) v = Wf[((size_t)l * s.PLD + (size_t)first * DN_MULT + n) * H + cc * 16 + kk];
    }
    const int o = (kk / 8) * (N * 8) + (n / 8) * 64 + (n % 8) * 8 + (kk % 8);
    if (split) {   // fp16 hi | lo of 16 W (a power of two: exact; keeps the lo halves of typical weights out of the
      __half* d16 = reinterpret_cast<__half*>(dst);   // fp16 subnormal range; the epilogues fold the 1/16 back in)
      const float vs = fminf(fmaxf(v * DN_WSCALE, -60000.f), 60000.f);
      const __half hi = __float2half_rn(vs);
      d16[o] = hi;
      d16[N * 16 + o] = __float2half_rn(vs - __half2float(hi));
    } else {
      dst[o] = __float2bfloat16(v);
    }
  }
}

// two floats -> packed bf16x2 with relu fused into the conversion (lo in bits 0..15)
__device__ __forceinline__ uint32_t pack_bf16_relu(float lo, float hi) {
  uint32_t d;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
// Split tier: (a, b) -> packed fp16 hi pair (optionally relu'd) and the packed fp16 pair of what the hi half misses:
// a = hi + lo to ~22 significand bits (11 + 11).  fp16 rather than bf16 (8 + 8 bits): three MMAs per product then
// carry fp32-grade operands.  Range: activations are O(1..1e3), clamped at +-6e4 (fp16 max 65504).
__device__ __forceinline__ uint32_t pack_f16(float lo, float hi) {
  uint32_t d;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
__device__ __forceinline__ uint32_t pack_f16_relu(float lo, float hi) {
  uint32_t d;
  asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
template <bool RELU>
__device__ __forceinline__ void split_pack(float a, float b, uint32_t& hi, uint32_t& lo) {
  a = RELU ? fminf(a, 60000.f) : fminf(fmaxf(a, -60000.f), 60000.f);
  b = RELU ? fminf(b, 60000.f) : fminf(fmaxf(b, -60000.f), 60000.f);
  hi = RELU ? pack_f16_relu(a, b) : pack_f16(a, b);
  const float2 h = __half22float2(*reinterpret_cast<const __half2*>(&hi));
  const float ar = RELU ? fmaxf(a, 0.f) : a, br = RELU ? fmaxf(b, 0.f) : b;
  lo = pack_f16(ar - h.x, br - h.y);
}
// (a, b) -> packed bf16 hi pair and bf16 lo pair (saved pre-gate values: range-safe, ~16 bits)
__device__ __forceinline__ void split_pack_bf16(float a, float b, uint32_t& hi, uint32_t& lo) {
  hi = pack_bf16(a, b);
  lo = pack_bf16(a - __uint_as_float(hi << 16), b - __uint_as_float(hi & 0xffff0000u));
}
__device__ __forceinline__ float dn_finite(float v) { return fminf(fmaxf(v, -3.0e38f), 3.0e38f); }   // NaN -> -3e38
__device__ __forceinline__ float dn_relu(float v) { return fminf(fmaxf(v, 0.f), 3.0e38f); }          // NaN -> 0

// Forward rational-quadratic spline of one (row, dim) from its 3*NB-1 unnormalised parameters (Appendix A.5-A.7,
// nflows unconstrained_rational_quadratic_spline with linear tails), hardware ex2 / lg2 / rcp.
__device__ __forceinline__ float dn_rqs_forward(float* q, float x, float B, float& lad) {
  constexpr int NB = DN_NB;
  if (!(x >= -B && x <= B)) { lad = 0.f; return x; }
  float mw = q[0], mh = q[NB];
#pragma unroll
  for (int i = 1; i < NB; ++i) { mw = fmaxf(mw, q[i]); mh = fmaxf(mh, q[NB + i]); }
  float sw = 0.f, sh = 0.f;
#pragma unroll
  for (int i = 0; i < NB; ++i) {
    q[i] = fx_ex2((q[i] - mw) * 1.4426950408889634f);
    q[NB + i] = fx_ex2((q[NB + i] - mh) * 1.4426950408889634f);
    sw += q[i]; sh += q[NB + i];
  }
  const float cwn = (2.f * B) * (1.f - kMinBin * NB) * fx_rcp(sw);
  const float chn = (2.f * B) * (1.f - kMinBin * NB) * fx_rcp(sh);
  const float base = (2.f * B) * kMinBin;
  int k = 0;
  float xk = -B, wk = 0.f, px = -B, cx = 0.f;
#pragma unroll
  for (int i = 0; i < NB; ++i) {
    cx += fmaf(cwn, q[i], base);
    const float nx = (i == NB - 1) ? B : cx - B;
    if (i == 0 || x >= px) { k = i; xk = px; wk = nx - px; }
    px = nx;
  }
  float yk = -B, hk = 0.f, py = -B, cy = 0.f;
#pragma unroll
  for (int i = 0; i < NB; ++i) {
    cy += fmaf(chn, q[NB + i], base);
    const float ny = (i == NB - 1) ? B : cy - B;
    if (i == k) { yk = py; hk = ny - py; }
    py = ny;
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
  const float rw = fx_rcp(wk);
  const float sl = hk * rw;
  const float th = (x - xk) * rw;
  const float omt = 1.f - th;
  const float tt = th * omt;
  const float den = fmaf(dk + dk1 - 2.f * sl, tt, sl);
  const float rden = fx_rcp(den);
  const float dnum = sl * sl * (dk1 * th * th + 2.f * sl * tt + dk * omt * omt);
  lad = fx_log(dnum * rden * rden);
  return fmaf(hk * (sl * th * th + dk * tt), rden, yk);
}

// shared-memory carve (bytes)
constexpr int DN_CST_BIAS = 0;                         // floats: b0 | ba0 | bb0 | ba1 | bb1
constexpr int DN_CST_BF = 5 * DN_H;                    // [16 dims][48]
constexpr int DN_CST_CTX = DN_CST_BF + 16 * 48;        // [2 events][5][256]: relu(c0)+b0 | g1 | g1*bb0 | g2 | g2*bb1
constexpr int DN_CST_FLOATS = DN_CST_CTX + 2 * 5 * DN_H;
template <int SPLIT>
struct DnL {
  static constexpr int NSLOT = SPLIT == 3 ? 3 : 6;           // weight ring depth (16 KB slots)
  static constexpr int OFF_ACT = 0;                          // [128 x 256] bf16, UMMA layout (hi half when SPLIT == 3)
  static constexpr int OFF_ACTLO = 65536;                    // lo half (SPLIT == 3)
  static constexpr int OFF_XT = SPLIT == 3 ? 131072 : 65536; // [128 x 32] bf16 (x hi | x lo), UMMA layout
  static constexpr int OFF_RING = OFF_XT + 8192;
  static constexpr int OFF_CST = OFF_RING + NSLOT * DN_SLOT;
  static constexpr int OFF_Y = OFF_CST + DN_CST_FLOATS * 4;        // [2][128][12] fp32 layer inputs / outputs
  static constexpr int OFF_RED = OFF_Y + 2 * DN_ROWS * 12 * 4;     // [128] logdet partial of half 1, [2][128] bad flags
  static constexpr int OFF_BAR = OFF_RED + 3 * DN_ROWS * 4;
  static constexpr int SMEM = OFF_BAR + 256;
  static_assert(SMEM <= 232448, "shared-memory budget of one CTA");
};

struct DnRegQ {
  const float* q;
  __device__ __forceinline__ float operator()(int i) const { return q[i]; }
};

// SPLIT = 1: bf16 operands (PF_TIER_BF16 inference).  SPLIT = 3: every product as hi*hi + lo*hi + hi*lo of the bf16
// hi / lo halves of BOTH operands (~16 significand bits each, fp32 accumulate) and the spline on the exact expf / logf /
// division path: the tensor-core tier that meets the FP32 tolerance (PF_TIER_BF16X3), and the forward of the fused
// training step (SAVE: additionally stores what the backward needs, see DnSave).
template <int SPLIT, bool SAVE>
__global__ void __launch_bounds__(DN_THREADS, 1) tc_dense_forward_kernel(const DnArgs a) {
  static_assert(!SAVE || SPLIT == 3, "the training forward runs on the split operands");
  using LY = DnL<SPLIT>;
  constexpr int DN_NSLOT = LY::NSLOT;
  extern __shared__ __align__(128) uint8_t smem[];
  const DnShape& s = a.s;
  const int D = s.D;
  uint8_t* sACT = smem + LY::OFF_ACT;
  uint8_t* sACTLO = smem + LY::OFF_ACTLO;
  uint8_t* sXT = smem + LY::OFF_XT;
  uint8_t* ring = smem + LY::OFF_RING;
  float* cst = reinterpret_cast<float*>(smem + LY::OFF_CST);
  float* ybuf = reinterpret_cast<float*>(smem + LY::OFF_Y);
  float* red = reinterpret_cast<float*>(smem + LY::OFF_RED);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + LY::OFF_BAR);
  uint64_t* full = bars;                     // [NSLOT]
  uint64_t* empty = bars + DN_NSLOT;         // [NSLOT]
  uint64_t* acc_bar = bars + 2 * DN_NSLOT;   // accumulator of the current stage complete
  uint64_t* accb_bar = acc_bar + 1;          // second final-layer group complete
  uint64_t* op_bar = acc_bar + 2;            // operands of the next stage written (8 warp arrivals)
  volatile int* abort_s = reinterpret_cast<volatile int*>(acc_bar + 3);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 4);

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  if (tid == 0) {
    for (int i = 0; i < DN_NSLOT; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(acc_bar, 1);
    mbar_init(accb_bar, 1);
    mbar_init(op_bar, 8);
    *abort_s = 0;
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc(tmem_slot, 512);
  for (int i = tid; i < (LY::OFF_XT + 8192) / 16; i += DN_THREADS) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  auto wait = [&](uint64_t* bar, uint32_t parity, int code) -> bool {
    for (uint32_t i = 0; i < (1u << 22); ++i) {
      if (mbar_try_wait(bar, parity)) return true;
      if ((i & 255) == 255 && *abort_s) return false;
    }
    *abort_s = 1;
    raise_flag(a.err, code);
    return false;
  };

  if (warp < 8) {
    // ============================ row warps ============================
    const int half = warp >> 2, q4 = warp & 3;
    const int r = q4 * 32 + lane;
    const int rt = half * 128 + r;                              // index among the 256 row threads
    const uint32_t tl = tmem + ((uint32_t)(q4 * 32) << 16);
    const uint32_t rowoff = (uint32_t)(r / 8) * 128 + (uint32_t)(r % 8) * 16;
    const int ch = half * 128;                                  // this thread's column half in the hidden stages
    const int jfirst = half ? 1 + s.ga : 0;                     // this thread's dimensions in the spline stage
    const int jcount = half ? s.gb : 1 + s.ga;
    uint32_t acc_par = 0, accb_par = 0;
    bool ok = true;
    const bool prof = a.prof && blockIdx.x == 0 && tid == 0;
    long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // 0 consts, 1 hidden wait, 2 hidden compute, 3 F wait, 4 spline, 5 total
    long long tmark = clock64();
    const long long t_begin = tmark;
    auto mark = [&](int k) { if (prof) { const long long n = clock64(); pc[k] += n - tmark; tmark = n; } };
    auto arrive_ops = [&]() {
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(op_bar);
    };
    // write one value of the next layer's x tile: column jn (bf16 hi) and 16 + jn (bf16 lo)
    auto put_x = [&](int jn, float v, int l, int tile) {
      (void)l; (void)tile;
      const uint32_t o0 = (uint32_t)(jn / 8) * 2048 + rowoff + (jn % 8) * 2;
      if constexpr (SPLIT == 3) {
        const float xc = fminf(fmaxf(v, -60000.f), 60000.f);       // NaN -> -6e4 (the row is flagged `bad` elsewhere)
        const __half hi = __float2half_rn(xc);
        *reinterpret_cast<__half*>(sXT + o0) = hi;
        *reinterpret_cast<__half*>(sXT + 4096 + o0) = __float2half_rn(xc - __half2float(hi));
        if constexpr (SAVE)   // bf16 x tile of this layer for the initial layer's weight gradient
          *reinterpret_cast<__nv_bfloat16*>(a.sv.xt + ((size_t)l * a.sv.tiles + tile) * 4096 + o0) = __float2bfloat16(dn_finite(v));
      } else {
        const float xc = dn_finite(v);
        const __nv_bfloat16 hi = __float2bfloat16(xc);
        const __nv_bfloat16 lo = __float2bfloat16(xc - __bfloat162float(hi));
        *reinterpret_cast<__nv_bfloat16*>(sXT + o0) = hi;
        *reinterpret_cast<__nv_bfloat16*>(sXT + 4096 + o0) = lo;
      }
    };
    constexpr float ASC = SPLIT == 3 ? 1.f / DN_WSCALE : 1.f;    // accumulators of the split tier carry 16 W
    int tile = blockIdx.x;
    for (; tile < a.n_tiles && ok; tile += gridDim.x) {
      const int64_t grow = (int64_t)tile * DN_ROWS + r;
      const bool valid = grow < a.R;
      const int64_t first_row = (int64_t)tile * DN_ROWS;
      const int64_t last_row = first_row + DN_ROWS - 1 < a.R ? first_row + DN_ROWS - 1 : a.R - 1;
      const bool ctx_staged = !a.cm.row2ctx && a.cm.n_per_ctx >= DN_ROWS;
      const int64_t e0 = a.cm(first_row), e1 = a.cm(last_row);
      const int64_t cidx = valid ? a.cm(grow) : e0;
      const int esel = (ctx_staged && cidx != e0) ? 1 : 0;
      const float* projr = a.proj + cidx * a.ldproj;
      float* yin = ybuf + r * 12;
      float* yout = ybuf + DN_ROWS * 12 + r * 12;
      float ld_acc = 0.f;
      bool bad = false;
      if (half == 0) {
        for (int j = 0; j < D; ++j) {
          const float v = valid ? a.x[grow * D + j] : 0.f;
          yin[j] = v;
          put_x(D - 1 - j, v, 0, tile);
          if constexpr (SAVE) a.sv.XS[((size_t)tile * D + j) * DN_ROWS + r] = v;
        }
      }
      arrive_ops();                                             // x tile of layer 0
      for (int l = 0; l < s.L && ok; ++l) {
        if constexpr (SAVE) {   // training: per-row gates come from global memory -- ask the L2 for the NEXT layer's block
          if (rt == 0 && l + 1 < s.L) {
            const char* pj = reinterpret_cast<const char*>(a.proj + ((size_t)tile * a.ldproj + (size_t)(l + 1) * 3 * DN_H) * DN_ROWS);
            for (int o = 0; o < 3 * DN_H * DN_ROWS * 4; o += 65536) bulk_prefetch_l2(pj + o, 65536);
          }
        }
        // ---- this layer's constants -> shared memory (all 256 row threads) ----
        bar_sync(1, 256);                                       // previous layer's readers are done
        {
          const float* srcs[5] = {a.b0 + (size_t)l * DN_H, a.ba + ((size_t)l * 2) * DN_H, a.bb + ((size_t)l * 2) * DN_H,
                                  a.ba + ((size_t)l * 2 + 1) * DN_H, a.bb + ((size_t)l * 2 + 1) * DN_H};
          float v[5], c[6];
#pragma unroll
          for (int i = 0; i < 5; ++i) v[i] = __ldg(srcs[i] + rt);
          if (ctx_staged) {
#pragma unroll
            for (int i = 0; i < 6; ++i)
              c[i] = __ldg(a.proj + (i < 3 ? e0 : e1) * a.ldproj + ((size_t)l * 3 + (i % 3)) * DN_H + rt);
          }
          float bfv[3];
#pragma unroll
          for (int i = 0; i < 3; ++i) {
            const int e = rt + 256 * i, j = e / 48, k = e % 48;
            bfv[i] = (j < D && k < DN_MULT) ? __ldg(a.bf + (size_t)l * s.PLD + j * DN_MULT + k) : 0.f;
          }
#pragma unroll
          for (int i = 0; i < 5; ++i) cst[DN_CST_BIAS + i * DN_H + rt] = v[i];
          if (ctx_staged) {   // fold the biases into the event's context terms: one add / one fma per element later
#pragma unroll
            for (int ev = 0; ev < 2; ++ev) {
              float* d = cst + DN_CST_CTX + ev * 5 * DN_H + rt;
              d[0] = c[3 * ev] + v[0];
              d[DN_H] = c[3 * ev + 1];
              d[2 * DN_H] = c[3 * ev + 1] * v[2];
              d[3 * DN_H] = c[3 * ev + 2];
              d[4 * DN_H] = c[3 * ev + 2] * v[4];
            }
          }
#pragma unroll
          for (int i = 0; i < 3; ++i) cst[DN_CST_BF + rt + 256 * i] = bfv[i];
        }
        bar_sync(1, 256);
        mark(0);
        const float* cctx = cst + DN_CST_CTX + esel * 5 * DN_H;
        const float* pl = projr + (size_t)l * 3 * DN_H;
        // SAVE (training): the projection is tile-transposed [tile][L*3*H][128]: column c of this row at plt[c * 128]
        const float* plt = a.proj + ((size_t)tile * a.ldproj + (size_t)l * 3 * DN_H) * DN_ROWS + r;
        (void)plt;
        // ---- five hidden stages: I, A1, B1, A2, B2 ----
#pragma unroll 1
        for (int st = 0; st < 5 && ok; ++st) {
          if (!(ok = wait(acc_bar, acc_par, 10 + st))) break;
          mark(1);
          acc_par ^= 1;
          tc_fence_after();
          const bool res = (st == 0 || st == 2 || st == 4);          // stage writes the residual stream h
          const float* bias = cst + DN_CST_BIAS + st * DN_H;          // b0 | ba0 | bb0 | ba1 | bb1
          const int sec = st == 0 ? 0 : (st == 2 ? 1 : 2);            // context section used by this stage
          // staged: add = relu(c0)+b0 (stage I) or gate*bias (B stages), mul = gate
          const float* cadd = cctx + (st == 0 ? 0 : (st == 2 ? 2 : 4)) * DN_H;
          const float* cmul = cctx + (st == 2 ? 1 : 3) * DN_H;
#pragma unroll 1
          for (int b = 0; b < 4; ++b) {
            const int c0 = ch + 32 * b;
            uint32_t v[32];
            float o[32];
            tmem_ld32(tl + DN_HACC + c0, v);
            if (!res) {                       // A stages: u = acc + ba
              tmem_wait_ld();
#pragma unroll
              for (int i = 0; i < 32; i += 4) {
                const float4 t4 = *reinterpret_cast<const float4*>(bias + c0 + i);
                o[i] = fmaf(__uint_as_float(v[i]), ASC, t4.x); o[i + 1] = fmaf(__uint_as_float(v[i + 1]), ASC, t4.y);
                o[i + 2] = fmaf(__uint_as_float(v[i + 2]), ASC, t4.z); o[i + 3] = fmaf(__uint_as_float(v[i + 3]), ASC, t4.w);
              }
            } else {
              if (st == 0) {                  // h0 = acc + b0 + relu(c0)
                tmem_wait_ld();
                if (ctx_staged) {
#pragma unroll
                  for (int i = 0; i < 32; i += 4) {
                    const float4 t4 = *reinterpret_cast<const float4*>(cadd + c0 + i);
                    o[i] = fmaf(__uint_as_float(v[i]), ASC, t4.x); o[i + 1] = fmaf(__uint_as_float(v[i + 1]), ASC, t4.y);
                    o[i + 2] = fmaf(__uint_as_float(v[i + 2]), ASC, t4.z); o[i + 3] = fmaf(__uint_as_float(v[i + 3]), ASC, t4.w);
                  }
                } else {
#pragma unroll
                  for (int i = 0; i < 32; i += 4) {
                    const float4 t4 = *reinterpret_cast<const float4*>(bias + c0 + i);
                    float4 g4;
                    if constexpr (SAVE) {
                      g4 = make_float4(__ldg(plt + (size_t)(c0 + i) * DN_ROWS), __ldg(plt + (size_t)(c0 + i + 1) * DN_ROWS),
                                       __ldg(plt + (size_t)(c0 + i + 2) * DN_ROWS), __ldg(plt + (size_t)(c0 + i + 3) * DN_ROWS));
                    } else {
                      g4 = __ldg(reinterpret_cast<const float4*>(pl + c0 + i));
                    }
                    o[i] = fmaf(__uint_as_float(v[i]), ASC, t4.x) + g4.x; o[i + 1] = fmaf(__uint_as_float(v[i + 1]), ASC, t4.y) + g4.y;
                    o[i + 2] = fmaf(__uint_as_float(v[i + 2]), ASC, t4.z) + g4.z; o[i + 3] = fmaf(__uint_as_float(v[i + 3]), ASC, t4.w) + g4.w;
                  }
                }
              } else {                        // h += gate * (acc + bb)
                uint32_t hr[32];
                tmem_ld32(tl + DN_HRES + c0, hr);
                tmem_wait_ld();
                if (ctx_staged) {
#pragma unroll
                  for (int i = 0; i < 32; i += 4) {
                    const float4 g4 = *reinterpret_cast<const float4*>(cmul + c0 + i);
                    const float4 a4 = *reinterpret_cast<const float4*>(cadd + c0 + i);
                    o[i] = fmaf(g4.x * ASC, __uint_as_float(v[i]), __uint_as_float(hr[i]) + a4.x);
                    o[i + 1] = fmaf(g4.y * ASC, __uint_as_float(v[i + 1]), __uint_as_float(hr[i + 1]) + a4.y);
                    o[i + 2] = fmaf(g4.z * ASC, __uint_as_float(v[i + 2]), __uint_as_float(hr[i + 2]) + a4.z);
                    o[i + 3] = fmaf(g4.w * ASC, __uint_as_float(v[i + 3]), __uint_as_float(hr[i + 3]) + a4.w);
                  }
                } else {
#pragma unroll
                  for (int i = 0; i < 32; i += 4) {
                    const float4 t4 = *reinterpret_cast<const float4*>(bias + c0 + i);
                    float4 g4;
                    if constexpr (SAVE) {
                      const float* pg = plt + (size_t)(sec * DN_H + c0 + i) * DN_ROWS;
                      g4 = make_float4(__ldg(pg), __ldg(pg + DN_ROWS), __ldg(pg + 2 * DN_ROWS), __ldg(pg + 3 * DN_ROWS));
                    } else {
                      g4 = __ldg(reinterpret_cast<const float4*>(pl + sec * DN_H + c0 + i));
                    }
                    o[i] = fmaf(g4.x, fmaf(__uint_as_float(v[i]), ASC, t4.x), __uint_as_float(hr[i]));
                    o[i + 1] = fmaf(g4.y, fmaf(__uint_as_float(v[i + 1]), ASC, t4.y), __uint_as_float(hr[i + 1]));
                    o[i + 2] = fmaf(g4.z, fmaf(__uint_as_float(v[i + 2]), ASC, t4.z), __uint_as_float(hr[i + 2]));
                    o[i + 3] = fmaf(g4.w, fmaf(__uint_as_float(v[i + 3]), ASC, t4.w), __uint_as_float(hr[i + 3]));
                  }
                }
              }
              uint32_t ow[32];
#pragma unroll
              for (int i = 0; i < 32; ++i) ow[i] = __float_as_uint(o[i]);
              tmem_st16(tl + DN_HRES + c0, ow);
              tmem_st16(tl + DN_HRES + c0 + 16, ow + 16);
            }
            // next stage's A operand: relu fused into the bf16 conversion, except after the last block, where the
            // final layer takes h itself.  (Rows are independent in the MMA: a non-finite value stays in its row,
            // which the `bad` flag turns into NaN as the reference does.)
            if constexpr (SPLIT == 3) {
              uint8_t* gimg = nullptr;
              if constexpr (SAVE) gimg = a.sv.img + dn_img_off(l, st, tile, a.sv.tiles);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                uint4 w4, l4;
                if (st == 4) {
                  split_pack<false>(o[8 * i], o[8 * i + 1], w4.x, l4.x); split_pack<false>(o[8 * i + 2], o[8 * i + 3], w4.y, l4.y);
                  split_pack<false>(o[8 * i + 4], o[8 * i + 5], w4.z, l4.z); split_pack<false>(o[8 * i + 6], o[8 * i + 7], w4.w, l4.w);
                } else {
                  split_pack<true>(o[8 * i], o[8 * i + 1], w4.x, l4.x); split_pack<true>(o[8 * i + 2], o[8 * i + 3], w4.y, l4.y);
                  split_pack<true>(o[8 * i + 4], o[8 * i + 5], w4.z, l4.z); split_pack<true>(o[8 * i + 6], o[8 * i + 7], w4.w, l4.w);
                }
                const size_t io = (size_t)(c0 / 8 + i) * 2048 + rowoff;
                *reinterpret_cast<uint4*>(sACT + io) = w4;
                *reinterpret_cast<uint4*>(sACTLO + io) = l4;
                if constexpr (SAVE) {   // bf16 image of the same values: relu mask + weight-gradient operand of the backward
                  uint4 b4;
                  if (st == 4) {
                    b4.x = pack_bf16(o[8 * i], o[8 * i + 1]); b4.y = pack_bf16(o[8 * i + 2], o[8 * i + 3]);
                    b4.z = pack_bf16(o[8 * i + 4], o[8 * i + 5]); b4.w = pack_bf16(o[8 * i + 6], o[8 * i + 7]);
                  } else {
                    b4.x = pack_bf16_relu(o[8 * i], o[8 * i + 1]); b4.y = pack_bf16_relu(o[8 * i + 2], o[8 * i + 3]);
                    b4.z = pack_bf16_relu(o[8 * i + 4], o[8 * i + 5]); b4.w = pack_bf16_relu(o[8 * i + 6], o[8 * i + 7]);
                  }
                  *reinterpret_cast<uint4*>(gimg + io) = b4;
                }
              }
              if constexpr (SAVE) {
                if (st == 2 || st == 4) {   // v = acc + bb: the pre-gate block output (d gate = gh * v in the backward)
                  uint8_t* vhi = a.sv.img + dn_img_off(l, st == 2 ? DN_IMG_V1 : DN_IMG_V2, tile, a.sv.tiles);
                  uint8_t* vlo = a.sv.img + dn_img_off(l, st == 2 ? DN_IMG_V1L : DN_IMG_V2L, tile, a.sv.tiles);
#pragma unroll
                  for (int i = 0; i < 4; ++i) {
                    const float4 b0 = *reinterpret_cast<const float4*>(bias + c0 + 8 * i);
                    const float4 b1 = *reinterpret_cast<const float4*>(bias + c0 + 8 * i + 4);
                    uint4 w4, l4;
                    split_pack_bf16(fmaf(__uint_as_float(v[8 * i]), ASC, b0.x), fmaf(__uint_as_float(v[8 * i + 1]), ASC, b0.y), w4.x, l4.x);
                    split_pack_bf16(fmaf(__uint_as_float(v[8 * i + 2]), ASC, b0.z), fmaf(__uint_as_float(v[8 * i + 3]), ASC, b0.w), w4.y, l4.y);
                    split_pack_bf16(fmaf(__uint_as_float(v[8 * i + 4]), ASC, b1.x), fmaf(__uint_as_float(v[8 * i + 5]), ASC, b1.y), w4.z, l4.z);
                    split_pack_bf16(fmaf(__uint_as_float(v[8 * i + 6]), ASC, b1.z), fmaf(__uint_as_float(v[8 * i + 7]), ASC, b1.w), w4.w, l4.w);
                    const size_t io = (size_t)(c0 / 8 + i) * 2048 + rowoff;
                    *reinterpret_cast<uint4*>(vhi + io) = w4;
                    *reinterpret_cast<uint4*>(vlo + io) = l4;
                  }
                }
              }
            } else {
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                uint4 w4;
                if (st == 4) {
                  w4.x = pack_bf16(o[8 * i], o[8 * i + 1]); w4.y = pack_bf16(o[8 * i + 2], o[8 * i + 3]);
                  w4.z = pack_bf16(o[8 * i + 4], o[8 * i + 5]); w4.w = pack_bf16(o[8 * i + 6], o[8 * i + 7]);
                } else {
                  w4.x = pack_bf16_relu(o[8 * i], o[8 * i + 1]); w4.y = pack_bf16_relu(o[8 * i + 2], o[8 * i + 3]);
                  w4.z = pack_bf16_relu(o[8 * i + 4], o[8 * i + 5]); w4.w = pack_bf16_relu(o[8 * i + 6], o[8 * i + 7]);
                }
                *reinterpret_cast<uint4*>(sACT + (size_t)(c0 / 8 + i) * 2048 + rowoff) = w4;
              }
            }
          }
          if (res) tmem_wait_st();
          arrive_ops();
          mark(2);
        }
        if (!ok) break;
        // ---- final layer: 3*NB-1 spline parameters per dimension -> forward spline ----
        if (!(ok = wait(acc_bar, acc_par, 16))) break;       // group A (dims 1..ga)
        acc_par ^= 1;
        if (half == 1) {                                       // group B (dims ga+1..): only half 1 reads it
          if (!(ok = wait(accb_bar, accb_par, 17))) break;
        }
        tc_fence_after();
        mark(3);
        const float* bfc = cst + DN_CST_BF;
#pragma unroll 1
        for (int jj = 0; jj < jcount; ++jj) {
          const int j = jfirst + jj;
          float q[48];
#pragma unroll
          for (int i = 0; i < 48; i += 4) {
            const float4 t4 = *reinterpret_cast<const float4*>(bfc + j * 48 + i);
            q[i] = t4.x; q[i + 1] = t4.y; q[i + 2] = t4.z; q[i + 3] = t4.w;
          }
          if (j > 0) {
            const int col = half ? s.NA + (j - 1 - s.ga) * DN_MULT : (j - 1) * DN_MULT;
            uint32_t v[48];
            tmem_ld16(tl + col, v);
            tmem_ld16(tl + col + 16, v + 16);
            tmem_ld16(tl + col + 32, v + 32);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < DN_MULT; ++i) q[i] = fmaf(__uint_as_float(v[i]), ASC, q[i]);
          }
          const float xin = yin[D - 1 - j];
          bad = bad || !(fabsf(xin) <= 3.0e38f);
          if constexpr (SAVE) {   // row-minor layout: the 32 rows of a warp store one 128-byte line per parameter
            float* pr = a.sv.P + (((size_t)l * a.sv.tiles + tile) * s.PLD + j * DN_MULT) * DN_ROWS + r;
#pragma unroll
            for (int i = 0; i < DN_MULT; ++i) pr[(size_t)i * DN_ROWS] = q[i];
          }
          float lad;
          float y;
          // split tier, inference (log q at the FP32 tolerance): exact expf / logf / division.  bf16 tier and the training
          // forward (whose parameters are saved before this point and differentiated by the backward): hardware ex2 / lg2 / rcp
          if constexpr (SPLIT == 3 && !SAVE) y = rqs_apply<DN_NB, false>(DnRegQ{q}, xin, a.tail, lad);
          else y = dn_rqs_forward(q, xin, a.tail, lad);
          ld_acc += lad;
          yout[j] = y;
          if (l + 1 < s.L) {
            put_x(D - 1 - j, y, l + 1, tile);
            if constexpr (SAVE) a.sv.XS[(((size_t)(l + 1) * a.sv.tiles + tile) * D + j) * DN_ROWS + r] = y;
          }
        }
        if (half == 0) {                                       // keep the group-B barrier phase in step
          if (!(ok = wait(accb_bar, accb_par, 18))) break;
        }
        accb_par ^= 1;
        if (l + 1 < s.L) arrive_ops();                          // x tile of the next layer
        mark(4);
        float* tsw = yin; yin = yout; yout = tsw;
      }
      if (!ok) break;
      // ---- z, logdet: combine the two halves of every row ----
      if (half == 1) red[r] = ld_acc;
      red[(1 + half) * DN_ROWS + r] = bad ? 1.f : 0.f;
      bar_sync(2, 256);
      if (valid) {
        // a non-finite value that entered the conditioner turns the whole row into NaN, as the reference's does
        const bool anybad = red[DN_ROWS + r] != 0.f || red[2 * DN_ROWS + r] != 0.f;
        for (int jj = 0; jj < jcount; ++jj) {
          const int j = jfirst + jj;
          a.z[grow * D + j] = anybad ? __int_as_float(0x7fc00000) : yin[j];
        }
        if (half == 0 && a.logdet) a.logdet[grow] = anybad ? __int_as_float(0x7fc00000) : ld_acc + red[r];
      }
      bar_sync(2, 256);                                         // red / ybuf are reused by the next tile
    }
    if (!ok && half == 0) {   // barrier timeout (flag raised): abandoned and unreached tiles come back as NaN rows
      for (; tile < a.n_tiles; tile += gridDim.x) {
        const int64_t grow = (int64_t)tile * DN_ROWS + r;
        if (grow >= a.R) continue;
        for (int jj = 0; jj < D; ++jj) a.z[grow * D + jj] = __int_as_float(0x7fc00000);
        if (a.logdet) a.logdet[grow] = __int_as_float(0x7fc00000);
      }
    }
    if (prof) {
      for (int i = 0; i < 5; ++i) a.prof[i] = pc[i];
      a.prof[5] = clock64() - t_begin;
    }
  } else if (warp == 8) {
    // ============================ MMA issuer ============================
    uint32_t op_par = 0, slot = 0, full_par = 0;
    // split tier: fp16 x fp16 -> fp32 (a_format = b_format = 0); bf16 tier: bf16 x bf16 -> fp32
    constexpr uint32_t FMT = SPLIT == 3 ? ((1u << 7) | (1u << 10)) : 0u;   // bits to clear from the bf16 descriptor
    const uint32_t idesc_h = umma_idesc_bf16(DN_ROWS, 256) & ~FMT;
    const uint32_t idesc_a = umma_idesc_bf16(DN_ROWS, s.NA) & ~FMT, idesc_b = umma_idesc_bf16(DN_ROWS, s.NBn) & ~FMT;
    const uint32_t aX = smem_u32(sXT), aACT = smem_u32(sACT), aACTLO = smem_u32(sACTLO), aRing = smem_u32(ring);
    (void)aACTLO;
    bool ok = true;
    for (int tile = blockIdx.x; tile < a.n_tiles && ok; tile += gridDim.x) {
      for (int l = 0; l < s.L && ok; ++l) {
        for (int st = 0; st < 7 && ok; ++st) {               // I, A1, B1, A2, B2, FA, FB
          const int N = st < 5 ? 256 : (st == 5 ? s.NA : s.NBn);
          const int nch = st == 0 ? 2 : 16;
          const uint32_t idesc = st < 5 ? idesc_h : (st == 5 ? idesc_a : idesc_b);
          const uint32_t acc = tmem + (st == 6 ? (uint32_t)s.NA : 0u);
          const uint32_t abase = st == 0 ? aX : aACT;
          if (st != 6) {                                       // FB reads the same operand as FA
            if (!(ok = wait(&full[slot], full_par, 30))) break;   // weights first: prefetched long ago
            if (!(ok = wait(op_bar, op_par, 20 + st))) break;
            op_par ^= 1;
            tc_fence_after();
          }
          if constexpr (SPLIT == 3) {
            // one ring slot per K = 16 chunk: bf16 hi | lo of the weights; three MMAs: hi*hi + lo*hi + hi*lo
            const int nk = st == 0 ? 1 : 16;
            const uint32_t abase_lo = st == 0 ? aX + 4096u : aACTLO;
            for (int c0 = 0; c0 < nk && ok; ++c0) {
              if (!(ok = wait(&full[slot], full_par, 31))) break;
              tc_fence_after();
              const uint32_t bbase = aRing + slot * DN_SLOT;
              if (elect_one_sync()) {
                const uint64_t bh = umma_smem_desc(bbase, N * 16, 128);
                const uint64_t bl = umma_smem_desc(bbase + (uint32_t)N * 32u, N * 16, 128);
                const uint64_t ah = umma_smem_desc(abase + (uint32_t)c0 * 4096u, 2048, 128);
                const uint64_t al = umma_smem_desc(abase_lo + (uint32_t)c0 * 4096u, 2048, 128);
                umma_ss(acc, ah, bh, idesc, c0 > 0 ? 1u : 0u);
                umma_ss(acc, al, bh, idesc, 1u);
                umma_ss(acc, ah, bl, idesc, 1u);
                umma_commit(&empty[slot]);
                if (c0 + 1 >= nk) umma_commit(st == 6 ? accb_bar : acc_bar);
              }
              __syncwarp();
              if (++slot == DN_NSLOT) { slot = 0; full_par ^= 1; }
            }
          } else {
          for (int c0 = 0; c0 < nch && ok; c0 += 2) {
            if (!(ok = wait(&full[slot], full_par, 31))) break;
            tc_fence_after();
            const uint32_t bbase = aRing + slot * DN_SLOT;
            if (elect_one_sync()) {
              uint64_t bd = umma_smem_desc(bbase, N * 16, 128);
              const uint64_t bstep = (uint64_t)((2u * (uint32_t)N * 16u) >> 4);
              uint64_t ad = umma_smem_desc(abase + (uint32_t)c0 * 4096u, 2048, 128);
              umma_ss(acc, ad, bd, idesc, c0 > 0 ? 1u : 0u);
              ad += 256u; bd += bstep;
              umma_ss(acc, ad, bd, idesc, 1u);
              umma_commit(&empty[slot]);
              if (c0 + 2 >= nch) umma_commit(st == 6 ? accb_bar : acc_bar);
            }
            __syncwarp();
            if (++slot == DN_NSLOT) { slot = 0; full_par ^= 1; }
          }
          }
        }
      }
    }
  } else if (warp == 9) {
    // ============================ weight loader ============================
    uint32_t slot = 0, empty_par = 1;   // a fresh barrier passes a wait on parity 1
    bool ok = true;
    for (int tile = blockIdx.x; tile < a.n_tiles && ok; tile += gridDim.x) {
      for (int l = 0; l < s.L && ok; ++l) {
        const uint8_t* src = a.image + (size_t)l * (SPLIT == 3 ? s.layer_bytes3 : s.layer_bytes);
        for (int st = 0; st < 7 && ok; ++st) {
          const int N = st < 5 ? 256 : (st == 5 ? s.NA : s.NBn);
          // slots of this stage: SPLIT 1 = two K chunks per slot, SPLIT 3 = one (hi, lo) pair per slot; 2 * N * 32 bytes either way
          const int nch = SPLIT == 3 ? (st == 0 ? 2 : 32) : (st == 0 ? 2 : 16);
          const uint32_t bytes = 2u * (uint32_t)N * 32u;
          for (int c0 = 0; c0 < nch; c0 += 2) {
            if (!(ok = wait(&empty[slot], empty_par, 40))) break;
            if (elect_one_sync()) {
              mbar_arrive_expect_tx(&full[slot], bytes);
              bulk_g2s(ring + (size_t)slot * DN_SLOT, src, bytes, &full[slot]);
            }
            __syncwarp();
            src += bytes;
            if (++slot == DN_NSLOT) { slot = 0; empty_par ^= 1; }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc(tmem, 512);
}

#include "pf_tc_pair.cuh"
#include "pf_tc_pair2.cuh"

}  // namespace pf
using namespace pf;

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
#include "pf_gemm.cuh"

namespace pf {
// proj[e, n] = relu | sigmoid (acc + b[n]) in the handle's sorted hidden order (same as the FP32 tier's projection)
struct EpiCtxProjDense {
  float* out; int ld; const float* bias; int H;
  __device__ __forceinline__ void operator()(int m, int n, float acc) const {
    float v = acc + bias[n];
    int sec = (n / H) % 3;
    out[(int64_t)m * ld + n] = sec == 0 ? fmaxf(v, 0.f) : 1.f / (1.f + expf(-v));
  }
};
}  // namespace pf

static bool dn_plan(const pf_flow* f, DnShape& s) {
  memset(&s, 0, sizeof(s));
  if (f->H != DN_H || f->NB != DN_NB || f->D < 2 || f->D > 11 || f->desc.num_blocks != 2) return false;
  s.D = f->D; s.L = f->L; s.PLD = f->PLD; s.DP = f->DP;
  s.ga = (f->D - 1 + 1) / 2;
  s.gb = f->D - 1 - s.ga;
  s.NA = round_up(s.ga * DN_MULT, 16);
  s.NBn = s.gb > 0 ? round_up(s.gb * DN_MULT, 16) : 16;
  if (s.NA > 256 || s.NBn > 256 || s.NA + s.NBn > 512) return false;
  s.layer_bytes = 66u * 8192u + 16u * 32u * (uint32_t)(s.NA + s.NBn);
  s.layer_bytes3 = 65u * 16384u + 16u * 64u * (uint32_t)(s.NA + s.NBn);
  return true;
}

static bool dn_tier(const pf_flow* f) { return f->desc.tier == PF_TIER_BF16 || f->desc.tier == PF_TIER_BF16X3; }

bool pf_dn_supported(const pf_flow* f) {
  DnShape s;
  return dn_tier(f) && dn_plan(f, s);
}

pf_status pf_dn_pack(pf_flow* f, cudaStream_t st) {
  DnShape s;
  if (!dn_tier(f) || !dn_plan(f, s)) return PF_OK;   // the stepwise kernel (or the FP32 tier) serves other shapes
  DnState* t = reinterpret_cast<DnState*>(f->dense);
  if (!t) {
    t = new DnState();
    memset(t, 0, sizeof(*t));
    t->s = s;
    cudaError_t e = cudaMalloc(&t->image, (size_t)s.L * s.layer_bytes);
    if (e == cudaSuccess) e = cudaMalloc(&t->image3, (size_t)s.L * s.layer_bytes3);
    // CTA-pair form of the training forward (LeanNPE's D = 11; PF_TRAIN_PAIR=0 keeps the one-CTA kernel)
    const char* pp = getenv("PF_TRAIN_PAIR");
    const bool pair_on = s.D == 11 && f->sm_count >= 2 && !(pp && pp[0] == '0');
    if (e == cudaSuccess && pair_on) e = cudaMalloc(&t->imagep, (size_t)s.L * 2 * DP_LAYER_BYTES);
    const char* p2 = getenv("PF_DENSE_PAIR");
    const bool pair2_on = s.D == 11 && f->sm_count >= 2 && p2 && p2[0] == '1';   // opt-in: +4 % only (profiles/experiments/r02_dense_pair2.md)
    if (e == cudaSuccess && pair2_on) e = cudaMalloc(&t->image2, (size_t)s.L * 2 * D2_LAYER_BYTES);
    if (e != cudaSuccess) {
      if (t->image) cudaFree(t->image);
      if (t->image3) cudaFree(t->image3);
      if (t->imagep) cudaFree(t->imagep);
      delete t;
      return cuda_fail(e, "dense density image allocation", __FILE__, __LINE__);
    }
    const char* pr = getenv("PF_TC_PROF");
    if (pr && pr[0] == '1' && cudaMalloc(&t->prof, 16 * sizeof(long long)) != cudaSuccess) t->prof = nullptr;
    f->dense = t;
  }
  dim3 grid(98 + 97, s.L);
  dn_pack_kernel<<<grid, 256, 0, st>>>(s, t->image, t->image3, f->W0, f->Wa, f->Wb, f->Wf);
  PF_LAUNCH_CHECK(f);
  if (t->imagep) {
    dp_pack_kernel<<<dim3(DP_PACK_SLOTS, s.L, 2), 256, 0, st>>>(s, t->imagep, f->W0, f->Wa, f->Wb, f->Wf);
    PF_LAUNCH_CHECK(f);
  }
  if (t->image2) {
    d2_pack_kernel<<<dim3(2 + 7 * 16, s.L, 2), 256, 0, st>>>(s, t->image2, f->W0, f->Wa, f->Wb, f->Wf);
    PF_LAUNCH_CHECK(f);
  }
  PF_TRY(pf_tt_pack(f, st));   // backward images of the fused training step (pf_tc_train.cu)
  return PF_OK;
}

void pf_dn_destroy(pf_flow* f) {
  DnState* t = reinterpret_cast<DnState*>(f->dense);
  if (!t) return;
  pf_tt_destroy(f);
  cudaFree(t->image);
  cudaFree(t->image3);
  if (t->imagep) cudaFree(t->imagep);
  if (t->image2) cudaFree(t->image2);
  if (t->prof) cudaFree(t->prof);
  delete t;
  f->dense = nullptr;
}

template <int SPLIT, bool SAVE>
static pf_status dn_launch(pf_flow* f, cudaStream_t st, const DnArgs& a) {
  static bool attr_done[64] = {};
  const int dev = f->desc.device;
  if (dev < 64 && !attr_done[dev]) {
    PF_CUDA(cudaFuncSetAttribute(tc_dense_forward_kernel<SPLIT, SAVE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 DnL<SPLIT>::SMEM));
    attr_done[dev] = true;
  }
  const int grid = a.n_tiles < f->sm_count ? a.n_tiles : f->sm_count;
  tc_dense_forward_kernel<SPLIT, SAVE><<<grid, DN_THREADS, DnL<SPLIT>::SMEM, st>>>(a);
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}

static void dn_prof_report(pf_flow* f, DnState* t, cudaStream_t st) {
  if (!t->prof) return;
  long long h[16];
  if (cudaStreamSynchronize(st) != cudaSuccess) return;
  if (cudaMemcpy(h, t->prof, sizeof(h), cudaMemcpyDeviceToHost) != cudaSuccess) return;
  fprintf(stderr, "[pf_dn prof] CTA0: total %lld cycles: constants %lld | hidden wait %lld compute %lld | final wait %lld "
          "spline %lld | MMA warp (pair form): operands %lld own weights %lld peer weights %lld issue %lld\n", h[5], h[0], h[1],
          h[2], h[3], h[4], h[8], h[9], h[10], h[11]);
}

// x [R, D] -> z [R, D], logdet [R].  `err` = the handle's error flag.  split = 1 (bf16 operands) | 3 (split operands,
// exact spline: the FP32-tolerance tier).
pf_status pf_dn_forward(pf_flow* f, cudaStream_t st, const float* x, const float* ctx, const int64_t* row2ctx,
                        int64_t n_per_ctx, int64_t R, int64_t E, float* z, float* logdet, uint32_t flags,
                        void* workspace, size_t workspace_bytes, int* err, int split) {
  DnState* t = reinterpret_cast<DnState*>(f->dense);
  if (!t) return PF_ERR_NOT_PACKED;
  const DnShape& s = t->s;
  const bool is_proj = flags & PF_FLAG_CTX_IS_PROJ;
  const bool dense = !row2ctx && n_per_ctx == 1;
  const int64_t ldproj = (int64_t)f->L * 3 * DN_H;
  const int64_t chunk = dense ? (R < PF_ROW_CHUNK ? R : PF_ROW_CHUNK) : R;
  const int64_t prow = is_proj ? 0 : (dense ? chunk : E);
  if ((size_t)prow * ldproj * sizeof(float) > workspace_bytes || (prow && !workspace)) return PF_ERR_WORKSPACE;
  float* wproj = reinterpret_cast<float*>(workspace);
  auto project = [&](const float* c, int64_t rows) -> pf_status {
    const int N = f->L * 3 * DN_H;
    EpiCtxProjDense epi{wproj, N, f->bctx, DN_H};
    if (flags & PF_FLAG_SANITIZE_CTX)
      return sgemm_nt<EpiCtxProjDense, A_SANITIZE, 64>(f, st, c, f->C, f->Wctx, f->C, (int)rows, N, f->C, epi);
    return sgemm_nt<EpiCtxProjDense, A_PLAIN, 64>(f, st, c, f->C, f->Wctx, f->C, (int)rows, N, f->C, epi);
  };
  const float* proj_all = nullptr;
  if (is_proj) proj_all = ctx;
  else if (!dense) {
    PF_TRY(project(ctx, E));
    proj_all = wproj;
  }
  for (int64_t r0 = 0; r0 < R; r0 += chunk) {
    const int64_t rows = (R - r0) < chunk ? (R - r0) : chunk;
    DnArgs a;
    memset(&a, 0, sizeof(a));
    a.s = s;
    a.image = split == 3 ? t->image3 : t->image;
    a.b0 = f->b0; a.ba = f->ba; a.bb = f->bb; a.bf = f->bf;
    a.ldproj = ldproj;
    if (proj_all) {
      a.proj = proj_all;
      a.cm = CtxMap{row2ctx, n_per_ctx, r0};
    } else {
      PF_TRY(project(ctx + r0 * f->C, rows));
      a.proj = wproj;
      a.cm = CtxMap{nullptr, 1, 0};
    }
    a.x = x + r0 * s.D;
    a.R = rows;
    a.n_tiles = (int)((rows + DN_ROWS - 1) / DN_ROWS);
    a.z = z + r0 * s.D;
    a.logdet = logdet ? logdet + r0 : nullptr;
    a.tail = f->desc.tail_bound;
    a.err = err;
    a.prof = t->prof;
    if (split == 3) PF_TRY((dn_launch<3, false>(f, st, a)));
    else if (t->image2) {     // bf16 operands: CTA pairs, two tiles in flight per pair
      static bool attr_done[64] = {};
      const int dev = f->desc.device;
      if (dev < 64 && !attr_done[dev]) {
        PF_CUDA(cudaFuncSetAttribute(tc_dense_pair2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, D2_SMEM));
        attr_done[dev] = true;
      }
      a.image = t->image2;
      const int n_super = (a.n_tiles + 1) / 2;
      const int pairs = n_super < f->sm_count / 2 ? n_super : f->sm_count / 2;
      tc_dense_pair2_kernel<<<2 * pairs, D2_THREADS, D2_SMEM, st>>>(a);
      PF_LAUNCH_CHECK(f);
    } else PF_TRY((dn_launch<1, false>(f, st, a)));
  }
  dn_prof_report(f, t, st);
  return PF_OK;
}

// Forward of the fused training step: all R rows in ONE launch on the split operands, `proj` = the context projection of
// every row ([R][L*3*H], already computed), activations kept in `sv` for pf_tt_backward.
pf_status pf_dn_forward_save(pf_flow* f, cudaStream_t st, const float* x, const float* proj, int64_t R, float* z,
                             float* logdet, uint8_t* img, uint8_t* xt, float* P, float* XS) {
  DnState* t = reinterpret_cast<DnState*>(f->dense);
  if (!t) return PF_ERR_NOT_PACKED;
  const DnShape& s = t->s;
  DnArgs a;
  memset(&a, 0, sizeof(a));
  a.s = s;
  a.image = t->image3;
  a.b0 = f->b0; a.ba = f->ba; a.bb = f->bb; a.bf = f->bf;
  a.ldproj = (int64_t)f->L * 3 * DN_H;
  a.proj = proj;
  a.cm = CtxMap{nullptr, 1, 0};
  a.x = x;
  a.R = R;
  a.n_tiles = (int)((R + DN_ROWS - 1) / DN_ROWS);
  a.z = z;
  a.logdet = logdet;
  a.tail = f->desc.tail_bound;
  a.err = f->tg_err;
  a.prof = t->prof;
  a.sv = DnSave{img, xt, P, XS, R, a.n_tiles};
  if (t->imagep) {     // CTA pairs: two SMs per 128-row tile
    static bool attr_done[64] = {};
    const int dev = f->desc.device;
    if (dev < 64 && !attr_done[dev]) {
      PF_CUDA(cudaFuncSetAttribute(tc_dense_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DP_SMEM));
      attr_done[dev] = true;
    }
    a.image = t->imagep;
    const int pairs = a.n_tiles < f->sm_count / 2 ? a.n_tiles : f->sm_count / 2;
    tc_dense_pair_kernel<<<2 * pairs, DP_THREADS, DP_SMEM, st>>>(a);
    PF_LAUNCH_CHECK(f);
    dn_prof_report(f, t, st);
    return PF_OK;
  }
  PF_TRY((dn_launch<3, true>(f, st, a)));
  dn_prof_report(f, t, st);
  return PF_OK;
}

const uint8_t* pf_dn_image3(const pf_flow* f) {
  const DnState* t = reinterpret_cast<const DnState*>(f->dense);
  return t ? t->image3 : nullptr;
}
