// Fused training step of the tensor-core tiers, backward side (A12 of SURVEY.md 8a: what loss.backward() does through
// nflows at experiments/train_lean_npe.py:365, for the flow head):
//
//   tt_backward_kernel   ONE persistent kernel for the whole data-gradient chain.  A CTA keeps a 128-row tile on chip from
//                        the last layer to the first: per layer the spline VJP (fp32, in the epilogue) and seven tcgen05
//                        MMA stages  d h2 = d p . Wf  (two groups) -> d relu(u2) = (g2 d h2) . Wb2 -> d h1 += [h1>0] (. Wa2)
//                        -> block 1 likewise -> d x = d h0 . W0,  every product as hi*hi + lo*hi + hi*lo of bf16 pairs
//                        (gradients span many decades: bf16 keeps fp32's exponent range), fp32 accumulation and the fp32
//                        residual gradient in tensor memory.  Relu masks, gates and pre-gate values come from the images
//                        the forward saved; the gradient operand tiles it builds in shared memory are written out once
//                        (bulk copy shared -> global) as the images the weight-gradient kernel contracts.
//   tt_wgrad_kernel      one CTA per (layer, weight matrix) -- both 128-row blocks of the matrix in two accumulators against
//                        ONE fetch of the activations; the 288-wide context matrices stay one block per CTA --:
//                        D[out, in] = sum over ALL flow rows
//                        of G[row, out] A[row, in] -- both operands are the saved row-major tile images, fed to tcgen05.mma
//                        as MN-major operands (the contraction index is the tile row), bulk loads through two rings; bias
//                        gradients are the column sums of the same G tiles.  Plain read-modify-write stores into nflows'
//                        own weight layout: every output entry has exactly one owner CTA (no atomics).
#include <cuda_fp16.h>
#include <stdlib.h>

#include "pf_common.cuh"
#include "pf_sm100.cuh"
#include "pf_spline.cuh"
#include "pf_train.cuh"

namespace pf {
using namespace pf::sm100;

constexpr int TT_THREADS = 320;
constexpr int TT_NSLOT = 5;      // weight ring depth (16 KB slots): what shared memory leaves beside the two operand tiles
constexpr int TT_SLOT = 16384;
constexpr int TT_H = 256;
constexpr int TT_NB = 16;
constexpr int TT_MULT = 3 * TT_NB - 1;
constexpr int TT_PCOL = 48;      // columns a dimension's 47 parameter gradients take in an operand tile: 8-aligned, so a
                                 // thread writes them as six conflict-free 16-byte stores (the 48th is a zero pad)

struct TtShape {
  int D, L, PLD, DP;
  int ga, gb;        // dims 1..ga in parameter group A, ga+1..ga+gb in group B (dim 0 has no hidden input: bias only)
  int NA, NBn;       // padded parameter counts of the groups (multiples of 16, <= 256)
  uint32_t layer_bytes;
};

struct TtState {
  TtShape s;
  uint8_t* wimg;     // [L][layer_bytes] backward weight image
  uint8_t* wimgp;    // [L][2][slots][8 KB] pair form (pf_tc_train_pair.cuh), NULL when the pair kernels are off
  uint8_t* cfimg;    // context projection, forward: [L*3 column blocks][C/16 chunks][fp16 hi | lo of 16 W, 16 KB]
  uint8_t* cbimg;    // context projection, d ctx:   [L*3 blocks][16 chunks][bf16 hi | lo of W^T, (C x 16) x 2 x 2 B]
  int C, CK;         // context width, K = 16 chunks of it
};
constexpr int TT_CB_CHUNK = TT_CIMG_COLS * 16 * 2 * 2;    // 18 432 bytes: [288 x 16] bf16 hi | lo

// ------------------------------------------------------------------------------------------------
// backward weight image: 16 KB slots, each the bf16 (hi | lo) pair of one K = 16 chunk of a TRANSPOSED matrix in the
// UMMA K-major no-swizzle layout, in the order the backward consumes them:
//   FA^T  NA/16 slots   B[n = hidden unit][k = parameter of dims 1..ga]
//   FB^T  NBn/16 slots
//   Wb2^T, Wa2^T, Wb1^T, Wa1^T   16 slots each   B[n = input unit][k = output unit]
//   W0^T  1 slot = 16 chunks of ([16 x 16] hi 512 B | lo 512 B)   B[n = x column][k = hidden unit]
// ------------------------------------------------------------------------------------------------
__global__ void tt_pack_kernel(TtShape s, uint8_t* __restrict__ image, const float* __restrict__ WfT,
                               const float* __restrict__ WaT, const float* __restrict__ WbT, const float* __restrict__ W0T) {
  const int l = blockIdx.y, slot = blockIdx.x;
  const int nA = s.NA / 16, nB = s.NBn / 16;
  __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(image + (size_t)l * s.layer_bytes + (size_t)slot * TT_SLOT);
  const int H = TT_H;
  if (slot < nA + nB + 64) {
    int kind, c;   // 0 FA, 1 FB, 2.. hidden
    if (slot < nA) { kind = 0; c = slot; }
    else if (slot < nA + nB) { kind = 1; c = slot - nA; }
    else { kind = 2 + (slot - nA - nB) / 16; c = (slot - nA - nB) % 16; }
    for (int e = threadIdx.x; e < 256 * 16; e += blockDim.x) {
      const int n = e / 16, kk = e % 16, k = c * 16 + kk;
      float v = 0.f;
      if (kind <= 1) {
        const int first = kind == 0 ? 1 : 1 + s.ga, cnt = kind == 0 ? s.ga : s.gb;
        const int dim = k / TT_PCOL, i = k % TT_PCOL;
        if (dim < cnt && i < TT_MULT) v = WfT[((size_t)l * H + n) * s.PLD + (size_t)(first + dim) * TT_MULT + i];
      } else {
        const int hs = kind - 2;                    // 0 Wb2, 1 Wa2, 2 Wb1, 3 Wa1
        const int blk = hs < 2 ? 1 : 0;
        const float* W = (hs & 1) ? WaT : WbT;
        v = W[(((size_t)l * 2 + blk) * H + n) * H + k];
      }
      const int o = (kk / 8) * (256 * 8) + (n / 8) * 64 + (n % 8) * 8 + (kk % 8);
      const __nv_bfloat16 hi = __float2bfloat16(v);
      dst[o] = hi;
      dst[256 * 16 + o] = __float2bfloat16(v - __bfloat162float(hi));
    }
  } else {
    for (int e = threadIdx.x; e < 16 * 16 * 16; e += blockDim.x) {
      const int c = e / 256, n = (e / 16) % 16, kk = e % 16;
      const float v = n < s.D ? W0T[((size_t)l * s.DP + n) * H + c * 16 + kk] : 0.f;
      const int o = c * 512 + (kk / 8) * (16 * 8) + (n / 8) * 64 + (n % 8) * 8 + (kk % 8);
      const __nv_bfloat16 hi = __float2bfloat16(v);
      dst[o] = hi;
      dst[o + 256] = __float2bfloat16(v - __bfloat162float(hi));
    }
  }
}

// context-layer weight images.  Wctx is the handle's [L*3*H][C] array (rows = sorted hidden units, sections c0 | gate 1 |
// gate 2 per layer).  blockIdx.y = column block nb = l * 3 + section (256 rows of Wctx).
//   forward (blockIdx.x <  CK):      chunk kc of B[n = hidden unit][k = context column]: fp16 hi | lo of 16 W   (16 KB)
//   d ctx   (blockIdx.x >= CK, 16):  chunk kc of B[n = context column][k = hidden unit]: bf16 hi | lo of W      (18 KB)
__global__ void tt_pack_ctx_kernel(const float* __restrict__ Wctx, int C, int CK, uint8_t* __restrict__ cfimg,
                                   uint8_t* __restrict__ cbimg) {
  const int nb = blockIdx.y;
  if ((int)blockIdx.x < CK) {
    const int kc = blockIdx.x;
    __half* dst = reinterpret_cast<__half*>(cfimg + ((size_t)nb * CK + kc) * TT_SLOT);
    for (int e = threadIdx.x; e < 256 * 16; e += blockDim.x) {
      const int n = e / 16, kk = e % 16, c = kc * 16 + kk;
      const float v = c < C ? Wctx[((size_t)nb * TT_H + n) * C + c] * 16.f : 0.f;
      const float vs = fminf(fmaxf(v, -60000.f), 60000.f);
      const int o = (kk / 8) * (256 * 8) + (n / 8) * 64 + (n % 8) * 8 + (kk % 8);
      const __half hi = __float2half_rn(vs);
      dst[o] = hi;
      dst[256 * 16 + o] = __float2half_rn(vs - __half2float(hi));
    }
  } else {
    const int kc = blockIdx.x - CK;
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(cbimg + ((size_t)nb * 16 + kc) * TT_CB_CHUNK);
    for (int e = threadIdx.x; e < TT_CIMG_COLS * 16; e += blockDim.x) {
      const int n = e / 16, kk = e % 16;                     // n = context column, k = hidden unit kc * 16 + kk
      const float v = n < C ? Wctx[((size_t)nb * TT_H + kc * 16 + kk) * C + n] : 0.f;
      const int o = (kk / 8) * (TT_CIMG_COLS * 8) + (n / 8) * 64 + (n % 8) * 8 + (kk % 8);
      const __nv_bfloat16 hi = __float2bfloat16(v);
      dst[o] = hi;
      dst[TT_CIMG_COLS * 16 + o] = __float2bfloat16(v - __bfloat162float(hi));
    }
  }
}

// ------------------------------------------------------------------------------------------------
// the backward chain kernel
// ------------------------------------------------------------------------------------------------
struct TtArgs {
  TtShape s;
  const uint8_t* wimg;
  const float* proj; int64_t ldproj;    // [tiles][L*3*H][128]  tile-transposed (pf_train.cuh)
  const float* P; const float* XS;      // [L][tiles][PLD][128], [L][tiles][D][128]  saved by the forward
  const uint8_t* img;                   // [L][TT_NA][tiles][TT_IMG]
  const float* gz; const float* gl;     // upstream gradients [R, D], [R] (either may be NULL = 0)
  uint8_t* gimg;                        // [L][TT_NG][tiles][TT_IMG]
  float* gx;                            // [R, D] or NULL
  float* dbf[TT_MAX_L];                 // final_layer.bias gradient of each layer (dim 0's 47 entries) or NULL
  int64_t R; int n_tiles; float tail; int* err;
  long long* prof;                      // NULL, or 8 cycle counters filled by CTA 0 (PF_TC_PROF=1)
};

constexpr int TT_OFF_ACT = 0;
constexpr int TT_OFF_ACTLO = 65536;
constexpr int TT_OFF_RING = 131072;
constexpr int TT_OFF_GY = TT_OFF_RING + TT_NSLOT * TT_SLOT;     // [128][12] d loss / d layer output
constexpr int TT_OFF_GXD = TT_OFF_GY + TT_ROWS * 12 * 4;        // [128][12] direct d/dx of the spline
constexpr int TT_OFF_BAR = TT_OFF_GXD + TT_ROWS * 12 * 4;
constexpr int TT_SMEM = TT_OFF_BAR + 256;

__device__ __forceinline__ void bulk_s2g(void* gdst, const void* ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void split_bf16(float a, float b, uint32_t& hi, uint32_t& lo) {
  hi = pack_bf16(a, b);
  lo = pack_bf16(a - __uint_as_float(hi << 16), b - __uint_as_float(hi & 0xffff0000u));
}
__device__ __forceinline__ uint4 ldg128(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

struct TtRegQ {
  const float* q;
  __device__ __forceinline__ float operator()(int i) const { return q[i]; }
};
struct TtGradOut {
  float* g;
  __device__ __forceinline__ void operator()(int i, float v) const { g[i] = v; }
};

__global__ void __launch_bounds__(TT_THREADS, 1) tt_backward_kernel(const TtArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  const TtShape& s = a.s;
  const int D = s.D;
  uint8_t* sACT = smem + TT_OFF_ACT;
  uint8_t* sACTLO = smem + TT_OFF_ACTLO;
  uint8_t* ring = smem + TT_OFF_RING;
  float* gy_s = reinterpret_cast<float*>(smem + TT_OFF_GY);
  float* gxd_s = reinterpret_cast<float*>(smem + TT_OFF_GXD);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + TT_OFF_BAR);
  uint64_t* full = bars;                     // [NSLOT]
  uint64_t* empty = bars + TT_NSLOT;         // [NSLOT]
  uint64_t* acc_bar = bars + 2 * TT_NSLOT;   // MMAs of the current stage complete
  uint64_t* op_bar = acc_bar + 1;            // operand tile of the next stage written (8 warp arrivals)
  volatile int* abort_s = reinterpret_cast<volatile int*>(acc_bar + 2);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 3);

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  if (tid == 0) {
    for (int i = 0; i < TT_NSLOT; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(acc_bar, 1);
    mbar_init(op_bar, 8);
    *abort_s = 0;
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc(tmem_slot, 512);
  for (int i = tid; i < 131072 / 16; i += TT_THREADS) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
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
    // ============================ row warps: two threads per flow row ============================
    const int half = warp >> 2, q4 = warp & 3;
    const int r = q4 * 32 + lane;
    const int rt = half * 128 + r;
    const uint32_t tl = tmem + ((uint32_t)(q4 * 32) << 16);
    const uint32_t rowoff = (uint32_t)(r / 8) * 128 + (uint32_t)(r % 8) * 16;
    const int ch = half * 128;
    uint32_t acc_par = 0;
    bool ok = true;
    // cycle profile of CTA 0 (PF_TC_PROF=1): 0 spline VJP | 1 wait before the second parameter tile | 2 hidden waits |
    // 3 hidden epilogues | 4 wait + d x epilogue | 5 publish / reopen | 6 total
    const bool prof = a.prof && blockIdx.x == 0 && tid == 0;
    long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long tmark = clock64();
    const long long t_begin = tmark;
    auto mark = [&](int k) { if (prof) { const long long n = clock64(); pc[k] += n - tmark; tmark = n; } };
    // operand tile complete: make it visible to the async proxy, start the bulk copy of its hi half to `gdst` (the
    // gradient image the weight-gradient kernel reads), hand it to the MMA warp
    auto publish = [&](uint8_t* gdst) {
      fence_proxy_async_smem();
      bar_sync(3, 256);
      if (rt == 0) {
        bulk_s2g(gdst, sACT, TT_IMG);
        bulk_commit();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(op_bar);
    };
    // before the operand tile is overwritten: the MMAs reading it are complete (the caller waited on acc_bar) and the
    // bulk copy has finished reading shared memory
    auto reopen = [&]() {
      if (rt == 0) bulk_wait_read0();
      bar_sync(3, 256);
    };
    // 32 columns of this thread's row -> bf16 hi / lo tile images in global memory (the context-layer gradients: operands
    // of the weight-gradient kernel and of the d ctx product)
    auto store_image = [&](uint8_t* ghi, uint8_t* glo, int c0, const float* vals) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        uint4 w4, l4;
        split_bf16(vals[8 * i], vals[8 * i + 1], w4.x, l4.x); split_bf16(vals[8 * i + 2], vals[8 * i + 3], w4.y, l4.y);
        split_bf16(vals[8 * i + 4], vals[8 * i + 5], w4.z, l4.z); split_bf16(vals[8 * i + 6], vals[8 * i + 7], w4.w, l4.w);
        const size_t io = (size_t)(c0 / 8 + i) * 2048 + rowoff;
        *reinterpret_cast<uint4*>(ghi + io) = w4;
        *reinterpret_cast<uint4*>(glo + io) = l4;
      }
    };
    for (int tile = blockIdx.x; tile < a.n_tiles && ok; tile += gridDim.x) {
      const int64_t grow = (int64_t)tile * TT_ROWS + r;
      const bool valid = grow < a.R;
      const float glr = (valid && a.gl) ? a.gl[grow] : 0.f;
      if (half == 0)
        for (int j = 0; j < D; ++j) gy_s[r * 12 + j] = (valid && a.gz) ? a.gz[grow * D + j] : 0.f;
      bar_sync(3, 256);
      for (int l = s.L - 1; l >= 0 && ok; --l) {
        // row-minor layouts: column c of this thread's row sits at base[c * 128] (one 128-byte line per warp access)
        const float* projr = a.proj + ((size_t)tile * a.ldproj + (size_t)l * 3 * TT_H) * TT_ROWS + r;
        // one thread asks the L2 for what this tile reads next (contiguous blocks of the saved arrays): the spline
        // parameters of the layer below, and this layer's gates / pre-gate values / relu-mask images
        if (rt == 0) {
          if (l > 0) {
            const float* pn = a.P + ((size_t)(l - 1) * a.n_tiles + tile) * s.PLD * TT_ROWS;
            for (uint32_t o = 0; o < (uint32_t)s.PLD * TT_ROWS * 4u; o += 65536u) {
              const uint32_t n = (uint32_t)s.PLD * TT_ROWS * 4u - o;
              bulk_prefetch_l2(reinterpret_cast<const char*>(pn) + o, n < 65536u ? n : 65536u);
            }
            bulk_prefetch_l2(a.XS + ((size_t)(l - 1) * a.n_tiles + tile) * D * TT_ROWS, (uint32_t)D * TT_ROWS * 4u);
          }
          const char* pj = reinterpret_cast<const char*>(a.proj + ((size_t)tile * a.ldproj + (size_t)l * 3 * TT_H) * TT_ROWS);
          for (int o = 0; o < 3 * TT_H * TT_ROWS * 4; o += 65536) bulk_prefetch_l2(pj + o, 65536);
          for (int w = 0; w < TT_NA; ++w) {
            if (w == TT_A_H2) continue;
            bulk_prefetch_l2(a.img + tt_img_off(l, w, TT_NA, tile, a.n_tiles), TT_IMG);
          }
        }
        // ---------------- E0: spline VJP -> d params (two operand tiles: dims 1..ga, then ga+1..) ----------------
#pragma unroll 1
        for (int ps = 0; ps < 2 && ok; ++ps) {
          const int gfirst = ps == 0 ? 1 : 1 + s.ga, gcnt = ps == 0 ? s.ga : s.gb;
          const int h0n = (gcnt + 1) / 2;
          const int jbeg = gfirst + (half ? h0n : 0);
          const int jcnt = (half ? gcnt - h0n : h0n) + ((ps == 0 && half == 1) ? 1 : 0);   // + dim 0 (bias only)
          bool opened = false;
#pragma unroll 1
          for (int jj = 0; jj < jcnt; ++jj) {
            const bool is0 = (ps == 0 && half == 1 && jj == jcnt - 1);
            const int j = is0 ? 0 : jbeg + jj;
            float q[TT_MULT], gp[TT_MULT];
            const float* pr = a.P + (((size_t)l * a.n_tiles + tile) * s.PLD + j * TT_MULT) * TT_ROWS + r;
#pragma unroll
            for (int i = 0; i < TT_MULT; ++i) q[i] = __ldg(pr + (size_t)i * TT_ROWS);
            const float xin = a.XS[(((size_t)l * a.n_tiles + tile) * D + (D - 1 - j)) * TT_ROWS + r];
            const float gyj = gy_s[r * 12 + j];
            const float gxd = rqs_forward_backward_fast<TT_NB>(TtRegQ{q}, xin, a.tail, gyj, glr, TtGradOut{gp});
            gxd_s[r * 12 + j] = valid ? gxd : 0.f;
            if (is0) {
              // dim 0's parameters are bias only: their gradient goes straight to final_layer.bias[0:47]
              float* dst = a.dbf[l];
              float keep0 = 0.f, keep1 = 0.f;
#pragma unroll
              for (int i = 0; i < TT_MULT; ++i) {
                float v = valid ? gp[i] : 0.f;
#pragma unroll
                for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == (i & 31)) { if (i < 32) keep0 = v; else keep1 = v; }
              }
              if (dst) {
                atomicAdd(dst + lane, keep0);
                if (lane + 32 < TT_MULT) atomicAdd(dst + lane + 32, keep1);
              }
              continue;
            }
            if (!opened) {
              mark(0);
              if (ps == 1) {   // the MMAs of group A must be done with the operand tile
                if (!(ok = wait(acc_bar, acc_par, 50))) break;
                acc_par ^= 1;
              }
              mark(1);
              reopen();
              mark(5);
              opened = true;
            }
            const int cg0 = (j - gfirst) * (TT_PCOL / 8);       // first 8-column group of this dimension
#pragma unroll
            for (int c8 = 0; c8 < TT_PCOL / 8; ++c8) {
              float e[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) e[i] = (valid && 8 * c8 + i < TT_MULT) ? gp[8 * c8 + i < TT_MULT ? 8 * c8 + i : 0] : 0.f;
              uint4 w4, l4;
              split_bf16(e[0], e[1], w4.x, l4.x); split_bf16(e[2], e[3], w4.y, l4.y);
              split_bf16(e[4], e[5], w4.z, l4.z); split_bf16(e[6], e[7], w4.w, l4.w);
              const size_t io = (size_t)(cg0 + c8) * 2048 + rowoff;
              *reinterpret_cast<uint4*>(sACT + io) = w4;
              *reinterpret_cast<uint4*>(sACTLO + io) = l4;
            }
          }
          if (!ok) break;
          if (!opened) {     // a half without a parameter dim in this group (tiny D): still take part in the protocol
            if (ps == 1) {
              if (!(ok = wait(acc_bar, acc_par, 51))) break;
              acc_par ^= 1;
            }
            reopen();
          }
          mark(0);
          publish(a.gimg + tt_img_off(l, ps == 0 ? TT_G_PA : TT_G_PB, TT_NG, tile, a.n_tiles));
          mark(5);
        }
        if (!ok) break;
        // ---------------- E2 .. E6: the five hidden data-gradient epilogues ----------------
        // k = 0: d h2 complete -> g2 d h2 (operand), gate-2 gradient, residual := d h2
        // k = 1: [u2 > 0] acc  (operand)
        // k = 2: d h1 = residual + [h1 > 0] acc -> g1 d h1 (operand), gate-1 gradient, residual := d h1
        // k = 3: [u1 > 0] acc  (operand)
        // k = 4: d h0 = residual + [h0 > 0] acc (operand), relu(context) gradient
#pragma unroll 1
        for (int k = 0; k < 5 && ok; ++k) {
          if (!(ok = wait(acc_bar, acc_par, 52 + k))) break;
          acc_par ^= 1;
          tc_fence_after();
          mark(2);
          reopen();
          mark(5);
          const bool res_in = (k == 2 || k == 4), gate = (k == 0 || k == 2), has_mask = k >= 1;
          const int mimg = k == 1 ? TT_A_U2 : (k == 2 ? TT_A_H1 : (k == 3 ? TT_A_U1 : TT_A_H0));
          const uint8_t* mbase = a.img + tt_img_off(l, mimg, TT_NA, tile, a.n_tiles);
          const int vimg = k == 0 ? TT_A_V2 : TT_A_V1;
          const uint8_t* vbase = a.img + tt_img_off(l, vimg, TT_NA, tile, a.n_tiles);
          const uint8_t* vlbase = a.img + tt_img_off(l, vimg + 1, TT_NA, tile, a.n_tiles);
          const int sec = k == 0 ? 2 : (k == 2 ? 1 : 0);
#pragma unroll 1
          for (int b = 0; b < 4; ++b) {
            const int c0 = ch + 32 * b;
            uint32_t v[32], hr[32];
            tmem_ld32(tl + c0, v);
            if (res_in) tmem_ld32(tl + 256 + c0, hr);
            uint32_t mkw[16], vhw[16], vlw[16];      // 32 bf16 each: relu mask source, pre-gate value hi / lo
            float g[32];
            if (has_mask) {
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint4 t4 = ldg128(mbase + (size_t)(c0 / 8 + i) * 2048 + rowoff);
                mkw[4 * i] = t4.x; mkw[4 * i + 1] = t4.y; mkw[4 * i + 2] = t4.z; mkw[4 * i + 3] = t4.w;
              }
            }
            if (gate) {
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint4 h4 = ldg128(vbase + (size_t)(c0 / 8 + i) * 2048 + rowoff);
                const uint4 l4 = ldg128(vlbase + (size_t)(c0 / 8 + i) * 2048 + rowoff);
                vhw[4 * i] = h4.x; vhw[4 * i + 1] = h4.y; vhw[4 * i + 2] = h4.z; vhw[4 * i + 3] = h4.w;
                vlw[4 * i] = l4.x; vlw[4 * i + 1] = l4.y; vlw[4 * i + 2] = l4.z; vlw[4 * i + 3] = l4.w;
              }
            }
            if (gate || k == 4) {
#pragma unroll
              for (int i = 0; i < 32; ++i) g[i] = __ldg(projr + (size_t)(sec * TT_H + c0 + i) * TT_ROWS);
            }
            tmem_wait_ld();
            float o[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              float x = __uint_as_float(v[i]);
              if (has_mask) {
                const uint32_t w = mkw[i / 2];
                const uint32_t hbits = (i & 1) ? (w >> 16) : (w & 0xffffu);
                if ((hbits & 0x7fffu) == 0u) x = 0.f;
              }
              if (res_in) x += __uint_as_float(hr[i]);
              o[i] = x;
            }
            if (k == 0 || k == 2) {          // residual gradient stays in tensor memory
              uint32_t ow[32];
#pragma unroll
              for (int i = 0; i < 32; ++i) ow[i] = __float_as_uint(o[i]);
              tmem_st16(tl + 256 + c0, ow);
              tmem_st16(tl + 256 + c0 + 16, ow + 16);
            }
            if (gate) {                      // d gate (pre-sigmoid) = d h * v * g (1 - g); operand = g * d h
              float gp4[32];
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                const uint32_t wh = vhw[i / 2], wl = vlw[i / 2];
                const float vv = (i & 1) ? __uint_as_float(wh & 0xffff0000u) + __uint_as_float(wl & 0xffff0000u)
                                         : __uint_as_float(wh << 16) + __uint_as_float(wl << 16);
                gp4[i] = o[i] * vv * g[i] * (1.f - g[i]);
                o[i] *= g[i];
              }
              store_image(a.gimg + tt_img_off(l, TT_G_C0 + sec, TT_NG, tile, a.n_tiles),
                          a.gimg + tt_img_off(l, TT_G_C0L + sec, TT_NG, tile, a.n_tiles), c0, gp4);
            } else if (k == 4) {             // relu(context_layer) backward: g[] holds the relu'd context term
              float gc[32];
#pragma unroll
              for (int i = 0; i < 32; ++i) gc[i] = g[i] > 0.f ? o[i] : 0.f;
              store_image(a.gimg + tt_img_off(l, TT_G_C0, TT_NG, tile, a.n_tiles),
                          a.gimg + tt_img_off(l, TT_G_C0L, TT_NG, tile, a.n_tiles), c0, gc);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              uint4 w4, l4;
              split_bf16(o[8 * i], o[8 * i + 1], w4.x, l4.x); split_bf16(o[8 * i + 2], o[8 * i + 3], w4.y, l4.y);
              split_bf16(o[8 * i + 4], o[8 * i + 5], w4.z, l4.z); split_bf16(o[8 * i + 6], o[8 * i + 7], w4.w, l4.w);
              const size_t io = (size_t)(c0 / 8 + i) * 2048 + rowoff;
              *reinterpret_cast<uint4*>(sACT + io) = w4;
              *reinterpret_cast<uint4*>(sACTLO + io) = l4;
            }
          }
          if (k == 0 || k == 2) tmem_wait_st();
          mark(3);
          publish(a.gimg + tt_img_off(l, TT_G_V2 + k, TT_NG, tile, a.n_tiles));
          mark(5);
        }
        if (!ok) break;
        // ---------------- E7: d x of this layer -> d output of the layer below ----------------
        if (!(ok = wait(acc_bar, acc_par, 58))) break;
        acc_par ^= 1;
        tc_fence_after();
        if (half == 0) {
          uint32_t v[16];
          tmem_ld16(tl, v);
          tmem_wait_ld();
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            if (k < D) {
              const float gxin = __uint_as_float(v[k]) + gxd_s[r * 12 + k];
              gy_s[r * 12 + (D - 1 - k)] = gxin;          // undo the layer's ReversePermutation
              if (l == 0 && a.gx && valid) a.gx[grow * D + (D - 1 - k)] = gxin;
            }
          }
        }
        tc_fence_before();
        bar_sync(3, 256);
        mark(4);
      }
    }
    if (rt == 0) bulk_wait0();      // the last gradient image is in global memory before the CTA exits
    if (prof) {
      for (int i = 0; i < 6; ++i) a.prof[i] = pc[i];
      a.prof[6] = clock64() - t_begin;
    }
  } else if (warp == 8) {
    // ============================ MMA issuer ============================
    uint32_t op_par = 0, slot = 0, full_par = 0;
    const uint32_t idesc256 = umma_idesc_bf16(TT_ROWS, 256), idesc16 = umma_idesc_bf16(TT_ROWS, 16);
    const uint32_t aACT = smem_u32(sACT), aACTLO = smem_u32(sACTLO), aRing = smem_u32(ring);
    bool ok = true;
    for (int tile = blockIdx.x; tile < a.n_tiles && ok; tile += gridDim.x) {
      for (int l = s.L - 1; l >= 0 && ok; --l) {
        for (int m = 0; m < 7 && ok; ++m) {          // FA^T, FB^T, Wb2^T, Wa2^T, Wb1^T, Wa1^T, W0^T
          const int nk = m == 0 ? s.NA / 16 : (m == 1 ? s.NBn / 16 : 16);
          if (!(ok = wait(&full[slot], full_par, 30))) break;
          if (!(ok = wait(op_bar, op_par, 20 + m))) break;
          op_par ^= 1;
          tc_fence_after();
          if (m < 6) {
            for (int c = 0; c < nk && ok; ++c) {
              if (!(ok = wait(&full[slot], full_par, 31))) break;
              tc_fence_after();
              const uint32_t bbase = aRing + slot * TT_SLOT;
              if (elect_one_sync()) {
                const uint64_t bh = umma_smem_desc(bbase, 256 * 16, 128);
                const uint64_t bl = umma_smem_desc(bbase + 8192u, 256 * 16, 128);
                const uint64_t ah = umma_smem_desc(aACT + (uint32_t)c * 4096u, 2048, 128);
                const uint64_t al = umma_smem_desc(aACTLO + (uint32_t)c * 4096u, 2048, 128);
                umma_ss(tmem, ah, bh, idesc256, (m == 1 || c > 0) ? 1u : 0u);
                umma_ss(tmem, al, bh, idesc256, 1u);
                umma_ss(tmem, ah, bl, idesc256, 1u);
                umma_commit(&empty[slot]);
                if (c + 1 >= nk) umma_commit(acc_bar);
              }
              __syncwarp();
              if (++slot == TT_NSLOT) { slot = 0; full_par ^= 1; }
            }
          } else {
            const uint32_t bbase = aRing + slot * TT_SLOT;
            if (elect_one_sync()) {
              for (int c = 0; c < 16; ++c) {
                const uint64_t bh = umma_smem_desc(bbase + (uint32_t)c * 1024u, 16 * 16, 128);
                const uint64_t bl = umma_smem_desc(bbase + (uint32_t)c * 1024u + 512u, 16 * 16, 128);
                const uint64_t ah = umma_smem_desc(aACT + (uint32_t)c * 4096u, 2048, 128);
                const uint64_t al = umma_smem_desc(aACTLO + (uint32_t)c * 4096u, 2048, 128);
                umma_ss(tmem, ah, bh, idesc16, c > 0 ? 1u : 0u);
                umma_ss(tmem, al, bh, idesc16, 1u);
                umma_ss(tmem, ah, bl, idesc16, 1u);
              }
              umma_commit(&empty[slot]);
              umma_commit(acc_bar);
            }
            __syncwarp();
            if (++slot == TT_NSLOT) { slot = 0; full_par ^= 1; }
          }
        }
      }
    }
  } else if (warp == 9) {
    // ============================ weight loader ============================
    uint32_t slot = 0, empty_par = 1;
    bool ok = true;
    const int nslots = s.NA / 16 + s.NBn / 16 + 65;
    for (int tile = blockIdx.x; tile < a.n_tiles && ok; tile += gridDim.x) {
      for (int l = s.L - 1; l >= 0 && ok; --l) {
        const uint8_t* src = a.wimg + (size_t)l * s.layer_bytes;
        for (int i = 0; i < nslots; ++i) {
          if (!(ok = wait(&empty[slot], empty_par, 40))) break;
          if (elect_one_sync()) {
            mbar_arrive_expect_tx(&full[slot], TT_SLOT);
            bulk_g2s(ring + (size_t)slot * TT_SLOT, src, TT_SLOT, &full[slot]);
          }
          __syncwarp();
          src += TT_SLOT;
          if (++slot == TT_NSLOT) { slot = 0; empty_par ^= 1; }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc(tmem, 512);
}

#include "pf_tc_train_pair.cuh"

// ------------------------------------------------------------------------------------------------
// context rows [R, C] fp32 -> bf16 tile images [tiles][C/8 column groups][16 row groups][8 rows x 16 B]
// ------------------------------------------------------------------------------------------------
__global__ void tt_ctx_image_kernel(const float* __restrict__ ctx, int C, int64_t R, int tiles, uint8_t* __restrict__ cimg) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;      // (tile, column group, row in tile)
  const int ncg = TT_CIMG_COLS / 8;
  if (i >= (int64_t)tiles * ncg * TT_ROWS) return;
  const int rr = (int)(i % TT_ROWS), cg = (int)((i / TT_ROWS) % ncg), tile = (int)(i / ((int64_t)TT_ROWS * ncg));
  const int64_t row = (int64_t)tile * TT_ROWS + rr;
  float v[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = (row < R && cg * 8 + k < C) ? ctx[row * C + cg * 8 + k] : 0.f;
  uint4 w;
  w.x = pack_bf16(v[0], v[1]); w.y = pack_bf16(v[2], v[3]); w.z = pack_bf16(v[4], v[5]); w.w = pack_bf16(v[6], v[7]);
  *reinterpret_cast<uint4*>(cimg + (size_t)tile * TT_CIMG + (size_t)cg * 2048 + (rr / 8) * 128 + (rr % 8) * 16) = w;
}

// ------------------------------------------------------------------------------------------------
// the weight-gradient kernel
// ------------------------------------------------------------------------------------------------
struct TtWgArgs {
  TtShape s;
  const uint8_t* img;      // forward images
  const uint8_t* xt;       // x-tile images
  const uint8_t* gimg;     // gradient images
  const uint8_t* cimg;     // context images
  int C;
  int tiles;
  int64_t R;
  const int* perm;         // [H] sorted slot -> nflows unit
  int mn_swap;             // descriptor convention switch (PF_MN_SWAP; bring-up aid)
  int wide;                // job table: 1 = WG_JOBS_WIDE per layer (both gradient halves against ONE fetch of the activations)
  int* err;
  pf_layer_grads g[TT_MAX_L];
};

constexpr int WG_THREADS = 192;       // 4 epilogue / column-sum warps, MMA issuer, loader
constexpr int WG_NA = 2;              // ring of second operands (activation / x / context image of a tile)
constexpr int WG_NG = 3;              // ring of gradient blocks (two of them when a job has one block per tile)
constexpr int WG_GBYTES = 32768;      // [128 rows x 128 columns] of a gradient image
constexpr int WG_ABYTES = 65536;      // [128 rows x 256 columns] activation image
// wide job: 2 x 64 KB + 3 x 32 KB; context job: 2 x 72 KB + 2 x 32 KB
constexpr int WG_OFF_BAR = (WG_NA * WG_ABYTES + WG_NG * WG_GBYTES) > (WG_NA * TT_CIMG + 2 * WG_GBYTES)
                               ? (WG_NA * WG_ABYTES + WG_NG * WG_GBYTES) : (WG_NA * TT_CIMG + 2 * WG_GBYTES);
constexpr int WG_OFF_EPI = 128 * (TT_CIMG_COLS + 1) * 4;   // epilogue: row pointers behind the staged [128][N + 1] block
static_assert(WG_OFF_EPI + 128 * 12 <= WG_NA * TT_CIMG + 2 * WG_GBYTES, "epilogue staging must fit in the operand rings");
constexpr int WG_SMEM = WG_OFF_BAR + 256;
static_assert(WG_SMEM <= 232448, "shared memory of the weight-gradient kernel");
constexpr int WG_JOBS = 20;           // per layer: final 4 | 4 hidden matrices x 2 | initial 2 | 3 context layers x 2
// Wide table: final PA | final PB | 4 hidden matrices | initial (each: BOTH 128-column gradient blocks, two accumulators)
// | 3 context layers x 2.  96 KB of operands per 1 024 tensor cycles in the narrow form, 128 KB per 2 048 in the wide one, and
// 130 CTAs are one wave instead of 200 in two: 177 -> 159 us at 7 608 rows (PF_WGRAD_WIDE=0 keeps the narrow table).
constexpr int WG_JOBS_WIDE = 13;

// instruction descriptor, kind::f16: bf16 x bf16 -> fp32, BOTH operands MN-major (bits 15 / 16), dense
__host__ __device__ constexpr uint32_t umma_idesc_bf16_mn(int M, int N) {
  return umma_idesc_bf16(M, N) | (1u << 15) | (1u << 16);
}

__global__ void __launch_bounds__(WG_THREADS, 1) tt_wgrad_kernel(const TtWgArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  const TtShape& s = a.s;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + WG_OFF_BAR);
  uint64_t* afull = bars;                   // [WG_NA] second operand of a tile landed
  uint64_t* aempty = bars + WG_NA;          // [WG_NA] every MMA of the tile done (1 commit)
  uint64_t* gfull = bars + 2 * WG_NA;       // [WG_NG] gradient block landed
  uint64_t* gempty = gfull + WG_NG;         // [WG_NG] MMAs + column sums done with the block (1 commit + 4 warps)
  uint64_t* acc_bar = gempty + WG_NG;
  volatile int* abort_s = reinterpret_cast<volatile int*>(acc_bar + 1);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 2);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  // ---- decode the job ----
  // mat: 0 final | 1 Wb2 | 2 Wa2 | 3 Wb1 | 4 Wa1 | 5 W0 | 6 context_layer | 7 blocks.0.context_layer | 8 blocks.1.context_layer
  // mt0: first 128-column block of the gradient image, nh: blocks of this job (accumulator h at TMEM column 256 h)
  int l, mat, mt0, nh;
  bool is_pa = false;
  if (a.wide) {
    l = blockIdx.x / WG_JOBS_WIDE;
    const int j = blockIdx.x % WG_JOBS_WIDE;
    if (j < 2) { mat = 0; is_pa = j == 0; mt0 = 0; nh = 2; }
    else if (j < 7) { mat = j - 1; mt0 = 0; nh = 2; }
    else { mat = 6 + (j - 7) / 2; mt0 = (j - 7) & 1; nh = 1; }
  } else {
    l = blockIdx.x / WG_JOBS;
    const int jk = blockIdx.x % WG_JOBS;
    mat = jk < 4 ? 0 : 1 + (jk - 4) / 2;
    mt0 = jk < 4 ? (jk & 1) : ((jk - 4) & 1);
    is_pa = jk < 2;
    nh = 1;
  }
  const int gwhich = mat == 0 ? (is_pa ? TT_G_PA : TT_G_PB) : (mat <= 5 ? TT_G_V2 + (mat - 1) : TT_G_C0 + (mat - 6));
  const int awhich = mat == 0 ? TT_A_H2 : (mat == 1 ? TT_A_U2 : (mat == 2 ? TT_A_H1 : (mat == 3 ? TT_A_U1 : TT_A_H0)));
  const int N = mat == 5 ? 16 : (mat >= 6 ? TT_CIMG_COLS : 256);
  const uint32_t abytes = mat == 5 ? TT_XT : (mat >= 6 ? TT_CIMG : WG_ABYTES);
  const uint8_t* gsrc = a.gimg + tt_img_off(l, gwhich, TT_NG, 0, a.tiles) + (size_t)mt0 * WG_GBYTES;
  const uint8_t* asrc = mat == 5 ? a.xt + (size_t)l * a.tiles * TT_XT
                                 : (mat >= 6 ? a.cimg : a.img + tt_img_off(l, awhich, TT_NA, 0, a.tiles));
  const size_t astride = mat == 5 ? TT_XT : (mat >= 6 ? TT_CIMG : TT_IMG);
  const uint32_t aslot = mat >= 6 ? TT_CIMG : WG_ABYTES;      // ring geometry of this job
  const int ng = nh == 2 ? WG_NG : 2;
  uint8_t* const gring = smem + WG_NA * aslot;
  const int units = a.tiles * nh;                             // (tile, gradient block) pairs, tile-major

  if (tid == 0) {
    for (int i = 0; i < WG_NA; ++i) { mbar_init(&afull[i], 1); mbar_init(&aempty[i], 1); }
    for (int i = 0; i < WG_NG; ++i) { mbar_init(&gfull[i], 1); mbar_init(&gempty[i], 5); }
    mbar_init(acc_bar, 1);
    *abort_s = 0;
    fence_mbar_init();
  }
  if (warp == 4) tmem_alloc(tmem_slot, 512);
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

  if (warp < 4) {
    // ---- column sums of the gradient tiles (bias gradients), then the epilogue ----
    const int m = warp * 32 + lane;                 // column of this CTA's 128-column block = row of the weight block
    // column sums: thread (cg = m / 8, r8 = m % 8) reads row r8 of every 8-row core matrix of column group cg as one
    // 16-byte load (a quarter warp covers one 128-byte core matrix: conflict-free), keeps 8 partial column sums, and the
    // 8 threads of a column group are reduced by shuffles once per tile
    const int cg = m >> 3, r8 = m & 7;
    float bs[2][8];
#pragma unroll
    for (int i = 0; i < 8; ++i) bs[0][i] = bs[1][i] = 0.f;
    uint32_t st = 0, par = 0;
    bool ok = true;
    for (int u = 0; u < units && ok; ++u) {
      if (!(ok = wait(&gfull[st], par, 70))) break;
      const uint8_t* gb = gring + st * WG_GBYTES + (size_t)cg * 2048 + r8 * 16;
      float ps[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
      for (int rg = 0; rg < 16; ++rg) {
        const uint4 w = *reinterpret_cast<const uint4*>(gb + rg * 128);
        ps[0] += __uint_as_float(w.x << 16); ps[1] += __uint_as_float(w.x & 0xffff0000u);
        ps[2] += __uint_as_float(w.y << 16); ps[3] += __uint_as_float(w.y & 0xffff0000u);
        ps[4] += __uint_as_float(w.z << 16); ps[5] += __uint_as_float(w.z & 0xffff0000u);
        ps[6] += __uint_as_float(w.w << 16); ps[7] += __uint_as_float(w.w & 0xffff0000u);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&gempty[st]);        // shared memory is no longer read: the block may be refilled
      const bool second = nh == 2 && (u & 1);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float v = ps[i];
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        if (second) bs[1][i] += v; else bs[0][i] += v;
      }
      if (++st == (uint32_t)ng) { st = 0; par ^= 1; }
    }
    // thread m needs the sum of column m: column group m / 8, element m % 8 -- held (identically) by the 8 threads of
    // that group: pick it with a shuffle from lane (m / 8 % 4) * 8 of this warp ... the group lives in this very warp
    float bsum2[2] = {0.f, 0.f};
    {
      const int src_lane = ((m >> 3) & 3) * 8;      // any lane of column group m / 8 (all eight hold the same sums)
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float v = __shfl_sync(0xffffffffu, bs[h][i], src_lane);
          if ((m & 7) == i) bsum2[h] = v;
        }
    }
    if (ok) ok = wait(acc_bar, 0, 71);
    tc_fence_after();
    // ---- epilogue: accumulator -> shared memory (transposed to nflows' column order) -> coalesced read-modify-write ----
    // A thread owns one ROW of the weight block (its TMEM lane); written straight from there, every store of a warp hit 32
    // different rows of the gradient (32 sectors per instruction, and a permuted column on top).  The block is staged in
    // the (now idle) operand stages as S[row][column in nflows' order], row pitch N + 1 floats (conflict-free both ways),
    // then each warp walks whole rows with consecutive lanes on consecutive columns.
    const bool go = ok && !*abort_s;
    float* S = reinterpret_cast<float*>(smem);
    const int ldS = N + 1;
    float** rowptr = reinterpret_cast<float**>(smem + WG_OFF_EPI);
    int* roww = reinterpret_cast<int*>(smem + WG_OFF_EPI + 128 * sizeof(float*));
    for (int h = 0; h < nh; ++h) {
    const int mt = mt0 + h;
    const float bsum = h == 0 ? bsum2[0] : bsum2[1];
    bar_sync(1, 128);         // every epilogue warp is done reading the last operand block / the previous block's staging
    if (go) {
      const pf_layer_grads& g = a.g[l];
      const int gm = mt * 128 + m;                  // column within the gradient image
      // destination row / bias slot of this thread
      float* wrow = nullptr;
      float* bdst = nullptr;
      int wr = -1;                                  // row index in nflows' space (mask rule)
      if (mat == 0) {
        const int first = is_pa ? 1 : 1 + s.ga, cnt = is_pa ? s.ga : s.gb;
        if (gm / TT_PCOL < cnt && gm % TT_PCOL < TT_MULT) {
          wr = (first + gm / TT_PCOL) * TT_MULT + gm % TT_PCOL;
          if (g.final_w) wrow = g.final_w + (size_t)wr * TT_H;
          if (g.final_b) bdst = g.final_b + wr;
        }
      } else if (mat <= 4) {
        const int blk = mat <= 2 ? 1 : 0, j = (mat & 1) ? 1 : 0;      // Wb = linear_layers.1, Wa = linear_layers.0
        wr = a.perm[gm];
        if (g.blk_lin_w[blk][j]) wrow = g.blk_lin_w[blk][j] + (size_t)wr * TT_H;
        if (g.blk_lin_b[blk][j]) bdst = g.blk_lin_b[blk][j] + wr;
      } else if (mat == 5) {
        wr = a.perm[gm];
        if (g.initial_w) wrow = g.initial_w + (size_t)wr * s.D;
        if (g.initial_b) bdst = g.initial_b + wr;
      } else {
        wr = a.perm[gm];
        float* w = mat == 6 ? g.ctx_w : g.blk_ctx_w[mat - 7];
        float* b = mat == 6 ? g.ctx_b : g.blk_ctx_b[mat - 7];
        if (w) wrow = w + (size_t)wr * a.C;
        if (b) bdst = b + wr;
      }
      if (bdst) *bdst += bsum;
      rowptr[m] = wrow;
      // threshold of the mask comparison (see the write-back below); rows without a destination never pass it
      roww[m] = mat >= 6 ? 0 : (mat == 0 ? (wr >= 0 ? wr / TT_MULT : -1) : hidden_degree(wr, s.D));
      const uint32_t tl = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)h * 256u;
      float* srow = S + (size_t)m * ldS;
      for (int c0 = 0; c0 < N; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(tl + c0, v);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int n = c0 + i;
          const int dc = (mat <= 4) ? a.perm[n] : n;   // hidden input units: sorted slot -> nflows column
          srow[dc] = __uint_as_float(v[i]);
        }
      }
    }
    bar_sync(1, 128);
    if (go) {
      const int ncols = mat >= 6 ? a.C : (mat == 5 ? s.D : TT_H);
      // four rows per pass, every load of the pass issued before the first add: the read-modify-write is latency-bound
      // (one L2 / HBM round trip per pass instead of one per 32 columns)
      constexpr int RB = 4, KC = TT_CIMG_COLS / 32;       // 9 column steps cover the widest block (288)
      // mask rule as one comparison thr[row] >= cv[column] (nflows' MADE masks, pf_common.cuh rule_mask):
      //   hidden  deg(row) >= deg(col) | final  row / MULT >= deg(col) | initial  deg(row) >= col + 1 | context: no mask
      // cv depends on (lane, k) only: computed once, so no division and no branch sits between the loads of a pass
      int cv[KC];
#pragma unroll
      for (int k = 0; k < KC; ++k) {
        const int c = lane + 32 * k;
        cv[k] = c >= ncols ? 0x7fffffff : (mat >= 6 ? -1 : (mat == 5 ? c + 1 : hidden_degree(c, s.D)));
      }
      for (int r0 = warp * 32; r0 < warp * 32 + 32; r0 += RB) {
        float old[RB][KC];
        float* wrow[RB];
#pragma unroll
        for (int q = 0; q < RB; ++q) {
          wrow[q] = rowptr[r0 + q];
          const int thr = wrow[q] ? roww[r0 + q] : (int)0x80000000;
#pragma unroll
          for (int k = 0; k < KC; ++k) {
            const bool on = thr >= cv[k];
            old[q][k] = on ? __ldcg(wrow[q] + lane + 32 * k) : __int_as_float(0x7fc00000);
          }
        }
#pragma unroll
        for (int q = 0; q < RB; ++q) {
          const float* srow = S + (size_t)(r0 + q) * ldS;
          const int thr = wrow[q] ? roww[r0 + q] : (int)0x80000000;
#pragma unroll
          for (int k = 0; k < KC; ++k)
            if (thr >= cv[k]) wrow[q][lane + 32 * k] = old[q][k] + srow[lane + 32 * k];
        }
      }
    }
    }      // blocks of the job
  } else if (warp == 4) {
    // ---- MMA issuer ----
    const int NH = mat >= 6 ? N / 2 : N;              // the context image is 288 columns wide: two N = 144 instructions
    const uint32_t idesc = umma_idesc_bf16_mn(128, NH);
    const uint32_t lbo = a.mn_swap ? 2048u : 128u, sbo = a.mn_swap ? 128u : 2048u;
    uint32_t st = 0, par = 0;
    bool ok = true;
    for (int t = 0; t < a.tiles && ok; ++t) {
      const uint32_t as = (uint32_t)t & 1u;
      if (!(ok = wait(&afull[as], ((uint32_t)t >> 1) & 1u, 72))) break;
      const uint32_t ab = smem_u32(smem + as * aslot);
      for (int h = 0; h < nh && ok; ++h) {
        if (!(ok = wait(&gfull[st], par, 74))) break;
        tc_fence_after();
        const uint32_t gb = smem_u32(gring + st * WG_GBYTES);
        if (elect_one_sync()) {
          const uint32_t acc = tmem + (uint32_t)h * 256u;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint64_t ad = umma_smem_desc(gb + (uint32_t)ks * 256u, lbo, sbo);
            const uint64_t bd = umma_smem_desc(ab + (uint32_t)ks * 256u, lbo, sbo);
            umma_ss(acc, ad, bd, idesc, (t > 0 || ks > 0) ? 1u : 0u);
            if (mat >= 6) {     // columns NH .. 2 NH - 1: NH / 8 column groups further into the image
              const uint64_t bd2 = umma_smem_desc(ab + (uint32_t)(NH / 8) * 2048u + (uint32_t)ks * 256u, lbo, sbo);
              umma_ss(acc + (uint32_t)NH, ad, bd2, idesc, (t > 0 || ks > 0) ? 1u : 0u);
            }
          }
          umma_commit(&gempty[st]);
          if (h + 1 == nh) {
            umma_commit(&aempty[as]);
            if (t + 1 == a.tiles) umma_commit(acc_bar);
          }
        }
        __syncwarp();
        if (++st == (uint32_t)ng) { st = 0; par ^= 1; }
      }
    }
  } else {
    // ---- loader ----
    uint32_t st = 0, par = 1;
    bool ok = true;
    for (int t = 0; t < a.tiles && ok; ++t) {
      const uint32_t as = (uint32_t)t & 1u;
      if (!(ok = wait(&aempty[as], (((uint32_t)t >> 1) & 1u) ^ 1u, 73))) break;
      if (elect_one_sync()) {
        mbar_arrive_expect_tx(&afull[as], abytes);
        bulk_g2s(smem + as * aslot, asrc + (size_t)t * astride, abytes, &afull[as]);
      }
      __syncwarp();
      for (int h = 0; h < nh && ok; ++h) {
        if (!(ok = wait(&gempty[st], par, 75))) break;
        if (elect_one_sync()) {
          mbar_arrive_expect_tx(&gfull[st], WG_GBYTES);
          bulk_g2s(gring + st * WG_GBYTES, gsrc + (size_t)t * TT_IMG + (size_t)h * WG_GBYTES, WG_GBYTES, &gfull[st]);
        }
        __syncwarp();
        if (++st == (uint32_t)ng) { st = 0; par ^= 1; }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------
// context projection of the training rows on tcgen05 (replaces a [R, C] x [C, L*3*H] GEMM whose fp32 operands were
// converted on the fly): out[tile][n][row] = relu | sigmoid (ctx . Wctx^T + b), split fp16 operands as in the forward.
// One CTA per (tile, group of column blocks): the tile's context image stays in shared memory, the weight chunks stream
// through a ring, two accumulators alternate so that the epilogue of block j runs under the MMAs of block j + 1.
// ------------------------------------------------------------------------------------------------
struct TtCpArgs {
  const uint8_t* cimg16;   // [tiles][2][TT_CIMG] fp16 hi | lo context images (K-major operand layout)
  const uint8_t* cfimg;
  const float* bias;       // [L*3*H] sorted order
  float* out;              // [tiles][NB*256][128]
  int CK, NB, nb_per_cta, tiles;
  int* err;
};
constexpr int CP_THREADS = 320;      // 8 epilogue warps (two per TMEM lane quarter: column halves), MMA issuer, loader
constexpr int CP_OFF_RING = 2 * TT_CIMG;
constexpr int CP_NSLOT = 3;
constexpr int CP_OFF_BAR = CP_OFF_RING + CP_NSLOT * TT_SLOT;
constexpr int CP_SMEM = CP_OFF_BAR + 128;

__global__ void tt_ctx_image16_kernel(const float* __restrict__ ctx, int C, int64_t R, int tiles, uint8_t* __restrict__ cimg16) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int ncg = TT_CIMG_COLS / 8;
  if (i >= (int64_t)tiles * ncg * TT_ROWS) return;
  const int rr = (int)(i % TT_ROWS), cg = (int)((i / TT_ROWS) % ncg), tile = (int)(i / ((int64_t)TT_ROWS * ncg));
  const int64_t row = (int64_t)tile * TT_ROWS + rr;
  uint32_t h[4], l[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    float a = (row < R && cg * 8 + 2 * k < C) ? ctx[row * C + cg * 8 + 2 * k] : 0.f;
    float b = (row < R && cg * 8 + 2 * k + 1 < C) ? ctx[row * C + cg * 8 + 2 * k + 1] : 0.f;
    a = fminf(fmaxf(a, -60000.f), 60000.f);
    b = fminf(fmaxf(b, -60000.f), 60000.f);
    const __half2 hh = __floats2half2_rn(a, b);
    const float2 hf = __half22float2(hh);
    const __half2 ll = __floats2half2_rn(a - hf.x, b - hf.y);
    h[k] = *reinterpret_cast<const uint32_t*>(&hh);
    l[k] = *reinterpret_cast<const uint32_t*>(&ll);
  }
  const size_t o = (size_t)cg * 2048 + (rr / 8) * 128 + (rr % 8) * 16;
  *reinterpret_cast<uint4*>(cimg16 + ((size_t)tile * 2) * TT_CIMG + o) = make_uint4(h[0], h[1], h[2], h[3]);
  *reinterpret_cast<uint4*>(cimg16 + ((size_t)tile * 2 + 1) * TT_CIMG + o) = make_uint4(l[0], l[1], l[2], l[3]);
}

__global__ void __launch_bounds__(CP_THREADS, 1) tt_ctxproj_kernel(const TtCpArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* sA = smem;                       // hi image | lo image
  uint8_t* ring = smem + CP_OFF_RING;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + CP_OFF_BAR);
  uint64_t* full = bars;                    // [NSLOT]
  uint64_t* empty = bars + CP_NSLOT;        // [NSLOT]
  uint64_t* a_bar = bars + 2 * CP_NSLOT;    // context images landed
  uint64_t* acc_full = a_bar + 1;           // [2] accumulator complete
  uint64_t* acc_free = a_bar + 3;           // [2] accumulator drained by the epilogue (4 warp arrivals)
  volatile int* abort_s = reinterpret_cast<volatile int*>(a_bar + 5);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(a_bar + 6);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int tile = blockIdx.x;
  const int nb0 = blockIdx.y * a.nb_per_cta;
  const int nb1 = min(a.NB, nb0 + a.nb_per_cta);
  if (tid == 0) {
    for (int i = 0; i < CP_NSLOT; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(a_bar, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_free[i], 8); }
    *abort_s = 0;
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc(tmem_slot, 512);
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
    // ---- epilogue: two threads per row of the tile (column halves) ----
    const int q4 = warp & 3, half = warp >> 2;
    const int r = q4 * 32 + lane;
    const uint32_t tl = tmem + ((uint32_t)(q4 * 32) << 16);
    uint32_t par[2] = {0, 0};
    bool ok = true;
    for (int nb = nb0; nb < nb1 && ok; ++nb) {
      const int ab = (nb - nb0) & 1;
      if (!(ok = wait(&acc_full[ab], par[ab], 80))) break;
      par[ab] ^= 1;
      tc_fence_after();
      const int sec = nb % 3;
      float* dst = a.out + ((size_t)tile * a.NB * 256 + (size_t)nb * 256) * TT_ROWS + r;
      const float* bias = a.bias + (size_t)nb * 256;
#pragma unroll 1
      for (int c0 = half * 128; c0 < half * 128 + 128; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(tl + ab * 256 + c0, v);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float x = fmaf(__uint_as_float(v[i]), 1.f / 16.f, __ldg(bias + c0 + i));
          // sigmoid through the hardware ex2 / rcp (~2^-22 relative: far below the tier's tolerance)
          dst[(size_t)(c0 + i) * TT_ROWS] = sec == 0 ? fmaxf(x, 0.f) : fx_rcp(1.f + fx_ex2(-1.4426950408889634f * x));
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_free[ab]);
    }
  } else if (warp == 8) {
    // ---- MMA issuer ----
    const uint32_t idesc = umma_idesc_bf16(TT_ROWS, 256) & ~((1u << 7) | (1u << 10));     // fp16 operands
    const uint32_t aHi = smem_u32(sA), aLo = aHi + TT_CIMG, aRing = smem_u32(ring);
    uint32_t slot = 0, full_par = 0, free_par[2] = {1, 1};
    bool ok = wait(a_bar, 0, 81);
    tc_fence_after();
    for (int nb = nb0; nb < nb1 && ok; ++nb) {
      const int ab = (nb - nb0) & 1;
      if (!(ok = wait(&acc_free[ab], free_par[ab], 82))) break;
      free_par[ab] ^= 1;
      tc_fence_after();
      for (int kc = 0; kc < a.CK && ok; ++kc) {
        if (!(ok = wait(&full[slot], full_par, 83))) break;
        tc_fence_after();
        const uint32_t bbase = aRing + slot * TT_SLOT;
        if (elect_one_sync()) {
          const uint64_t bh = umma_smem_desc(bbase, 256 * 16, 128), bl = umma_smem_desc(bbase + 8192u, 256 * 16, 128);
          const uint64_t ah = umma_smem_desc(aHi + (uint32_t)kc * 4096u, 2048, 128);
          const uint64_t al = umma_smem_desc(aLo + (uint32_t)kc * 4096u, 2048, 128);
          const uint32_t acc = tmem + (uint32_t)ab * 256u;
          umma_ss(acc, ah, bh, idesc, kc > 0 ? 1u : 0u);
          umma_ss(acc, al, bh, idesc, 1u);
          umma_ss(acc, ah, bl, idesc, 1u);
          umma_commit(&empty[slot]);
          if (kc + 1 == a.CK) umma_commit(&acc_full[ab]);
        }
        __syncwarp();
        if (++slot == CP_NSLOT) { slot = 0; full_par ^= 1; }
      }
    }
  } else {
    // ---- loader ----
    if (elect_one_sync()) {
      mbar_arrive_expect_tx(a_bar, 2 * TT_CIMG);
      bulk_g2s(sA, a.cimg16 + (size_t)tile * 2 * TT_CIMG, 2 * TT_CIMG, a_bar);
    }
    __syncwarp();
    uint32_t slot = 0, empty_par = 1;
    bool ok = true;
    for (int nb = nb0; nb < nb1 && ok; ++nb) {
      const uint8_t* src = a.cfimg + (size_t)nb * a.CK * TT_SLOT;
      for (int kc = 0; kc < a.CK; ++kc) {
        if (!(ok = wait(&empty[slot], empty_par, 84))) break;
        if (elect_one_sync()) {
          mbar_arrive_expect_tx(&full[slot], TT_SLOT);
          bulk_g2s(ring + (size_t)slot * TT_SLOT, src + (size_t)kc * TT_SLOT, TT_SLOT, &full[slot]);
        }
        __syncwarp();
        if (++slot == CP_NSLOT) { slot = 0; empty_par ^= 1; }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------
// d ctx [R, C] = sum over (layer, context Linear) of G [R, 256] . W [256, C] on tcgen05: the gradient images (bf16
// hi / lo, written by the backward chain kernel) are the A operand, chunk by chunk; the transposed weight chunks stream
// beside them through one ring.  One CTA per (tile, slice of the L*3 images); slices accumulate atomically.
// ------------------------------------------------------------------------------------------------
struct TtDcArgs {
  const uint8_t* gimg;     // gradient images
  const uint8_t* cbimg;
  float* gctx;             // [R, C], zero-filled by the caller
  int L, C, tiles, imgs_per_cta;
  int64_t R;
  int* err;
};
constexpr int DC_THREADS = 192;
constexpr int DC_STAGE = 8192 + TT_CB_CHUNK;        // A hi 4 KB | A lo 4 KB | B 18 KB
constexpr int DC_NSTAGE = 8;
constexpr int DC_OFF_BAR = DC_NSTAGE * DC_STAGE;
constexpr int DC_SMEM = DC_OFF_BAR + 256;

__global__ void __launch_bounds__(DC_THREADS, 1) tt_dctx_kernel(const TtDcArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + DC_OFF_BAR);
  uint64_t* full = bars;
  uint64_t* empty = bars + DC_NSTAGE;
  uint64_t* acc_bar = bars + 2 * DC_NSTAGE;
  volatile int* abort_s = reinterpret_cast<volatile int*>(acc_bar + 1);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 2);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int tile = blockIdx.x;
  const int i0 = blockIdx.y * a.imgs_per_cta, i1 = min(a.L * 3, i0 + a.imgs_per_cta);
  if (tid == 0) {
    for (int i = 0; i < DC_NSTAGE; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(acc_bar, 1);
    *abort_s = 0;
    fence_mbar_init();
  }
  if (warp == 4) tmem_alloc(tmem_slot, 512);
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
  constexpr int NH = TT_CIMG_COLS / 2;      // two N = 144 instructions per chunk
  if (warp < 4) {
    // A thread owns one row of the tile (its TMEM lane): adding from there would touch 32 different rows of d ctx per
    // instruction.  The [128 x C] block is staged in the idle operand stages (row pitch C_pad + 1 floats) and added row by
    // row with consecutive lanes on consecutive columns.
    const bool ok = i0 < i1 && wait(acc_bar, 0, 90);
    tc_fence_after();
    bar_sync(1, 128);
    const bool go = ok && !*abort_s;
    float* S = reinterpret_cast<float*>(smem);
    constexpr int ldS = TT_CIMG_COLS + 1;
    static_assert(128 * ldS * 4 <= DC_OFF_BAR, "epilogue staging must fit in the operand stages");
    if (go) {
      const int r = warp * 32 + lane;
      const uint32_t tl = tmem + ((uint32_t)(warp * 32) << 16);
      float* srow = S + (size_t)r * ldS;
      for (int c0 = 0; c0 < TT_CIMG_COLS; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(tl + c0, v);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 16; ++i) srow[c0 + i] = __uint_as_float(v[i]);
      }
    }
    bar_sync(1, 128);
    if (go) {
      for (int r = warp; r < TT_ROWS; r += 4) {
        const int64_t row = (int64_t)tile * TT_ROWS + r;
        if (row >= a.R) break;
        const float* srow = S + (size_t)r * ldS;
        float* dst = a.gctx + row * a.C;
        for (int c = lane; c < a.C; c += 32) atomicAdd(dst + c, srow[c]);
      }
    }
  } else if (warp == 4) {
    const uint32_t idesc = umma_idesc_bf16(TT_ROWS, NH);
    uint32_t st = 0, par = 0;
    bool ok = true;
    for (int im = i0; im < i1 && ok; ++im) {
      for (int kc = 0; kc < 16 && ok; ++kc) {
        if (!(ok = wait(&full[st], par, 91))) break;
        tc_fence_after();
        const uint32_t base = smem_u32(smem + st * DC_STAGE);
        if (elect_one_sync()) {
          const uint64_t ah = umma_smem_desc(base, 2048, 128), al = umma_smem_desc(base + 4096u, 2048, 128);
          const uint32_t bb = base + 8192u;
          const uint32_t first = (im == i0 && kc == 0) ? 0u : 1u;
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const uint64_t bh = umma_smem_desc(bb + (uint32_t)h * (NH / 8) * 128u, TT_CIMG_COLS * 16, 128);
            const uint64_t bl = umma_smem_desc(bb + (uint32_t)TT_CIMG_COLS * 32u + (uint32_t)h * (NH / 8) * 128u, TT_CIMG_COLS * 16, 128);
            const uint32_t acc = tmem + (uint32_t)h * NH;
            umma_ss(acc, ah, bh, idesc, first);
            umma_ss(acc, al, bh, idesc, 1u);
            umma_ss(acc, ah, bl, idesc, 1u);
          }
          umma_commit(&empty[st]);
          if (im + 1 == i1 && kc == 15) umma_commit(acc_bar);
        }
        __syncwarp();
        if (++st == DC_NSTAGE) { st = 0; par ^= 1; }
      }
    }
  } else {
    uint32_t st = 0, par = 1;
    bool ok = true;
    for (int im = i0; im < i1 && ok; ++im) {
      const int l = im / 3, sec = im % 3;
      const uint8_t* ghi = a.gimg + tt_img_off(l, TT_G_C0 + sec, TT_NG, tile, a.tiles);
      const uint8_t* glo = a.gimg + tt_img_off(l, TT_G_C0L + sec, TT_NG, tile, a.tiles);
      const uint8_t* wsrc = a.cbimg + (size_t)im * 16 * TT_CB_CHUNK;
      for (int kc = 0; kc < 16; ++kc) {
        if (!(ok = wait(&empty[st], par, 92))) break;
        if (elect_one_sync()) {
          uint8_t* d = smem + st * DC_STAGE;
          mbar_arrive_expect_tx(&full[st], DC_STAGE);
          bulk_g2s(d, ghi + (size_t)kc * 4096, 4096, &full[st]);
          bulk_g2s(d + 4096, glo + (size_t)kc * 4096, 4096, &full[st]);
          bulk_g2s(d + 8192, wsrc + (size_t)kc * TT_CB_CHUNK, TT_CB_CHUNK, &full[st]);
        }
        __syncwarp();
        if (++st == DC_NSTAGE) { st = 0; par ^= 1; }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem, 512);
}

}  // namespace pf
using namespace pf;

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static bool tt_plan(const pf_flow* f, TtShape& s) {
  memset(&s, 0, sizeof(s));
  if (f->H != TT_H || f->NB != TT_NB || f->D < 2 || f->D > 11 || f->desc.num_blocks != 2 || f->L > TT_MAX_L) return false;
  if (f->desc.tier != PF_TIER_BF16 && f->desc.tier != PF_TIER_BF16X3) return false;
  s.D = f->D; s.L = f->L; s.PLD = f->PLD; s.DP = f->DP;
  s.ga = (f->D - 1 + 1) / 2;
  s.gb = f->D - 1 - s.ga;
  s.NA = round_up(s.ga * TT_PCOL, 16);
  s.NBn = s.gb > 0 ? round_up(s.gb * TT_PCOL, 16) : 16;
  if (s.NA > 256 || s.NBn > 256) return false;
  s.layer_bytes = (uint32_t)(s.NA / 16 + s.NBn / 16 + 65) * TT_SLOT;
  return true;
}

bool pf_tt_supported(const pf_flow* f) { return f->train != nullptr; }

pf_status pf_tt_pack(pf_flow* f, cudaStream_t st) {
  TtShape s;
  if (!tt_plan(f, s)) return PF_OK;
  TtState* t = reinterpret_cast<TtState*>(f->train);
  if (!t) {
    t = new TtState();
    memset(t, 0, sizeof(*t));
    t->s = s;
    t->C = f->C;
    t->CK = (f->C + 15) / 16;
    cudaError_t e = cudaMalloc(&t->wimg, (size_t)s.L * s.layer_bytes);
    if (e == cudaSuccess) e = cudaMalloc(&t->cfimg, (size_t)s.L * 3 * t->CK * TT_SLOT);
    if (e == cudaSuccess) e = cudaMalloc(&t->cbimg, (size_t)s.L * 3 * 16 * TT_CB_CHUNK);
    const char* pp = getenv("PF_TRAIN_PAIR");
    const bool pair_on = f->sm_count >= 2 && !(pp && pp[0] == '0');
    if (e == cudaSuccess && pair_on) e = cudaMalloc(&t->wimgp, (size_t)s.L * 2 * bp_layer_slots(s) * BP_SLOT);
    if (e != cudaSuccess) {
      if (t->wimg) cudaFree(t->wimg);
      if (t->cfimg) cudaFree(t->cfimg);
      if (t->cbimg) cudaFree(t->cbimg);
      delete t;
      return cuda_fail(e, "backward weight image allocation", __FILE__, __LINE__);
    }
    f->train = t;
  }
  dim3 grid(s.NA / 16 + s.NBn / 16 + 65, s.L);
  tt_pack_kernel<<<grid, 256, 0, st>>>(s, t->wimg, f->WfT, f->WaT, f->WbT, f->W0T);
  PF_LAUNCH_CHECK(f);
  if (t->wimgp) {
    tt_pack_pair_kernel<<<dim3(bp_layer_slots(s), s.L, 2), 256, 0, st>>>(s, t->wimgp, f->WfT, f->WaT, f->WbT, f->W0T);
    PF_LAUNCH_CHECK(f);
  }
  if (f->C <= TT_CIMG_COLS) {
    dim3 g2(t->CK + 16, s.L * 3);
    tt_pack_ctx_kernel<<<g2, 256, 0, st>>>(f->Wctx, f->C, t->CK, t->cfimg, t->cbimg);
    PF_LAUNCH_CHECK(f);
  }
  return PF_OK;
}

// context projection of R training rows into the tile-transposed layout; `scratch` holds the fp16 context images
pf_status pf_tt_ctxproj(pf_flow* f, cudaStream_t st, const float* ctx, int64_t R, float* out, uint8_t* scratch) {
  TtState* t = reinterpret_cast<TtState*>(f->train);
  if (!t || f->C > TT_CIMG_COLS) return PF_ERR_UNSUPPORTED;
  static bool attr_done[64] = {};
  const int dev = f->desc.device;
  if (dev < 64 && !attr_done[dev]) {
    PF_CUDA(cudaFuncSetAttribute(tt_ctxproj_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CP_SMEM));
    PF_CUDA(cudaFuncSetAttribute(tt_dctx_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DC_SMEM));
    attr_done[dev] = true;
  }
  const int tiles = (int)((R + TT_ROWS - 1) / TT_ROWS);
  {
    const int64_t n = (int64_t)tiles * (TT_CIMG_COLS / 8) * TT_ROWS;
    tt_ctx_image16_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(ctx, f->C, R, tiles, scratch);
    PF_LAUNCH_CHECK(f);
  }
  TtCpArgs a;
  memset(&a, 0, sizeof(a));
  a.cimg16 = scratch;
  a.cfimg = t->cfimg;
  a.bias = f->bctx;
  a.out = out;
  a.CK = t->CK;
  a.NB = f->L * 3;
  a.tiles = tiles;
  // Column blocks per CTA: one CTA per SM (shared memory), so the launch runs in ceil(CTAs / SMs) waves of
  // (blocks per CTA + a fixed part: context image load, pipeline fill, last epilogue ~ 0.3 blocks) each.  "About two
  // waves" used to give 300 CTAs for 60 tiles (and for 20): 2.03 waves = three, the last one with 4 CTAs.  Take the
  // split with the lowest modelled time instead.
  int best_nb = a.NB;
  {
    double best = 1e30;
    for (int nb = 1; nb <= a.NB; ++nb) {
      const int g = (a.NB + nb - 1) / nb;
      const int64_t ctas = (int64_t)tiles * g;
      const int64_t waves = (ctas + f->sm_count - 1) / f->sm_count;
      const double cost = (double)waves * ((double)nb + 0.3);
      if (cost < best - 1e-9) { best = cost; best_nb = nb; }
    }
  }
  a.nb_per_cta = best_nb;
  const int groups = (a.NB + a.nb_per_cta - 1) / a.nb_per_cta;
  a.err = f->tg_err;
  tt_ctxproj_kernel<<<dim3(tiles, groups), CP_THREADS, CP_SMEM, st>>>(a);
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}
size_t pf_tt_ctxproj_scratch(int64_t R) { return (size_t)((R + TT_ROWS - 1) / TT_ROWS) * 2 * TT_CIMG; }

pf_status pf_tt_dctx(pf_flow* f, cudaStream_t st, const TtWork& ws, int64_t R, float* gctx) {
  TtState* t = reinterpret_cast<TtState*>(f->train);
  if (!t || f->C > TT_CIMG_COLS) return PF_ERR_UNSUPPORTED;
  TtDcArgs a;
  memset(&a, 0, sizeof(a));
  a.gimg = ws.gimg;
  a.cbimg = t->cbimg;
  a.gctx = gctx;
  a.L = f->L; a.C = f->C;
  a.tiles = (int)((R + TT_ROWS - 1) / TT_ROWS);
  a.R = R;
  int slices = f->sm_count / a.tiles;                      // one wave: at most one CTA per SM
  if (slices < 1) slices = 1;
  if (slices > f->L * 3) slices = f->L * 3;
  a.imgs_per_cta = (f->L * 3 + slices - 1) / slices;
  slices = (f->L * 3 + a.imgs_per_cta - 1) / a.imgs_per_cta;
  a.err = f->tg_err;
  tt_dctx_kernel<<<dim3(a.tiles, slices), DC_THREADS, DC_SMEM, st>>>(a);
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}

void pf_tt_destroy(pf_flow* f) {
  TtState* t = reinterpret_cast<TtState*>(f->train);
  if (!t) return;
  cudaFree(t->wimg);
  if (t->wimgp) cudaFree(t->wimgp);
  cudaFree(t->cfimg);
  cudaFree(t->cbimg);
  delete t;
  f->train = nullptr;
}

pf_status pf_tt_backward(pf_flow* f, cudaStream_t st, const TtSaved& sv, const TtWork& ws, int64_t R, const float* gz,
                         const float* glogdet, float* gx, const pf_layer_grads* grads) {
  TtState* t = reinterpret_cast<TtState*>(f->train);
  if (!t) return PF_ERR_NOT_PACKED;
  static bool attr_done[64] = {};
  const int dev = f->desc.device;
  if (dev < 64 && !attr_done[dev]) {
    PF_CUDA(cudaFuncSetAttribute(tt_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TT_SMEM));
    PF_CUDA(cudaFuncSetAttribute(tt_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM));
    attr_done[dev] = true;
  }
  TtArgs a;
  memset(&a, 0, sizeof(a));
  a.s = t->s;
  a.wimg = t->wimg;
  a.proj = sv.proj;
  a.ldproj = (int64_t)f->L * 3 * TT_H;
  a.P = sv.P; a.XS = sv.XS; a.img = sv.img;
  a.gz = gz; a.gl = glogdet;
  a.gimg = ws.gimg;
  a.gx = gx;
  for (int l = 0; l < f->L; ++l) a.dbf[l] = grads ? grads[l].final_b : nullptr;
  a.R = R;
  a.n_tiles = (int)((R + TT_ROWS - 1) / TT_ROWS);
  a.tail = f->desc.tail_bound;
  a.err = f->tg_err;
  static long long* d_prof = nullptr;
  {
    const char* pr = getenv("PF_TC_PROF");
    if (pr && pr[0] == '1' && !d_prof && cudaMalloc(&d_prof, 8 * sizeof(long long)) != cudaSuccess) d_prof = nullptr;
  }
  a.prof = d_prof;
  if (t->wimgp) {      // CTA pairs: two SMs per 128-row tile
    static bool pattr_done[64] = {};
    if (dev < 64 && !pattr_done[dev]) {
      PF_CUDA(cudaFuncSetAttribute(tt_backward_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BP_SMEM));
      pattr_done[dev] = true;
    }
    a.wimg = t->wimgp;
    const int pairs = a.n_tiles < f->sm_count / 2 ? a.n_tiles : f->sm_count / 2;
    tt_backward_pair_kernel<<<2 * pairs, BP_THREADS, BP_SMEM, st>>>(a);
  } else {
    const int grid = a.n_tiles < f->sm_count ? a.n_tiles : f->sm_count;
    tt_backward_kernel<<<grid, TT_THREADS, TT_SMEM, st>>>(a);
  }
  PF_LAUNCH_CHECK(f);
  if (d_prof) {
    long long h[8];
    PF_CUDA(cudaStreamSynchronize(st));
    PF_CUDA(cudaMemcpy(h, d_prof, sizeof(h), cudaMemcpyDeviceToHost));
    fprintf(stderr, "[pf_tt prof] backward CTA0: total %lld cycles: spline VJP %lld | wait group A %lld | hidden wait %lld "
            "compute %lld | d x %lld | publish/reopen %lld\n", h[6], h[0], h[1], h[2], h[3], h[4], h[5]);
  }
  return PF_OK;
}

pf_status pf_tt_wgrad(pf_flow* f, cudaStream_t st, const TtSaved& sv, const TtWork& ws, int64_t R, const float* ctx,
                      const pf_layer_grads* grads) {
  TtState* t = reinterpret_cast<TtState*>(f->train);
  if (!t) return PF_ERR_NOT_PACKED;
  if (!grads) return PF_OK;
  if (f->C > TT_CIMG_COLS || !ctx) return PF_ERR_UNSUPPORTED;
  TtWgArgs a;
  memset(&a, 0, sizeof(a));
  a.s = t->s;
  a.img = sv.img; a.xt = sv.xt; a.gimg = ws.gimg; a.cimg = ws.cimg;
  a.C = f->C;
  a.tiles = (int)((R + TT_ROWS - 1) / TT_ROWS);
  {
    const int64_t n = (int64_t)a.tiles * (TT_CIMG_COLS / 8) * TT_ROWS;
    tt_ctx_image_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(ctx, f->C, R, a.tiles, ws.cimg);
    PF_LAUNCH_CHECK(f);
  }
  a.R = R;
  a.perm = f->d_perm;
  const char* sw = getenv("PF_MN_SWAP");
  a.mn_swap = sw && sw[0] == '1';
  a.err = f->tg_err;
  for (int l = 0; l < f->L; ++l) a.g[l] = grads[l];
  static const bool wide = [] { const char* e = getenv("PF_WGRAD_WIDE"); return !(e && e[0] == '0'); }();
  a.wide = wide ? 1 : 0;
  tt_wgrad_kernel<<<f->L * (wide ? WG_JOBS_WIDE : WG_JOBS), WG_THREADS, WG_SMEM, st>>>(a);
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}
