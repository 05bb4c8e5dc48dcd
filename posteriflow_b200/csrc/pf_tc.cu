// Tensor-core tier (PF_TIER_BF16) of the sampling direction: ONE persistent kernel per call runs the whole
// incremental autoregressive inverse of NSFPosteriorFlow.inverse (flows.py:620-655) + the LeanNPE epilogue
// (lean_npe.py:326-332) for tiles of 128 flow rows, with every conditioner matrix product on tcgen05.mma
// (bf16 operands, fp32 accumulation in tensor memory) and the rational-quadratic spline in fp32 registers.
//
// Per CTA (one per SM, 352 threads = 11 warps):
//   warps 0-7  "row" warps, TWO threads per flow row (warps 0-3 = half 0, 4-7 = half 1; both halves cover TMEM lanes
//              0..127, lane window = warp % 4).  They read accumulators with tcgen05.ld, apply bias / context / GLU /
//              residual / relu or the spline inverse, and write the next stage's A operand (bf16) back to tensor
//              memory (tcgen05.st) or shared memory.  Hidden stages: each half owns half of the group's bf16 words;
//              spline stage: half 1 turns the width logits into knots while half 0 searches the bin and solves.
//   warp 8     MMA issuer: stays converged, one elect.sync lane issues every tcgen05.mma + tcgen05.commit; the warp
//              also owns the TMEM allocation.
//   warp 9     weight loader: streams the pre-laid-out bf16 weight blobs (UMMA canonical K-major layout, built by
//              pf_tc_pack) from L2 into a 3-slot shared-memory ring with cp.async.bulk + mbarrier complete_tx.
//   warp 10    constants stager: up to 3 steps ahead, copies the step's biases and the tile's (<= 2) events' context
//              terms into shared memory so that no global-load latency sits on the dependency chain.
//
// Hidden units are sorted by autoregressive degree and every degree group is padded to GSP (even) slots, so
// a masked matrix restricted to "rows of degree d" is a dense [GSP x GSP*d] block: step d of a layer needs
// five skinny MMAs (N = 32, K = GSP*d) for the hidden group of degree d and one (N = 48, K = GSP*(d-1)) for
// the 3K-1 spline parameters of dimension d -- one conditioner pass of work per layer in total instead of
// nflows' D passes.  The five activation arrays a step must keep for later steps
// (relu(h0), relu(u1), relu(h1), relu(u2), h2: 5 x [128, 260] bf16) do not fit in shared memory, so three
// of them live in tensor memory and are fed to the MMA as the TMEM A operand; two live in shared memory.
#include <stdlib.h>

#include <vector>

#include "pf_common.cuh"
#include "pf_gemm.cuh"
#include "pf_sm100.cuh"
#include "pf_spline.cuh"
#include "pf_tail.cuh"

namespace pf {
using namespace pf::sm100;

constexpr int TC_ROWS = 128;      // rows per tile = UMMA M = TMEM lanes
constexpr int TC_NSLOT = 3;       // weight ring depth
constexpr int TC_THREADS = 352;   // 8 row warps (2 threads per row) + MMA warp + weight-loader warp + constants-stager warp
constexpr int TC_CB_BF = 48;      // per-step constants buffer (floats): bf[48] | 5 bias groups | W0[:, d-1] | 2 events x 3 ctx groups
__host__ __device__ constexpr int tc_gs4(int gsp) { return (gsp + 3) / 4 * 4; }
__host__ __device__ constexpr int tc_cb_floats(int gsp) { return TC_CB_BF + 12 * tc_gs4(gsp); }
constexpr int TC_NCB = 4;         // constants buffers in flight (the stager runs up to 3 steps ahead)

// stage order inside step d of a layer (d = 1..D; the hidden stages exist for d < D)
enum { ST_F = 0, ST_I = 1, ST_A1 = 2, ST_B1 = 3, ST_A2 = 4, ST_B2 = 5, ST_COUNT = 6 };

struct TcShape {
  int D, G, GSP, KTOT, KA, MULT, NF, L, PLD;
  int pc_f;              // K=16 chunks per ring item of the final-layer blobs
  uint32_t slot_bytes;   // ring slot size
  uint32_t layer_bytes;  // weight image bytes per layer
  int off[PF_MAX_D + 1]; // off[g] = #hidden units of degree <= g (unpadded, degree-sorted order)
};
__host__ __device__ inline int tc_nchunks(const TcShape& s, int d, int st) {
  if (st == ST_F) return (s.GSP * (d - 1) + 15) / 16;
  if (st == ST_I) return (d >= 2 && d <= s.G) ? 2 : 0;   // I(1) has no MMA: its only input x_1 is added in fp32
  return (s.GSP * d + 15) / 16;
}
__host__ __device__ inline int tc_n(const TcShape& s, int st) { return st == ST_F ? s.NF : 32; }
__host__ __device__ inline int tc_piece(const TcShape& s, int st, int nch) { return st == ST_F ? s.pc_f : nch; }

struct TcItem { int d, st, c0, nc, N; uint32_t off; };

struct TcState {
  TcShape shape;
  uint8_t* image;        // [L][layer_bytes] bf16 blobs in program order
  float* arena;          // padded-order fp32 side arrays
  float *b0p, *bap, *bbp;   // [L][KTOT], [L][2][KTOT], [L][2][KTOT]
  float *Wctxp, *bctxp;     // [L*3*KTOT][C], [L*3*KTOT]
  TcItem* d_items; int n_items;
  int kernel_smem;
  bool debug_sync;
  long long* prof;       // device, 16 counters (PF_TC_PROF=1)
};

// ------------------------------------------------------------------------------------------------
// packing
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ int tc_slot_to_sorted(const TcShape& s, int slot) {
  int g = slot / s.GSP, i = slot % s.GSP;  // g = degree - 1
  if (g >= s.G) return -1;
  int n = s.off[g + 1] - s.off[g];
  return i < n ? s.off[g] + i : -1;
}

// padded-order side arrays of one layer: dst[row][slot] = src[row][sorted(slot)] or 0
struct TcPadArgs {
  const float *b0, *ba, *bb, *bctx, *Wctx;   // packed (degree-sorted) fp32 arrays of the handle
  float *b0p, *bap, *bbp, *bctxp, *Wctxp;    // padded-slot copies
  int H, C;
};
constexpr int TC_PAD_BLOCKS = 48;

// blockIdx.y = layer.  Blocks [0, n_items): one weight blob each (UMMA canonical K-major layout, program order).
// Blocks [n_items, n_items + TC_PAD_BLOCKS): the layer's slices of the padded side arrays (biases, context weights).
__global__ void tc_pack_image_kernel(TcShape s, const TcItem* __restrict__ items, int n_items, uint8_t* __restrict__ image,
                                     const float* __restrict__ W0, int DP, const float* __restrict__ Wa,
                                     const float* __restrict__ Wb, const float* __restrict__ Wf, int H, TcPadArgs pa) {
  const int l = blockIdx.y;
  if ((int)blockIdx.x >= n_items) {
    const int KT = s.KTOT, C = pa.C;
    // vector rows of this layer: b0 (1 row), ba (2), bb (2), bctx (3) = 8 rows of KT; then 3*KT context-weight rows of C
    const int nvec = 8 * KT;
    const int64_t total = (int64_t)nvec + (int64_t)3 * KT * C;
    for (int64_t i = (int64_t)(blockIdx.x - n_items) * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)TC_PAD_BLOCKS * blockDim.x) {
      if (i < nvec) {
        const int row = (int)(i / KT), col = tc_slot_to_sorted(s, (int)(i % KT));
        const float* src; float* dst; int srow, nrow;
        if (row < 1) { src = pa.b0; dst = pa.b0p; srow = row; nrow = 1; }
        else if (row < 3) { src = pa.ba; dst = pa.bap; srow = row - 1; nrow = 2; }
        else if (row < 5) { src = pa.bb; dst = pa.bbp; srow = row - 3; nrow = 2; }
        else { src = pa.bctx; dst = pa.bctxp; srow = row - 5; nrow = 3; }
        const size_t r = (size_t)l * nrow + srow;
        dst[r * KT + i % KT] = col >= 0 ? src[r * pa.H + col] : 0.f;
      } else {
        const int64_t k = i - nvec;
        const int c = (int)(k % C);
        const int rs = (int)(k / C);                   // sec * KT + slot
        const int sec = rs / KT, col = tc_slot_to_sorted(s, rs % KT);
        const size_t srow = ((size_t)l * 3 + sec);
        pa.Wctxp[(srow * KT + rs % KT) * C + c] = col >= 0 ? pa.Wctx[(srow * pa.H + col) * C + c] : 0.f;
      }
    }
    return;
  }
  const TcItem it = items[blockIdx.x];
  __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(image + (size_t)l * s.layer_bytes + it.off);
  const int K = it.nc * 16, N = it.N;
  const int gsz = it.d <= s.G ? s.off[it.d] - s.off[it.d - 1] : 0;
  for (int e = threadIdx.x; e < N * K; e += blockDim.x) {
    int n = e / K, k = e % K;
    float v = 0.f;
    if (it.st == ST_F) {
      int col = tc_slot_to_sorted(s, it.c0 * 16 + k);
      if (n < s.MULT && col >= 0) v = Wf[((size_t)l * s.PLD + (size_t)(it.d - 1) * s.MULT + n) * H + col];
    } else if (it.st == ST_I) {
      int j = k < 16 ? k : k - 16;  // columns 16.. multiply the low-order bf16 half of x
      // column d-1 (x_d itself) is left out: the epilogue adds W0[g, d-1] * x_d in fp32 (hoisted I stage)
      if (n < gsz && j < s.D && j != it.d - 1) v = W0[((size_t)l * H + s.off[it.d - 1] + n) * DP + j];
    } else {
      int blk = (it.st - ST_A1) / 2;
      const float* W = ((it.st - ST_A1) & 1) ? Wb : Wa;
      int col = tc_slot_to_sorted(s, it.c0 * 16 + k);
      if (n < gsz && col >= 0) v = W[(((size_t)l * 2 + blk) * H + s.off[it.d - 1] + n) * H + col];
    }
    dst[(k / 8) * (N * 8) + (n / 8) * 64 + (n % 8) * 8 + (k % 8)] = __float2bfloat16(v);
  }
}

// ------------------------------------------------------------------------------------------------
// the sampler kernel
// ------------------------------------------------------------------------------------------------
struct TcArgs {
  TcShape s;
  const uint8_t* image;
  const float *b0p, *bap, *bbp, *bf;
  const float* W0; int DP, H;   // fp32 initial-layer weights [L][H sorted][DP] (x_d column of the hoisted I stage)
  const float* proj;   // [ctx rows][L*3*KTOT] padded order: relu(c0) | sigmoid gate 1 | sigmoid gate 2
  int64_t ldproj;
  CtxMap cm;
  const float* z;      // [R, D] or NULL -> Philox
  uint64_t seed;
  int64_t R;
  int64_t row_base;    // global index of local row 0 (Philox counter)
  int n_tiles;
  float* x; float* logdet; uint8_t* railed;
  uint32_t flags; float tail, clamp;
  ScalerDev sc;
  int* err;
  int dbg_stop;        // >= 0: leave after this many MMA stages (bring-up aid, PF_TC_STOP)
  long long* prof;     // NULL, or 16 cycle counters filled by CTA 0 (PF_TC_PROF=1)
};

// TMEM column map (512 columns allocated)
template <int GSP>
struct TcTmem {
  static constexpr int NW = GSP / 2;         // 32-bit columns per group
  // columns per activation array (G <= 10 groups).  The TMEM A operand of tcgen05.mma must start at a
  // multiple of 4 columns (measured: tests/tmem_align_probe.cu), tcgen05.ld/st accept any column.
  static constexpr int ARR = (10 * NW + 3) / 4 * 4;
  static constexpr int A0 = 0, A1 = ARR, A2 = 2 * ARR, PAD = 3 * ARR;
  static constexpr int KCOLS = (10 * GSP + 15) / 16 * 8;  // columns the widest K loop reads from one array
  static constexpr int ACC_F = (A2 + KCOLS + 15) / 16 * 16;  // zero pad columns between the last array and ACC_F
  static constexpr int ACC_H = ACC_F + 48;
  static constexpr int ACC_I = ACC_H + 32;
  static_assert(ACC_I + 32 <= 512, "TMEM budget");
};

__device__ __forceinline__ float finite_clamp(float v) { return fminf(fmaxf(v, -3.0e38f), 3.0e38f); }  // NaN -> -3e38
__device__ __forceinline__ float relu_finite(float v) { return fminf(fmaxf(v, 0.f), 3.0e38f); }       // NaN -> 0

template <int NW>
__device__ __forceinline__ void tmem_store_words(uint32_t dst, const uint32_t* w) {
  static_assert(NW == 13 || NW == 10, "instantiated group widths");
  tmem_st8(dst, w);
  if (NW == 13) {
    tmem_st4(dst + 8, w + 8);
    tmem_st1(dst + 12, w + 12);
  } else {
    tmem_st2(dst + 8, w + 8);
  }
  tmem_wait_st();
}

// NWO consecutive bf16x2 words of one row into an operand array in the UMMA K-major no-swizzle layout (16 bytes of a row per
// 8-column group, groups 2048 bytes apart), starting at word A4 of a 16-byte row chunk: 16- / 8-byte stores wherever the
// alignment allows.  Word-by-word, every store of a warp hit 8 banks four times over (rows are 16 bytes apart): 7 stores =
// 28 shared-memory wavefronts per thread group against ~8 here.  All indices are compile-time (no local memory).
template <int A4, int NWO>
__device__ __forceinline__ void store_words_umma(uint8_t* p /* first word's address */, const uint32_t* w) {
  constexpr int HEAD = (4 - A4) < NWO ? (4 - A4) : NWO;
  if (HEAD == 4) *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
  else if (HEAD == 3) { *reinterpret_cast<uint32_t*>(p) = w[0]; *reinterpret_cast<uint2*>(p + 4) = make_uint2(w[1], w[2]); }
  else if (HEAD == 2) *reinterpret_cast<uint2*>(p) = make_uint2(w[0], w[1]);
  else *reinterpret_cast<uint32_t*>(p) = w[0];
  uint8_t* q = p - 4 * A4 + 2048;           // word 0 of the next column group
  constexpr int FULL = (NWO - HEAD) / 4, REM = (NWO - HEAD) % 4;
#pragma unroll
  for (int c = 0; c < FULL; ++c)
    *reinterpret_cast<uint4*>(q + 2048 * c) = make_uint4(w[HEAD + 4 * c], w[HEAD + 4 * c + 1], w[HEAD + 4 * c + 2], w[HEAD + 4 * c + 3]);
  uint8_t* r = q + 2048 * FULL;
  constexpr int I = HEAD + 4 * FULL;
  if (REM == 3) { *reinterpret_cast<uint2*>(r) = make_uint2(w[I], w[I + 1]); *reinterpret_cast<uint32_t*>(r + 8) = w[I + 2]; }
  else if (REM == 2) *reinterpret_cast<uint2*>(r) = make_uint2(w[I], w[I + 1]);
  else if (REM == 1) *reinterpret_cast<uint32_t*>(r) = w[I];
}

struct RegQ {
  const float* q;
  __device__ __forceinline__ float operator()(int i) const { return q[i]; }
};

// INV = true: sampling direction (z -> x, NSFPosteriorFlow.inverse).  INV = false: density direction
// (x -> z, NSFPosteriorFlow.forward, flows.py:610-618) on the SAME stepwise schedule: with x known, step d
// evaluates dim d's spline forward and then the hidden group of degree d -- the masks make that identical to one
// dense MADE pass, and the weight image, staging and barrier protocol are shared between the directions.
template <int NB, int GSP, bool INV>
__global__ void __launch_bounds__(TC_THREADS, 1) tc_sampler_kernel(const TcArgs a) {
  using TM = TcTmem<GSP>;
  constexpr int NW = GSP / 2;
  constexpr int MULT = 3 * NB - 1;
  static_assert(MULT <= 48 && GSP <= 32 && GSP % 2 == 0, "shape limits of this kernel");
  const TcShape& s = a.s;
  extern __shared__ __align__(128) uint8_t smem[];
  const int KA = s.KA, D = s.D;
  uint8_t* sA3 = smem;
  uint8_t* sA4 = sA3 + (size_t)KA * 256;
  uint8_t* sX = sA4 + (size_t)KA * 256;
  uint8_t* ring = sX + 8192;
  float* ybuf = reinterpret_cast<float*>(ring + (size_t)TC_NSLOT * s.slot_bytes);
  constexpr int GS4 = tc_gs4(GSP), CBF = tc_cb_floats(GSP);
  float* cbuf = ybuf + (2 * TC_ROWS * D + 3) / 4 * 4;   // [TC_NCB][CBF] per-step constants, staged by warp 6
  float* xwk = cbuf + TC_NCB * CBF;                      // [128][17] width knots, half 1 -> half 0 of a row pair
  uint64_t* bars = reinterpret_cast<uint64_t*>(xwk + TC_ROWS * 17 + (TC_ROWS * 17 & 1));
  uint64_t* full = bars;                 // [NSLOT] weights landed
  uint64_t* empty = bars + TC_NSLOT;     // [NSLOT] MMAs reading the slot completed
  uint64_t* acc_bar = bars + 2 * TC_NSLOT;      // accumulator of the current stage complete
  uint64_t* op_bar = bars + 2 * TC_NSLOT + 1;   // operands of the next stage written (4 warp arrivals)
  uint64_t* cfull = bars + 2 * TC_NSLOT + 2;              // [TC_NCB] constants of a step staged
  uint64_t* cempty = bars + 2 * TC_NSLOT + 2 + TC_NCB;    // [TC_NCB] constants of a step consumed (4 warp arrivals)
  volatile int* abort_s = reinterpret_cast<volatile int*>(bars + 2 * TC_NSLOT + 2 + 2 * TC_NCB);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * TC_NSLOT + 3 + 2 * TC_NCB);

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // provably warp-uniform role index

  if (tid == 0) {
    for (int i = 0; i < TC_NSLOT; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(acc_bar, 1);
    mbar_init(op_bar, 8);
    for (int i = 0; i < TC_NCB; ++i) { mbar_init(&cfull[i], 1); mbar_init(&cempty[i], 8); }
    *abort_s = 0;
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc(tmem_slot, 512);
  // zero the shared-memory operand arrays (stale columns are multiplied by zero weights: must be finite)
  for (int i = tid; i < (2 * KA * 256 + 8192) / 16; i += TC_THREADS) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  // bounded wait; on timeout raise the CTA-wide abort flag so every role leaves its loops
  auto wait = [&](uint64_t* bar, uint32_t parity, int code) -> bool {
    for (uint32_t i = 0; i < (1u << 21); ++i) {
      if (mbar_try_wait(bar, parity)) return true;
      if ((i & 255) == 255 && *abort_s) return false;
    }
    *abort_s = 1;
    raise_flag(a.err, code);
    return false;
  };

  const int lvl = a.dbg_stop;
  if (warp < 8 && lvl != -11) {
    // ============================ row warps ============================
    // Two threads per flow row: warps 0-3 ("half 0") and 4-7 ("half 1") both cover rows 0..127 (TMEM lane
    // window = warp % 4).  In the hidden stages half 0 owns bf16 words 0..6 of the step's 13-word group and
    // half 1 words 7..12; in the spline stage half 1 turns the 16 width logits into the 17 width knots while
    // half 0 does heights, bin search, derivatives and the closing solve.  Two warps per scheduler instead of
    // one is what hides the dependent-instruction latency of these epilogues.
    const int half = warp >> 2, q4 = warp & 3;
    const int r = q4 * 32 + lane;
    const uint32_t tl = tmem + ((uint32_t)(q4 * 32) << 16);   // this warp's TMEM lane window
    constexpr int NWH = (NW + 1) / 2;                            // words of half 0 (half 1 has NW - NWH)
    const int wo = half ? NWH : 0;                               // first word of this thread's share
    const int nwo = half ? NW - NWH : NWH;
    if (half == 0) {  // zero the TMEM activation arrays + pad columns
      uint32_t zz[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) zz[i] = 0u;
      for (int c = 0; c < TM::ACC_F; c += 16) tmem_st16(tl + c, zz);
      tmem_wait_st();
    }
    uint32_t acc_par = 0, stepc = 0;
    const bool prof = a.prof && blockIdx.x == 0 && tid == 0;
    long long pw[6] = {0, 0, 0, 0, 0, 0}, pc[6] = {0, 0, 0, 0, 0, 0};
    long long tmark = clock64();
    const long long t_begin = tmark;
    int pcur = ST_F;   // stage whose epilogue is running
    auto mark_c = [&](int) { if (prof) { long long n = clock64(); pc[pcur] += n - tmark; tmark = n; } };
    auto mark_w = [&](int st) { if (prof) { long long n = clock64(); pw[st] += n - tmark; tmark = n; pcur = st; } };
    const uint32_t rowoff = (uint32_t)(r / 8) * 128 + (uint32_t)(r % 8) * 16;
    float* xr = xwk + r * 17;                                    // this row's 17 width knots (half 1 -> half 0)
    // store this thread's words of the group starting at word column `wbase` / element slot `slot0`
    auto store_tmem = [&](uint32_t wbase, const uint32_t* w) {
      const uint32_t dst = wbase + wo;
      tmem_st4(dst, w);
      if (nwo >= 6) tmem_st2(dst + 4, w + 4);
      else tmem_st1(dst + 4, w + 4);
      if (nwo == 7) tmem_st1(dst + 6, w + 6);
      tmem_wait_st();
    };
    auto store_smem = [&](uint8_t* arr, int slot0, const uint32_t* w) {
      const uint32_t b0 = (uint32_t)slot0 * 2 + 4 * wo;
      uint8_t* p = arr + (b0 >> 4) * 2048 + rowoff + (b0 & 15);
      constexpr int N0 = NWH, N1 = NW - NWH;
      if (half == 0) {
        switch ((b0 >> 2) & 3) {
          case 0: store_words_umma<0, N0>(p, w); break;
          case 1: store_words_umma<1, N0>(p, w); break;
          case 2: store_words_umma<2, N0>(p, w); break;
          default: store_words_umma<3, N0>(p, w); break;
        }
      } else {
        switch ((b0 >> 2) & 3) {
          case 0: store_words_umma<0, N1>(p, w); break;
          case 1: store_words_umma<1, N1>(p, w); break;
          case 2: store_words_umma<2, N1>(p, w); break;
          default: store_words_umma<3, N1>(p, w); break;
        }
      }
      fence_proxy_async_smem();
    };
    // 2*NWH constants of this thread's share from group `g` of the step's constants
    auto load_consts = [&](const float* grp, float* t) {
#pragma unroll
      for (int i = 0; i < NWH; ++i) {
        const float2 c2 = *reinterpret_cast<const float2*>(grp + 2 * wo + 2 * i);
        t[2 * i] = c2.x; t[2 * i + 1] = c2.y;
      }
    };
    bool ok = !(lvl == -12 || lvl == -13);
    int tile = blockIdx.x;
    for (; tile < a.n_tiles && ok; tile += gridDim.x) {
      const int64_t grow = (int64_t)tile * TC_ROWS + r;
      const bool valid = grow < a.R;
      float* y0 = ybuf + r * D;                 // ybuf[0][r][:]
      float* y1 = ybuf + TC_ROWS * D + r * D;   // ybuf[1][r][:]
      if (half == 0) {
        float zr[PF_MAX_D];
#pragma unroll
        for (int j = 0; j < PF_MAX_D; ++j) zr[j] = 0.f;
        if (valid) {
          if (a.z) {
            for (int j = 0; j < D; ++j) zr[j] = a.z[grow * D + j];
          } else {
            philox_normal_row((uint64_t)(a.row_base + grow), a.seed, D, zr);
          }
        }
        for (int j = 0; j < D; ++j) y0[j] = zr[j];
      }
      // context constants: staged in shared memory for the (at most two) events of this tile when rows map
      // to contexts in contiguous runs of >= 128; otherwise each row reads its own projection from L2
      const bool ctx_staged = !a.cm.row2ctx && a.cm.n_per_ctx >= TC_ROWS;
      const int64_t cidx = valid ? a.cm(grow) : a.cm((int64_t)tile * TC_ROWS);
      const int esel = (ctx_staged && cidx != a.cm((int64_t)tile * TC_ROWS)) ? 1 : 0;
      const float* projr = a.proj + cidx * a.ldproj;
      float ld_acc = 0.f;
      bool bad = false;
      float* yin = y0;
      float* yout = y1;
      for (int li = 0; li < s.L && ok; ++li) {
        const int l = INV ? s.L - 1 - li : li;
        const bool reversed = (l != s.L - 1);
        const float* pl = projr + (size_t)l * 3 * s.KTOT;
        for (int d = 1; d <= D && ok; ++d) {
          const int j = d - 1;
          // this step's constants (biases, context gates), staged by the stager warp ahead of time
          const int cbi = stepc & (TC_NCB - 1);
          const float* cb = cbuf + cbi * CBF;
          if (!(ok = wait(&cfull[cbi], (stepc / TC_NCB) & 1, 16))) break;
          const float* cbb = cb + TC_CB_BF;                      // 5 bias groups of GS4 floats
          const float* cbw = cbb + 5 * GS4;                      // initial-layer weights of x_d for this group
          const float* cbc = cbw + GS4 + esel * 3 * GS4;          // 3 context groups of this row's event
          float xd;
          if (d == 1) { mark_c(0); pcur = ST_F; }
          // ---------------- F(d): spline parameters of dim d -> inverse spline, split over the thread pair -------
          if (half == 1) {
            // ---- the axis that is NOT searched (widths when sampling, heights for the density direction):
            // 16 logits -> its 17 knots, handed to half 0 through shared memory ----
            constexpr int OTHER = INV ? 0 : NB;
            float qw[NB];
#pragma unroll
            for (int i = 0; i < NB; i += 4) {
              const float4 v4 = *reinterpret_cast<const float4*>(cb + OTHER + i);
              qw[i] = v4.x; qw[i + 1] = v4.y; qw[i + 2] = v4.z; qw[i + 3] = v4.w;
            }
            if (d > 1) {
              if (!(ok = wait(acc_bar, acc_par, 10))) break;
              acc_par ^= 1;
              tc_fence_after();
              uint32_t v[16];
              tmem_ld16(tl + TM::ACC_F + OTHER, v);
              tmem_wait_ld();
#pragma unroll
              for (int i = 0; i < NB; ++i) qw[i] += __uint_as_float(v[i]);
            }
            float mw = qw[0];
#pragma unroll
            for (int i = 1; i < NB; ++i) mw = fmaxf(mw, qw[i]);
            float s0 = 0.f, s1 = 0.f;
#pragma unroll
            for (int i = 0; i < NB; i += 2) {
              qw[i] = fx_ex2((qw[i] - mw) * 1.4426950408889634f);
              qw[i + 1] = fx_ex2((qw[i + 1] - mw) * 1.4426950408889634f);
              s0 += qw[i]; s1 += qw[i + 1];
            }
            const float cwn = (2.f * a.tail) * (1.f - kMinBin * NB) * fx_rcp(s0 + s1);
            const float base = (2.f * a.tail) * kMinBin;
            float cx = 0.f;
            xr[0] = -a.tail;
#pragma unroll
            for (int i = 0; i < NB; ++i) {
              cx += fmaf(cwn, qw[i], base);
              xr[i + 1] = (i == NB - 1) ? a.tail : cx - a.tail;
            }
            bar_arrive(1 + q4, 64);
            bar_sync(5 + q4, 64);          // half 0 has read the knots and published x_d
            xd = INV ? yout[j] : yin[D - 1 - j];   // the conditioner input of this dimension
          } else {
            // ---- searched axis (heights when sampling, widths for the density direction) + derivatives:
            // bin search, then the closing solve.  q[0,16) = logits of the searched axis, q[16,31) = derivatives ----
            constexpr int SEARCH = INV ? NB : 0;
            float q[32];
#pragma unroll
            for (int i = 0; i < NB; i += 4) {
              const float4 v4 = *reinterpret_cast<const float4*>(cb + SEARCH + i);
              const float4 d4 = *reinterpret_cast<const float4*>(cb + 2 * NB + i);
              q[i] = v4.x; q[i + 1] = v4.y; q[i + 2] = v4.z; q[i + 3] = v4.w;
              q[NB + i] = d4.x; q[NB + i + 1] = d4.y; q[NB + i + 2] = d4.z; q[NB + i + 3] = d4.w;
            }
            const float y = INV ? yin[reversed ? D - 1 - j : j] : yin[D - 1 - j];
            if (d > 1) {
              mark_c(ST_F);
              if (!(ok = wait(acc_bar, acc_par, 10))) break;
              mark_w(ST_F);
              acc_par ^= 1;
              tc_fence_after();
              uint32_t v[32];
              tmem_ld16(tl + TM::ACC_F + SEARCH, v);
              tmem_ld16(tl + TM::ACC_F + 2 * NB, v + 16);
              tmem_wait_ld();
#pragma unroll
              for (int i = 0; i < 2 * NB - 1; ++i) q[i] += __uint_as_float(v[i]);
            }
            const float B = a.tail;
            float mh = q[0];
#pragma unroll
            for (int i = 1; i < NB; ++i) mh = fmaxf(mh, q[i]);
            float s0 = 0.f, s1 = 0.f;
#pragma unroll
            for (int i = 0; i < NB; i += 2) {
              q[i] = fx_ex2((q[i] - mh) * 1.4426950408889634f);
              q[i + 1] = fx_ex2((q[i + 1] - mh) * 1.4426950408889634f);
              s0 += q[i]; s1 += q[i + 1];
            }
            const float chn = (2.f * B) * (1.f - kMinBin * NB) * fx_rcp(s0 + s1);
            const float base = (2.f * B) * kMinBin;
            int k = 0;
            float yk = -B, hk = 0.f, py = -B, cy = 0.f;
#pragma unroll
            for (int i = 0; i < NB; ++i) {
              cy += fmaf(chn, q[i], base);
              const float ny = (i == NB - 1) ? B : cy - B;
              if (i == 0 || y >= py) { k = i; yk = py; hk = ny - py; }
              py = ny;
            }
            float u0 = kEdgeLogit, u1 = kEdgeLogit;
#pragma unroll
            for (int i = 0; i < NB - 1; ++i) {
              const float u = q[NB + i];
              if (i + 1 == k) u0 = u;
              if (i == k) u1 = u;
            }
            const float dk = kMinDeriv + (u0 > 20.f ? u0 : fx_log(1.f + fx_exp(u0)));
            const float dk1 = kMinDeriv + (u1 > 20.f ? u1 : fx_log(1.f + fx_exp(u1)));
            bar_sync(1 + q4, 64);          // the other axis' knots of this row are in shared memory
            const float ok_ = xr[k];                  // other-axis knot and bin size at the bin found
            const float os_ = xr[k + 1] - ok_;
            const bool inside = y >= -B && y <= B;       // outside (or NaN): linear tails, identity
            float res, lad;
            if (INV) {
              const float xk = ok_, wk = os_;
              const float sl = hk * fx_rcp(wk);
              const float dy = y - yk;
              const float t = dk + dk1 - 2.f * sl;
              const float qa = fmaf(dy, t, hk * (sl - dk));
              const float qb = fmaf(hk, dk, -dy * t);
              const float qc = -sl * dy;
              const float disc = fmaxf(fmaf(qb, qb, -4.f * qa * qc), 0.f);
              const float root = 2.f * qc * fx_rcp(-qb - fx_sqrt(disc));
              const float tt = root * (1.f - root);
              const float den = fmaf(t, tt, sl);
              const float omr = 1.f - root;
              const float dnum = sl * sl * (dk1 * root * root + 2.f * sl * tt + dk * omr * omr);
              res = fmaf(root, wk, xk);
              lad = -fx_log(dnum * fx_rcp(den * den));
            } else {   // forward map: (yk, hk) here are the searched x-axis knot / width, (ok_, os_) the y axis
              const float xk = yk, wk = hk, y0k = ok_, h0k = os_;
              const float rw = fx_rcp(wk);
              const float sl = h0k * rw;
              const float th = (y - xk) * rw;
              const float omt = 1.f - th;
              const float tt = th * omt;
              const float den = fmaf(dk + dk1 - 2.f * sl, tt, sl);
              const float rden = fx_rcp(den);
              res = fmaf(h0k * (sl * th * th + dk * tt), rden, y0k);
              const float dnum = sl * sl * (dk1 * th * th + 2.f * sl * tt + dk * omt * omt);
              lad = fx_log(dnum * rden * rden);
            }
            const float xo = inside ? res : y;
            ld_acc += inside ? lad : 0.f;
            yout[j] = xo;
            const float xin = INV ? xo : y;              // what the conditioner of the later dimensions sees
            bad = bad || !(fabsf(xin) <= 3.0e38f);
            const float xc = finite_clamp(xin);
            const __nv_bfloat16 hi = __float2bfloat16(xc);
            const __nv_bfloat16 lo = __float2bfloat16(xc - __bfloat162float(hi));
            *reinterpret_cast<__nv_bfloat16*>(sX + (j / 8) * 2048 + rowoff + (j % 8) * 2) = hi;
            *reinterpret_cast<__nv_bfloat16*>(sX + (2 + j / 8) * 2048 + rowoff + (j % 8) * 2) = lo;
            bar_arrive(5 + q4, 64);
            xd = xin;
          }
          if (d == D) {       // last dim: no hidden group follows, and the next stage (F(1)) has no MMA
            mark_c(ST_F);
            __syncwarp();
            if (lane == 0) mbar_arrive(&cempty[cbi]);
            ++stepc;
            break;
          }
          const int slot0 = j * GSP;   // first padded slot of the degree-d group
          float h[2 * NWH];
          // ---- I(d), hoisted: W0[g, <d] x[<d] was issued together with F(d) (its x_d column is zero in the
          // blob); the x_d term is added here in fp32: h0 = acc + W0[g, d] x_d + b0 + relu(Wc0 ctx + b) ----
          {
            float t[2 * NWH], wc[2 * NWH];
            load_consts(cbb, t);
            load_consts(cbw, wc);
#pragma unroll
            for (int i = 0; i < 2 * NWH; ++i) t[i] = fmaf(wc[i], xd, t[i]);
            if (ctx_staged) {
              load_consts(cbc, wc);
#pragma unroll
              for (int i = 0; i < 2 * NWH; ++i) t[i] += wc[i];
            } else {
#pragma unroll
              for (int i = 0; i < 2 * NWH; ++i)
                if (i < 2 * nwo) t[i] += __ldg(pl + slot0 + 2 * wo + i);
            }
            if (d > 1) {
              uint32_t v[16];
              tmem_ld16(tl + TM::ACC_I + 2 * wo, v);
              tmem_wait_ld();
#pragma unroll
              for (int i = 0; i < 2 * NWH; ++i) h[i] = __uint_as_float(v[i]) + t[i];
            } else {
#pragma unroll
              for (int i = 0; i < 2 * NWH; ++i) h[i] = t[i];
            }
            uint32_t w[NWH];
#pragma unroll
            for (int i = 0; i < NWH; ++i) w[i] = pack_bf16(relu_finite(h[2 * i]), relu_finite(h[2 * i + 1]));
            store_tmem(tl + TM::A0 + j * NW, w);
          }
          fence_proxy_async_smem();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(op_bar);
          // ---------------- residual blocks ----------------
#pragma unroll
          for (int blk = 0; blk < 2; ++blk) {
            {  // A: u = Wa relu(h) + ba -> relu(u)
              float t[2 * NWH];
              load_consts(cbb + (1 + 2 * blk) * GS4, t);
              mark_c(ST_A1 + 2 * blk);
              if (!(ok = wait(acc_bar, acc_par, 12 + blk))) break;
              mark_w(ST_A1 + 2 * blk);
              acc_par ^= 1;
              tc_fence_after();
              uint32_t v[16];
              tmem_ld16(tl + TM::ACC_H + 2 * wo, v);
              tmem_wait_ld();
              uint32_t w[NWH];
#pragma unroll
              for (int i = 0; i < NWH; ++i)
                w[i] = pack_bf16(relu_finite(__uint_as_float(v[2 * i]) + t[2 * i]),
                                 relu_finite(__uint_as_float(v[2 * i + 1]) + t[2 * i + 1]));
              if (blk == 0) store_tmem(tl + TM::A1 + j * NW, w);
              else store_smem(sA3, slot0, w);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(op_bar);
            {  // B: h += gate * (Wb relu(u) + bb)
              float t[2 * NWH], g[2 * NWH];
              load_consts(cbb + (2 + 2 * blk) * GS4, t);
              if (ctx_staged) {
                load_consts(cbc + (blk + 1) * GS4, g);
              } else {
                const float* gp = pl + (size_t)(blk + 1) * s.KTOT + slot0 + 2 * wo;
#pragma unroll
                for (int i = 0; i < 2 * NWH; ++i) g[i] = i < 2 * nwo ? __ldg(gp + i) : 0.f;
              }
              mark_c(ST_B1 + 2 * blk);
              if (!(ok = wait(acc_bar, acc_par, 14 + blk))) break;
              mark_w(ST_B1 + 2 * blk);
              acc_par ^= 1;
              tc_fence_after();
              uint32_t v[16];
              tmem_ld16(tl + TM::ACC_H + 2 * wo, v);
              tmem_wait_ld();
#pragma unroll
              for (int i = 0; i < 2 * NWH; ++i) h[i] = fmaf(g[i], __uint_as_float(v[i]) + t[i], h[i]);
              uint32_t w[NWH];
              if (blk == 0) {
#pragma unroll
                for (int i = 0; i < NWH; ++i) w[i] = pack_bf16(relu_finite(h[2 * i]), relu_finite(h[2 * i + 1]));
                store_tmem(tl + TM::A2 + j * NW, w);
              } else {
#pragma unroll
                for (int i = 0; i < NWH; ++i) w[i] = pack_bf16(finite_clamp(h[2 * i]), finite_clamp(h[2 * i + 1]));
                store_smem(sA4, slot0, w);
              }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(op_bar);
          }
          if (lane == 0) mbar_arrive(&cempty[cbi]);   // (after the __syncwarp above: all lanes are done reading)
          ++stepc;
        }
        float* tsw = yin; yin = yout; yout = tsw;
      }
      if (!ok) break;
      // ---------------- tail of inverse + ParamScaler (shared with the FP32 tier) ----------------
      if (valid && half == 0) {
        float v[PF_MAX_D];
        if (INV) {
#pragma unroll
          for (int jj = 0; jj < PF_MAX_D; ++jj) v[jj] = jj < D ? yin[D - 1 - jj] : 0.f;
          if (bad) {  // a non-finite value entered the conditioner: the reference turns the row into NaN
            for (int jj = 0; jj < D; ++jj) v[jj] = __int_as_float(0x7fc00000);
            ld_acc = __int_as_float(0x7fc00000);
          }
          const bool rail = finish_row(v, D, a.flags, a.clamp, a.sc);
          if (a.railed) a.railed[grow] = rail ? 1 : 0;
        } else {      // density direction: z = output of the last layer as it is (flows.py:615-618)
#pragma unroll
          for (int jj = 0; jj < PF_MAX_D; ++jj) v[jj] = jj < D ? yin[jj] : 0.f;
          if (bad) {
            for (int jj = 0; jj < D; ++jj) v[jj] = __int_as_float(0x7fc00000);
            ld_acc = __int_as_float(0x7fc00000);
          }
        }
        for (int jj = 0; jj < D; ++jj) a.x[grow * D + jj] = v[jj];
        if (a.logdet) a.logdet[grow] = ld_acc;
      }
    }
    if (!ok && half == 0) {
      // The CTA gave up (barrier timeout, flag raised): a timeout must never pass as data -- every row of the tile it
      // abandoned and of the tiles it will not reach is written as NaN.
      for (; tile < a.n_tiles; tile += gridDim.x) {
        const int64_t grow = (int64_t)tile * TC_ROWS + r;
        if (grow >= a.R) continue;
        for (int jj = 0; jj < D; ++jj) a.x[grow * D + jj] = __int_as_float(0x7fc00000);
        if (a.logdet) a.logdet[grow] = __int_as_float(0x7fc00000);
        if (a.railed) a.railed[grow] = 0;
      }
    }
    if (prof) {
      for (int i = 0; i < 6; ++i) { a.prof[i] = pw[i]; a.prof[6 + i] = pc[i]; }
      a.prof[15] = clock64() - t_begin;
    }
  } else if (warp == 8) {
    // ============================ MMA issuer ============================
    // the whole warp runs the (warp-uniform) loops; one elected lane issues tcgen05.mma / commit
    if (lvl <= -13) *abort_s = 1;
    if (lvl > -11) {
      uint32_t op_par = 0, slot = 0, full_par = 0;
      int nstage = 0;
      const bool mprof = a.prof && blockIdx.x == 0;
      long long m_op = 0, m_full = 0, m_issue = 0, m_loop = 0, m_commit = 0;
      long long n_mma = 0;
      const uint32_t idesc_h = umma_idesc_bf16(TC_ROWS, 32), idesc_f = umma_idesc_bf16(TC_ROWS, s.NF);
      const uint32_t aX = smem_u32(sX), aA3 = smem_u32(sA3), aA4 = smem_u32(sA4), aRing = smem_u32(ring);
      bool ok = true;
      for (int tile = blockIdx.x; tile < a.n_tiles && ok; tile += gridDim.x) {
        for (int li = 0; li < s.L && ok; ++li) {
        const int l = INV ? s.L - 1 - li : li;
          for (int d = 1; d <= D && ok; ++d) {
            // MMA stages of step d: [F(d) + I(d)] (one commit), A1, B1, A2, B2
            for (int ms = 0; ms < 5 && ok; ++ms) {
              if (ms > 0 && d == D) break;
              const int st0 = ms == 0 ? ST_F : ms + 1;
              const int ngrp = ms == 0 ? 2 : 1;            // stage 0 has two item groups: F then I
              const int n0 = tc_nchunks(s, d, st0), n1 = ms == 0 ? tc_nchunks(s, d, ST_I) : 0;
              if (n0 + n1 == 0) continue;                 // step 1: bias-only F, fp32-only I
              if (nstage++ == a.dbg_stop) {      // PF_TC_STOP: forced protocol failure (tests, bring-up)
                *abort_s = 1;
                ok = false;
                if (elect_one_sync()) raise_flag(a.err, 99);
                __syncwarp();
                break;
              }
              long long t0 = mprof ? clock64() : 0;
              if (!(ok = wait(&full[slot], full_par, 30))) break;   // first item's weights (prefetched long ago)
              if (!(ok = wait(op_bar, op_par, 20 + st0))) break;
              if (mprof) { long long n = clock64(); m_op += n - t0; t0 = n; }
              op_par ^= 1;
              tc_fence_after();
              bool first_item = true;
              for (int grp = 0; grp < ngrp && ok; ++grp) {
                const int st = grp == 0 ? st0 : ST_I;
                const int nch = grp == 0 ? n0 : n1;
                if (nch == 0) continue;
                const bool last_grp = (grp == ngrp - 1) || (grp == 0 && n1 == 0);
                const int N = tc_n(s, st);
                const uint32_t idesc = st == ST_F ? idesc_f : idesc_h;
                const uint32_t acc = tmem + (st == ST_F ? TM::ACC_F : (st == ST_I ? TM::ACC_I : TM::ACC_H));
                const int pc = tc_piece(s, st, nch);
                for (int c0 = 0; c0 < nch && ok; c0 += pc) {
                  const int nc = min(pc, nch - c0);
                  long long t1 = mprof ? clock64() : 0;
                  if (!first_item && !(ok = wait(&full[slot], full_par, 30))) break;
                  first_item = false;
                  if (mprof) m_full += clock64() - t1;
                  tc_fence_after();
                  const uint32_t bbase = aRing + slot * s.slot_bytes;
                  const long long t2 = mprof ? clock64() : 0;
                  if (elect_one_sync()) {
                    // descriptors advance by adding to their low word (start address field, 16-byte units)
                    uint64_t bd = umma_smem_desc(bbase, N * 16, 128);
                    const uint64_t bstep = (uint64_t)((2u * (uint32_t)N * 16u) >> 4);
                    uint32_t accf = c0 > 0 ? 1u : 0u;
                    if (st == ST_F || st == ST_I || st == ST_B2) {
                      const uint32_t abase = st == ST_F ? aA4 : (st == ST_I ? aX : aA3);
                      uint64_t ad = umma_smem_desc(abase + (uint32_t)c0 * 4096u, 2048, 128);
#pragma unroll 4
                      for (int i = 0; i < nc; ++i) {
                        umma_ss(acc, ad, bd, idesc, accf);
                        accf = 1u; ad += 256u; bd += bstep;
                      }
                    } else {
                      uint32_t at = tmem + (st == ST_A1 ? TM::A0 : (st == ST_B1 ? TM::A1 : TM::A2)) + c0 * 8;
#pragma unroll 4
                      for (int i = 0; i < nc; ++i) {
                        umma_ts(acc, at, bd, idesc, accf);
                        accf = 1u; at += 8u; bd += bstep;
                      }
                    }
                    const long long t3 = mprof ? clock64() : 0;
                    umma_commit(&empty[slot]);
                    if (last_grp && c0 + nc >= nch) umma_commit(acc_bar);
                    if (mprof) { const long long t4 = clock64(); m_loop += t3 - t2; m_commit += t4 - t3; n_mma += nc; }
                  }
                  __syncwarp();
                  if (++slot == TC_NSLOT) { slot = 0; full_par ^= 1; }
                }
              }
              if (mprof) m_issue += clock64() - t0;
            }
          }
        }
      }
      if (mprof && elect_one_sync()) {
        a.prof[12] = m_op; a.prof[13] = m_full; a.prof[14] = m_issue;
        a.prof[16] = m_loop; a.prof[17] = m_commit; a.prof[18] = n_mma;
      }
    }
  } else if (warp == 9) {
    // ============================ weight loader ============================
    if (lvl != -11 && lvl != -12) {
      uint32_t slot = 0, empty_par = 1;   // a fresh barrier passes a wait on parity 1
      bool ok = true;
      for (int tile = blockIdx.x; tile < a.n_tiles && ok; tile += gridDim.x) {
        for (int li = 0; li < s.L && ok; ++li) {
        const int l = INV ? s.L - 1 - li : li;
          const uint8_t* src = a.image + (size_t)l * s.layer_bytes;
          for (int d = 1; d <= D && ok; ++d) {
            for (int st = 0; st < ST_COUNT && ok; ++st) {
              if (st > 0 && d == D) break;
              const int nch = tc_nchunks(s, d, st);
              if (nch == 0) continue;
              const int N = tc_n(s, st);
              const int pc = tc_piece(s, st, nch);
              for (int c0 = 0; c0 < nch; c0 += pc) {
                const uint32_t bytes = (uint32_t)min(pc, nch - c0) * (uint32_t)N * 32u;
                if (!(ok = wait(&empty[slot], empty_par, 40))) break;
                if (elect_one_sync()) {
                  mbar_arrive_expect_tx(&full[slot], bytes);
                  bulk_g2s(ring + (size_t)slot * s.slot_bytes, src, bytes, &full[slot]);
                }
                __syncwarp();
                src += bytes;
                if (++slot == TC_NSLOT) { slot = 0; empty_par ^= 1; }
              }
            }
          }
        }
      }
    }
  } else if (warp == 10) {
    // ============================ constants stager ============================
    // One step ahead of the row warps: the biases of the step's six stages and the context terms of the
    // tile's (at most two) events go to shared memory, so no global-load latency sits on the serial chain.
    if (lvl != -11 && lvl != -12) {
      uint32_t stepc = 0;
      bool ok = true;
      for (int tile = blockIdx.x; tile < a.n_tiles && ok; tile += gridDim.x) {
        const int64_t r_first = (int64_t)tile * TC_ROWS;
        const int64_t r_last = (r_first + TC_ROWS - 1 < a.R ? r_first + TC_ROWS - 1 : a.R - 1);
        const bool ctx_staged = !a.cm.row2ctx && a.cm.n_per_ctx >= TC_ROWS;
        const float* p0 = a.proj + a.cm(r_first) * a.ldproj;
        const float* p1 = a.proj + a.cm(r_last) * a.ldproj;
        for (int li = 0; li < s.L && ok; ++li) {
        const int l = INV ? s.L - 1 - li : li;
          for (int d = 1; d <= D && ok; ++d) {
            const int buf = stepc & (TC_NCB - 1);
            if (!(ok = wait(&cempty[buf], ((stepc / TC_NCB) & 1) ^ 1, 41))) break;
            float* cb = cbuf + buf * CBF;
            const int j = d - 1, slot0 = j * GSP;
            // address phase, then all loads in flight together, then the stores
            constexpr int PER = (CBF + 31) / 32;
            const float* src[PER];
#pragma unroll
            for (int k = 0; k < PER; ++k) {
              const int i = lane + 32 * k;
              const float* ptr = nullptr;
              if (i < TC_CB_BF) {
                if (i < MULT) ptr = a.bf + (size_t)l * s.PLD + (size_t)j * MULT + i;
              } else if (i < CBF && d < D) {
                const int kk = i - TC_CB_BF, grp = kk / GS4, e = kk % GS4;
                if (e < GSP) {
                  if (grp == 0) ptr = a.b0p + (size_t)l * s.KTOT + slot0 + e;
                  else if (grp < 5) ptr = (((grp - 1) & 1) ? a.bbp : a.bap) + ((size_t)l * 2 + (grp - 1) / 2) * s.KTOT + slot0 + e;
                  else if (grp == 5) {
                    if (e < s.off[d] - s.off[d - 1]) ptr = a.W0 + ((size_t)l * a.H + s.off[d - 1] + e) * a.DP + j;
                  } else if (ctx_staged) {
                    const int ev = (grp - 6) / 3, sec = (grp - 6) % 3;
                    ptr = (ev ? p1 : p0) + ((size_t)l * 3 + sec) * s.KTOT + slot0 + e;
                  }
                }
              }
              src[k] = ptr;
            }
            float val[PER];
#pragma unroll
            for (int k = 0; k < PER; ++k) val[k] = src[k] ? __ldg(src[k]) : 0.f;
#pragma unroll
            for (int k = 0; k < PER; ++k)
              if (lane + 32 * k < CBF) cb[lane + 32 * k] = val[k];
            __syncwarp();
            if (lane == 0) mbar_arrive(&cfull[buf]);
            ++stepc;
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc(tmem, 512);
}

}  // namespace pf
using namespace pf;

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static bool tc_plan(const pf_flow* f, TcShape& s, std::vector<TcItem>& items) {
  memset(&s, 0, sizeof(s));
  s.D = f->D; s.G = f->G; s.L = f->L; s.MULT = f->MULT; s.PLD = f->PLD;
  if (f->NB != 16 || f->D > 16 || f->G > 10 || f->desc.num_blocks != 2) return false;
  int gsp = (f->max_group + 1) / 2 * 2;
  if (gsp != 26 && gsp != 20) return false;   // instantiated group sizes (H=256 | 192 at D=11)
  s.GSP = gsp;
  s.KTOT = s.G * gsp;
  s.KA = round_up(s.KTOT, 16);
  s.NF = round_up(s.MULT, 16);
  for (int g = 0; g <= PF_MAX_D; ++g) s.off[g] = f->off[g];
  s.slot_bytes = 32u * (uint32_t)s.KA * 2u;
  s.pc_f = (int)(s.slot_bytes / ((uint32_t)s.NF * 32u));
  uint32_t off = 0;
  items.clear();
  for (int d = 1; d <= s.D; ++d)
    for (int st = 0; st < ST_COUNT; ++st) {
      if (st > 0 && d == s.D) break;
      int nch = tc_nchunks(s, d, st);
      if (nch == 0) continue;
      int pc = tc_piece(s, st, nch), N = tc_n(s, st);
      for (int c0 = 0; c0 < nch; c0 += pc) {
        int nc = std::min(pc, nch - c0);
        items.push_back(TcItem{d, st, c0, nc, N, off});
        off += (uint32_t)nc * N * 32u;
      }
    }
  s.layer_bytes = off;
  return true;
}

static int tc_smem_bytes(const TcShape& s) {
  size_t y = ((size_t)2 * TC_ROWS * s.D + 3) / 4 * 4 + TC_NCB * (size_t)tc_cb_floats(s.GSP) + TC_ROWS * 17;
  return (int)((size_t)2 * s.KA * 256 + 8192 + (size_t)TC_NSLOT * s.slot_bytes + y * 4 + (2 * TC_NSLOT + 2 * TC_NCB + 4) * 8);
}

template <int NB, int GSP, bool INV>
static pf_status tc_launch(pf_flow* f, cudaStream_t st, const TcArgs& a, int smem) {
  static bool attr_done[64] = {};
  int dev = f->desc.device;
  if (dev < 64 && !attr_done[dev]) {
    PF_CUDA(cudaFuncSetAttribute(tc_sampler_kernel<NB, GSP, INV>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_done[dev] = true;
  }
  int grid = a.n_tiles < f->sm_count ? a.n_tiles : f->sm_count;
  tc_sampler_kernel<NB, GSP, INV><<<grid, TC_THREADS, smem, st>>>(a);
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}

pf_status pf_tc_pack(pf_flow* f, cudaStream_t st) {
  if (f->desc.tier != PF_TIER_BF16) return PF_OK;
  TcShape s;
  std::vector<TcItem> items;
  if (!tc_plan(f, s, items)) return PF_ERR_UNSUPPORTED;
  TcState* t = reinterpret_cast<TcState*>(f->tc);
  const size_t L = f->L, KT = s.KTOT, C = f->C;
  if (!t) {
    t = new TcState();
    memset(t, 0, sizeof(*t));
    t->shape = s;
    t->n_items = (int)items.size();
    t->kernel_smem = tc_smem_bytes(s);
    const char* dbg = getenv("PF_TC_DEBUG");
    t->debug_sync = dbg && dbg[0] == '1';
    const char* pr = getenv("PF_TC_PROF");
    if (pr && pr[0] == '1' && cudaMalloc(&t->prof, 24 * sizeof(long long)) != cudaSuccess) t->prof = nullptr;
    size_t nf = L * KT + 2 * (L * 2 * KT) + L * 3 * KT * C + L * 3 * KT;
    cudaError_t e = cudaMalloc(&t->image, L * s.layer_bytes);
    if (e == cudaSuccess) e = cudaMalloc(&t->arena, nf * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&t->d_items, items.size() * sizeof(TcItem));
    if (e == cudaSuccess) e = cudaMemcpy(t->d_items, items.data(), items.size() * sizeof(TcItem), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
      if (t->image) cudaFree(t->image);
      if (t->arena) cudaFree(t->arena);
      if (t->d_items) cudaFree(t->d_items);
      delete t;
      return cuda_fail(e, "tensor-core tier allocation", __FILE__, __LINE__);
    }
    float* p = t->arena;
    t->b0p = p; p += L * KT;
    t->bap = p; p += L * 2 * KT;
    t->bbp = p; p += L * 2 * KT;
    t->Wctxp = p; p += L * 3 * KT * C;
    t->bctxp = p;
    f->tc = t;
  }
  dim3 grid(t->n_items + TC_PAD_BLOCKS, f->L);
  TcPadArgs pa{f->b0, f->ba, f->bb, f->bctx, f->Wctx, t->b0p, t->bap, t->bbp, t->bctxp, t->Wctxp, f->H, f->C};
  tc_pack_image_kernel<<<grid, 256, 0, st>>>(s, t->d_items, t->n_items, t->image, f->W0, f->DP, f->Wa, f->Wb, f->Wf, f->H, pa);
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}

void pf_tc_destroy(pf_flow* f) {
  TcState* t = reinterpret_cast<TcState*>(f->tc);
  if (!t) return;
  cudaFree(t->image); cudaFree(t->arena); cudaFree(t->d_items);
  if (t->prof) cudaFree(t->prof);
  delete t;
  f->tc = nullptr;
}

// proj[e, (l*3 + sec)*KTOT + slot] in the padded slot order of this tier
struct EpiCtxProjPad {
  float* out; int ld; const float* bias; int KT;
  __device__ __forceinline__ void operator()(int m, int n, float acc) const {
    float v = acc + bias[n];
    int sec = (n / KT) % 3;
    out[(int64_t)m * ld + n] = sec == 0 ? fmaxf(v, 0.f) : 1.f / (1.f + expf(-v));
  }
};

pf_status pf_tc_ctxproj(pf_flow* f, cudaStream_t st, const float* ctx, int64_t E, uint32_t flags, float* proj) {
  TcState* t = reinterpret_cast<TcState*>(f->tc);
  if (!t) return PF_ERR_NOT_PACKED;
  const int N = f->L * 3 * t->shape.KTOT;
  EpiCtxProjPad epi{proj, N, t->bctxp, t->shape.KTOT};
  if (flags & PF_FLAG_SANITIZE_CTX)
    return sgemm_nt<EpiCtxProjPad, A_SANITIZE, 64>(f, st, ctx, f->C, t->Wctxp, f->C, (int)E, N, f->C, epi);
  return sgemm_nt<EpiCtxProjPad, A_PLAIN, 64>(f, st, ctx, f->C, t->Wctxp, f->C, (int)E, N, f->C, epi);
}

size_t pf_tc_workspace_bytes(const pf_flow* f, int64_t R, int64_t E) {
  const TcState* t = reinterpret_cast<const TcState*>(f->tc);
  int KT = t ? t->shape.KTOT : f->G * 32;
  const int64_t chunk = R < PF_ROW_CHUNK ? R : PF_ROW_CHUNK;
  const int64_t prow = E > chunk ? E : chunk;
  return (size_t)prow * f->L * 3 * KT * sizeof(float) + 512;
}

// both directions: `z` = input rows ([R, D]; NULL = Philox noise, sampling direction only), `x` = output rows
static pf_status tc_run(pf_flow* f, cudaStream_t st, bool inverse, const float* z, uint64_t seed, const float* ctx,
                        const int64_t* row2ctx, int64_t n_per_ctx, int64_t R, int64_t E, float* x, float* logdet,
                        uint8_t* railed, uint32_t flags, const pf_scaler* scaler, void* workspace,
                        size_t workspace_bytes) {
  TcState* t = reinterpret_cast<TcState*>(f->tc);
  if (!t) return PF_ERR_NOT_PACKED;
  if (t->debug_sync) {
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
      snprintf(g_cuda_err, sizeof(g_cuda_err), "error pending on entry of inverse: %s", cudaGetErrorString(e));
      return PF_ERR_CUDA;
    }
  }
  const TcShape& s = t->shape;
  const bool is_proj = flags & PF_FLAG_CTX_IS_PROJ;
  const bool dense = !row2ctx && n_per_ctx == 1;
  const int64_t ldproj = (int64_t)f->L * 3 * s.KTOT;
  const int64_t chunk = dense ? (R < PF_ROW_CHUNK ? R : PF_ROW_CHUNK) : R;
  const int64_t prow = is_proj ? 0 : (dense ? chunk : E);
  if ((size_t)prow * ldproj * sizeof(float) > workspace_bytes || (prow && !workspace)) return PF_ERR_WORKSPACE;
  float* wproj = reinterpret_cast<float*>(workspace);
  ScalerDev sc{nullptr, nullptr, nullptr, nullptr};
  if (flags & (PF_FLAG_WRAP | PF_FLAG_DENORMALIZE)) {
    if (!scaler || !scaler->lo || !scaler->hi || !scaler->log_mask || !scaler->circ_mask) return PF_ERR_INVALID;
    sc = ScalerDev{scaler->lo, scaler->hi, scaler->log_mask, scaler->circ_mask};
  }
  const float* proj_all = nullptr;
  if (is_proj) proj_all = ctx;
  else if (!dense) {
    PF_TRY(pf_tc_ctxproj(f, st, ctx, E, flags, wproj));
    proj_all = wproj;
    if (t->debug_sync) {
      cudaError_t e = cudaStreamSynchronize(st);
      if (e != cudaSuccess) {
        snprintf(g_cuda_err, sizeof(g_cuda_err), "ctxproj in inverse: %s (ctx %p E %lld wproj %p)", cudaGetErrorString(e),
                 (const void*)ctx, (long long)E, (void*)wproj);
        return PF_ERR_CUDA;
      }
    }
  }
  for (int64_t r0 = 0; r0 < R; r0 += chunk) {
    const int64_t rows = (R - r0) < chunk ? (R - r0) : chunk;
    TcArgs a;
    memset(&a, 0, sizeof(a));
    a.s = s;
    a.image = t->image;
    a.b0p = t->b0p; a.bap = t->bap; a.bbp = t->bbp; a.bf = f->bf;
    a.W0 = f->W0; a.DP = f->DP; a.H = f->H;
    a.ldproj = ldproj;
    if (proj_all) {
      a.proj = proj_all;
      a.cm = CtxMap{row2ctx, n_per_ctx, r0};
    } else {
      PF_TRY(pf_tc_ctxproj(f, st, ctx + r0 * f->C, rows, flags, wproj));
      a.proj = wproj;
      a.cm = CtxMap{nullptr, 1, 0};
    }
    a.z = z ? z + r0 * s.D : nullptr;
    a.seed = seed;
    a.R = rows;
    a.row_base = r0;
    a.n_tiles = (int)((rows + TC_ROWS - 1) / TC_ROWS);
    a.x = x + r0 * s.D;
    a.logdet = logdet ? logdet + r0 : nullptr;
    a.railed = railed ? railed + r0 : nullptr;
    a.flags = flags; a.tail = f->desc.tail_bound; a.clamp = f->desc.clamp_bound;
    a.sc = sc;
    a.err = f->tg_err;
    a.prof = t->prof;
    {
      const char* stop = getenv("PF_TC_STOP");
      a.dbg_stop = stop ? atoi(stop) : -1;
    }
    if (a.dbg_stop == -10) continue;
    if (inverse) {
      if (s.GSP == 26) PF_TRY((tc_launch<16, 26, true>(f, st, a, t->kernel_smem)));
      else PF_TRY((tc_launch<16, 20, true>(f, st, a, t->kernel_smem)));
    } else {
      if (s.GSP == 26) PF_TRY((tc_launch<16, 26, false>(f, st, a, t->kernel_smem)));
      else PF_TRY((tc_launch<16, 20, false>(f, st, a, t->kernel_smem)));
    }
  }
  if (t->prof) {
    long long h[24];
    PF_CUDA(cudaStreamSynchronize(st));
    PF_CUDA(cudaMemcpy(h, t->prof, sizeof(h), cudaMemcpyDeviceToHost));
    const char* nm[6] = {"F", "I", "A1", "B1", "A2", "B2"};
    fprintf(stderr, "[pf_tc prof] CTA0 rows: total %lld cycles\n", h[15]);
    for (int i = 0; i < 6; ++i) fprintf(stderr, "[pf_tc prof]   stage %-2s wait %10lld  compute %10lld\n", nm[i], h[i], h[6 + i]);
    fprintf(stderr, "[pf_tc prof] MMA thread: wait-operands %lld  wait-weights %lld  stage total (incl. waits) %lld\n", h[12],
            h[13], h[14]);
    fprintf(stderr, "[pf_tc prof] MMA thread: issue loops %lld cycles for %lld MMAs, commits %lld\n", h[16], h[18], h[17]);
  }
  if (t->debug_sync) {
    PF_CUDA(cudaStreamSynchronize(st));
    PF_TRY(pf_check_flag(f, "tensor-core sampler"));
  }
  return PF_OK;
}

pf_status pf_tc_inverse(pf_flow* f, cudaStream_t st, const float* z, uint64_t seed, const float* ctx,
                        const int64_t* row2ctx, int64_t n_per_ctx, int64_t R, int64_t E, float* x, float* logdet,
                        uint8_t* railed, uint32_t flags, const pf_scaler* scaler, void* workspace,
                        size_t workspace_bytes) {
  return tc_run(f, st, true, z, seed, ctx, row2ctx, n_per_ctx, R, E, x, logdet, railed, flags, scaler, workspace,
                workspace_bytes);
}

// Density direction (NSFPosteriorFlow.forward, flows.py:610-618) without saved activations: x [R, D] -> z, logdet.
pf_status pf_tc_forward(pf_flow* f, cudaStream_t st, const float* x, const float* ctx, const int64_t* row2ctx,
                        int64_t n_per_ctx, int64_t R, int64_t E, float* z, float* logdet, uint32_t flags,
                        void* workspace, size_t workspace_bytes) {
  if (!x || !z) return PF_ERR_INVALID;
  const uint32_t fl = flags & (PF_FLAG_SANITIZE_CTX | PF_FLAG_CTX_IS_PROJ);
  // dense per-layer schedule (pf_tc_dense.cu) when the shape is covered; a pre-projected context is in this tier's
  // padded layout and stays with the stepwise kernel.  PF_TC_DENSE=0 forces the stepwise kernel (A/B, tests).
  TcState* t = reinterpret_cast<TcState*>(f->tc);
  const char* dn = getenv("PF_TC_DENSE");
  if (f->dense && t && !(fl & PF_FLAG_CTX_IS_PROJ) && !(dn && dn[0] == '0')) {
    PF_TRY(pf_dn_forward(f, st, x, ctx, row2ctx, n_per_ctx, R, E, z, logdet, fl, workspace, workspace_bytes, f->tg_err, 1));
    if (t->debug_sync) {
      PF_CUDA(cudaStreamSynchronize(st));
      PF_TRY(pf_check_flag(f, "dense density kernel"));
    }
    return PF_OK;
  }
  return tc_run(f, st, false, x, 0, ctx, row2ctx, n_per_ctx, R, E, z, logdet, nullptr, fl, nullptr, workspace,
                workspace_bytes);
}
