// CTA-PAIR form of the training forward (tc_dense_forward_kernel<3, true>): one 128-row tile is shared by a cluster of
// two CTAs on tcgen05.mma.cta_group::2 with M = 128.  Each CTA owns 64 rows of the tile (its own operand tile, x tile,
// residual stream, spline) and HALF of every weight chunk; one elected thread of the leader CTA issues the MMAs of the
// pair.  Measured (tests/umma_2cta_probe.cu): the pair instruction costs 71 cycles at N = 256 against 133 for the
// one-CTA form, and in each CTA the accumulator fills all 128 TMEM lanes x N / 2 columns (lanes 0..63 = its rows x
// outputs [0, N/2), lanes 64..127 = the same rows x outputs [N/2, N)) -- so FOUR threads share a row, the tensor time
// and the epilogue work per thread both halve, and nothing is exchanged between the two CTAs but barrier arrivals.
// A fused training step is one tile's latency on 8 .. 60 of the 148 SMs: this form doubles the SMs a
// tile uses.  Included by pf_tc_dense.cu (inside namespace pf, after the one-CTA kernel: shares its helpers and DnArgs).
//
// Same results layout as the one-CTA kernel: images / P / XS / xt are written at the rows' places in the 128-row tile,
// so the backward and weight-gradient kernels read them unchanged.
#pragma once

constexpr int DP_ROWS = 64;            // rows of the tile owned by one CTA
constexpr int DP_THREADS = 320;        // 8 row warps (4 threads per row), MMA issuer / relay, weight loader
constexpr int DP_NSLOT = 16;           // weight ring depth
constexpr int DP_SLOT = 8192;          // one K = 16 chunk of 128 outputs: fp16 hi (2 x 2 KB) | lo (2 x 2 KB)
constexpr int DP_NHF = 96, DP_NHC = 48;   // outputs per CTA of the final stages FA / FB (2 dims x 47 -> 96) and FC (1 dim -> 48)
constexpr uint32_t DP_LAYER_BYTES = 65u * 8192u + 16u * (2u * 64u * DP_NHF + 64u * DP_NHC);   // per layer and CTA rank
constexpr int DP_PACK_SLOTS = 65 + 48;

constexpr int DP_OFF_ACT = 0;                              // [64 x 256] fp16 hi, 1024 B per 8 columns
constexpr int DP_OFF_ACTLO = 32768;
constexpr int DP_OFF_XT = 65536;                           // [64 x 16] fp16 hi (2 KB) | lo (2 KB)
constexpr int DP_OFF_RING = DP_OFF_XT + 4096;
constexpr int DP_OFF_CST = DP_OFF_RING + DP_NSLOT * DP_SLOT;   // floats: 5 x 256 biases | [16 dims][48] final-layer biases
constexpr int DP_CST_FLOATS = 5 * DN_H + 16 * 48;
constexpr int DP_OFF_Y = DP_OFF_CST + DP_CST_FLOATS * 4;       // [2][64][12] fp32 layer inputs / outputs
constexpr int DP_OFF_RED = DP_OFF_Y + 2 * DP_ROWS * 12 * 4;    // [4][64] logdet partials, [4][64] bad flags
constexpr int DP_OFF_BAR = DP_OFF_RED + 8 * DP_ROWS * 4;
constexpr int DP_SMEM = DP_OFF_BAR + 512;

// final stage f = 0, 1 (FA, FB): output half h of the pair holds dims 1 + 4 f + 2 h and the next one (47 parameters each,
// 2 zero pads); f = 2 (FC): half h holds dim 9 + h (1 zero pad).  D = 11 only (LeanNPE).
__device__ __forceinline__ int dp_final_dim(int f, int h, int n, int& p) {
  if (f < 2) {
    if (n >= 2 * DN_MULT) return -1;
    p = n % DN_MULT;
    return 1 + 4 * f + 2 * h + n / DN_MULT;
  }
  if (n >= DN_MULT) return -1;
  p = n;
  return 9 + h;
}

// pair weight image: [L][2 ranks][DP_LAYER_BYTES], slots in the order the MMA warp consumes them:
//   I (1 slot) | A1 B1 A2 B2 (16 slots each) : NH = 128 outputs of this rank   8 KB
//   FA FB (16 each) : NH = 96   6 KB      FC (16) : NH = 48   3 KB
// a slot = fp16(16 W) of one K = 16 chunk as [2 K-slabs][NH][8] (UMMA K-major no-swizzle) followed by the lo halves
__global__ void dp_pack_kernel(DnShape s, uint8_t* __restrict__ imagep, const float* __restrict__ W0,
                               const float* __restrict__ Wa, const float* __restrict__ Wb, const float* __restrict__ Wf) {
  const int l = blockIdx.y, rank = blockIdx.z, c = blockIdx.x;
  int stage, cc, NH;
  uint32_t off;
  if (c < 1) { stage = 0; cc = 0; NH = 128; off = 0; }
  else if (c < 65) { stage = 1 + (c - 1) / 16; cc = (c - 1) % 16; NH = 128; off = (uint32_t)c * 8192u; }
  else {
    const int f = (c - 65) / 16;
    cc = (c - 65) % 16;
    stage = 5 + f;
    NH = f < 2 ? DP_NHF : DP_NHC;
    off = 65u * 8192u + (uint32_t)(f < 2 ? f * 16 + cc : 32) * 64u * DP_NHF + (f < 2 ? 0u : (uint32_t)cc * 64u * DP_NHC);
  }
  __half* dst = reinterpret_cast<__half*>(imagep + ((size_t)l * 2 + rank) * DP_LAYER_BYTES + off);
  const int H = DN_H;
  for (int e = threadIdx.x; e < NH * 16; e += blockDim.x) {
    const int n = e / 16, kk = e % 16;
    float v = 0.f;
    if (stage == 0) {
      if (kk < s.D) v = W0[((size_t)l * H + rank * 128 + n) * s.DP + kk];
    } else if (stage <= 4) {
      const int blk = (stage - 1) / 2;
      const float* W = ((stage - 1) & 1) ? Wb : Wa;
      v = W[(((size_t)l * 2 + blk) * H + rank * 128 + n) * H + cc * 16 + kk];
    } else {
      int p = 0;
      const int d = dp_final_dim(stage - 5, rank, n, p);
      if (d >= 0 && d < s.D) v = Wf[((size_t)l * s.PLD + (size_t)d * DN_MULT + p) * H + cc * 16 + kk];
    }
    const int o = (kk / 8) * (NH * 8) + (n / 8) * 64 + (n % 8) * 8 + (kk % 8);
    const float vs = fminf(fmaxf(v * DN_WSCALE, -60000.f), 60000.f);
    const __half hi = __float2half_rn(vs);
    dst[o] = hi;
    dst[NH * 16 + o] = __float2half_rn(vs - __half2float(hi));
  }
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(DP_THREADS, 1) tc_dense_pair_kernel(const DnArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  const DnShape& s = a.s;
  const int D = s.D;
  uint8_t* sACT = smem + DP_OFF_ACT;
  uint8_t* sACTLO = smem + DP_OFF_ACTLO;
  uint8_t* sXT = smem + DP_OFF_XT;
  uint8_t* ring = smem + DP_OFF_RING;
  float* cst = reinterpret_cast<float*>(smem + DP_OFF_CST);
  float* ybuf = reinterpret_cast<float*>(smem + DP_OFF_Y);
  float* red = reinterpret_cast<float*>(smem + DP_OFF_RED);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + DP_OFF_BAR);
  uint64_t* full = bars;                       // [NSLOT] weight chunk landed: this CTA's half -- and, in the leader, the peer's
                                               // (second arrival, relayed by the peer's warp 8: one barrier to poll per chunk)
  uint64_t* empty = bars + DP_NSLOT;           // [NSLOT] the pair's MMAs reading the slot completed (multicast commit)
  uint64_t* acc_bar = bars + 3 * DP_NSLOT;     // hidden-stage accumulator complete (multicast commit)
  uint64_t* accf_bar = acc_bar + 1;            // [3] final stages FA, FB, FC
  uint64_t* op_bar = acc_bar + 4;              // leader only: operands of the next stage written, 8 + 8 warp arrivals
  volatile int* abort_s = reinterpret_cast<volatile int*>(acc_bar + 5);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 6);

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const uint32_t rank = cluster_cta_rank();
  const int n_pairs = (int)(gridDim.x >> 1), pair = (int)(blockIdx.x >> 1);
  if (tid == 0) {
    for (int i = 0; i < DP_NSLOT; ++i) { mbar_init(&full[i], rank == 0 ? 2 : 1); mbar_init(&empty[i], 1); }
    mbar_init(acc_bar, 1);
    for (int i = 0; i < 3; ++i) mbar_init(&accf_bar[i], 1);
    mbar_init(op_bar, 16);
    *abort_s = 0;
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc2(tmem_slot, 256);
  for (int i = tid; i < DP_OFF_RING / 16; i += DP_THREADS) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                          // both CTAs' barriers exist before anything arrives on them
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  auto wait = [&](uint64_t* bar, uint32_t parity, int code) -> bool {
    for (uint32_t i = 0; i < (1u << 22); ++i) {            // bounded as in the one-CTA kernels: a protocol bug must not hang the GPU
      if (mbar_try_wait(bar, parity)) return true;            // suspends in hardware: a spinning probe loop would take
      if ((i & 1023) == 1023 && *abort_s) return false;          // issue slots from the row warps of the same scheduler
    }
    *abort_s = 1;
    raise_flag(a.err, code);
    return false;
  };

  if (warp < 8) {
    // ============================ row warps: four threads per flow row ============================
    const int q4 = warp & 3, wh = warp >> 2, nh = q4 >> 1;
    const int rr = (q4 & 1) * 32 + lane;                        // row within this CTA's 64
    const int rit = (int)rank * DP_ROWS + rr;                   // row within the 128-row tile
    const int rt = wh * 128 + q4 * 32 + lane;                   // index among the 256 row threads (constants staging)
    const int part = nh * 2 + wh;                               // which of the row's four threads
    const uint32_t tl = tmem + ((uint32_t)(q4 * 32) << 16);
    const uint32_t rowoff = (uint32_t)(rr / 8) * 128 + (uint32_t)(rr % 8) * 16;      // operand tile (1024 B per 8 columns)
    const uint32_t growoff = (uint32_t)(rit / 8) * 128 + (uint32_t)(rit % 8) * 16;   // saved images (2048 B per 8 columns)
    const int gcol0 = 128 * nh + 64 * wh;                       // first hidden unit of this thread's 64
    const int tcol0 = 64 * wh;                                  // ... and its TMEM column
    uint32_t acc_par = 0, accf_par = 0;
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
      if (lane == 0) { if (rank == 0) mbar_arrive(op_bar); else mbar_arrive_remote(op_bar, 0); }
    };
    auto put_x = [&](int jn, float v, int l, int tile) {
      const uint32_t o0 = (uint32_t)(jn / 8) * 1024 + rowoff + (jn % 8) * 2;
      const float xc = fminf(fmaxf(v, -60000.f), 60000.f);
      const __half hi = __float2half_rn(xc);
      *reinterpret_cast<__half*>(sXT + o0) = hi;
      *reinterpret_cast<__half*>(sXT + 2048 + o0) = __float2half_rn(xc - __half2float(hi));
      *reinterpret_cast<__nv_bfloat16*>(a.sv.xt + ((size_t)l * a.sv.tiles + tile) * 4096 + (uint32_t)(jn / 8) * 2048 + growoff +
                                        (jn % 8) * 2) = __float2bfloat16(dn_finite(v));
    };
    constexpr float ASC = 1.f / DN_WSCALE;
    int tile = pair;
    for (; tile < a.n_tiles && ok; tile += n_pairs) {
      const int64_t grow = (int64_t)tile * DN_ROWS + rit;
      const bool valid = grow < a.R;
      float* yin = ybuf + rr * 12;
      float* yout = ybuf + DP_ROWS * 12 + rr * 12;
      float ld_acc = 0.f;
      bool bad = false;
      if (part == 0) {
        for (int j = 0; j < D; ++j) {
          const float v = valid ? a.x[grow * D + j] : 0.f;
          yin[j] = v;
          put_x(D - 1 - j, v, 0, tile);
          a.sv.XS[((size_t)tile * D + j) * DN_ROWS + rit] = v;
        }
      }
      arrive_ops();                                             // x tile of layer 0
      for (int l = 0; l < s.L && ok; ++l) {
        if (rt == 0 && l + 1 < s.L) {   // ask the L2 for the next layer's gates of this CTA's rows (the whole tile block)
          const char* pj = reinterpret_cast<const char*>(a.proj + ((size_t)tile * a.ldproj + (size_t)(l + 1) * 3 * DN_H) * DN_ROWS);
          for (int o = (int)rank * 65536; o < 3 * DN_H * DN_ROWS * 4; o += 131072) bulk_prefetch_l2(pj + o, 65536);
        }
        // ---- this layer's constants -> shared memory (all 256 row threads) ----
        bar_sync(1, 256);
        {
          const float* srcs[5] = {a.b0 + (size_t)l * DN_H, a.ba + ((size_t)l * 2) * DN_H, a.bb + ((size_t)l * 2) * DN_H,
                                  a.ba + ((size_t)l * 2 + 1) * DN_H, a.bb + ((size_t)l * 2 + 1) * DN_H};
          float v[5], bfv[3];
#pragma unroll
          for (int i = 0; i < 5; ++i) v[i] = __ldg(srcs[i] + rt);
#pragma unroll
          for (int i = 0; i < 3; ++i) {
            const int e = rt + 256 * i, j = e / 48, k = e % 48;
            bfv[i] = (j < D && k < DN_MULT) ? __ldg(a.bf + (size_t)l * s.PLD + j * DN_MULT + k) : 0.f;
          }
#pragma unroll
          for (int i = 0; i < 5; ++i) cst[i * DN_H + rt] = v[i];
#pragma unroll
          for (int i = 0; i < 3; ++i) cst[5 * DN_H + rt + 256 * i] = bfv[i];
        }
        bar_sync(1, 256);
        mark(0);
        // the projection is tile-transposed [tile][L*3*H][128]: column c of this row at plt[c * 128]
        const float* plt = a.proj + ((size_t)tile * a.ldproj + (size_t)l * 3 * DN_H) * DN_ROWS + rit;
        // ---- five hidden stages: I, A1, B1, A2, B2 ----
#pragma unroll 1
        for (int st = 0; st < 5 && ok; ++st) {
          const bool res = (st == 0 || st == 2 || st == 4);
          const float* bias = cst + st * DN_H;
          const int sec = st == 0 ? 0 : (st == 2 ? 1 : 2);
          uint8_t* gimg = a.sv.img + dn_img_off(l, st, tile, a.sv.tiles);
          // the row's context terms / gates of a 32-column block (global memory, one 128-byte line per column and warp):
          // independent of the accumulator, so the first block's are loaded BEFORE the wait on the tensor pipe and the
          // second block's while the first block's operand words are being stored
          float g[32];
          auto load_gates = [&](int c0) {
            const float* pg = plt + (size_t)(sec * DN_H + c0) * DN_ROWS;
#pragma unroll
            for (int i = 0; i < 32; ++i) g[i] = __ldg(pg + (size_t)i * DN_ROWS);
          };
          if (res) load_gates(gcol0);
          if (!(ok = wait(acc_bar, acc_par, 10 + st))) break;
          mark(1);
          acc_par ^= 1;
          tc_fence_after();
#pragma unroll 1
          for (int b = 0; b < 2; ++b) {
            const int c0 = gcol0 + 32 * b;                       // hidden units c0 .. c0 + 31
            const uint32_t tc = (uint32_t)(tcol0 + 32 * b);      // their accumulator columns
            uint32_t v[32];
            float o[32];
            tmem_ld32(tl + tc, v);
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
#pragma unroll
                for (int i = 0; i < 32; ++i) o[i] = fmaf(__uint_as_float(v[i]), ASC, bias[c0 + i]) + g[i];
              } else {                        // h += gate * (acc + bb)
                uint32_t hr[32];
                tmem_ld32(tl + 128 + tc, hr);
                tmem_wait_ld();
#pragma unroll
                for (int i = 0; i < 32; ++i)
                  o[i] = fmaf(g[i], fmaf(__uint_as_float(v[i]), ASC, bias[c0 + i]), __uint_as_float(hr[i]));
              }
              uint32_t ow[32];
#pragma unroll
              for (int i = 0; i < 32; ++i) ow[i] = __float_as_uint(o[i]);
              tmem_st16(tl + 128 + tc, ow);
              tmem_st16(tl + 128 + tc + 16, ow + 16);
              if (b == 0) load_gates(c0 + 32);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              uint4 w4, l4, b4;
              if (st == 4) {
                split_pack<false>(o[8 * i], o[8 * i + 1], w4.x, l4.x); split_pack<false>(o[8 * i + 2], o[8 * i + 3], w4.y, l4.y);
                split_pack<false>(o[8 * i + 4], o[8 * i + 5], w4.z, l4.z); split_pack<false>(o[8 * i + 6], o[8 * i + 7], w4.w, l4.w);
                b4.x = pack_bf16(o[8 * i], o[8 * i + 1]); b4.y = pack_bf16(o[8 * i + 2], o[8 * i + 3]);
                b4.z = pack_bf16(o[8 * i + 4], o[8 * i + 5]); b4.w = pack_bf16(o[8 * i + 6], o[8 * i + 7]);
              } else {
                split_pack<true>(o[8 * i], o[8 * i + 1], w4.x, l4.x); split_pack<true>(o[8 * i + 2], o[8 * i + 3], w4.y, l4.y);
                split_pack<true>(o[8 * i + 4], o[8 * i + 5], w4.z, l4.z); split_pack<true>(o[8 * i + 6], o[8 * i + 7], w4.w, l4.w);
                b4.x = pack_bf16_relu(o[8 * i], o[8 * i + 1]); b4.y = pack_bf16_relu(o[8 * i + 2], o[8 * i + 3]);
                b4.z = pack_bf16_relu(o[8 * i + 4], o[8 * i + 5]); b4.w = pack_bf16_relu(o[8 * i + 6], o[8 * i + 7]);
              }
              const size_t io = (size_t)(c0 / 8 + i) * 1024 + rowoff;
              *reinterpret_cast<uint4*>(sACT + io) = w4;
              *reinterpret_cast<uint4*>(sACTLO + io) = l4;
              *reinterpret_cast<uint4*>(gimg + (size_t)(c0 / 8 + i) * 2048 + growoff) = b4;
            }
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
                const size_t io = (size_t)(c0 / 8 + i) * 2048 + growoff;
                *reinterpret_cast<uint4*>(vhi + io) = w4;
                *reinterpret_cast<uint4*>(vlo + io) = l4;
              }
            }
          }
          if (res) tmem_wait_st();
          arrive_ops();
          mark(2);
        }
        if (!ok) break;
        // ---- final layer in three stages of <= 4 dims (two per output half): every thread of a row owns one dim ----
        const float* bfc = cst + 5 * DN_H;
#pragma unroll 1
        for (int f = 0; f < 3 && ok; ++f) {
          if (!(ok = wait(&accf_bar[f], accf_par, 16 + f))) break;
          mark(3);
          tc_fence_after();
          const int j = f < 2 ? 1 + 4 * f + 2 * nh + wh : (wh == 0 ? 9 + nh : (nh == 0 ? 0 : -1));
          if (j < 0 || j >= D) continue;
          float q[48];
#pragma unroll
          for (int i = 0; i < 48; i += 4) {
            const float4 t4 = *reinterpret_cast<const float4*>(bfc + j * 48 + i);
            q[i] = t4.x; q[i + 1] = t4.y; q[i + 2] = t4.z; q[i + 3] = t4.w;
          }
          if (j > 0) {
            const uint32_t col = f < 2 ? (uint32_t)(DP_NHF * f + DN_MULT * wh) : (uint32_t)(2 * DP_NHF);
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
          {
            float* pr = a.sv.P + (((size_t)l * a.sv.tiles + tile) * s.PLD + j * DN_MULT) * DN_ROWS + rit;
#pragma unroll
            for (int i = 0; i < DN_MULT; ++i) pr[(size_t)i * DN_ROWS] = q[i];
          }
          float lad;
          const float y = dn_rqs_forward(q, xin, a.tail, lad);
          ld_acc += lad;
          yout[j] = y;
          if (l + 1 < s.L) {
            put_x(D - 1 - j, y, l + 1, tile);
            a.sv.XS[(((size_t)(l + 1) * a.sv.tiles + tile) * D + j) * DN_ROWS + rit] = y;
          }
          mark(4);
        }
        if (!ok) break;
        accf_par ^= 1;
        if (l + 1 < s.L) arrive_ops();                          // x tile of the next layer
        float* tsw = yin; yin = yout; yout = tsw;
      }
      if (!ok) break;
      // ---- z, logdet: combine the four threads of every row ----
      red[part * DP_ROWS + rr] = ld_acc;
      red[(4 + part) * DP_ROWS + rr] = bad ? 1.f : 0.f;
      bar_sync(2, 256);
      if (valid && wh == 0) {
        // a non-finite value that entered the conditioner turns the whole row into NaN, as the reference's does
        bool anybad = false;
        float ld = 0.f;
#pragma unroll
        for (int p = 0; p < 4; ++p) { anybad = anybad || red[(4 + p) * DP_ROWS + rr] != 0.f; ld += red[p * DP_ROWS + rr]; }
        const int j0 = nh ? (D + 1) / 2 : 0, j1 = nh ? D : (D + 1) / 2;
        for (int j = j0; j < j1; ++j) a.z[grow * D + j] = anybad ? __int_as_float(0x7fc00000) : yin[j];
        if (nh == 0 && a.logdet) a.logdet[grow] = anybad ? __int_as_float(0x7fc00000) : ld;
      }
      bar_sync(2, 256);                                         // red / ybuf are reused by the next tile
    }
    if (!ok && part == 0) {   // barrier timeout (flag raised): abandoned and unreached tiles come back as NaN rows
      for (; tile < a.n_tiles; tile += n_pairs) {
        const int64_t grow = (int64_t)tile * DN_ROWS + rit;
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
    // ============================ MMA issuer (leader) / weight relay (peer) ============================
    bool ok = true;
    if (rank != 0) {
      // peer: tell the leader when this CTA's half of a chunk has landed (the leader's tensor instruction reads both halves)
      uint32_t slot = 0, full_par = 0;
      for (int tile = pair; tile < a.n_tiles && ok; tile += n_pairs)
        for (int i = 0; i < s.L * DP_PACK_SLOTS && ok; ++i) {
          if (!(ok = wait(&full[slot], full_par, 31))) break;
          if (elect_one_sync()) mbar_arrive_remote(&full[slot], 0);
          __syncwarp();
          if (++slot == DP_NSLOT) { slot = 0; full_par ^= 1; }
        }
    } else {
      uint32_t op_par = 0, slot = 0;
      constexpr uint32_t FMT = (1u << 7) | (1u << 10);            // fp16 x fp16 -> fp32
      const uint32_t idesc_h = umma_idesc_bf16(128, 256) & ~FMT;
      const uint32_t idesc_f = umma_idesc_bf16(128, 2 * DP_NHF) & ~FMT, idesc_c = umma_idesc_bf16(128, 2 * DP_NHC) & ~FMT;
      // descriptors advance by adding to their start-address field (16-byte units; shared addresses stay below 2^18)
      const uint64_t adX = umma_smem_desc(smem_u32(sXT), 1024, 128), adA = umma_smem_desc(smem_u32(sACT), 1024, 128);
      const uint64_t bd128 = umma_smem_desc(smem_u32(ring), 128 * 16, 128), bdF = umma_smem_desc(smem_u32(ring), DP_NHF * 16, 128),
                     bdC = umma_smem_desc(smem_u32(ring), DP_NHC * 16, 128);
      // chunks whose two halves are known to have landed, ahead of the issue pointer: confirmed while the warp would
      // otherwise idle on the operand barrier, so that the issue loop itself polls nothing (a poll costs ~70 cycles, the
      // three MMAs of a chunk 195: polling inline made the issuing thread, not the tensor pipe, the pacer)
      uint32_t cslot = 0, cpar = 0;
      int ready = 0;
      auto confirm = [&]() {
        while (ready < DP_NSLOT) {
          if (!mbar_test_wait(&full[cslot], cpar)) break;
          ++ready;
          if (++cslot == DP_NSLOT) { cslot = 0; cpar ^= 1; }
        }
      };
      const bool mprof = a.prof && blockIdx.x == 0;
      long long m_op = 0, m_w = 0, m_issue = 0;
      for (int tile = pair; tile < a.n_tiles && ok; tile += n_pairs) {
        for (int l = 0; l < s.L && ok; ++l) {
          for (int st = 0; st < 8 && ok; ++st) {               // I, A1, B1, A2, B2, FA, FB, FC
            const int NH = st < 5 ? 128 : (st < 7 ? DP_NHF : DP_NHC);
            const int nk = st == 0 ? 1 : 16;
            const uint32_t idesc = st < 5 ? idesc_h : (st < 7 ? idesc_f : idesc_c);
            const uint64_t bd0 = st < 5 ? bd128 : (st < 7 ? bdF : bdC);
            const uint32_t acc = tmem + (st < 6 ? 0u : (st == 6 ? (uint32_t)DP_NHF : (uint32_t)(2 * DP_NHF)));
            uint64_t* done = st < 5 ? acc_bar : &accf_bar[st - 5];
            if (st <= 5) {                                        // FB, FC read the same operand tile as FA
              const long long t0 = mprof ? clock64() : 0;
              bool got = false;
              for (uint32_t i = 0; i < (1u << 26); ++i) {
                if (mbar_test_wait(op_bar, op_par)) { got = true; break; }
                confirm();
                if ((i & 1023) == 1023 && *abort_s) break;
              }
              if (!got) { if (!*abort_s) { *abort_s = 1; raise_flag(a.err, 20 + st); } ok = false; break; }
              if (mprof) m_op += clock64() - t0;
              op_par ^= 1;
            }
            for (int c0 = 0; c0 < nk && ok; ++c0) {
              const long long t1 = mprof ? clock64() : 0;
              if (ready == 0) {
                if (!(ok = wait(&full[cslot], cpar, 31))) break;
                ++ready;
                if (++cslot == DP_NSLOT) { cslot = 0; cpar ^= 1; }
              }
              --ready;
              const long long t2 = mprof ? clock64() : 0;
              tc_fence_after();
              if (elect_one_sync()) {
                const uint64_t bh = bd0 + (uint64_t)((slot * DP_SLOT) >> 4);
                const uint64_t bl = bh + (uint64_t)((NH * 32) >> 4);
                const uint64_t ah = st == 0 ? adX : adA + (uint64_t)(c0 * (2048 >> 4));
                const uint64_t al = ah + (uint64_t)((st == 0 ? 2048 : 32768) >> 4);
                umma_ss2(acc, ah, bh, idesc, c0 > 0 ? 1u : 0u);
                umma_ss2(acc, al, bh, idesc, 1u);
                umma_ss2(acc, ah, bl, idesc, 1u);
                umma_commit2(&empty[slot], 3);
                if (c0 + 1 >= nk) umma_commit2(done, 3);
              }
              __syncwarp();
              if (mprof) { const long long t3 = clock64(); m_w += t2 - t1; m_issue += t3 - t2; }
              if (++slot == DP_NSLOT) slot = 0;
            }
          }
        }
      }
      if (mprof && elect_one_sync()) { a.prof[8] = m_op; a.prof[9] = m_w; a.prof[10] = 0; a.prof[11] = m_issue; }
    }
  } else if (warp == 9) {
    // ============================ weight loader (each CTA: its half of every chunk) ============================
    uint32_t slot = 0, empty_par = 1;
    bool ok = true;
    for (int tile = pair; tile < a.n_tiles && ok; tile += n_pairs) {
      for (int l = 0; l < s.L && ok; ++l) {
        const uint8_t* src = a.image + ((size_t)l * 2 + rank) * DP_LAYER_BYTES;
        for (int st = 0; st < 8 && ok; ++st) {
          const uint32_t bytes = 64u * (uint32_t)(st < 5 ? 128 : (st < 7 ? DP_NHF : DP_NHC));
          const int nk = st == 0 ? 1 : 16;
          for (int c0 = 0; c0 < nk; ++c0) {
            if (!(ok = wait(&empty[slot], empty_par, 40))) break;
            if (elect_one_sync()) {
              mbar_arrive_expect_tx(&full[slot], bytes);
              bulk_g2s(ring + (size_t)slot * DP_SLOT, src, bytes, &full[slot]);
            }
            __syncwarp();
            src += bytes;
            if (++slot == DP_NSLOT) { slot = 0; empty_par ^= 1; }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                          // nobody leaves while the other CTA may still signal it
  if (warp == 8) tmem_dealloc2(tmem, 256);
}
