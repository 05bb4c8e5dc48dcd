// Density direction for inference (log-prob, bf16 operands) on CTA pairs with TWO tiles in flight per pair.
// The one-CTA kernel (tc_dense_forward_kernel<1, false>) keeps the tensor pipe 38 % busy: accumulator + fp32 residual
// stream fill the 512 TMEM columns, so a tile's epilogue and its MMAs cannot overlap.  On a CTA pair (cta_group::2,
// M = 128: see pf_tc_pair.cuh) a 128-row tile needs only 128 + 128 TMEM columns and a 32 KB operand tile per CTA, which
// leaves room for a SECOND tile: the row warps run the epilogue of stream 0 while the MMAs of stream 1 execute, and vice
// versa.  Both streams are at the same stage of the same layer (stream 1 half a stage behind), so a stage's weights are
// loaded ONCE per pair of tiles into one of two whole-stage buffers (64 KB) instead of once per tile through a ring:
// half the L2 -> shared-memory weight traffic per row of the one-CTA kernel.
// Included by pf_tc_dense.cu (inside namespace pf): shares DnArgs, the spline and the pack helpers.
#pragma once

constexpr int D2_ROWS = 64;              // rows of a tile owned by one CTA
constexpr int D2_THREADS = 320;
constexpr int D2_WBUF = 65536;           // one stage of weights of this CTA: 16 chunks x (128 outputs x 16 K x 2 B)
constexpr int D2_NSTAGE = 8;             // I, A1, B1, A2, B2, FA, FB, FC
constexpr int D2_NHF = 96, D2_NHC = 48;  // outputs per CTA of FA / FB (2 dims) and FC (1 dim)
__host__ __device__ constexpr uint32_t d2_stage_bytes(int st) {
  return st == 0 ? 2u * 4096u : (st < 5 ? 16u * 4096u : (st < 7 ? 16u * 32u * D2_NHF : 16u * 32u * D2_NHC));
}
constexpr uint32_t D2_LAYER_BYTES = 2u * 4096u + 4u * 65536u + 2u * 16u * 32u * D2_NHF + 16u * 32u * D2_NHC;   // per layer and rank
__host__ __device__ constexpr uint32_t d2_stage_off(int st) {
  return st == 0 ? 0u : (st < 6 ? 8192u + (uint32_t)(st - 1) * 65536u
                                : 8192u + 4u * 65536u + (uint32_t)(st - 5) * 16u * 32u * D2_NHF);
}

constexpr int D2_OFF_W = 0;                                   // [2 buffers][64 KB]
constexpr int D2_OFF_ACT = 2 * D2_WBUF;                       // [2 streams][64 x 256] bf16 (1024 B per 8 columns)
constexpr int D2_OFF_XT = D2_OFF_ACT + 2 * 32768;             // [2 streams][64 x 32] bf16: x hi (2 KB) | x lo (2 KB)
constexpr int D2_OFF_CST = D2_OFF_XT + 2 * 4096;              // floats: 5 x 256 biases | [16 dims][48] final-layer biases
constexpr int D2_CST_FLOATS = 5 * DN_H + 16 * 48;
constexpr int D2_OFF_Y = D2_OFF_CST + D2_CST_FLOATS * 4;      // [2 streams][2][64][12] fp32 layer inputs / outputs
constexpr int D2_OFF_RED = D2_OFF_Y + 2 * 2 * D2_ROWS * 12 * 4;   // [2 streams][8][64]: logdet partials, bad flags
constexpr int D2_OFF_BAR = D2_OFF_RED + 2 * 8 * D2_ROWS * 4;
constexpr int D2_SMEM = D2_OFF_BAR + 512;
static_assert(D2_SMEM <= 232448, "shared-memory budget of one CTA");

// bf16 pair image, [L][2 ranks][D2_LAYER_BYTES]: the stages in consumption order, each the rank's half of every K = 16
// chunk as [2 K-slabs][NH][8] (UMMA K-major no-swizzle).  Stage I holds W0 twice (for the hi and the lo half of x).
__global__ void d2_pack_kernel(DnShape s, uint8_t* __restrict__ imagep, const float* __restrict__ W0,
                               const float* __restrict__ Wa, const float* __restrict__ Wb, const float* __restrict__ Wf) {
  const int l = blockIdx.y, rank = blockIdx.z;
  int st, cc;     // stage, chunk within the stage
  if (blockIdx.x < 2) { st = 0; cc = blockIdx.x; }
  else { st = 1 + (blockIdx.x - 2) / 16; cc = (blockIdx.x - 2) % 16; }
  const int NH = st < 5 ? 128 : (st < 7 ? D2_NHF : D2_NHC);
  __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(imagep + ((size_t)l * 2 + rank) * D2_LAYER_BYTES + d2_stage_off(st) +
                                                        (size_t)cc * NH * 32);
  const int H = DN_H;
  for (int e = threadIdx.x; e < NH * 16; e += blockDim.x) {
    const int n = e / 16, kk = e % 16;
    float v = 0.f;
    if (st == 0) {
      if (kk < s.D) v = W0[((size_t)l * H + rank * 128 + n) * s.DP + kk];
    } else if (st <= 4) {
      const int blk = (st - 1) / 2;
      const float* W = ((st - 1) & 1) ? Wb : Wa;
      v = W[(((size_t)l * 2 + blk) * H + rank * 128 + n) * H + cc * 16 + kk];
    } else {
      int p = 0;
      const int d = dp_final_dim(st - 5, rank, n, p);
      if (d >= 0 && d < s.D) v = Wf[((size_t)l * s.PLD + (size_t)d * DN_MULT + p) * H + cc * 16 + kk];
    }
    dst[(kk / 8) * (NH * 8) + (n / 8) * 64 + (n % 8) * 8 + (kk % 8)] = __float2bfloat16(v);
  }
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(D2_THREADS, 1) tc_dense_pair2_kernel(const DnArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  const DnShape& s = a.s;
  const int D = s.D;
  uint8_t* sW = smem + D2_OFF_W;
  uint8_t* sACT = smem + D2_OFF_ACT;
  uint8_t* sXT = smem + D2_OFF_XT;
  float* cst = reinterpret_cast<float*>(smem + D2_OFF_CST);
  float* ybuf = reinterpret_cast<float*>(smem + D2_OFF_Y);
  float* red = reinterpret_cast<float*>(smem + D2_OFF_RED);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + D2_OFF_BAR);
  uint64_t* wfull = bars;              // [2] this CTA's half of a stage's weights landed
  uint64_t* wempty = bars + 2;         // [2] both streams' MMAs of the stage completed (multicast commit)
  uint64_t* pwfull = bars + 4;         // [2] leader only: the peer's half landed (relay)
  uint64_t* acc_bar = bars + 6;        // [2 streams] hidden-stage accumulator complete (multicast commit)
  uint64_t* accf_bar = bars + 8;       // [2 streams][3] final stages FA, FB, FC
  uint64_t* op_bar = bars + 14;        // [2 streams] leader only: operands of the stream's next stage written (8 + 8 warps)
  volatile int* abort_s = reinterpret_cast<volatile int*>(bars + 16);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 17);

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const uint32_t rank = cluster_cta_rank();
  const int n_pairs = (int)(gridDim.x >> 1), pair = (int)(blockIdx.x >> 1);
  const int n_super = (a.n_tiles + 1) / 2;     // a pair works on tiles 2 i and 2 i + 1 at a time
  if (tid == 0) {
    for (int i = 0; i < 2; ++i) { mbar_init(&wfull[i], 1); mbar_init(&wempty[i], 1); mbar_init(&pwfull[i], 1); mbar_init(&acc_bar[i], 1); }
    for (int i = 0; i < 6; ++i) mbar_init(&accf_bar[i], 1);
    for (int i = 0; i < 2; ++i) mbar_init(&op_bar[i], 16);
    *abort_s = 0;
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc2(tmem_slot, 512);
  for (int i = tid; i < (2 * 32768 + 2 * 4096) / 16; i += D2_THREADS) reinterpret_cast<uint4*>(sACT)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  auto wait = [&](uint64_t* bar, uint32_t parity, int code) -> bool {
    for (uint32_t i = 0; i < (1u << 26); ++i) {
      if (mbar_test_wait(bar, parity)) return true;
      if ((i & 1023) == 1023 && *abort_s) return false;
    }
    *abort_s = 1;
    raise_flag(a.err, code);
    return false;
  };

  if (warp < 8) {
    // ============================ row warps: four threads per flow row, two tiles (streams) interleaved ============================
    const int q4 = warp & 3, wh = warp >> 2, nh = q4 >> 1;
    const int rr = (q4 & 1) * 32 + lane;
    const int rit = (int)rank * D2_ROWS + rr;
    const int rt = wh * 128 + q4 * 32 + lane;
    const int part = nh * 2 + wh;
    const uint32_t tl = tmem + ((uint32_t)(q4 * 32) << 16);
    const uint32_t rowoff = (uint32_t)(rr / 8) * 128 + (uint32_t)(rr % 8) * 16;
    const int gcol0 = 128 * nh + 64 * wh;
    const int tcol0 = 64 * wh;
    uint32_t acc_par = 0, accf_par = 0;          // both streams flip their barriers equally often: one parity each
    bool ok = true;
    const bool prof = a.prof && blockIdx.x == 0 && tid == 0;
    long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // 0 consts, 1 hidden wait, 2 hidden compute, 3 F wait, 4 spline, 5 total
    long long tmark = clock64();
    const long long t_begin = tmark;
    auto mark = [&](int k) { if (prof) { const long long n = clock64(); pc[k] += n - tmark; tmark = n; } };
    auto arrive_ops = [&](int t) {
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { if (rank == 0) mbar_arrive(&op_bar[t]); else mbar_arrive_remote(&op_bar[t], 0); }
    };
    auto put_x = [&](int t, int jn, float v) {
      uint8_t* xt = sXT + t * 4096;
      const uint32_t o0 = (uint32_t)(jn / 8) * 1024 + rowoff + (jn % 8) * 2;
      const float xc = dn_finite(v);
      const __nv_bfloat16 hi = __float2bfloat16(xc);
      *reinterpret_cast<__nv_bfloat16*>(xt + o0) = hi;
      *reinterpret_cast<__nv_bfloat16*>(xt + 2048 + o0) = __float2bfloat16(xc - __bfloat162float(hi));
    };
    int sup = pair;
    for (; sup < n_super && ok; sup += n_pairs) {
      // per-stream row state in scalars (selected with t ? .. : ..): arrays indexed by the runtime stream id would live
      // in local memory
      const int tile0 = 2 * sup, tile1 = 2 * sup + 1;
      const int64_t grow0 = (int64_t)tile0 * DN_ROWS + rit, grow1 = (int64_t)tile1 * DN_ROWS + rit;
      const bool valid0 = grow0 < a.R, valid1 = tile1 < a.n_tiles && grow1 < a.R;
      const float* projr0 = a.proj + a.cm(valid0 ? grow0 : 0) * a.ldproj;
      const float* projr1 = a.proj + a.cm(valid1 ? grow1 : 0) * a.ldproj;
      float ld0 = 0.f, ld1 = 0.f;
      bool bad0 = false, bad1 = false;
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        const bool vd = t ? valid1 : valid0;
        const int64_t gr = t ? grow1 : grow0;
        float* yin = ybuf + (t * 2) * D2_ROWS * 12 + rr * 12;
        if (part == 0) {
          for (int j = 0; j < D; ++j) {
            const float v = vd ? a.x[gr * D + j] : 0.f;
            yin[j] = v;
            put_x(t, D - 1 - j, v);
          }
        }
        arrive_ops(t);                                          // x tile of layer 0
      }
      int ysel = 0;                                             // which y buffer holds the layer's input
      for (int l = 0; l < s.L && ok; ++l) {
        // ---- this layer's constants -> shared memory (all 256 row threads; shared by both streams) ----
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
        // The gates / context terms of this layer are read per row from the projection (row-major, one 3 KB block per
        // event and layer).  A CTA's 64 rows of a stream belong to at most two events when rows map to contexts in runs of
        // >= 64: pull those blocks into L1 now (24 lines each), so that the epilogues' loads are L1 hits instead of one
        // exposed L2 round trip per 32-column block (measured: 825 -> cycles per block against 475 with staged terms).
        if (!a.cm.row2ctx && a.cm.n_per_ctx >= D2_ROWS && rt < 96) {
          const int t = rt / 48, ev = (rt / 24) & 1, line = rt % 24;
          const int64_t base_row = (int64_t)(2 * sup + t) * DN_ROWS + (int64_t)rank * D2_ROWS;
          int64_t r = base_row + (ev ? D2_ROWS - 1 : 0);
          if (r >= a.R) r = a.R - 1;
          if (base_row < a.R) {
            const float* pb = a.proj + a.cm(r) * a.ldproj + (size_t)l * 3 * DN_H + line * 32;
            asm volatile("prefetch.global.L1 [%0];" ::"l"(pb));
          }
        }
        mark(0);
        // ---- five hidden stages, stream 0 then stream 1: one stream's epilogue runs under the other's MMAs ----
#pragma unroll 1
        for (int st = 0; st < 5 && ok; ++st) {
          const bool res = (st == 0 || st == 2 || st == 4);
          const float* bias = cst + st * DN_H;
          const int sec = st == 0 ? 0 : (st == 2 ? 1 : 2);
#pragma unroll 1
          for (int t = 0; t < 2 && ok; ++t) {
            if (!(ok = wait(&acc_bar[t], acc_par, 10 + st))) break;
            mark(1);
            tc_fence_after();
            const uint32_t tb = tl + 256u * t;                   // the stream's accumulator | residual columns
            uint8_t* act = sACT + t * 32768;
            const float* pl = (t ? projr1 : projr0) + (size_t)l * 3 * DN_H + sec * DN_H;
#pragma unroll 1
            for (int b = 0; b < 2; ++b) {
              const int c0 = gcol0 + 32 * b;
              const uint32_t tc = (uint32_t)(tcol0 + 32 * b);
              uint32_t v[32];
              float o[32];
              tmem_ld32(tb + tc, v);
              if (!res) {                       // A stages: u = acc + ba
                tmem_wait_ld();
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                  const float4 t4 = *reinterpret_cast<const float4*>(bias + c0 + i);
                  o[i] = __uint_as_float(v[i]) + t4.x; o[i + 1] = __uint_as_float(v[i + 1]) + t4.y;
                  o[i + 2] = __uint_as_float(v[i + 2]) + t4.z; o[i + 3] = __uint_as_float(v[i + 3]) + t4.w;
                }
              } else {
                if (st == 0) {                  // h0 = acc + b0 + relu(c0)
                  tmem_wait_ld();
#pragma unroll
                  for (int i = 0; i < 32; i += 4) {
                    const float4 t4 = *reinterpret_cast<const float4*>(bias + c0 + i);
                    const float4 g4 = __ldg(reinterpret_cast<const float4*>(pl + c0 + i));
                    o[i] = __uint_as_float(v[i]) + t4.x + g4.x; o[i + 1] = __uint_as_float(v[i + 1]) + t4.y + g4.y;
                    o[i + 2] = __uint_as_float(v[i + 2]) + t4.z + g4.z; o[i + 3] = __uint_as_float(v[i + 3]) + t4.w + g4.w;
                  }
                } else {                        // h += gate * (acc + bb)
                  uint32_t hr[32];
                  tmem_ld32(tb + 128 + tc, hr);
                  tmem_wait_ld();
#pragma unroll
                  for (int i = 0; i < 32; i += 4) {
                    const float4 t4 = *reinterpret_cast<const float4*>(bias + c0 + i);
                    const float4 g4 = __ldg(reinterpret_cast<const float4*>(pl + c0 + i));
                    o[i] = fmaf(g4.x, __uint_as_float(v[i]) + t4.x, __uint_as_float(hr[i]));
                    o[i + 1] = fmaf(g4.y, __uint_as_float(v[i + 1]) + t4.y, __uint_as_float(hr[i + 1]));
                    o[i + 2] = fmaf(g4.z, __uint_as_float(v[i + 2]) + t4.z, __uint_as_float(hr[i + 2]));
                    o[i + 3] = fmaf(g4.w, __uint_as_float(v[i + 3]) + t4.w, __uint_as_float(hr[i + 3]));
                  }
                }
                uint32_t ow[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) ow[i] = __float_as_uint(o[i]);
                tmem_st16(tb + 128 + tc, ow);
                tmem_st16(tb + 128 + tc + 16, ow + 16);
              }
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
                *reinterpret_cast<uint4*>(act + (size_t)(c0 / 8 + i) * 1024 + rowoff) = w4;
              }
            }
            if (res) tmem_wait_st();
            arrive_ops(t);
            mark(2);
          }
          acc_par ^= 1;
        }
        if (!ok) break;
        // ---- final layer in three stages of <= 4 dims: every thread of a row owns one dim per stage and stream ----
        const float* bfc = cst + 5 * DN_H;
#pragma unroll 1
        for (int t = 0; t < 2 && ok; ++t) {
          float* yin = ybuf + (t * 2 + ysel) * D2_ROWS * 12 + rr * 12;
          float* yout = ybuf + (t * 2 + (ysel ^ 1)) * D2_ROWS * 12 + rr * 12;
          const uint32_t tb = tl + 256u * t;
#pragma unroll 1
          for (int f = 0; f < 3 && ok; ++f) {
            if (!(ok = wait(&accf_bar[t * 3 + f], accf_par, 16 + f))) break;
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
              const uint32_t col = f < 2 ? (uint32_t)(D2_NHF * f + DN_MULT * wh) : (uint32_t)(2 * D2_NHF);
              uint32_t v[48];
              tmem_ld16(tb + col, v);
              tmem_ld16(tb + col + 16, v + 16);
              tmem_ld16(tb + col + 32, v + 32);
              tmem_wait_ld();
#pragma unroll
              for (int i = 0; i < DN_MULT; ++i) q[i] += __uint_as_float(v[i]);
            }
            const float xin = yin[D - 1 - j];
            const bool isbad = !(fabsf(xin) <= 3.0e38f);
            float lad;
            const float y = dn_rqs_forward(q, xin, a.tail, lad);
            if (t) { ld1 += lad; bad1 = bad1 || isbad; } else { ld0 += lad; bad0 = bad0 || isbad; }
            yout[j] = y;
            if (l + 1 < s.L) put_x(t, D - 1 - j, y);
            mark(4);
          }
          if (!ok) break;
          if (l + 1 < s.L) arrive_ops(t);                       // x tile of the stream's next layer
        }
        if (!ok) break;
        accf_par ^= 1;
        ysel ^= 1;
      }
      if (!ok) break;
      // ---- z, logdet: combine the four threads of every row, stream by stream ----
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        float* rd = red + t * 8 * D2_ROWS;
        rd[part * D2_ROWS + rr] = t ? ld1 : ld0;
        rd[(4 + part) * D2_ROWS + rr] = (t ? bad1 : bad0) ? 1.f : 0.f;
      }
      bar_sync(2, 256);
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        const bool vd = t ? valid1 : valid0;
        const int64_t gr = t ? grow1 : grow0;
        if (vd && wh == 0) {
          const float* rd = red + t * 8 * D2_ROWS;
          const float* yfin = ybuf + (t * 2 + ysel) * D2_ROWS * 12 + rr * 12;
          bool anybad = false;
          float ld = 0.f;
#pragma unroll
          for (int p = 0; p < 4; ++p) { anybad = anybad || rd[(4 + p) * D2_ROWS + rr] != 0.f; ld += rd[p * D2_ROWS + rr]; }
          const int j0 = nh ? (D + 1) / 2 : 0, j1 = nh ? D : (D + 1) / 2;
          for (int j = j0; j < j1; ++j) a.z[gr * D + j] = anybad ? __int_as_float(0x7fc00000) : yfin[j];
          if (nh == 0 && a.logdet) a.logdet[gr] = anybad ? __int_as_float(0x7fc00000) : ld;
        }
      }
      bar_sync(2, 256);                                         // red / ybuf are reused by the next pair of tiles
    }
    if (!ok && part == 0) {   // barrier timeout (flag raised): abandoned and unreached tiles come back as NaN rows
      for (; sup < n_super; sup += n_pairs)
        for (int t = 0; t < 2; ++t) {
          const int64_t gr = (int64_t)(2 * sup + t) * DN_ROWS + rit;
          if (gr >= a.R) continue;
          for (int jj = 0; jj < D; ++jj) a.z[gr * D + jj] = __int_as_float(0x7fc00000);
          if (a.logdet) a.logdet[gr] = __int_as_float(0x7fc00000);
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
      uint32_t buf = 0, par = 0;
      for (int sup = pair; sup < n_super && ok; sup += n_pairs)
        for (int i = 0; i < s.L * D2_NSTAGE && ok; ++i) {
          if (!(ok = wait(&wfull[buf], par, 31))) break;
          if (elect_one_sync()) mbar_arrive_remote(&pwfull[buf], 0);    // this CTA's half of the stage has landed
          __syncwarp();
          if (++buf == 2) { buf = 0; par ^= 1; }
        }
    } else {
      uint32_t op_par = 0, buf = 0, wpar = 0;
      const uint32_t idesc_h = umma_idesc_bf16(128, 256);
      const uint32_t idesc_f = umma_idesc_bf16(128, 2 * D2_NHF), idesc_c = umma_idesc_bf16(128, 2 * D2_NHC);
      const uint64_t adX = umma_smem_desc(smem_u32(sXT), 1024, 128), adA = umma_smem_desc(smem_u32(sACT), 1024, 128);
      const uint64_t bd128 = umma_smem_desc(smem_u32(sW), 128 * 16, 128), bdF = umma_smem_desc(smem_u32(sW), D2_NHF * 16, 128),
                     bdC = umma_smem_desc(smem_u32(sW), D2_NHC * 16, 128);
      for (int sup = pair; sup < n_super && ok; sup += n_pairs) {
        for (int l = 0; l < s.L && ok; ++l) {
          for (int st = 0; st < D2_NSTAGE && ok; ++st) {
            const int NH = st < 5 ? 128 : (st < 7 ? D2_NHF : D2_NHC);
            const int nk = st == 0 ? 2 : 16;
            const uint32_t idesc = st < 5 ? idesc_h : (st < 7 ? idesc_f : idesc_c);
            const uint64_t bd0 = (st < 5 ? bd128 : (st < 7 ? bdF : bdC)) + (uint64_t)((buf * D2_WBUF) >> 4);
            const uint32_t accc = st < 6 ? 0u : (st == 6 ? (uint32_t)D2_NHF : (uint32_t)(2 * D2_NHF));
            if (!(ok = wait(&wfull[buf], wpar, 30))) break;      // the stage's weights: this CTA's half and the peer's
            if (!(ok = wait(&pwfull[buf], wpar, 32))) break;
#pragma unroll 1
            for (int t = 0; t < 2 && ok; ++t) {
              if (st <= 5) {                                      // FB, FC read the same operand tile as FA
                if (!(ok = wait(&op_bar[t], op_par, 20 + st))) break;
              }
              tc_fence_after();
              if (elect_one_sync()) {
                const uint32_t acc = tmem + 256u * t + accc;
                uint64_t* done = st < 5 ? &acc_bar[t] : &accf_bar[t * 3 + (st - 5)];
                if (st == 0) {                                    // x hi . W0 + x lo . W0
                  const uint64_t ax = adX + (uint64_t)((t * 4096) >> 4);
                  umma_ss2(acc, ax, bd0, idesc, 0u);
                  umma_ss2(acc, ax + (2048 >> 4), bd0 + (4096 >> 4), idesc, 1u);
                } else {
                  const uint64_t aa = adA + (uint64_t)((t * 32768) >> 4);
#pragma unroll 4
                  for (int c0 = 0; c0 < 16; ++c0)
                    umma_ss2(acc, aa + (uint64_t)(c0 * (2048 >> 4)), bd0 + (uint64_t)((c0 * NH * 32) >> 4), idesc, c0 > 0 ? 1u : 0u);
                }
                umma_commit2(done, 3);
                if (t == 1) umma_commit2(&wempty[buf], 3);        // both streams are through with the stage's weights
              }
              __syncwarp();
            }
            if (st <= 5) op_par ^= 1;
            (void)nk;
            if (++buf == 2) { buf = 0; wpar ^= 1; }
          }
        }
      }
    }
  } else if (warp == 9) {
    // ============================ weight loader: whole stages, this CTA's half ============================
    uint32_t buf = 0, epar = 1;
    bool ok = true;
    for (int sup = pair; sup < n_super && ok; sup += n_pairs) {
      for (int l = 0; l < s.L && ok; ++l) {
        const uint8_t* lsrc = a.image + ((size_t)l * 2 + rank) * D2_LAYER_BYTES;
        for (int st = 0; st < D2_NSTAGE && ok; ++st) {
          const uint32_t bytes = d2_stage_bytes(st);
          if (!(ok = wait(&wempty[buf], epar, 40))) break;
          if (elect_one_sync()) {
            mbar_arrive_expect_tx(&wfull[buf], bytes);
            const uint8_t* src = lsrc + d2_stage_off(st);
            for (uint32_t o = 0; o < bytes; o += 16384u)
              bulk_g2s(sW + (size_t)buf * D2_WBUF + o, src + o, bytes - o < 16384u ? bytes - o : 16384u, &wfull[buf]);
          }
          __syncwarp();
          if (++buf == 2) { buf = 0; epar ^= 1; }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 8) tmem_dealloc2(tmem, 512);
}
