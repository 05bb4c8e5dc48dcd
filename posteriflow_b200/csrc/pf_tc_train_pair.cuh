// CTA-PAIR form of the backward chain kernel (tt_backward_kernel): one 128-row tile on a cluster of two CTAs through
// tcgen05.mma.cta_group::2 with M = 128, exactly as the pair form of the forward (pf_tc_pair.cuh): each CTA owns 64 rows
// (operand tiles, residual gradient, spline VJP) and half of every transposed weight chunk, four threads share a row,
// the leader's elected thread issues the pair's MMAs, the peer's warp 8 relays "my half of the chunk has landed".
// Differences from the one-CTA kernel that the halved operand tile pays for: the two parameter-gradient tiles (dims 1..5,
// 6..10) have their own buffers, so the spline VJP of all dims runs in one go and FA^T / FB^T are issued back to back; the
// gradient images go to global memory with the epilogue's own 16-byte stores (no publish / reopen hand-shake).
// Included by pf_tc_train.cu (inside namespace pf, after the one-CTA kernel: shares TtArgs and its helpers).
#pragma once

constexpr int BP_ROWS = 64;
constexpr int BP_THREADS = 320;
constexpr int BP_NSLOT = 11;
constexpr int BP_SLOT = 8192;           // one K = 16 chunk of 128 outputs: bf16 hi (4 KB) | lo (4 KB)
constexpr int BP_OFF_T0 = 0;            // [64 x 256] bf16 hi (32 KB) | lo (32 KB): parameter gradients A, then every hidden operand
constexpr int BP_OFF_T1 = 65536;        // parameter gradients B
constexpr int BP_OFF_RING = 131072;
constexpr int BP_OFF_GY = BP_OFF_RING + BP_NSLOT * BP_SLOT;   // [64][12] d loss / d layer output
constexpr int BP_OFF_GXD = BP_OFF_GY + BP_ROWS * 12 * 4;      // [64][12] direct d/dx of the spline
constexpr int BP_OFF_BAR = BP_OFF_GXD + BP_ROWS * 12 * 4;
constexpr int BP_SMEM = BP_OFF_BAR + 512;
__host__ __device__ inline int bp_layer_slots(const TtShape& s) { return s.NA / 16 + s.NBn / 16 + 65; }

// pair backward weight image: [L][2 ranks][slots][8 KB], slots in consumption order FA^T | FB^T | Wb2^T Wa2^T Wb1^T Wa1^T | W0^T.
// A slot = the rank's 128 input units n of one K = 16 chunk: bf16 hi [2 K-slabs][128][8] | lo; the W0^T slot = 16 chunks of
// ([8 x columns of this rank] x 16 hidden units: hi 256 B | lo 256 B).
__global__ void tt_pack_pair_kernel(TtShape s, uint8_t* __restrict__ image, const float* __restrict__ WfT,
                                    const float* __restrict__ WaT, const float* __restrict__ WbT, const float* __restrict__ W0T) {
  const int l = blockIdx.y, slot = blockIdx.x, rank = blockIdx.z;
  const int nA = s.NA / 16, nB = s.NBn / 16;
  __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(image + (((size_t)l * 2 + rank) * bp_layer_slots(s) + slot) * BP_SLOT);
  const int H = TT_H;
  if (slot < nA + nB + 64) {
    int kind, c;   // 0 FA, 1 FB, 2.. hidden
    if (slot < nA) { kind = 0; c = slot; }
    else if (slot < nA + nB) { kind = 1; c = slot - nA; }
    else { kind = 2 + (slot - nA - nB) / 16; c = (slot - nA - nB) % 16; }
    for (int e = threadIdx.x; e < 128 * 16; e += blockDim.x) {
      const int nl = e / 16, kk = e % 16, k = c * 16 + kk, n = rank * 128 + nl;
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
      const int o = (kk / 8) * (128 * 8) + (nl / 8) * 64 + (nl % 8) * 8 + (kk % 8);
      const __nv_bfloat16 hi = __float2bfloat16(v);
      dst[o] = hi;
      dst[128 * 16 + o] = __float2bfloat16(v - __bfloat162float(hi));
    }
  } else {
    for (int e = threadIdx.x; e < 16 * 8 * 16; e += blockDim.x) {
      const int c = e / 128, nl = (e / 16) % 8, kk = e % 16, n = rank * 8 + nl;
      const float v = n < s.D ? W0T[((size_t)l * s.DP + n) * H + c * 16 + kk] : 0.f;
      const int o = c * 256 + (kk / 8) * 64 + nl * 8 + (kk % 8);
      const __nv_bfloat16 hi = __float2bfloat16(v);
      dst[o] = hi;
      dst[o + 128] = __float2bfloat16(v - __bfloat162float(hi));
    }
  }
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(BP_THREADS, 1) tt_backward_pair_kernel(const TtArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  const TtShape& s = a.s;
  const int D = s.D;
  uint8_t* sT0 = smem + BP_OFF_T0;
  uint8_t* sT1 = smem + BP_OFF_T1;
  uint8_t* ring = smem + BP_OFF_RING;
  float* gy_s = reinterpret_cast<float*>(smem + BP_OFF_GY);
  float* gxd_s = reinterpret_cast<float*>(smem + BP_OFF_GXD);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + BP_OFF_BAR);
  uint64_t* full = bars;                       // [NSLOT] weight chunk landed: this CTA's half -- and, in the leader, the peer's
                                               // (second arrival, relayed by the peer's warp 8: one barrier to poll per chunk)
  uint64_t* empty = bars + BP_NSLOT;           // [NSLOT] the pair's MMAs reading the slot completed (multicast commit)
  uint64_t* acc_bar = bars + 3 * BP_NSLOT;     // MMAs of the current stage complete (multicast commit)
  uint64_t* op_bar = acc_bar + 1;              // leader only: operand tile(s) of the next stage written, 8 + 8 warp arrivals
  volatile int* abort_s = reinterpret_cast<volatile int*>(acc_bar + 2);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 3);

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const uint32_t rank = cluster_cta_rank();
  const int n_pairs = (int)(gridDim.x >> 1), pair = (int)(blockIdx.x >> 1);
  if (tid == 0) {
    for (int i = 0; i < BP_NSLOT; ++i) { mbar_init(&full[i], rank == 0 ? 2 : 1); mbar_init(&empty[i], 1); }
    mbar_init(acc_bar, 1);
    mbar_init(op_bar, 16);
    *abort_s = 0;
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc2(tmem_slot, 256);
  for (int i = tid; i < BP_OFF_RING / 16; i += BP_THREADS) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  auto wait = [&](uint64_t* bar, uint32_t parity, int code) -> bool {
    for (uint32_t i = 0; i < (1u << 22); ++i) {
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
    const int rr = (q4 & 1) * 32 + lane;
    const int rit = (int)rank * BP_ROWS + rr;
    const int rt = wh * 128 + q4 * 32 + lane;
    const int part = nh * 2 + wh;
    const uint32_t tl = tmem + ((uint32_t)(q4 * 32) << 16);
    const uint32_t rowoff = (uint32_t)(rr / 8) * 128 + (uint32_t)(rr % 8) * 16;      // operand tiles: 1024 B per 8 columns
    const uint32_t growoff = (uint32_t)(rit / 8) * 128 + (uint32_t)(rit % 8) * 16;   // images: 2048 B per 8 columns
    const int gcol0 = 128 * nh + 64 * wh;
    const int tcol0 = 64 * wh;
    uint32_t acc_par = 0;
    bool ok = true;
    // cycle profile of CTA 0 (PF_TC_PROF=1): 0 spline VJP | 2 hidden waits | 3 hidden epilogues | 4 wait + d x | 6 total
    const bool prof = a.prof && blockIdx.x == 0 && tid == 0;
    long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long tmark = clock64();
    const long long t_begin = tmark;
    auto mark = [&](int k) { if (prof) { const long long n = clock64(); pc[k] += n - tmark; tmark = n; } };
    auto arrive_ops = [&]() {
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { if (rank == 0) mbar_arrive(op_bar); else mbar_arrive_remote(op_bar, 0); }
    };
    // 32 columns of this thread's row -> bf16 hi / lo tile images in global memory
    auto store_image = [&](uint8_t* ghi, uint8_t* glo, int c0, const float* vals) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        uint4 w4, l4;
        split_bf16(vals[8 * i], vals[8 * i + 1], w4.x, l4.x); split_bf16(vals[8 * i + 2], vals[8 * i + 3], w4.y, l4.y);
        split_bf16(vals[8 * i + 4], vals[8 * i + 5], w4.z, l4.z); split_bf16(vals[8 * i + 6], vals[8 * i + 7], w4.w, l4.w);
        const size_t io = (size_t)(c0 / 8 + i) * 2048 + growoff;
        *reinterpret_cast<uint4*>(ghi + io) = w4;
        *reinterpret_cast<uint4*>(glo + io) = l4;
      }
    };
    for (int tile = pair; tile < a.n_tiles && ok; tile += n_pairs) {
      const int64_t grow = (int64_t)tile * TT_ROWS + rit;
      const bool valid = grow < a.R;
      const float glr = (valid && a.gl) ? a.gl[grow] : 0.f;
      if (part == 0)
        for (int j = 0; j < D; ++j) gy_s[rr * 12 + j] = (valid && a.gz) ? a.gz[grow * D + j] : 0.f;
      bar_sync(3, 256);
      for (int l = s.L - 1; l >= 0 && ok; --l) {
        const float* projr = a.proj + ((size_t)tile * a.ldproj + (size_t)l * 3 * TT_H) * TT_ROWS + rit;
        if (rt == 0 && rank == 0) {   // L2 prefetch of what the tile reads next (as the one-CTA kernel)
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
        // ---------------- E0: spline VJP of dims part, part + 4, part + 8 -> the two parameter-gradient tiles ----------------
        uint8_t* gpa = a.gimg + tt_img_off(l, TT_G_PA, TT_NG, tile, a.n_tiles);
        uint8_t* gpb = a.gimg + tt_img_off(l, TT_G_PB, TT_NG, tile, a.n_tiles);
#pragma unroll 1
        for (int j = part; j < D; j += 4) {
          float q[TT_MULT], gp[TT_MULT];
          const float* pr = a.P + (((size_t)l * a.n_tiles + tile) * s.PLD + j * TT_MULT) * TT_ROWS + rit;
#pragma unroll
          for (int i = 0; i < TT_MULT; ++i) q[i] = __ldg(pr + (size_t)i * TT_ROWS);
          const float xin = a.XS[(((size_t)l * a.n_tiles + tile) * D + (D - 1 - j)) * TT_ROWS + rit];
          const float gyj = gy_s[rr * 12 + j];
          const float gxd = rqs_forward_backward_fast<TT_NB>(TtRegQ{q}, xin, a.tail, gyj, glr, TtGradOut{gp});
          gxd_s[rr * 12 + j] = valid ? gxd : 0.f;
          if (j == 0) {
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
          const bool ga_dim = j <= s.ga;
          uint8_t* tile_hi = ga_dim ? sT0 : sT1;
          uint8_t* gimg = ga_dim ? gpa : gpb;
          const int cg0 = (ga_dim ? j - 1 : j - 1 - s.ga) * (TT_PCOL / 8);
#pragma unroll
          for (int c8 = 0; c8 < TT_PCOL / 8; ++c8) {
            float e[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) e[i] = (valid && 8 * c8 + i < TT_MULT) ? gp[8 * c8 + i < TT_MULT ? 8 * c8 + i : 0] : 0.f;
            uint4 w4, l4;
            split_bf16(e[0], e[1], w4.x, l4.x); split_bf16(e[2], e[3], w4.y, l4.y);
            split_bf16(e[4], e[5], w4.z, l4.z); split_bf16(e[6], e[7], w4.w, l4.w);
            const size_t io = (size_t)(cg0 + c8) * 1024 + rowoff;
            *reinterpret_cast<uint4*>(tile_hi + io) = w4;
            *reinterpret_cast<uint4*>(tile_hi + 32768 + io) = l4;
            *reinterpret_cast<uint4*>(gimg + (size_t)(cg0 + c8) * 2048 + growoff) = w4;   // weight-gradient operand (hi)
          }
        }
        if (part == 3) {   // columns 240 .. 255 of both parameter-gradient images are padding: the weight-gradient kernel reads them
          const uint4 z4 = make_uint4(0, 0, 0, 0);
#pragma unroll
          for (int cg = 30; cg < 32; ++cg) {
            *reinterpret_cast<uint4*>(gpa + (size_t)cg * 2048 + growoff) = z4;
            *reinterpret_cast<uint4*>(gpb + (size_t)cg * 2048 + growoff) = z4;
          }
        }
        arrive_ops();
        mark(0);
        // ---------------- E2 .. E6: the five hidden data-gradient epilogues (as the one-CTA kernel) ----------------
        // k = 0: d h2 complete -> g2 d h2 (operand), gate-2 gradient, residual := d h2
        // k = 1: [u2 > 0] acc  (operand)
        // k = 2: d h1 = residual + [h1 > 0] acc -> g1 d h1 (operand), gate-1 gradient, residual := d h1
        // k = 3: [u1 > 0] acc  (operand)
        // k = 4: d h0 = residual + [h0 > 0] acc (operand), relu(context) gradient
#pragma unroll 1
        for (int k = 0; k < 5 && ok; ++k) {
          const bool res_in = (k == 2 || k == 4), gate = (k == 0 || k == 2), has_mask = k >= 1;
          const int mimg = k == 1 ? TT_A_U2 : (k == 2 ? TT_A_H1 : (k == 3 ? TT_A_U1 : TT_A_H0));
          const uint8_t* mbase = a.img + tt_img_off(l, mimg, TT_NA, tile, a.n_tiles);
          const int vimg = k == 0 ? TT_A_V2 : TT_A_V1;
          const uint8_t* vbase = a.img + tt_img_off(l, vimg, TT_NA, tile, a.n_tiles);
          const uint8_t* vlbase = a.img + tt_img_off(l, vimg + 1, TT_NA, tile, a.n_tiles);
          const int sec = k == 0 ? 2 : (k == 2 ? 1 : 0);
          uint8_t* gout = a.gimg + tt_img_off(l, TT_G_V2 + k, TT_NG, tile, a.n_tiles);
          uint32_t mkw[16], vhw[16], vlw[16];      // 32 bf16 each: relu mask source, pre-gate value hi / lo
          float g[32];
          // what a 32-column block reads from the forward's saves: independent of the accumulator, so the FIRST block's
          // loads are issued before the wait on the tensor pipe (their L2 round trip runs under the MMAs)
          auto load_saved = [&](int c0) {
            if (has_mask) {
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint4 t4 = ldg128(mbase + (size_t)(c0 / 8 + i) * 2048 + growoff);
                mkw[4 * i] = t4.x; mkw[4 * i + 1] = t4.y; mkw[4 * i + 2] = t4.z; mkw[4 * i + 3] = t4.w;
              }
            }
            if (gate) {
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint4 h4 = ldg128(vbase + (size_t)(c0 / 8 + i) * 2048 + growoff);
                const uint4 l4 = ldg128(vlbase + (size_t)(c0 / 8 + i) * 2048 + growoff);
                vhw[4 * i] = h4.x; vhw[4 * i + 1] = h4.y; vhw[4 * i + 2] = h4.z; vhw[4 * i + 3] = h4.w;
                vlw[4 * i] = l4.x; vlw[4 * i + 1] = l4.y; vlw[4 * i + 2] = l4.z; vlw[4 * i + 3] = l4.w;
              }
            }
            if (gate || k == 4) {
#pragma unroll
              for (int i = 0; i < 32; ++i) g[i] = __ldg(projr + (size_t)(sec * TT_H + c0 + i) * TT_ROWS);
            }
          };
          load_saved(gcol0);
          if (!(ok = wait(acc_bar, acc_par, 52 + k))) break;
          acc_par ^= 1;
          tc_fence_after();
          mark(2);
#pragma unroll 1
          for (int b = 0; b < 2; ++b) {
            const int c0 = gcol0 + 32 * b;
            const uint32_t tc = (uint32_t)(tcol0 + 32 * b);
            uint32_t v[32], hr[32];
            tmem_ld32(tl + tc, v);
            if (res_in) tmem_ld32(tl + 128 + tc, hr);
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
              tmem_st16(tl + 128 + tc, ow);
              tmem_st16(tl + 128 + tc + 16, ow + 16);
            }
            if (gate) {                      // d gate (pre-sigmoid) = d h * v * g (1 - g); operand = g * d h
              float gp4[32];
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                const uint32_t wh2 = vhw[i / 2], wl2 = vlw[i / 2];
                const float vv = (i & 1) ? __uint_as_float(wh2 & 0xffff0000u) + __uint_as_float(wl2 & 0xffff0000u)
                                         : __uint_as_float(wh2 << 16) + __uint_as_float(wl2 << 16);
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
            if (b == 0) load_saved(c0 + 32);   // the second block's saves: in flight while this block's operand words are stored
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              uint4 w4, l4;
              split_bf16(o[8 * i], o[8 * i + 1], w4.x, l4.x); split_bf16(o[8 * i + 2], o[8 * i + 3], w4.y, l4.y);
              split_bf16(o[8 * i + 4], o[8 * i + 5], w4.z, l4.z); split_bf16(o[8 * i + 6], o[8 * i + 7], w4.w, l4.w);
              const size_t io = (size_t)(c0 / 8 + i) * 1024 + rowoff;
              *reinterpret_cast<uint4*>(sT0 + io) = w4;
              *reinterpret_cast<uint4*>(sT0 + 32768 + io) = l4;
              *reinterpret_cast<uint4*>(gout + (size_t)(c0 / 8 + i) * 2048 + growoff) = w4;   // weight-gradient operand (hi)
            }
          }
          if (k == 0 || k == 2) tmem_wait_st();
          arrive_ops();
          mark(3);
        }
        if (!ok) break;
        // ---------------- E7: d x of this layer -> d output of the layer below ----------------
        if (!(ok = wait(acc_bar, acc_par, 58))) break;
        acc_par ^= 1;
        tc_fence_after();
        if (wh == 0) {      // lanes 0..63 of the accumulator hold x columns 0..7, lanes 64..127 columns 8..15
          uint32_t v[8];
          tmem_ld8(tl, v);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int k = 8 * nh + i;
            if (k < D) {
              const float gxin = __uint_as_float(v[i]) + gxd_s[rr * 12 + k];
              gy_s[rr * 12 + (D - 1 - k)] = gxin;          // undo the layer's ReversePermutation
              if (l == 0 && a.gx && valid) a.gx[grow * D + (D - 1 - k)] = gxin;
            }
          }
        }
        tc_fence_before();
        bar_sync(3, 256);
        mark(4);
      }
    }
    if (prof) {
      for (int i = 0; i < 6; ++i) a.prof[i] = pc[i];
      a.prof[6] = clock64() - t_begin;
    }
  } else if (warp == 8) {
    // ============================ MMA issuer (leader) / weight relay (peer) ============================
    bool ok = true;
    if (rank != 0) {
      uint32_t slot = 0, full_par = 0;
      const int per_tile = s.L * bp_layer_slots(s);
      for (int tile = pair; tile < a.n_tiles && ok; tile += n_pairs)
        for (int i = 0; i < per_tile && ok; ++i) {
          if (!(ok = wait(&full[slot], full_par, 31))) break;
          if (elect_one_sync()) mbar_arrive_remote(&full[slot], 0);   // this CTA's half of the chunk has landed
          __syncwarp();
          if (++slot == BP_NSLOT) { slot = 0; full_par ^= 1; }
        }
    } else {
      uint32_t op_par = 0, slot = 0;
      const uint32_t idesc256 = umma_idesc_bf16(TT_ROWS, 256), idesc16 = umma_idesc_bf16(TT_ROWS, 16);
      const uint64_t ad0 = umma_smem_desc(smem_u32(sT0), 1024, 128), ad1 = umma_smem_desc(smem_u32(sT1), 1024, 128);
      const uint64_t bd128 = umma_smem_desc(smem_u32(ring), 128 * 16, 128), bd8 = umma_smem_desc(smem_u32(ring), 128, 128);
      // weights confirmed ahead of the issue pointer while the warp waits for operands (see pf_tc_pair.cuh)
      uint32_t cslot = 0, cpar = 0;
      int ready = 0;
      auto confirm = [&]() {
        while (ready < BP_NSLOT) {
          if (!mbar_test_wait(&full[cslot], cpar)) break;
          ++ready;
          if (++cslot == BP_NSLOT) { cslot = 0; cpar ^= 1; }
        }
      };
      for (int tile = pair; tile < a.n_tiles && ok; tile += n_pairs) {
        for (int l = s.L - 1; l >= 0 && ok; --l) {
          for (int m = 0; m < 7 && ok; ++m) {          // FA^T, FB^T, Wb2^T, Wa2^T, Wb1^T, Wa1^T, W0^T
            const int nk = m == 0 ? s.NA / 16 : (m == 1 ? s.NBn / 16 : (m == 6 ? 1 : 16));
            const uint64_t ad = m == 1 ? ad1 : ad0;
            if (m != 1) {                               // FB^T follows FA^T at once: both tiles were written by the same stage
              bool got = false;
              for (uint32_t i = 0; i < (1u << 26); ++i) {
                if (mbar_test_wait(op_bar, op_par)) { got = true; break; }
                confirm();
                if ((i & 1023) == 1023 && *abort_s) break;
              }
              if (!got) { if (!*abort_s) { *abort_s = 1; raise_flag(a.err, 20 + m); } ok = false; break; }
              op_par ^= 1;
            }
            for (int c = 0; c < nk && ok; ++c) {
              if (ready == 0) {
                if (!(ok = wait(&full[cslot], cpar, 31))) break;
                ++ready;
                if (++cslot == BP_NSLOT) { cslot = 0; cpar ^= 1; }
              }
              --ready;
              tc_fence_after();
              if (elect_one_sync()) {
                const uint64_t soff = (uint64_t)((slot * BP_SLOT) >> 4);
                if (m < 6) {
                  const uint64_t bh = bd128 + soff, bl = bh + (4096 >> 4);
                  const uint64_t ah = ad + (uint64_t)(c * (2048 >> 4)), al = ah + (32768 >> 4);
                  umma_ss2(tmem, ah, bh, idesc256, (m == 1 || c > 0) ? 1u : 0u);
                  umma_ss2(tmem, al, bh, idesc256, 1u);
                  umma_ss2(tmem, ah, bl, idesc256, 1u);
                  umma_commit2(&empty[slot], 3);
                  if (m != 0 && c + 1 >= nk) umma_commit2(acc_bar, 3);
                } else {
#pragma unroll 4
                  for (int cc = 0; cc < 16; ++cc) {
                    const uint64_t bh = bd8 + soff + (uint64_t)(cc * (512 >> 4)), bl = bh + (256 >> 4);
                    const uint64_t ah = ad0 + (uint64_t)(cc * (2048 >> 4)), al = ah + (32768 >> 4);
                    umma_ss2(tmem, ah, bh, idesc16, cc > 0 ? 1u : 0u);
                    umma_ss2(tmem, al, bh, idesc16, 1u);
                    umma_ss2(tmem, ah, bl, idesc16, 1u);
                  }
                  umma_commit2(&empty[slot], 3);
                  umma_commit2(acc_bar, 3);
                }
              }
              __syncwarp();
              if (++slot == BP_NSLOT) slot = 0;
            }
          }
        }
      }
    }
  } else if (warp == 9) {
    // ============================ weight loader (each CTA: its half of every chunk) ============================
    uint32_t slot = 0, empty_par = 1;
    bool ok = true;
    const int nslots = bp_layer_slots(s);
    for (int tile = pair; tile < a.n_tiles && ok; tile += n_pairs) {
      for (int l = s.L - 1; l >= 0 && ok; --l) {
        const uint8_t* src = a.wimg + ((size_t)l * 2 + rank) * nslots * BP_SLOT;
        for (int i = 0; i < nslots; ++i) {
          if (!(ok = wait(&empty[slot], empty_par, 40))) break;
          if (elect_one_sync()) {
            mbar_arrive_expect_tx(&full[slot], BP_SLOT);
            bulk_g2s(ring + (size_t)slot * BP_SLOT, src, BP_SLOT, &full[slot]);
          }
          __syncwarp();
          src += BP_SLOT;
          if (++slot == BP_NSLOT) { slot = 0; empty_par ^= 1; }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 8) tmem_dealloc2(tmem, 256);
}
