// Tensor-core GEMM building blocks of the TRAINING path under PF_TIER_BF16 (forward with saved activations +
// backward, pf_fp32.cu): the same functor interfaces as sgemm_nt / sgemm_tn (pf_gemm.cuh), but the products run
// on tcgen05.mma with fp32 accumulation in tensor memory.
//
// Accuracy: operands arrive as fp32 and are split on the fly into TWO bf16 terms, a = a_hi + a_lo
// (a_lo = bf16(a - a_hi): 16 significand bits survive), and every K=16 chunk issues three MMAs
//     acc += a_hi b_hi + a_hi b_lo + a_lo b_hi            (the a_lo b_lo term, ~2^-18 |ab|, is dropped)
// so a product keeps ~1e-5 relative accuracy against 2^-9 for a plain bf16 one.  Measured on the training step
// (tests/test_gpu_parity.py, tier="bf16"): median error of the gradients 5-7x the float32 reference's own distance from
// float64; gradients of x / context within 2e-3 of their scale, weight-gradient norms within 1e-3.
//
// Shape of one CTA (288 threads; 80-96 KB of shared memory and two CTAs per SM with 32- / 64-column tiles, 128 KB and
// one per SM with 128-column tiles): a 128-row x BN-column output tile, the
// accumulator is BN columns of tensor memory (lane = output row).  The contraction advances through a 4-stage ring
// of 32-element stages: eight producer warps convert operands into the UMMA canonical K-major no-swizzle layout
// (8-row x 16-byte core matrices; element (row, k) at (k/8) * ROWS*16 + row*16 + (k%8)*2, i.e. LBO = ROWS*16 along K,
// SBO = 128 along rows -- the layout verified by tests/umma_probe.cu; one 16-byte store per (row, 8 k), bank-conflict
// free), a ninth, converged warp elects one lane to issue each stage's six MMAs; the epilogue functor runs on values
// read back with tcgen05.ld and transposed through shared memory so its global accesses are coalesced.
#pragma once
#include <type_traits>
#include <utility>

#include "pf_common.cuh"
#include "pf_gemm.cuh"
#include "pf_sm100.cuh"

namespace pf {

constexpr int TG_ROWS = 128;       // output rows per CTA = UMMA M = TMEM lanes
constexpr int TG_KSTG = 32;        // contraction elements per ring stage (two K=16 MMA chunks)
constexpr int TG_STAGES = 4;       // ring depth
constexpr int TG_PRODUCERS = 256;  // warps 0-7: operand producers during the main loop, epilogue afterwards
constexpr int TG_THREADS = TG_PRODUCERS + 32;   // + warp 8: MMA issuer, owns the TMEM allocation
constexpr int TG_EU = 8;           // epilogue elements per thread whose global reads are in flight together
                                   // (16 for the 32-column tiles -- their whole share in one batch -- measured slower:
                                   //  2.61 vs 2.28 ms per 948-row training step)
// An epilogue functor may ask for the warp's lanes to run along the output ROW index (kLanesAlongRows = true): it
// then writes a row-minor ("tile-transposed") layout with coalesced stores.  Default: lanes along the column index.
template <class E, class = void>
struct tg_lanes_along_rows : std::false_type {};
template <class E>
struct tg_lanes_along_rows<E, std::enable_if_t<E::kLanesAlongRows>> : std::true_type {};
template <class E, class = void>
struct tg_has_fetch : std::false_type {};
template <class E>
struct tg_has_fetch<E, std::void_t<decltype(std::declval<const E&>().fetch(0, 0))>> : std::true_type {};
template <int BN>
__host__ __device__ constexpr int tg_stage_bytes() { return 2 * (TG_KSTG / 8) * (TG_ROWS + BN) * 16; }   // hi + lo of A and B
template <int BN>
__host__ __device__ constexpr int tg_smem_bytes() { return TG_STAGES * tg_stage_bytes<BN>() + 128; }

__device__ __forceinline__ void tg_split_store(uint8_t* hi_dst, uint8_t* lo_dst, const float* v) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __nv_bfloat16 h0 = __float2bfloat16(v[2 * i]), h1 = __float2bfloat16(v[2 * i + 1]);
    const float r0 = v[2 * i] - __bfloat162float(h0), r1 = v[2 * i + 1] - __bfloat162float(h1);
    h[i] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
    l[i] = sm100::pack_bf16(r0, r1);
  }
  *reinterpret_cast<uint4*>(hi_dst) = make_uint4(h[0], h[1], h[2], h[3]);
  *reinterpret_cast<uint4*>(lo_dst) = make_uint4(l[0], l[1], l[2], l[3]);
}

// ---- operand sources: 8 consecutive contraction elements of one output row / column ------------------------------
// NT (forward products, data gradients): both operands are K-contiguous fp32 matrices
template <int AMODE>
struct SrcNT {
  const float* A; int lda; const float* B; int ldb; bool a_vec, b_vec;
  __device__ __forceinline__ void load8(const float* p, bool vec, int k, int kend, float* v) const {
    if (vec && k + 8 <= kend) {
      const float4 x0 = __ldg(reinterpret_cast<const float4*>(p)), x1 = __ldg(reinterpret_cast<const float4*>(p + 4));
      v[0] = x0.x; v[1] = x0.y; v[2] = x0.z; v[3] = x0.w; v[4] = x1.x; v[5] = x1.y; v[6] = x1.z; v[7] = x1.w;
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = (k + i < kend) ? __ldg(p + i) : 0.f;
    }
  }
  __device__ __forceinline__ void loadA(int row, int k, int kend, float* v) const {
    load8(A + (int64_t)row * lda + k, a_vec, k, kend, v);
#pragma unroll
    for (int i = 0; i < 8; ++i)
      v[i] = AMODE == A_RELU ? fmaxf(v[i], 0.f) : (AMODE == A_SANITIZE ? sanitize_ctx(v[i]) : v[i]);
  }
  __device__ __forceinline__ void loadB(int row, int k, int kend, float* v) const {
    load8(B + (int64_t)row * ldb + k, b_vec, k, kend, v);
  }
};
// TN (weight gradients): the contraction index is the ROW index of both operands, so the source transposes: a thread
// owns one output row / column n and gathers 8 consecutive m (each of the 8 loads is coalesced across the warp's n)
template <class LoadA, class LoadB>
struct SrcTN {
  LoadA la; LoadB lb;
  __device__ __forceinline__ void loadA(int row, int k, int kend, float* v) const {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = (k + i < kend) ? la(k + i, row) : 0.f;
  }
  __device__ __forceinline__ void loadB(int row, int k, int kend, float* v) const {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = (k + i < kend) ? lb(k + i, row) : 0.f;
  }
};

// C[MA, NBc] (+)= epi( sum_k A(row, k) B(col, k) ) over k in this CTA's slice [blockIdx.z * k_per_slice, ...).
// Warp-specialised: warps 0-7 produce ring stages (global fp32 -> split bf16 -> UMMA layout in shared memory; the
// loads of stage s+1 are in flight in registers while stage s is converted and stored), warp 8 waits for a full stage,
// issues its six MMAs and commits them to the stage's `empty` barrier; when the last stage's MMAs have completed the
// producers turn into the epilogue.
// PD = ring stages whose global loads a producer thread keeps in flight in registers ahead of the one it converts.
// Every launch uses PD = 1: PD = 2 (one CTA per SM, twice the bytes in flight) was measured on the 64-CTA launches of a
// ~1 k-row call and changed nothing (2.357 vs 2.367 ms per training step) -- those kernels are bound by their fixed
// prologue / first-load / epilogue latencies and the launch gap, not by the main loop.
template <class Src, class Epi, int BN, int PD>
__global__ void __launch_bounds__(TG_THREADS, (BN >= 128 || PD > 1) ? 1 : 2)
tc_gemm_kernel(const Src src, int MA, int NBc, int K, int k_per_slice, Epi epi, int* __restrict__ err) {
  using namespace sm100;
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr int STG = tg_stage_bytes<BN>();
  constexpr int KG = TG_KSTG / 8;                       // 8-wide k groups per stage
  constexpr int A_HALF = KG * TG_ROWS * 16, B_HALF = KG * BN * 16;   // bytes of one (hi or lo) operand image per stage
  constexpr int ITEMS = (TG_ROWS + BN) * KG;
  constexpr int IT = (ITEMS + TG_PRODUCERS - 1) / TG_PRODUCERS;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + TG_STAGES * STG);
  uint64_t* empty = full + TG_STAGES;
  uint64_t* acc_bar = empty + TG_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 1);
  const int tid = threadIdx.x;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int m0 = blockIdx.y * TG_ROWS, n0 = blockIdx.x * BN;
  const int kbeg = blockIdx.z * k_per_slice;
  const int kend = min(K, kbeg + k_per_slice);
  const int nst = kend > kbeg ? (kend - kbeg + TG_KSTG - 1) / TG_KSTG : 0;
  if (tid == 0) {
    for (int i = 0; i < TG_STAGES; ++i) { mbar_init(&full[i], TG_PRODUCERS); mbar_init(&empty[i], 1); }
    mbar_init(acc_bar, 1);
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc(tmem_slot, BN < 32 ? 32 : BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  bool ok = true;
  // Programmatic dependent launch: this grid may have been started while its predecessor in the stream was still
  // running (its barrier setup and TMEM allocation above overlap the predecessor's tail and the launch latency);
  // nothing the predecessor wrote is read before this point.  No-op for a normally launched grid.
  asm volatile("griddepcontrol.wait;" ::: "memory");

  if (warp < 8) {
    // ============================ producers ============================
    constexpr int NBUF = PD + 1;
    float buf[NBUF][IT][8];
    auto fetch = [&](int s, float (*dst)[8]) {
      const int k0 = kbeg + s * TG_KSTG;
#pragma unroll
      for (int u = 0; u < IT; ++u) {
        const int idx = tid + u * TG_PRODUCERS;
#pragma unroll
        for (int i = 0; i < 8; ++i) dst[u][i] = 0.f;
        if (idx < ITEMS) {
          const bool isA = idx < TG_ROWS * KG;
          const int j = isA ? idx : idx - TG_ROWS * KG;
          const int r = isA ? (j & (TG_ROWS - 1)) : (j % BN), kg = isA ? (j / TG_ROWS) : (j / BN);
          const int row = (isA ? m0 : n0) + r, k = k0 + kg * 8;
          if (row < (isA ? MA : NBc) && k < kend) {
            if (isA) src.loadA(row, k, kend, dst[u]);
            else src.loadB(row, k, kend, dst[u]);
          }
        }
      }
    };
#pragma unroll
    for (int p = 0; p < PD; ++p)
      if (p < nst) fetch(p, buf[p]);
    for (int s0 = 0; s0 < nst && ok; s0 += NBUF) {
#pragma unroll
      for (int jb = 0; jb < NBUF; ++jb) {      // unrolled: the register buffers are indexed statically
        const int s = s0 + jb;
        if (s < nst && ok) {
          if (s + PD < nst) fetch(s + PD, buf[(jb + PD) % NBUF]);
          const int slot = s % TG_STAGES;
          ok = mbar_wait(&empty[slot], ((s / TG_STAGES) & 1) ^ 1);   // a fresh barrier passes a wait on parity 1
          uint8_t* sb = smem + slot * STG;
#pragma unroll
          for (int u = 0; u < IT; ++u) {
            const int idx = tid + u * TG_PRODUCERS;
            if (idx < ITEMS) {
              const bool isA = idx < TG_ROWS * KG;
              const int j = isA ? idx : idx - TG_ROWS * KG;
              const int r = isA ? (j & (TG_ROWS - 1)) : (j % BN), kg = isA ? (j / TG_ROWS) : (j / BN);
              uint8_t* hi = sb + (isA ? 0 : 2 * A_HALF) + (uint32_t)kg * ((isA ? TG_ROWS : BN) * 16) + (uint32_t)r * 16;
              tg_split_store(hi, hi + (isA ? A_HALF : B_HALF), buf[jb][u]);
            }
          }
          fence_proxy_async_smem();
          mbar_arrive(&full[slot]);
        }
      }
    }
  } else {
    // ============================ MMA issuer (converged warp, one elected lane) ============================
    const uint32_t idesc = umma_idesc_bf16(TG_ROWS, BN);
    const uint32_t base = smem_u32(smem);
    for (int s = 0; s < nst && ok; ++s) {
      const int slot = s % TG_STAGES;
      ok = mbar_wait(&full[slot], (s / TG_STAGES) & 1);
      tc_fence_after();
      if (ok && elect_one_sync()) {
        const uint32_t sa = base + slot * STG;
        uint64_t ah = umma_smem_desc(sa, TG_ROWS * 16, 128), al = umma_smem_desc(sa + A_HALF, TG_ROWS * 16, 128);
        uint64_t bh = umma_smem_desc(sa + 2 * A_HALF, BN * 16, 128), bl = umma_smem_desc(sa + 2 * A_HALF + B_HALF, BN * 16, 128);
        const uint64_t astep = (uint64_t)((2u * TG_ROWS * 16u) >> 4), bstep = (uint64_t)((2u * (uint32_t)BN * 16u) >> 4);
#pragma unroll
        for (int c = 0; c < TG_KSTG / 16; ++c) {
          umma_ss(tmem, ah, bh, idesc, (s > 0 || c > 0) ? 1u : 0u);
          umma_ss(tmem, ah, bl, idesc, 1u);
          umma_ss(tmem, al, bh, idesc, 1u);
          ah += astep; al += astep; bh += bstep; bl += bstep;
        }
        umma_commit(&empty[slot]);
        if (s == nst - 1) umma_commit(acc_bar);
      }
      __syncwarp();
    }
  }
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // the successor may set itself up during our epilogue
  if (nst > 0 && ok) ok = mbar_wait(acc_bar, 0);
  tc_fence_after();
  if (!ok && (tid & 31) == 0) raise_flag(err, 60 + (warp == 8));
  ok = __syncthreads_and(ok);
  if (ok && nst > 0 && warp < 8) {
    // ============================ epilogue ============================
    // TMEM -> registers (thread = output row, the only shape tcgen05.ld offers) -> shared memory (the ring is free: every
    // MMA has completed) -> back out with the warp's lanes along the COLUMN index, so the functor's global loads /
    // stores / atomics of one instruction fall into one or two 128-byte lines instead of 32 rows' worth of sectors.
    constexpr int LD = BN + 1, HALF = BN / 2;
    float* stage = reinterpret_cast<float*>(smem);
    const int row = (warp & 3) * 32 + (tid & 31);
    const uint32_t tl = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    const int cbeg = (warp >> 2) * HALF;     // warps 0-3 take columns [0, HALF), warps 4-7 [HALF, BN)
#pragma unroll
    for (int c = 0; c < HALF; c += 16) {
      uint32_t v[16];
      tmem_ld16(tl + cbeg + c, v);
      tmem_wait_ld();
#pragma unroll
      for (int i = 0; i < 16; ++i) stage[row * LD + cbeg + c + i] = __uint_as_float(v[i]);
    }
    bar_sync(1, TG_PRODUCERS);
    if constexpr (tg_has_fetch<Epi>::value) {
      // the functor's global reads of TG_EU elements are issued together, then applied (one memory latency per
      // batch instead of one per element: with a single resident CTA there are few other warps to hide it)
      using Fetched = decltype(epi.fetch(0, 0));
      for (int base = tid; base < TG_ROWS * BN; base += TG_PRODUCERS * TG_EU) {
        Fetched t[TG_EU];
#pragma unroll
        for (int u = 0; u < TG_EU; ++u) {
          const int idx = base + u * TG_PRODUCERS, r = idx / BN, c = idx % BN;
          if (idx < TG_ROWS * BN && m0 + r < MA && n0 + c < NBc) t[u] = epi.fetch(m0 + r, n0 + c);
        }
#pragma unroll
        for (int u = 0; u < TG_EU; ++u) {
          const int idx = base + u * TG_PRODUCERS, r = idx / BN, c = idx % BN;
          if (idx < TG_ROWS * BN && m0 + r < MA && n0 + c < NBc) epi.apply(m0 + r, n0 + c, stage[r * LD + c], t[u]);
        }
      }
    } else if constexpr (tg_lanes_along_rows<Epi>::value) {
#pragma unroll 4
      for (int idx = tid; idx < TG_ROWS * BN; idx += TG_PRODUCERS) {
        const int r = idx % TG_ROWS, c = idx / TG_ROWS;      // LD = BN + 1 is odd: conflict-free column reads
        if (m0 + r < MA && n0 + c < NBc) epi(m0 + r, n0 + c, stage[r * LD + c]);
      }
    } else {
#pragma unroll 4
      for (int idx = tid; idx < TG_ROWS * BN; idx += TG_PRODUCERS) {
        const int r = idx / BN, c = idx % BN;
        if (m0 + r < MA && n0 + c < NBc) epi(m0 + r, n0 + c, stage[r * LD + c]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc(tmem, BN < 32 ? 32 : BN);
}

// ---- host launchers ---------------------------------------------------------------------------------
template <auto Kernel>   // one flag array per kernel INSTANTIATION (a function-pointer type would be shared)
inline pf_status tg_set_smem(int bytes, int dev) {
  static bool done[64] = {};
  if (dev >= 0 && dev < 64 && done[dev]) return PF_OK;
  PF_CUDA(cudaFuncSetAttribute(Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  if (dev >= 0 && dev < 64) done[dev] = true;
  return PF_OK;
}

template <class Src, class Epi, int BN, int PD = 1>
inline pf_status tg_launch(pf_flow* f, cudaStream_t st, const Src& src, int MA, int NBc, int K, int kps, int slices,
                           const Epi& epi) {
  PF_TRY((tg_set_smem<tc_gemm_kernel<Src, Epi, BN, PD>>(tg_smem_bytes<BN>(), f->desc.device)));
  dim3 grid((NBc + BN - 1) / BN, (MA + TG_ROWS - 1) / TG_ROWS, slices);
  if (f->train_pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(TG_THREADS);
    cfg.dynamicSmemBytes = tg_smem_bytes<BN>();
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    PF_CUDA(cudaLaunchKernelEx(&cfg, tc_gemm_kernel<Src, Epi, BN, PD>, src, MA, NBc, K, kps, epi, f->tg_err));
    f->launches++;
    return PF_OK;
  }
  tc_gemm_kernel<Src, Epi, BN, PD><<<grid, TG_THREADS, tg_smem_bytes<BN>(), st>>>(src, MA, NBc, K, kps, epi, f->tg_err);
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}

template <class Epi, int AMODE>
inline pf_status tc_gemm_nt(pf_flow* f, cudaStream_t st, const float* A, int lda, const float* B, int ldb, int M, int N,
                            int K, const Epi& epi) {
  if (M <= 0 || N <= 0) return PF_OK;
  const SrcNT<AMODE> src{A, lda, B, ldb, (lda & 3) == 0 && (reinterpret_cast<uintptr_t>(A) & 15) == 0,
                         (ldb & 3) == 0 && (reinterpret_cast<uintptr_t>(B) & 15) == 0};
  const int mt = (M + TG_ROWS - 1) / TG_ROWS;
  // Tile width: every column tile re-reads (and re-converts) the A rows of its row tile, so wide tiles as soon as
  // they still give most SMs a CTA (7 515 rows x 256 columns: 118 CTAs of 128 x 128 read A twice, 472 CTAs of
  // 128 x 32 eight times = 61 MB of L2 traffic per product); narrow tiles for the ~1 k-row calls of a sharded batch.
  if (10 * mt * ((N + 127) / 128) >= 7 * f->sm_count) return tg_launch<SrcNT<AMODE>, Epi, 128>(f, st, src, M, N, K, K, 1, epi);
  if (mt * ((N + 63) / 64) >= 2 * f->sm_count) return tg_launch<SrcNT<AMODE>, Epi, 64>(f, st, src, M, N, K, K, 1, epi);
  return tg_launch<SrcNT<AMODE>, Epi, 32>(f, st, src, M, N, K, K, 1, epi);
}

template <class Epi, class LoadA, class LoadB>
inline pf_status tc_gemm_tn(pf_flow* f, cudaStream_t st, const LoadA& la, const LoadB& lb, int M, int N1, int N2,
                            const Epi& epi) {
  if (M <= 0 || N1 <= 0 || N2 <= 0) return PF_OK;
  const SrcTN<LoadA, LoadB> src{la, lb};
  // rows per slice: every slice adds a full tile of atomics, and these kernels share the GPU with the dependency
  // chain on the main stream: about one CTA per SM, never under 128 rows per slice.  128-column tiles for the big
  // batches (fewer re-reads of the gradient rows), 64-column ones otherwise.
  auto plan = [&](int BN, int& mps, int& slices) {
    const int tiles = ((N1 + TG_ROWS - 1) / TG_ROWS) * ((N2 + BN - 1) / BN);
    slices = (f->sm_count + tiles - 1) / tiles;
    mps = ((M + slices - 1) / slices + TG_KSTG - 1) / TG_KSTG * TG_KSTG;
    if (mps < 128) mps = 128;
    slices = (M + mps - 1) / mps;
  };
  int mps, slices;
  if (M >= 4096 && N2 >= 128) {
    plan(128, mps, slices);
    return tg_launch<SrcTN<LoadA, LoadB>, Epi, 128>(f, st, src, N1, N2, M, mps, slices, epi);
  }
  plan(64, mps, slices);
  return tg_launch<SrcTN<LoadA, LoadB>, Epi, 64>(f, st, src, N1, N2, M, mps, slices, epi);
}

// tier / size dispatch: the FP32 parity tier keeps its FP32-FMA kernels, and so do calls with fewer rows than
// PF_TRAIN_TC_MIN_ROWS (default 512: a 128-row tile of such a call is mostly padding).
template <class Epi, int AMODE, int BN>
inline pf_status gemm_nt(pf_flow* f, cudaStream_t st, const float* A, int lda, const float* B, int ldb, int M, int N,
                         int K, const Epi& epi) {
  if (f->train_tc && M >= f->train_tc_min_rows) return tc_gemm_nt<Epi, AMODE>(f, st, A, lda, B, ldb, M, N, K, epi);
  return sgemm_nt<Epi, AMODE, BN>(f, st, A, lda, B, ldb, M, N, K, epi);
}
template <class Epi, class LoadA, class LoadB>
inline pf_status gemm_tn(pf_flow* f, cudaStream_t st, const LoadA& la, const LoadB& lb, int M, int N1, int N2,
                         const Epi& epi) {
  if (f->train_tc && M >= f->train_tc_min_rows) return tc_gemm_tn<Epi, LoadA, LoadB>(f, st, la, lb, M, N1, N2, epi);
  return sgemm_tn<Epi, LoadA, LoadB>(f, st, la, lb, M, N1, N2, epi);
}

}  // namespace pf
