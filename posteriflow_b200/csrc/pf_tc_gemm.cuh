// Tensor-core GEMM building blocks of the TRAINING path under PF_TIER_BF16 (forward with saved activations +
// backward, pf_fp32.cu): the same functor interfaces as sgemm_nt / sgemm_tn (pf_gemm.cuh), but the products run
// on tcgen05.mma with fp32 accumulation in tensor memory.
//
// Accuracy: operands arrive as fp32 and are split on the fly into TWO bf16 terms, a = a_hi + a_lo
// (a_lo = bf16(a - a_hi): 16 significand bits survive), and every K=16 chunk issues three MMAs
//     acc += a_hi b_hi + a_hi b_lo + a_lo b_hi            (the a_lo b_lo term, ~2^-18 |ab|, is dropped)
// so gradients keep ~1e-5 relative accuracy: the training step can use the tensor pipe without giving up the
// weight-gradient tolerance that a plain bf16 product (2^-9) would.
//
// Shape of one CTA (256 threads, 80-96 KB of shared memory: two per SM, so one CTA's operand loads overlap the other's
// MMAs and epilogue): a 128-row x BN-column output tile, the accumulator is BN columns of
// tensor memory (lane = output row).  K advances in stages of KS = 128: all threads convert the stage's operands
// into the UMMA canonical K-major no-swizzle layout in shared memory (8-row x 16-byte core matrices; element
// (row, k) at (k/8) * ROWS*16 + row*16 + (k%8)*2, i.e. LBO = ROWS*16 along K, SBO = 128 along rows -- the layout
// verified by tests/umma_probe.cu), one 16-byte store per (row, 8 k) so stores are bank-conflict free; then a
// converged warp elects one lane to issue the stage's MMAs and a single commit.  The products here have K <= 768
// (hidden 256, final 517, context 3*256 / 288) or K = rows of a <= 256-row slice (weight gradients): two to six
// stages of load -> MMA -> wait, the overlap comes from the second resident CTA; the epilogue functor
// runs on values read back with tcgen05.ld and transposed through shared memory (tg_epilogue).
#pragma once
#include "pf_common.cuh"
#include "pf_gemm.cuh"
#include "pf_sm100.cuh"

namespace pf {

constexpr int TG_ROWS = 128, TG_KS = 128, TG_THREADS = 256;
constexpr int TG_UB = 8;   // items whose global loads are in flight together per thread
template <int BN>
constexpr int tg_smem_bytes() { return 2 * TG_KS * TG_ROWS * 2 + 2 * TG_KS * BN * 2 + 64; }

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

// issue the three split-bf16 MMAs of every K=16 chunk of one stage, then one commit (call from a converged warp)
template <int BN>
__device__ __forceinline__ void tg_issue_stage(uint32_t tmem, uint32_t aAh, uint32_t aAl, uint32_t aBh, uint32_t aBl,
                                               int nchunks, bool first_stage, uint64_t* bar) {
  using namespace sm100;
  if (elect_one_sync()) {
    const uint32_t idesc = umma_idesc_bf16(TG_ROWS, BN);
    uint64_t ah = umma_smem_desc(aAh, TG_ROWS * 16, 128), al = umma_smem_desc(aAl, TG_ROWS * 16, 128);
    uint64_t bh = umma_smem_desc(aBh, BN * 16, 128), bl = umma_smem_desc(aBl, BN * 16, 128);
    const uint64_t astep = (uint64_t)((2u * TG_ROWS * 16u) >> 4), bstep = (uint64_t)((2u * (uint32_t)BN * 16u) >> 4);
    uint32_t accf = first_stage ? 0u : 1u;
    for (int c = 0; c < nchunks; ++c) {
      umma_ss(tmem, ah, bh, idesc, accf);
      umma_ss(tmem, ah, bl, idesc, 1u);
      umma_ss(tmem, al, bh, idesc, 1u);
      accf = 1u;
      ah += astep; al += astep; bh += bstep; bl += bstep;
    }
    umma_commit(bar);
  }
  __syncwarp();
}

struct TgSmem {
  uint8_t *Ah, *Al, *Bh, *Bl;
  uint64_t* bar;
  uint32_t* tmem_slot;
};
template <int BN>
__device__ __forceinline__ TgSmem tg_carve(uint8_t* smem) {
  TgSmem s;
  s.Ah = smem;
  s.Al = s.Ah + TG_KS * TG_ROWS * 2;
  s.Bh = s.Al + TG_KS * TG_ROWS * 2;
  s.Bl = s.Bh + TG_KS * BN * 2;
  s.bar = reinterpret_cast<uint64_t*>(s.Bl + TG_KS * BN * 2);
  s.tmem_slot = reinterpret_cast<uint32_t*>(s.bar + 1);
  return s;
}

// prologue / epilogue shared by the two kernels
template <int BN>
__device__ __forceinline__ uint32_t tg_begin(const TgSmem& s) {
  using namespace sm100;
  if (threadIdx.x == 0) {
    mbar_init(s.bar, 1);
    fence_mbar_init();
  }
  if ((threadIdx.x >> 5) == 0) {
    __syncwarp();
    tmem_alloc(s.tmem_slot, BN < 32 ? 32 : BN);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  return *s.tmem_slot;
}

// Epilogue shared by the two kernels: the accumulator tile goes TMEM -> registers (thread = output row, the only
// shape tcgen05.ld offers) -> shared memory (the operand buffers are free once the last MMA has completed) -> back
// out with the warp's lanes along the COLUMN index, so the functor's global loads / stores / atomics of one
// instruction fall into one or two 128-byte lines instead of 32 rows' worth of sectors.
template <int BN, class F>
__device__ __forceinline__ void tg_epilogue(uint8_t* smem, uint32_t tmem, int warp, int tid, F&& f) {
  using namespace sm100;
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
  tc_fence_before();
  __syncthreads();
  for (int idx = tid; idx < TG_ROWS * BN; idx += TG_THREADS) {
    const int r = idx / BN, c = idx % BN;
    f(r, c, stage[r * LD + c]);
  }
}

// C[M,N] = epi( op(A)[M,K] . B[N,K]^T ), both operands K-contiguous fp32 (same contract as sgemm_nt)
template <class Epi, int AMODE, int BN>
__global__ void __launch_bounds__(TG_THREADS, 2)
tc_gemm_nt_kernel(const float* __restrict__ A, int lda, const float* __restrict__ B, int ldb, int M, int N, int K,
                  Epi epi, int* __restrict__ err) {
  using namespace sm100;
  extern __shared__ __align__(128) uint8_t smem[];
  const TgSmem s = tg_carve<BN>(smem);
  const uint32_t tmem = tg_begin<BN>(s);
  const int tid = threadIdx.x;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int m0 = blockIdx.y * TG_ROWS, n0 = blockIdx.x * BN;
  const bool a_vec = (lda & 3) == 0 && (reinterpret_cast<uintptr_t>(A) & 15) == 0;
  const bool b_vec = (ldb & 3) == 0 && (reinterpret_cast<uintptr_t>(B) & 15) == 0;
  uint32_t phase = 0;
  bool ok = true;
  for (int k0 = 0; k0 < K && ok; k0 += TG_KS) {
    const int kc = min(TG_KS, K - k0);
    const int ngrp = (kc + 15) / 16 * 2;   // 8-wide k groups of this stage (whole K=16 chunks)
    // A items [0, nA) then B items [nA, nA + nB); an item = one row x 8 consecutive k.  The global loads of TG_UB
    // items are all issued before the first conversion, so a thread has 2 * TG_UB 16-byte loads in flight (with
    // one CTA of 8 warps per SM, a load-convert-store loop would expose the full L2 latency per item).
    const int nA = TG_ROWS * ngrp, nB = BN * ngrp;
    for (int base = tid; base < nA + nB; base += TG_THREADS * TG_UB) {
      float v[TG_UB][8];
#pragma unroll
      for (int u = 0; u < TG_UB; ++u) {
        const int idx = base + u * TG_THREADS;
#pragma unroll
        for (int i = 0; i < 8; ++i) v[u][i] = 0.f;
        if (idx < nA + nB) {
          const bool isA = idx < nA;
          const int j = isA ? idx : idx - nA;
          const int r = isA ? (j & (TG_ROWS - 1)) : (j % BN), kg = isA ? (j >> 7) : (j / BN);
          const int row = (isA ? m0 : n0) + r, k = k0 + kg * 8;
          if (row < (isA ? M : N)) {
            const float* p = (isA ? A + (int64_t)row * lda : B + (int64_t)row * ldb) + k;
            if ((isA ? a_vec : b_vec) && k + 8 <= K) {
              const float4 x0 = __ldg(reinterpret_cast<const float4*>(p)), x1 = __ldg(reinterpret_cast<const float4*>(p + 4));
              v[u][0] = x0.x; v[u][1] = x0.y; v[u][2] = x0.z; v[u][3] = x0.w;
              v[u][4] = x1.x; v[u][5] = x1.y; v[u][6] = x1.z; v[u][7] = x1.w;
            } else {
#pragma unroll
              for (int i = 0; i < 8; ++i)
                if (k + i < K) v[u][i] = __ldg(p + i);
            }
          }
        }
      }
#pragma unroll
      for (int u = 0; u < TG_UB; ++u) {
        const int idx = base + u * TG_THREADS;
        if (idx < nA + nB) {
          const bool isA = idx < nA;
          const int j = isA ? idx : idx - nA;
          const int r = isA ? (j & (TG_ROWS - 1)) : (j % BN), kg = isA ? (j >> 7) : (j / BN);
          if (isA) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
              v[u][i] = AMODE == A_RELU ? fmaxf(v[u][i], 0.f) : (AMODE == A_SANITIZE ? sanitize_ctx(v[u][i]) : v[u][i]);
          }
          const uint32_t off = (uint32_t)kg * ((isA ? TG_ROWS : BN) * 16) + (uint32_t)r * 16;
          tg_split_store((isA ? s.Ah : s.Bh) + off, (isA ? s.Al : s.Bl) + off, v[u]);
        }
      }
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (warp == 4) {
      tc_fence_after();
      tg_issue_stage<BN>(tmem, smem_u32(s.Ah), smem_u32(s.Al), smem_u32(s.Bh), smem_u32(s.Bl), ngrp / 2, k0 == 0, s.bar);
    }
    ok = mbar_wait(s.bar, phase);   // MMAs done: shared memory may be rewritten, the accumulator read
    phase ^= 1;
    tc_fence_after();
  }
  if (!ok && tid == 0) atomicExch(err, 60);
  if (__syncthreads_and(ok)) {
    tg_epilogue<BN>(smem, tmem, warp, tid, [&](int r, int c, float acc) {
      const int m = m0 + r, n = n0 + c;
      if (m < M && n < N) epi(m, n, acc);
    });
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, BN < 32 ? 32 : BN);
}

// C[N1,N2] = epi( sum_m A(m,N1) . B(m,N2) ), split over m in slices (same contract as sgemm_tn: the epilogue
// accumulates atomically).  The contraction index is the ROW index of both operands, so the loaders transpose on
// the way into shared memory: a thread owns one n and gathers 8 consecutive m (each of the 8 loads is coalesced
// across the warp's 32 consecutive n).
template <class Epi, class LoadA, class LoadB, int BN>
__global__ void __launch_bounds__(TG_THREADS, 2)
tc_gemm_tn_kernel(LoadA la, LoadB lb, int M, int N1, int N2, int m_per_slice, Epi epi, int* __restrict__ err) {
  using namespace sm100;
  extern __shared__ __align__(128) uint8_t smem[];
  const TgSmem s = tg_carve<BN>(smem);
  const uint32_t tmem = tg_begin<BN>(s);
  const int tid = threadIdx.x;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int n10 = blockIdx.y * TG_ROWS, n20 = blockIdx.x * BN;
  const int mbeg = blockIdx.z * m_per_slice;
  const int mend = min(M, mbeg + m_per_slice);
  uint32_t phase = 0;
  bool ok = true;
  for (int k0 = mbeg; k0 < mend && ok; k0 += TG_KS) {
    const int kc = min(TG_KS, mend - k0);
    const int ngrp = (kc + 15) / 16 * 2;
    const int nA = TG_ROWS * ngrp, nB = BN * ngrp;
    for (int base = tid; base < nA + nB; base += TG_THREADS * TG_UB) {
      float v[TG_UB][8];
#pragma unroll
      for (int u = 0; u < TG_UB; ++u) {
        const int idx = base + u * TG_THREADS;
        const bool isA = idx < nA;
        const int j = isA ? idx : idx - nA;
        const int r = isA ? (j & (TG_ROWS - 1)) : (j % BN), kg = isA ? (j >> 7) : (j / BN);
        const int n = (isA ? n10 : n20) + r, m = k0 + kg * 8;
        const bool live = idx < nA + nB && n < (isA ? N1 : N2);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[u][i] = (live && m + i < mend) ? (isA ? la(m + i, n) : lb(m + i, n)) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < TG_UB; ++u) {
        const int idx = base + u * TG_THREADS;
        if (idx < nA + nB) {
          const bool isA = idx < nA;
          const int j = isA ? idx : idx - nA;
          const int r = isA ? (j & (TG_ROWS - 1)) : (j % BN), kg = isA ? (j >> 7) : (j / BN);
          const uint32_t off = (uint32_t)kg * ((isA ? TG_ROWS : BN) * 16) + (uint32_t)r * 16;
          tg_split_store((isA ? s.Ah : s.Bh) + off, (isA ? s.Al : s.Bl) + off, v[u]);
        }
      }
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (warp == 4) {
      tc_fence_after();
      tg_issue_stage<BN>(tmem, smem_u32(s.Ah), smem_u32(s.Al), smem_u32(s.Bh), smem_u32(s.Bl), ngrp / 2, k0 == mbeg, s.bar);
    }
    ok = mbar_wait(s.bar, phase);
    phase ^= 1;
    tc_fence_after();
  }
  if (!ok && tid == 0) atomicExch(err, 61);
  if (__syncthreads_and(ok) && mbeg < mend) {
    tg_epilogue<BN>(smem, tmem, warp, tid, [&](int r, int c, float acc) {
      const int n1 = n10 + r, n2 = n20 + c;
      if (n1 < N1 && n2 < N2) epi(n1, n2, acc);
    });
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, BN < 32 ? 32 : BN);
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

template <class Epi, int AMODE>
inline pf_status tc_gemm_nt(pf_flow* f, cudaStream_t st, const float* A, int lda, const float* B, int ldb, int M, int N,
                            int K, const Epi& epi) {
  if (M <= 0 || N <= 0) return PF_OK;
  const int mt = (M + TG_ROWS - 1) / TG_ROWS;
  // 32-column tiles while 64-column ones would leave most SMs without a CTA (the training step's ~1 k rows)
  if (mt * ((N + 63) / 64) < f->sm_count) {
    constexpr int BN = 32;
    PF_TRY((tg_set_smem<tc_gemm_nt_kernel<Epi, AMODE, BN>>(tg_smem_bytes<BN>(), f->desc.device)));
    dim3 grid((N + BN - 1) / BN, mt);
    tc_gemm_nt_kernel<Epi, AMODE, BN><<<grid, TG_THREADS, tg_smem_bytes<BN>(), st>>>(A, lda, B, ldb, M, N, K, epi, f->tg_err);
  } else {
    constexpr int BN = 64;
    PF_TRY((tg_set_smem<tc_gemm_nt_kernel<Epi, AMODE, BN>>(tg_smem_bytes<BN>(), f->desc.device)));
    dim3 grid((N + BN - 1) / BN, mt);
    tc_gemm_nt_kernel<Epi, AMODE, BN><<<grid, TG_THREADS, tg_smem_bytes<BN>(), st>>>(A, lda, B, ldb, M, N, K, epi, f->tg_err);
  }
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}

template <class Epi, class LoadA, class LoadB>
inline pf_status tc_gemm_tn(pf_flow* f, cudaStream_t st, const LoadA& la, const LoadB& lb, int M, int N1, int N2,
                            const Epi& epi) {
  if (M <= 0 || N1 <= 0 || N2 <= 0) return PF_OK;
  constexpr int BN = 64;
  const int tiles = ((N1 + TG_ROWS - 1) / TG_ROWS) * ((N2 + BN - 1) / BN);
  // rows per slice: 256 (one full stage), 128 when that leaves most of the GPU without a CTA.  Not finer: every
  // slice adds a full tile of atomics, and these kernels share the GPU with the dependency chain on the main stream.
  int mps = 256;
  if (tiles * ((M + mps - 1) / mps) < f->sm_count / 2) mps = 128;
  const int slices = (M + mps - 1) / mps;
  PF_TRY((tg_set_smem<tc_gemm_tn_kernel<Epi, LoadA, LoadB, BN>>(tg_smem_bytes<BN>(), f->desc.device)));
  dim3 grid((N2 + BN - 1) / BN, (N1 + TG_ROWS - 1) / TG_ROWS, slices);
  tc_gemm_tn_kernel<Epi, LoadA, LoadB, BN><<<grid, TG_THREADS, tg_smem_bytes<BN>(), st>>>(la, lb, M, N1, N2, mps, epi, f->tg_err);
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}

// tier / size dispatch: the FP32 parity tier keeps its FP32-FMA kernels, and so do calls with fewer rows than
// PF_TRAIN_TC_MIN_ROWS (default 2048).  Measured on B200 (scripts/prof_train.py, one CUDA-graph replay of forward +
// backward): 7 515 rows (config #4 on one GPU) 10.0 ms with these kernels against 13.9 ms with the FP32-FMA ones;
// 948 rows (its 8-GPU shard) 4.0 ms against 3.3 ms -- a CTA here is a serial load -> MMA -> epilogue chain and
// ~64 CTAs cannot hide it, while the FP32 kernels run three double-buffered CTAs per SM.
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
