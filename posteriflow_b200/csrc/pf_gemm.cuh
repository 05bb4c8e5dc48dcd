// FP32 FMA GEMM building blocks of the parity tier (PF_TIER_FP32).
//   sgemm_nt : C[M,N] = epi( op(A)[M,K] . B[N,K]^T )     both operands K-contiguous
//   sgemm_tn : C[N1,N2] = epi( sum_m A[m,N1] . B[m,N2] ) split over m (weight gradients)
// Epilogues are functors so bias / activation / GLU-residual / scatter-to-nflows-layout are fused
// into the producing kernel and no [R,H] tensor is re-read just to apply an elementwise op.
#pragma once
#include "pf_common.cuh"

namespace pf {

constexpr int NT_BM = 128, NT_BK = 16, NT_THREADS = 256;
enum { A_PLAIN = 0, A_RELU = 1, A_SANITIZE = 2 };  // transform applied to A on load

__device__ __forceinline__ float sanitize_ctx(float v) {  // flows.py:632 nan_to_num(nan=0, +inf=1e-3, -inf=-1e-3)
  if (v != v) return 0.f;
  if (isinf(v)) return v > 0.f ? 1e-3f : -1e-3f;
  return v;
}

// BN = 64 (thread tile TMx4) or 32 (thread tile TMx2: the sampler's 26-unit degree groups).
// BM = 128 (TM = 8) for row counts that fill the GPU; BM = 32 (TM = 2) for the training step's ~1 k rows, where
// 128-row tiles would leave three quarters of the SMs without a CTA.
template <class Epi, int AMODE, int BN, int BM = NT_BM>
__global__ void __launch_bounds__(NT_THREADS)
sgemm_nt_kernel(const float* __restrict__ A, int lda, const float* __restrict__ B, int ldb, int M, int N, int K,
                Epi epi) {
  constexpr int TN = BN / 16;
  constexpr int TM = BM / 16;
  static_assert(TM == 8 || TM == 2, "row tile");
  __shared__ __align__(16) float As[2][NT_BK][BM + 4];
  __shared__ __align__(16) float Bs[2][NT_BK][BN + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int lk = tid & 15;   // k within tile for loads
  const int lr = tid >> 4;   // 0..15 row (stride 16) for loads
  const int ty = tid >> 4, tx = tid & 15;

  float ra[TM], rb[TN];
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  auto gload = [&](int k0) {
#pragma unroll
    for (int i = 0; i < TM; ++i) {
      int m = m0 + lr + 16 * i, k = k0 + lk;
      float v = (m < M && k < K) ? A[(int64_t)m * lda + k] : 0.f;
      ra[i] = AMODE == A_RELU ? fmaxf(v, 0.f) : (AMODE == A_SANITIZE ? sanitize_ctx(v) : v);
    }
#pragma unroll
    for (int i = 0; i < TN; ++i) {
      int n = n0 + lr + 16 * i, k = k0 + lk;
      rb[i] = (n < N && k < K) ? B[(int64_t)n * ldb + k] : 0.f;
    }
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int i = 0; i < TM; ++i) As[buf][lk][lr + 16 * i] = ra[i];
#pragma unroll
    for (int i = 0; i < TN; ++i) Bs[buf][lk][lr + 16 * i] = rb[i];
  };

  const int nk = (K + NT_BK - 1) / NT_BK;
  gload(0);
  sstore(0);
  __syncthreads();
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) gload((kt + 1) * NT_BK);
#pragma unroll
    for (int k = 0; k < NT_BK; ++k) {
      float av[TM];
      if constexpr (TM == 8) {
        float4 a0 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 8]);
        float4 a1 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 8 + 4]);
        av[0] = a0.x; av[1] = a0.y; av[2] = a0.z; av[3] = a0.w;
        av[4] = a1.x; av[5] = a1.y; av[6] = a1.z; av[7] = a1.w;
      } else {
        float2 a0 = *reinterpret_cast<const float2*>(&As[buf][k][ty * 2]);
        av[0] = a0.x; av[1] = a0.y;
      }
      float bv[TN];
      if (TN == 4) {
        float4 b = *reinterpret_cast<const float4*>(&Bs[buf][k][tx * 4]);
        bv[0] = b.x; bv[1] = b.y; bv[TN - 2] = b.z; bv[TN - 1] = b.w;
      } else {
        float2 b = *reinterpret_cast<const float2*>(&Bs[buf][k][tx * 2]);
        bv[0] = b.x; bv[1] = b.y;
      }
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    if (kt + 1 < nk) {
      sstore(buf ^ 1);
      __syncthreads();
    }
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    int m = m0 + ty * TM + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      int n = n0 + tx * TN + j;
      if (n < N) epi(m, n, acc[i][j]);
    }
  }
}

template <class Epi, int AMODE, int BN>
inline pf_status sgemm_nt(pf_flow* f, cudaStream_t st, const float* A, int lda, const float* B, int ldb, int M, int N,
                          int K, const Epi& epi) {
  static_assert(BN == 64 || BN == 32, "tile");
  if (M <= 0 || N <= 0) return PF_OK;
  dim3 grid((N + BN - 1) / BN, (M + NT_BM - 1) / NT_BM);
  if ((int)(grid.x * grid.y) < 2 * f->sm_count) {   // too few 128-row tiles to fill the GPU: 32-row tiles
    grid.y = (M + 31) / 32;
    sgemm_nt_kernel<Epi, AMODE, BN, 32><<<grid, NT_THREADS, 0, st>>>(A, lda, B, ldb, M, N, K, epi);
  } else {
    sgemm_nt_kernel<Epi, AMODE, BN><<<grid, NT_THREADS, 0, st>>>(A, lda, B, ldb, M, N, K, epi);
  }
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}

// ---- TN (contraction over rows) with split-M and atomic accumulate --------------------------
constexpr int TN_B1 = 64, TN_B2 = 64, TN_BK = 16, TN_THREADS = 256;

// A(m, n1) = opA(A[m*lda + n1]) ; B(m, n2) = opB(...).  Operand transforms are functors so that
// relu(a), a*gate*..., etc. are applied on load.
template <class Epi, class LoadA, class LoadB>
__global__ void __launch_bounds__(TN_THREADS)
sgemm_tn_kernel(LoadA la, LoadB lb, int M, int N1, int N2, int m_per_slice, Epi epi) {
  __shared__ __align__(16) float As[2][TN_BK][TN_B1];
  __shared__ __align__(16) float Bs[2][TN_BK][TN_B2];
  const int tid = threadIdx.x;
  const int n10 = blockIdx.y * TN_B1, n20 = blockIdx.x * TN_B2;
  const int mbeg = blockIdx.z * m_per_slice;
  const int mend = min(M, mbeg + m_per_slice);
  const int lc = tid & 63, lrow = tid >> 6;  // 4 rows x 64 cols per pass
  const int ty = tid >> 4, tx = tid & 15;
  float ra[4], rb[4];
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  auto gload = [&](int mb) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int m = mb + lrow + 4 * i;
      ra[i] = (m < mend && n10 + lc < N1) ? la(m, n10 + lc) : 0.f;
      rb[i] = (m < mend && n20 + lc < N2) ? lb(m, n20 + lc) : 0.f;
    }
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      As[buf][lrow + 4 * i][lc] = ra[i];
      Bs[buf][lrow + 4 * i][lc] = rb[i];
    }
  };
  if (mbeg >= mend) return;
  const int nk = (mend - mbeg + TN_BK - 1) / TN_BK;
  gload(mbeg);
  sstore(0);
  __syncthreads();
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) gload(mbeg + (kt + 1) * TN_BK);
#pragma unroll
    for (int k = 0; k < TN_BK; ++k) {
      float4 a = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
      float4 b = *reinterpret_cast<const float4*>(&Bs[buf][k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    if (kt + 1 < nk) {
      sstore(buf ^ 1);
      __syncthreads();
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int n1 = n10 + ty * 4 + i;
    if (n1 >= N1) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n2 = n20 + tx * 4 + j;
      if (n2 < N2) epi(n1, n2, acc[i][j]);
    }
  }
}

template <class Epi, class LoadA, class LoadB>
inline pf_status sgemm_tn(pf_flow* f, cudaStream_t st, const LoadA& la, const LoadB& lb, int M, int N1, int N2,
                          const Epi& epi) {
  if (M <= 0 || N1 <= 0 || N2 <= 0) return PF_OK;
  int tiles = ((N1 + TN_B1 - 1) / TN_B1) * ((N2 + TN_B2 - 1) / TN_B2);
  int want = (2 * f->sm_count + tiles - 1) / tiles;            // ~2 CTAs per SM
  int max_slices = (M + 4 * TN_BK - 1) / (4 * TN_BK);
  int slices = want < 1 ? 1 : (want > max_slices ? max_slices : want);
  int mps = ((M + slices - 1) / slices + TN_BK - 1) / TN_BK * TN_BK;
  slices = (M + mps - 1) / mps;
  dim3 grid((N2 + TN_B2 - 1) / TN_B2, (N1 + TN_B1 - 1) / TN_B1, slices);
  sgemm_tn_kernel<Epi, LoadA, LoadB><<<grid, TN_THREADS, 0, st>>>(la, lb, M, N1, N2, mps, epi);
  PF_LAUNCH_CHECK(f);
  return PF_OK;
}

// ---- plain operand loaders for sgemm_tn -----------------------------------------------------
struct LoadPlain {
  const float* p; int ld;
  __device__ __forceinline__ float operator()(int m, int n) const { return p[(int64_t)m * ld + n]; }
};
struct LoadRelu {
  const float* p; int ld;
  __device__ __forceinline__ float operator()(int m, int n) const { return fmaxf(p[(int64_t)m * ld + n], 0.f); }
};

}  // namespace pf
