// Hardware probe (run on a B200 through gpurun): tcgen05.mma.cta_group::2 with M = 128, i.e. a CTA PAIR in which each
// CTA owns 64 rows of A and half of the N rows of B.  Questions it answers for a pair-split of the chain kernels:
//   (1) which TMEM (lane, column) of each CTA receives D[m, n]  (printed as a map + a one-line summary),
//   (2) cycles per instruction against the cta_group::1 M = 128 form at the same N.
// D[m, n] = (n + 1) + (m + 1) / 256: A[m][0] = 1, A[m][1] = m + 1 (global row m), B[n][0] = n + 1, B[n][1] = 1 / 256.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o umma_2cta_probe tests/umma_2cta_probe.cu
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "../posteriflow_b200/csrc/pf_sm100.cuh"
using namespace pf::sm100;

__device__ __forceinline__ uint32_t cta_rank() { return cluster_cta_rank(); }

// PAIR = 1: cluster of two CTAs, cta_group::2, M = 128 (64 rows per CTA), B split.  PAIR = 0: one CTA, M = 128, all of B.
template <int PAIR>
__global__ void __launch_bounds__(160) probe_kernel(const __nv_bfloat16* __restrict__ A, const __nv_bfloat16* __restrict__ B,
                                                    int N, float* __restrict__ out, long long* __restrict__ cyc, int reps,
                                                    int* err) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int rows = PAIR ? 64 : 128, nb = PAIR ? N / 2 : N;
  // NBUF copies of each operand: repetition i reads copy i % NBUF (a steady-state kernel never re-reads a tile)
  constexpr int NBUF = 8;
  uint8_t* sA = smem;                 // [NBUF][2 K-slabs][rows / 8][128 B]   4 KB apart
  uint8_t* sB = smem + NBUF * 4096;   // [NBUF][2 K-slabs][nb / 8][128 B]     8 KB apart
  __shared__ uint64_t bar_acc;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = PAIR ? cta_rank() : 0;
  if (tid == 0) {
    mbar_init(&bar_acc, 1);
    fence_mbar_init();
  }
  // operands -> shared memory (generic stores), K-major no-swizzle core matrices
  for (int e = tid; e < rows * 16; e += blockDim.x) {
    const int m = e / 16, k = e % 16;
    const int gm = (int)rank * rows + m;
    for (int b = 0; b < NBUF; ++b)
      *reinterpret_cast<__nv_bfloat16*>(sA + b * 4096 + (k / 8) * (rows * 16) + (m / 8) * 128 + (m % 8) * 16 + (k % 8) * 2) = A[gm * 16 + k];
  }
  for (int e = tid; e < nb * 16; e += blockDim.x) {
    const int n = e / 16, k = e % 16;
    const int gn = (int)rank * nb + n;
    for (int b = 0; b < NBUF; ++b)
      *reinterpret_cast<__nv_bfloat16*>(sB + b * 8192 + (k / 8) * (nb * 16) + (n / 8) * 128 + (n % 8) * 16 + (k % 8) * 2) = B[gn * 16 + k];
  }
  fence_proxy_async_smem();
  if (warp == 4) { if (PAIR) tmem_alloc2(&tmem_base_s, 512); else tmem_alloc(&tmem_base_s, 512); }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;
  // sentinel into every cell this CTA can see
  if (warp < 4) {
    uint32_t s16[16];
    for (int i = 0; i < 16; ++i) s16[i] = __float_as_uint(-1.f);
    const uint32_t tl = tmem + ((uint32_t)(warp * 32) << 16);
    for (int c = 0; c < 256; c += 16) tmem_st16(tl + c, s16);
    tmem_wait_st();
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();
  tc_fence_after();
  long long t0 = 0, t1 = 0;
  if (warp == 4 && rank == 0) {
    const uint32_t idesc = umma_idesc_bf16(128, N);
    t0 = clock64();
    if (elect_one_sync()) {
      for (int i = 0; i < reps; ++i) {
        const uint64_t ad = umma_smem_desc(smem_u32(sA) + (uint32_t)(i % NBUF) * 4096u, rows * 16, 128);
        const uint64_t bd = umma_smem_desc(smem_u32(sB) + (uint32_t)(i % NBUF) * 8192u, nb * 16, 128);
        if (PAIR) umma_ss2(tmem, ad, bd, idesc, i > 0 ? 1u : 0u);
        else umma_ss(tmem, ad, bd, idesc, i > 0 ? 1u : 0u);
      }
      if (PAIR) umma_commit2(&bar_acc, 3);
      else umma_commit(&bar_acc);
    }
    __syncwarp();
  }
  // every thread of every CTA waits for the accumulator barrier of ITS CTA (multicast commit in the pair form)
  if (!mbar_wait(&bar_acc, 0)) atomicExch(err, 1 + (int)rank);
  tc_fence_after();
  if (warp == 4 && rank == 0) {
    t1 = clock64();
    if (lane == 0) cyc[0] = t1 - t0;
  }
  if (warp < 4) {
    const uint32_t tl = tmem + ((uint32_t)(warp * 32) << 16);
    for (int c = 0; c < 256; c += 16) {
      uint32_t v[16];
      tmem_ld16(tl + c, v);
      tmem_wait_ld();
      for (int i = 0; i < 16; ++i) out[((size_t)rank * 128 + tid) * 256 + c + i] = __uint_as_float(v[i]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();
  if (warp == 4) { if (PAIR) tmem_dealloc2(tmem, 512); else tmem_dealloc(tmem, 512); }
}


template <int PAIR>
static int run(int N, int reps, bool print_map) {
  std::vector<__nv_bfloat16> A(128 * 16), B(256 * 16);
  for (int m = 0; m < 128; ++m)
    for (int k = 0; k < 16; ++k) A[m * 16 + k] = __float2bfloat16(k == 0 ? 1.f : (k == 1 ? (float)(m + 1) : 0.f));
  for (int n = 0; n < 256; ++n)
    for (int k = 0; k < 16; ++k) B[n * 16 + k] = __float2bfloat16(k == 0 ? (float)(n + 1) : (k == 1 ? 1.f / 256.f : 0.f));
  __nv_bfloat16 *dA, *dB;
  float* dO;
  long long* dC;
  int* dErr;
  cudaMalloc(&dA, A.size() * 2); cudaMalloc(&dB, B.size() * 2);
  cudaMalloc(&dO, 2 * 128 * 256 * 4); cudaMalloc(&dC, 8); cudaMalloc(&dErr, 4);
  cudaMemset(dErr, 0, 4); cudaMemset(dO, 0, 2 * 128 * 256 * 4); cudaMemset(dC, 0, 8);
  cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice);
  const size_t sm = 8 * (4096 + 8192);
  cudaFuncSetAttribute(probe_kernel<PAIR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(PAIR ? 2 : 1);
  cfg.blockDim = dim3(160);
  cfg.dynamicSmemBytes = sm;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = PAIR ? 2 : 1; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, probe_kernel<PAIR>, (const __nv_bfloat16*)dA, (const __nv_bfloat16*)dB, N, dO, dC, reps, dErr);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  int herr = 0;
  long long cyc = 0;
  std::vector<float> out(2 * 128 * 256);
  cudaMemcpy(&herr, dErr, 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(&cyc, dC, 8, cudaMemcpyDeviceToHost);
  cudaMemcpy(out.data(), dO, out.size() * 4, cudaMemcpyDeviceToHost);
  printf("probe PAIR=%d N=%d reps=%d: cuda=%s flag=%d issue+complete %lld cycles (%.1f per MMA)\n", PAIR, N, reps,
         cudaGetErrorString(e), herr, cyc, (double)cyc / reps);
  if (print_map && e == cudaSuccess) {
    // value = reps * ((n + 1) + (m + 1) / 256): decode (m, n) per (cta, lane, column)
    for (int cta = 0; cta < (PAIR ? 2 : 1); ++cta) {
      int written = 0, lane_lo = 999, lane_hi = -1, col_hi = -1;
      for (int lane = 0; lane < 128; ++lane)
        for (int c = 0; c < 256; ++c) {
          const float v = out[((size_t)cta * 128 + lane) * 256 + c] / reps;
          if (v < 0.f) continue;
          ++written;
          if (lane < lane_lo) lane_lo = lane;
          if (lane > lane_hi) lane_hi = lane;
          if (c > col_hi) col_hi = c;
        }
      printf("  cta %d: %d cells written, lanes %d..%d, columns 0..%d\n", cta, written, lane_lo, lane_hi, col_hi);
      for (int lane : {0, 1, 15, 16, 31, 32, 47, 48, 63, 64, 65, 95, 96, 127}) {
        printf("    lane %3d:", lane);
        for (int c : {0, 1, 15, 16, N / 2 - 1, N / 2, N - 1}) {
          if (c < 0 || c >= 256) continue;
          const float v = out[((size_t)cta * 128 + lane) * 256 + c] / reps;
          if (v < 0.f) { printf("  col %3d: -", c); continue; }
          const int n = (int)floorf(v) - 1;
          const int m = (int)lrintf((v - floorf(v)) * 256.f) - 1;
          printf("  col %3d: D[m=%3d,n=%3d]", c, m, n);
        }
        printf("\n");
      }
    }
  }
  cudaFree(dA); cudaFree(dB); cudaFree(dO); cudaFree(dC); cudaFree(dErr);
  return (e == cudaSuccess && herr == 0) ? 0 : 1;
}

int main() {
  int bad = 0;
  bad += run<1>(256, 1, true);
  bad += run<1>(32, 1, true);
  bad += run<0>(256, 1, false);
  bad += run<1>(256, 64, false);
  bad += run<1>(256, 512, false);
  bad += run<0>(256, 512, false);
  bad += run<0>(256, 64, false);
  bad += run<1>(128, 64, false);
  bad += run<0>(128, 64, false);
  printf("umma_2cta_probe: %s\n", bad ? "FAIL" : "done");
  return bad ? 1 : 0;
}
