// How long does a chain of small tcgen05.mma (M=128, K=16, bf16) take when every instruction accumulates
// into the SAME TMEM tile, versus round-robin over P independent tiles?  (B200; one CTA, one issuing thread)
//   mma_latency_probe            -> table over N in {32,48,64,128,256}, P in {1,2,4}, A from smem (SS) / TMEM (TS)
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#include "../posteriflow_b200/csrc/pf_sm100.cuh"
using namespace pf::sm100;

__global__ void __launch_bounds__(128) k(int N, int P, int ts, int n_mma, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;               // [2 k-chunks][128][16B] = 4 KB
  uint8_t* sB = smem + 4096;        // [2][N][16B]
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  if (tid == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (warp == 0) tmem_alloc(&slot, 512);
  for (int i = tid; i < (4096 + 2 * 256 * 16) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  {  // zero A region in TMEM (columns 448..455) and nothing else matters (values are zeros)
    uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    tmem_st8(tm + 448 + ((uint32_t)(warp * 32) << 16), z);
    tmem_wait_st();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {   // converged warp, one elected lane issues (see elect_one_sync in pf_sm100.cuh)
    tc_fence_after();
    const uint32_t idesc = umma_idesc_bf16(128, N);
    const uint64_t ad = umma_smem_desc(smem_u32(sA), 2048, 128);
    const uint64_t bd = umma_smem_desc(smem_u32(sB), N * 16, 128);
    // warm-up
    if (elect_one_sync()) {
      for (int i = 0; i < 8; ++i) {
        if (ts) umma_ts(tm, tm + 448, bd, idesc, 1); else umma_ss(tm, ad, bd, idesc, 1);
      }
      umma_commit(&bar);
    }
    __syncwarp();
    mbar_wait(&bar, 0);
    tc_fence_after();
    long long t0 = clock64();
    long long t1 = t0;
    if (elect_one_sync()) {
      for (int i = 0; i < n_mma; ++i) {
        const uint32_t d = tm + (uint32_t)((i & (P - 1)) * N);
        if (ts) umma_ts(d, tm + 448, bd, idesc, 1); else umma_ss(d, ad, bd, idesc, 1);
      }
      t1 = clock64();
      umma_commit(&bar);
    }
    __syncwarp();
    mbar_wait(&bar, 1);
    long long t2 = clock64();
    if ((tid & 31) == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 512);
}

int main() {
  long long* d;
  cudaMalloc(&d, 16);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 4096 + 2 * 256 * 16);
  const int n = 64;
  for (int ts = 0; ts < 2; ++ts)
    for (int N : {32, 48, 64, 128, 256})
      for (int P : {1, 2, 4}) {
        if (P * N > 448) continue;
        k<<<1, 128, 4096 + 2 * 256 * 16>>>(N, P, ts, n, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[2] = {0, 0};
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("%s N=%3d accumulators=%d: issue %6.1f cyc/mma, complete %6.1f cyc/mma  (floor %d)  %s\n", ts ? "TS" : "SS", N, P,
               (double)h[0] / n, (double)h[1] / n, 128 * N / 256, cudaGetErrorString(e));
      }
  return 0;
}
