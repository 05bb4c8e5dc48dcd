// Which TMEM column alignments do tcgen05.st / tcgen05.ld (32x32b.xN) and the TMEM A operand of tcgen05.mma
// accept?  One case per process (a fault poisons the context):  tmem_align_probe <st|ld|mma> <N> <column>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../posteriflow_b200/csrc/pf_sm100.cuh"
using namespace pf::sm100;

template <int N>
__device__ void stN(uint32_t addr, const uint32_t* v) {
  if (N == 1) tmem_st1(addr, v);
  if (N == 2) tmem_st2(addr, v);
  if (N == 4) tmem_st4(addr, v);
  if (N == 8) tmem_st8(addr, v);
  if (N == 16) tmem_st16(addr, v);
}
template <int N>
__device__ void ldN(uint32_t addr, uint32_t* v) {
  if (N == 1) tmem_ld1(addr, v);
  if (N == 2) tmem_ld2(addr, v);
  if (N == 4) tmem_ld4(addr, v);
  if (N == 8) tmem_ld8(addr, v);
  if (N == 16) tmem_ld16(addr, v);
  if (N == 32) tmem_ld32(addr, v);
}

template <int N>
__global__ void k_stld(int is_ld, int col, int* bad) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(&slot, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tl = slot + ((uint32_t)(warp * 32) << 16);
  uint32_t v[32], w[32];
  for (int i = 0; i < 32; ++i) v[i] = threadIdx.x * 1000u + col + i;
  if (!is_ld) {
    if (N <= 16) stN<N>(tl + col, v);
    tmem_wait_st();
    for (int i = 0; i < N; ++i) { tmem_ld1(tl + col + i, w + i); }
    tmem_wait_ld();
  } else {
    for (int i = 0; i < N; ++i) tmem_st1(tl + col + i, v + i);
    tmem_wait_st();
    ldN<N>(tl + col, w);
    tmem_wait_ld();
  }
  for (int i = 0; i < N; ++i) if (w[i] != v[i]) atomicAdd(bad, 1);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(slot, 128);
}

// TS MMA with the A operand ([128 x 16] bf16 = 8 columns) at column `col`, D at column 0, B = identity-ish
__global__ void k_mma(int col, int N, int dcol, int* bad) {
  __shared__ uint32_t slot;
  __shared__ uint64_t bar;
  __shared__ __align__(128) uint8_t sB[64 * 16 * 2];   // N<=64 rows x K=16
  const int warp = threadIdx.x >> 5, tid = threadIdx.x;
  if (tid == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (warp == 0) tmem_alloc(&slot, 512);
  // B[n][k] = (n % 16 == k): D[m][n] = A[m][n % 16]
  for (int e = tid; e < N * 16; e += blockDim.x) {
    int n = e / 16, k = e % 16;
    reinterpret_cast<__nv_bfloat16*>(sB)[(k / 8) * (N * 8) + (n / 8) * 64 + (n % 8) * 8 + (k % 8)] = __float2bfloat16(n % 16 == k ? 1.f : 0.f);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tl = slot + ((uint32_t)(warp * 32) << 16);
  for (int i = 0; i < 8; ++i) {
    uint32_t w = pack_bf16((float)(tid % 64) + 2 * i, (float)(tid % 64) + 2 * i + 1);
    tmem_st1(tl + col + i, &w);
  }
  tmem_wait_st();
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after();
    umma_ts(slot + dcol, slot + col, umma_smem_desc(smem_u32(sB), N * 16, 128), umma_idesc_bf16(128, N), 0);
    umma_commit(&bar);
  }
  if (!mbar_wait(&bar, 0)) atomicAdd(bad, 1000);
  tc_fence_after();
  for (int c = 0; c < N; c += 16) {
    uint32_t v[16];
    tmem_ld16(tl + dcol + c, v);
    tmem_wait_ld();
    for (int i = 0; i < 16; ++i) if (__uint_as_float(v[i]) != (float)(tid % 64) + i) atomicAdd(bad, 1);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(slot, 512);
}

int main(int argc, char** argv) {
  if (argc < 4) return 2;
  const char* op = argv[1];
  int N = atoi(argv[2]), col = atoi(argv[3]);
  int* bad;
  cudaMalloc(&bad, 4);
  cudaMemset(bad, 0, 4);
  if (!strcmp(op, "mma")) {
    int dcol = argc > 4 ? atoi(argv[4]) : 0;
    k_mma<<<1, 128>>>(col, N, dcol, bad);
    printf("(D col %d) ", dcol);
  } else {
    int is_ld = !strcmp(op, "ld");
    switch (N) {
      case 1: k_stld<1><<<1, 128>>>(is_ld, col, bad); break;
      case 2: k_stld<2><<<1, 128>>>(is_ld, col, bad); break;
      case 4: k_stld<4><<<1, 128>>>(is_ld, col, bad); break;
      case 8: k_stld<8><<<1, 128>>>(is_ld, col, bad); break;
      case 16: k_stld<16><<<1, 128>>>(is_ld, col, bad); break;
      case 32: k_stld<32><<<1, 128>>>(is_ld, col, bad); break;
    }
  }
  cudaError_t e = cudaDeviceSynchronize();
  int h = -1;
  cudaMemcpy(&h, bad, 4, cudaMemcpyDeviceToHost);
  printf("%s x%d col %d: %s mismatches=%d\n", op, N, col, cudaGetErrorString(e), h);
  return 0;
}
