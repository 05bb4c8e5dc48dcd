// Register <-> (TMEM lane, column) mapping of tcgen05.ld.16x256b (B200).  Pattern written with 32x32b stores
// (value = lane * 1000 + column), read back with 16x256b.x2 at lane offsets 0 and 16 of warp 0's sub-partition.
#include <cuda_runtime.h>
#include <stdio.h>
#include "../posteriflow_b200/csrc/pf_sm100.cuh"
using namespace pf::sm100;

__global__ void k(uint32_t* out) {
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (warp == 0) tmem_alloc(&slot, 32);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tl = slot + ((uint32_t)(warp * 32) << 16);
  uint32_t v[16];
  for (int i = 0; i < 16; ++i) v[i] = tid * 1000u + i;
  tmem_st16(tl, v);
  tmem_wait_st();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  for (int half = 0; half < 2; ++half) {
    uint32_t r[8];
    uint32_t addr = tl + ((uint32_t)(half * 16) << 16);
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(addr));
    tmem_wait_ld();
    for (int i = 0; i < 8; ++i) out[(half * 128 + tid) * 8 + i] = r[i];
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc(slot, 32);
}
int main() {
  uint32_t* d; cudaMalloc(&d, 2 * 128 * 8 * 4);
  k<<<1, 128>>>(d);
  cudaError_t e = cudaDeviceSynchronize();
  static uint32_t h[2 * 128 * 8];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  printf("%s\n", cudaGetErrorString(e));
  int bad = 0;
  for (int half = 0; half < 2; ++half)
    for (int t = 0; t < 128; ++t)
      for (int i = 0; i < 8; ++i) {
        uint32_t val = h[(half * 128 + t) * 8 + i];
        int lane = t & 31, w = t >> 5;
        // hypothesis: reg i -> chunk c = i/4, row = 32w + 16*half + lane/4 + 8*((i%4)/2), col = 8c + 2*(lane%4) + (i%2)
        int row = 32 * w + 16 * half + lane / 4 + 8 * ((i % 4) / 2), col = 8 * (i / 4) + 2 * (lane % 4) + (i % 2);
        if (val != (uint32_t)(row * 1000 + col)) { if (bad < 8) printf("t%d half%d r%d = %u (row %u col %u), expected row %d col %d\n", t, half, i, val, val / 1000, val % 1000, row, col); ++bad; }
      }
  printf("16x256b mapping hypothesis: %s (%d mismatches)\n", bad ? "FAIL" : "PASS", bad);
  return 0;
}
