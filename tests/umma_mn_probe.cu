// Hardware probe (run on a B200 through gpurun) for the convention the weight-gradient kernel relies on:
// tcgen05.mma with BOTH operands MN-major, no swizzle, fed from row-major tile IMAGES
//     image[(col / 8) * 2048 + (row / 8) * 128 + (row % 8) * 16 + (col % 8) * 2]      (128 rows, bf16)
// i.e. the K-major operand layout of the chain kernels re-read with the contraction index = tile row:
//     D[m, n] = sum_row G[row, m] * A[row, n]
// A core matrix is 8 K-rows x 16 bytes (8 MN elements) either way.  mode 0: LBO = 128 (next 8 rows along K),
// SBO = 2048 (next 8 columns along M/N)  -- cute's ((T,1,m),(8,k)):((1,T,SBO),(1T,LBO)).  mode 1: the two swapped.
// Prints which mode reproduces the host reference; exit code 0 iff exactly the documented one (mode 0) passes.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o umma_mn_probe tests/umma_mn_probe.cu
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "../posteriflow_b200/csrc/pf_sm100.cuh"
using namespace pf::sm100;

constexpr int ROWS = 128, MC = 128;

template <int N>
__global__ void __launch_bounds__(160) probe_kernel(const uint8_t* __restrict__ gimg, const uint8_t* __restrict__ aimg,
                                                    float* __restrict__ out, int mode, int* err) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sG = smem;                       // [16 colgroups][16 rowgroups][128 B]
  uint8_t* sA = smem + MC * ROWS * 2;       // [N/8 colgroups][16 rowgroups][128 B]
  __shared__ uint64_t bar_full, bar_acc;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) {
    mbar_init(&bar_full, 1);
    mbar_init(&bar_acc, 1);
    fence_mbar_init();
  }
  if (warp == 4) tmem_alloc(&tmem_base_s, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;
  if (warp == 4 && (tid & 31) == 0) {
    mbar_arrive_expect_tx(&bar_full, MC * ROWS * 2 + N * ROWS * 2);
    bulk_g2s(sG, gimg, MC * ROWS * 2, &bar_full);
    bulk_g2s(sA, aimg, N * ROWS * 2, &bar_full);
    if (!mbar_wait(&bar_full, 0)) atomicExch(err, 1);
    tc_fence_after();
    const uint32_t idesc = umma_idesc_bf16(128, N) | (1u << 15) | (1u << 16);
    const uint32_t lbo = mode == 0 ? 128u : 2048u, sbo = mode == 0 ? 2048u : 128u;
    for (int ks = 0; ks < ROWS / 16; ++ks) {
      uint64_t ad = umma_smem_desc(smem_u32(sG) + ks * 256, lbo, sbo);
      uint64_t bd = umma_smem_desc(smem_u32(sA) + ks * 256, lbo, sbo);
      umma_ss(tmem, ad, bd, idesc, ks > 0);
    }
    umma_commit(&bar_acc);
  }
  if (warp < 4) {
    if (!mbar_wait(&bar_acc, 0)) atomicExch(err, 2);
    tc_fence_after();
    const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
    for (int c = 0; c < N; c += 16) {
      uint32_t v[16];
      tmem_ld16(lane_addr + c, v);
      tmem_wait_ld();
      for (int i = 0; i < 16; ++i) out[tid * N + c + i] = __uint_as_float(v[i]);
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem, 256);
}

static float bf16_round(float x) { return __bfloat162float(__float2bfloat16(x)); }

template <int N>
static int run(int mode) {
  std::vector<float> G(ROWS * MC), A(ROWS * N), ref(MC * N), got(MC * N);
  srand(4321 + N);
  for (auto& v : G) v = bf16_round((rand() % 2001 - 1000) / 512.f);
  for (auto& v : A) v = bf16_round((rand() % 2001 - 1000) / 1024.f);
  for (int m = 0; m < MC; ++m)
    for (int n = 0; n < N; ++n) {
      double s = 0;
      for (int r = 0; r < ROWS; ++r) s += (double)G[r * MC + m] * A[r * N + n];
      ref[m * N + n] = (float)s;
    }
  auto image = [](const std::vector<float>& X, int cols) {
    std::vector<__nv_bfloat16> img(ROWS * cols);
    for (int r = 0; r < ROWS; ++r)
      for (int c = 0; c < cols; ++c)
        img[((c / 8) * 2048 + (r / 8) * 128 + (r % 8) * 16 + (c % 8) * 2) / 2] = __float2bfloat16(X[r * cols + c]);
    return img;
  };
  auto gi = image(G, MC), ai = image(A, N);
  uint8_t *dG, *dA;
  float* dO;
  int* dErr;
  cudaMalloc(&dG, gi.size() * 2);
  cudaMalloc(&dA, ai.size() * 2);
  cudaMalloc(&dO, got.size() * 4);
  cudaMalloc(&dErr, 4);
  cudaMemset(dErr, 0, 4);
  cudaMemset(dO, 0xFF, got.size() * 4);
  cudaMemcpy(dG, gi.data(), gi.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dA, ai.data(), ai.size() * 2, cudaMemcpyHostToDevice);
  size_t sm = MC * ROWS * 2 + N * ROWS * 2;
  cudaFuncSetAttribute(probe_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
  probe_kernel<N><<<1, 160, sm>>>(dG, dA, dO, mode, dErr);
  cudaError_t e = cudaDeviceSynchronize();
  int herr = 0;
  cudaMemcpy(&herr, dErr, 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(got.data(), dO, got.size() * 4, cudaMemcpyDeviceToHost);
  double mx = 0;
  for (int i = 0; i < MC * N; ++i) {
    double d = fabs((double)got[i] - ref[i]);
    if (!(d == d)) d = 1e30;
    if (d > mx) mx = d;
  }
  const bool pass = e == cudaSuccess && herr == 0 && mx < 1e-2;
  printf("mn_probe N=%d mode=%d (LBO=%s): cuda=%s flag=%d max_abs_err=%.3g %s\n", N, mode, mode == 0 ? "K step 128, SBO = MN step 2048" : "2048, SBO = 128",
         cudaGetErrorString(e), herr, mx, pass ? "PASS" : "FAIL");
  cudaFree(dG); cudaFree(dA); cudaFree(dO); cudaFree(dErr);
  return pass ? 0 : 1;
}

int main() {
  int m0 = run<256>(0) + run<16>(0);
  int m1 = run<256>(1) + run<16>(1);
  printf("umma_mn_probe: documented convention (mode 0) %s, swapped (mode 1) %s\n", m0 == 0 ? "PASS" : "FAIL", m1 == 0 ? "PASS" : "FAIL");
  return m0 == 0 ? 0 : 1;
}
