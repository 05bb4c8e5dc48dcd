// Hardware probe for the conventions the tensor-core tier relies on (run on a B200 through gpurun):
//   - K-major no-swizzle UMMA shared-memory descriptors (LBO = K-direction, SBO = 8-row-group stride)
//   - cp.async.bulk of a pre-laid-out weight blob + mbarrier complete_tx
//   - tcgen05.mma SS form (A in shared memory) and TS form (A in tensor memory, bf16 pairs per column)
//   - tcgen05.ld / tcgen05.st 32x32b row-per-thread access, tcgen05.commit -> mbarrier
// Prints one line per mode with the max abs error against a host reference; exit code 0 iff all pass.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o umma_probe tests/umma_probe.cu
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "../posteriflow_b200/csrc/pf_sm100.cuh"
using namespace pf::sm100;

constexpr int M = 128;

// mode 0: SS.  mode 1: TS, element k even in the low half of the column.  mode 2: TS, even k in the high half.
// mode 3 / 4: as 1 with the A operand at column base 65 / 66 written in 8+4+1-column pieces (unaligned bases).
template <int N, int K>
__global__ void __launch_bounds__(160) probe_kernel(const float* __restrict__ A, const __nv_bfloat16* __restrict__ Bblob,
                                                    float* __restrict__ out, int mode, int* err) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;                     // [K/8][128 rows][16 B]
  uint8_t* sB = smem + (K / 8) * M * 16;  // [K/8][N rows][16 B]
  __shared__ uint64_t bar_full, bar_acc;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) {
    mbar_init(&bar_full, 1);
    mbar_init(&bar_acc, 1);
    fence_mbar_init();
  }
  if (warp == 4) tmem_alloc(&tmem_base_s, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;
  const uint32_t accD = tmem;        // columns [0, N)
  const uint32_t accA = tmem + (mode == 3 ? 65 : (mode == 4 ? 66 : 64));   // columns [base, base + K/2)
  if (warp < 4) {
    const int r = tid;
    if (mode == 0) {
      for (int k = 0; k < K; k += 2) {
        uint32_t w = pack_bf16(A[r * K + k], A[r * K + k + 1]);
        *reinterpret_cast<uint32_t*>(sA + (k / 8) * (M * 16) + (r / 8) * 128 + (r % 8) * 16 + (k % 8) * 2) = w;
      }
      fence_proxy_async_smem();
    } else {
      const uint32_t lane_addr = accA + ((uint32_t)(warp * 32) << 16);
      for (int c = 0; c < K / 2; c += 8) {
        uint32_t v[8];
        for (int i = 0; i < 8; ++i) {
          float a0 = A[r * K + 2 * (c + i)], a1 = A[r * K + 2 * (c + i) + 1];
          v[i] = mode != 2 ? pack_bf16(a0, a1) : pack_bf16(a1, a0);
        }
        if (mode >= 3) {   // 8 columns as 4 + 1 + 2 + 1 (exercise the narrow shapes at odd bases)
          tmem_st4(lane_addr + c, v);
          tmem_st1(lane_addr + c + 4, v + 4);
          tmem_st2(lane_addr + c + 5, v + 5);
          tmem_st1(lane_addr + c + 7, v + 7);
        } else {
          tmem_st8(lane_addr + c, v);
        }
      }
      tmem_wait_st();
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 4 && (tid & 31) == 0) {
    tc_fence_after();
    mbar_arrive_expect_tx(&bar_full, N * K * 2);
    bulk_g2s(sB, Bblob, N * K * 2, &bar_full);
    if (!mbar_wait(&bar_full, 0)) atomicExch(err, 1);
    tc_fence_after();
    const uint32_t idesc = umma_idesc_bf16(M, N);
    for (int i = 0; i < K / 16; ++i) {
      uint64_t bd = umma_smem_desc(smem_u32(sB) + i * 2 * (N * 16), N * 16, 128);
      if (mode == 0) {
        uint64_t ad = umma_smem_desc(smem_u32(sA) + i * 2 * (M * 16), M * 16, 128);
        umma_ss(accD, ad, bd, idesc, i > 0);
      } else {
        umma_ts(accD, accA + i * 8, bd, idesc, i > 0);
      }
    }
    umma_commit(&bar_acc);
  }
  if (warp < 4) {
    if (!mbar_wait(&bar_acc, 0)) atomicExch(err, 2);
    tc_fence_after();
    const uint32_t lane_addr = accD + ((uint32_t)(warp * 32) << 16);
    for (int c = 0; c < N; c += 16) {
      uint32_t v[16];
      tmem_ld16(lane_addr + c, v);
      tmem_wait_ld();
      for (int i = 0; i < 16; ++i) out[tid * N + c + i] = __uint_as_float(v[i]);
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem, 128);
}

static float bf16_round(float x) { return __bfloat162float(__float2bfloat16(x)); }

template <int N, int K>
static int run(int mode) {
  std::vector<float> A(M * K), B(N * K), ref(M * N), got(M * N);
  srand(1234 + mode + N + K);
  for (auto& v : A) v = bf16_round((rand() % 2001 - 1000) / 512.f);
  for (auto& v : B) v = bf16_round((rand() % 2001 - 1000) / 1024.f);
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) {
      double s = 0;
      for (int k = 0; k < K; ++k) s += (double)A[m * K + k] * B[n * K + k];
      ref[m * N + n] = (float)s;
    }
  std::vector<__nv_bfloat16> blob(N * K);
  for (int n = 0; n < N; ++n)
    for (int k = 0; k < K; ++k) blob[(k / 8) * (N * 8) + (n / 8) * 64 + (n % 8) * 8 + (k % 8)] = __float2bfloat16(B[n * K + k]);
  float *dA, *dO;
  __nv_bfloat16* dB;
  int* dErr;
  cudaMalloc(&dA, A.size() * 4);
  cudaMalloc(&dO, got.size() * 4);
  cudaMalloc(&dB, blob.size() * 2);
  cudaMalloc(&dErr, 4);
  cudaMemset(dErr, 0, 4);
  cudaMemset(dO, 0xFF, got.size() * 4);
  cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, blob.data(), blob.size() * 2, cudaMemcpyHostToDevice);
  size_t sm = (K / 8) * M * 16 + (K / 8) * N * 16;
  probe_kernel<N, K><<<1, 160, sm>>>(dA, dB, dO, mode, dErr);
  cudaError_t e = cudaDeviceSynchronize();
  int herr = 0;
  cudaMemcpy(&herr, dErr, 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(got.data(), dO, got.size() * 4, cudaMemcpyDeviceToHost);
  double mx = 0;
  for (int i = 0; i < M * N; ++i) {
    double d = fabs((double)got[i] - ref[i]);
    if (!(d == d)) d = 1e30;
    if (d > mx) mx = d;
  }
  printf("probe N=%d K=%d mode=%d (%s): cuda=%s flag=%d max_abs_err=%.3g %s\n", N, K, mode,
         mode == 0 ? "SS" : (mode == 2 ? "TS even-k-high" : (mode == 1 ? "TS even-k-low" : "TS unaligned base")), cudaGetErrorString(e), herr, mx,
         (e == cudaSuccess && herr == 0 && mx < 1e-3) ? "PASS" : "FAIL");
  cudaFree(dA); cudaFree(dO); cudaFree(dB); cudaFree(dErr);
  return (e == cudaSuccess && herr == 0 && mx < 1e-3) ? 0 : 1;
}

int main() {
  int bad = 0;
  bad += run<48, 48>(0);
  bad += run<32, 32>(0);
  int ts1 = run<48, 48>(1);
  int ts2 = run<48, 48>(2);
  bad += (ts1 && ts2);
  bad += run<32, 64>(ts1 == 0 ? 1 : 2);
  bad += run<48, 48>(3);
  bad += run<32, 64>(4);
  printf("umma_probe: %s (TS packing: %s)\n", bad ? "FAIL" : "PASS", ts1 == 0 ? "even-k-low" : (ts2 == 0 ? "even-k-high" : "none"));
  return bad ? 1 : 0;
}
