// Test-only harness: runs the product's spline device functions (pf_spline.cuh, __host__ __device__)
// on the HOST so that `-m "not gpu"` tests can check them against the oracle without a GPU.
// Built by tests/test_spline_host.py with nvcc into a throw-away shared object.
#include "../posteriflow_b200/csrc/pf_spline.cuh"

namespace pf { thread_local char g_cuda_err[256] = ""; }

struct PtrParams {
  const float* p;
  __host__ __device__ float operator()(int i) const { return p[i]; }
};
struct PtrGrads {
  float* g;
  __host__ __device__ void operator()(int i, float v) const { g[i] = v; }
};

template <int NB>
static void apply_n(int inverse, const float* params, const float* v, float B, int n, float* out, float* lad) {
  for (int i = 0; i < n; ++i) {
    PtrParams P{params + (size_t)i * (3 * NB - 1)};
    out[i] = inverse ? pf::rqs_apply<NB, true>(P, v[i], B, lad[i]) : pf::rqs_apply<NB, false>(P, v[i], B, lad[i]);
  }
}
template <int NB>
static void bwd_n(const float* params, const float* x, float B, const float* gy, const float* gl, int n, float* gp,
                  float* gx) {
  for (int i = 0; i < n; ++i) {
    PtrParams P{params + (size_t)i * (3 * NB - 1)};
    PtrGrads G{gp + (size_t)i * (3 * NB - 1)};
    gx[i] = pf::rqs_forward_backward<NB>(P, x[i], B, gy[i], gl[i], G);
  }
}

extern "C" int h_rqs_apply(int nb, int inverse, const float* params, const float* v, float B, int n, float* out,
                           float* lad) {
  switch (nb) {
    case 8: apply_n<8>(inverse, params, v, B, n, out, lad); return 0;
    case 16: apply_n<16>(inverse, params, v, B, n, out, lad); return 0;
    case 24: apply_n<24>(inverse, params, v, B, n, out, lad); return 0;
    case 32: apply_n<32>(inverse, params, v, B, n, out, lad); return 0;
  }
  return -1;
}
extern "C" int h_rqs_backward(int nb, const float* params, const float* x, float B, const float* gy, const float* gl,
                              int n, float* gp, float* gx) {
  switch (nb) {
    case 8: bwd_n<8>(params, x, B, gy, gl, n, gp, gx); return 0;
    case 16: bwd_n<16>(params, x, B, gy, gl, n, gp, gx); return 0;
    case 24: bwd_n<24>(params, x, B, gy, gl, n, gp, gx); return 0;
    case 32: bwd_n<32>(params, x, B, gy, gl, n, gp, gx); return 0;
  }
  return -1;
}
