// pf_sample_fused_static: the fused posterior sampler (pf_sample_fused) for callers that cannot allocate -- code under
// CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  It has NO counterpart in the
// reference.
//
// The sampler's only scratch memory is the context projection it computes once per event (proj [E, L*3*KTOT] fp32,
// read by every tile of the event).  Here that scratch is a file-scope __device__ array of PF_STATIC_SCRATCH_BYTES:
// every GPU that loads the library holds it until the process exits, whether or not this entry point is ever called.
//
// In host code `pf_static_scratch` is the host-side stand-in of the symbol, NOT a device address; the device address
// comes from cudaGetSymbolAddress, asked once per device with that device current, and only that pointer is handed to
// the kernels.
//
// Sharing: every call on a device uses the same array and it keeps the last call's contents.  Nothing in it has to
// start at zero -- the projection kernel writes every element the sampler kernel reads (rows [0, E) of the padded
// layout) before the sampler starts, in stream order -- so nothing is cleared.  Two calls that OVERLAP on the GPU
// (different streams, or graph branches without an edge between them) race on the array: serialise them.
#include "pf_common.cuh"

namespace {

constexpr size_t PF_STATIC_SCRATCH_BYTES = (size_t)32 << 20;   // 32 MiB: 1075 events at the LeanNPE shape (31 200 B each)
static __device__ __align__(256) unsigned char pf_static_scratch[PF_STATIC_SCRATCH_BYTES];

constexpr int PF_STATIC_MAX_DEV = 64;
void* g_scratch_ptr[PF_STATIC_MAX_DEV] = {};   // device address of the array, per device (0 = not asked yet)

// Set-up, once per device: the device must be current.  cudaGetSymbolAddress loads this module on the device if it
// has not been loaded yet -- make the FIRST call on a device outside graph capture.
pf_status scratch_address(int device, void** out) {
  if (device < 0 || device >= PF_STATIC_MAX_DEV) return PF_ERR_INVALID;
  if (!g_scratch_ptr[device]) {
    void* p = nullptr;
    PF_CUDA(cudaSetDevice(device));
    PF_CUDA(cudaGetSymbolAddress(&p, pf_static_scratch));
    if (!p) return PF_ERR_CUDA;
    g_scratch_ptr[device] = p;
  }
  *out = g_scratch_ptr[device];
  return PF_OK;
}

}  // namespace

extern "C" pf_status pf_sample_fused_static(pf_flow* f, const float* ctx, const pf_scaler* scaler, const float* z,
                                            uint64_t seed, int64_t E, int64_t n, float* theta, float* logdet,
                                            uint8_t* railed, uint32_t flags, void* stream) {
  using namespace pf;
  if (!f || n < 1 || E < 0) return PF_ERR_INVALID;
  // scratch the existing entry point would ask its caller for.  Tensor-core tier: one projection row per event
  // (pf_tc_workspace_bytes with R = 1 is exactly E rows + padding); FP32 tier: its chunked activation workspace.
  const size_t need = f->desc.tier == PF_TIER_BF16 ? pf_tc_workspace_bytes(f, 1, E)
                                                   : pf_flow_workspace_bytes(f, PF_OP_SAMPLE, E * n, E);
  if (need > PF_STATIC_SCRATCH_BYTES) {
    snprintf(g_cuda_err, sizeof(g_cuda_err),
             "pf_sample_fused_static: %lld events x %lld samples need %zu bytes of scratch, the static array holds %zu "
             "(use pf_sample_fused with a caller-owned workspace, or split the events)",
             (long long)E, (long long)n, need, (size_t)PF_STATIC_SCRATCH_BYTES);
    return PF_ERR_WORKSPACE;
  }
  void* ws = nullptr;
  PF_TRY(scratch_address(f->desc.device, &ws));
  return pf_sample_fused(f, ctx, scaler, z, seed, E, n, theta, logdet, railed, flags, ws, PF_STATIC_SCRATCH_BYTES, stream);
}
