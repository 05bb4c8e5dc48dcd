// bf16 tensor-core tier (PF_TIER_BF16): tcgen05 sampler.  Placeholder until the kernel lands.
#include "pf_common.cuh"

pf_status pf_tc_pack(pf_flow*, cudaStream_t) { return PF_OK; }
pf_status pf_tc_inverse(pf_flow*, cudaStream_t, const float*, uint64_t, const float*, const int64_t*, int64_t, int64_t,
                        int64_t, float*, float*, uint8_t*, uint32_t, const pf_scaler*, void*, size_t) {
  return PF_ERR_UNSUPPORTED;
}
size_t pf_tc_workspace_bytes(const pf_flow*, int64_t, int64_t) { return 0; }
void pf_tc_destroy(pf_flow*) {}
