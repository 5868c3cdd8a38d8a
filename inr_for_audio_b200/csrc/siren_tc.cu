// siren_tc.cu -- bf16 tcgen05 SIREN path (placeholder until the tcgen05 kernels land; returns UNSUPPORTED).
#include "common.cuh"
namespace inr {
bool siren_tc_supported(const inr_siren_desc*) { return false; }
int64_t siren_tc_workspace_bytes(const inr_siren_desc*, int64_t, int) { return INR_E_UNSUPPORTED; }
int siren_tc_forward(const inr_siren_desc*, const float*, const float*, int64_t, int64_t, int64_t, float, float*, void*,
                     int64_t, cudaStream_t) { set_error("bf16 path not built"); return INR_E_UNSUPPORTED; }
int siren_tc_loss_backward(const inr_siren_desc*, const float*, const float*, const float*, int64_t, int64_t, int64_t,
                           float, int, float, const float*, float*, inr_opt_state*, float*, void*, int64_t,
                           cudaStream_t) { set_error("bf16 path not built"); return INR_E_UNSUPPORTED; }
}  // namespace inr
