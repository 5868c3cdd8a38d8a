/* inr_b200_debug.h -- test hooks and diagnostics of libinr_b200.so.  NOT part of the product ABI (include/inr_b200.h):
 * these entry points exist for tests/test_tc_gpu.py, tools/tc_probe.py, tools/fused_trace.py and the A/B notes under
 * profiles/; a reference-side integration never calls them. */
#ifndef INR_B200_DEBUG_H
#define INR_B200_DEBUG_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---- test hooks for the tcgen05 kernels (tests/test_tc_gpu.py): one layer GEMM / one dW GEMM on row-major fp32
 * inputs; `variant` bit 0 / bit 1 swap the LBO/SBO descriptor strides of operand A / B (convention probe). */
int64_t inr_dbg_tc_scratch_bytes(int H, int64_t rows);
int inr_dbg_tc_layer(int H, int mode, const float* A, const float* W, const float* bias, float omega,
                     const float* Cin, float* out0, float* out1, int64_t rows, int variant, void* scratch,
                     int64_t scratch_bytes, void* stream);
int inr_dbg_tc_dw(int H, const float* dZ, const float* S, float* gW, int64_t rows, int variant, void* scratch,
                  int64_t scratch_bytes, void* stream);

/* debug: clock64 stamps of the slot hand-offs of the fused forward kernel (pair 0), 4 per job */
int inr_dbg_fused_trace_enable(int on);
int inr_dbg_fused_trace_read(long long* host_out, int n);

#ifdef __cplusplus
}
#endif
#endif
