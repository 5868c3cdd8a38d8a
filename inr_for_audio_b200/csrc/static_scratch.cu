// static_scratch.cu -- inr_siren_forward_static: the network forward for callers that may not allocate device memory.
//
// Who needs this.  inr_siren_forward wants a caller-allocated workspace (the layer activations ping-pong through it).
// Some users of a fitted network cannot supply one:
//   - a real-time audio render callback (a plug-in or playback engine evaluating the fitted INR block by block):
//     cudaMalloc / cudaFree take driver-wide locks and may synchronise the device, so their latency is unbounded and they
//     are off limits on that thread for the same reason malloc is;
//   - a host process whose device memory is owned up front by somebody else's pool (a serving process that reserved the
//     whole HBM at start-up, an embedding application with a fixed budget): a late cudaMalloc there simply fails;
//   - plain-C / scripting-language hosts that bind only this library and have no CUDA runtime binding of their own to
//     allocate with.
// For them the workspace lives in the library image itself: one file-scope __device__ array of INR_STATIC_SCRATCH_BYTES.
// Every GPU on which the library's module is loaded holds that many bytes until the process exits, used or not -- that
// is why the size is modest and fixed.
//
// Rules that follow from the shared array:
//   - the whole call must fit: inr_siren_workspace_bytes(d, n, 0) <= INR_STATIC_SCRATCH_BYTES, else INR_E_WORKSPACE
//     (no silent chunking -- split the rows on the caller's side);
//   - all calls on one device use the same bytes: two calls whose GPU work can overlap (different streams, or different
//     host threads on unordered streams) race on them.  Keep the calls of a device in one stream order;
//   - the array keeps whatever the previous call left.  The forward needs nothing pre-zeroed: the packed weights and every
//     activation image are written in full before they are read, and the stale rows a last partial 128-row tile may see
//     only feed padded output rows that are never stored;
//   - the FIRST call on each device resolves the array's device address (cudaGetSymbolAddress, which also makes the
//     runtime load this module on that device).  Make that first call a warm-up outside the latency-critical section and
//     before any stream capture is begun; later calls on that device only read the cached pointer.
// Nothing on this path calls cudaMalloc, cudaFree or an allocating library routine (the one exception is the opt-in event
// profiler: between inr_profile_begin/inr_profile_end every launch creates two CUDA events).
#include <mutex>

#include "common.cuh"

namespace inr {

// In host code this name is only a handle for the runtime's symbol table -- it is NOT a device address and must never be
// handed to a kernel.  scratch_for_current_device() below asks the runtime for the real one.
static __device__ __align__(256) unsigned char g_static_scratch[INR_STATIC_SCRATCH_BYTES];

static constexpr int kMaxDevices = 64;
static void* g_scratch_addr[kMaxDevices] = {nullptr};   // per device ordinal, filled on first use
static std::mutex g_scratch_mu;

// device address of g_static_scratch on the CURRENT device (each device has its own copy at its own address)
static int scratch_for_current_device(void** out) {
  int dev = -1;
  INR_CUDA_CHECK(cudaGetDevice(&dev));
  INR_REQUIRE(dev >= 0 && dev < kMaxDevices, INR_E_UNSUPPORTED, "device ordinal %d outside [0,%d)", dev, kMaxDevices);
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  if (!g_scratch_addr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      set_error("cudaGetSymbolAddress(static scratch) on device %d: %s", dev, cudaGetErrorString(e));
      return INR_E_CUDA;
    }
    size_t sz = 0;
    e = cudaGetSymbolSize(&sz, g_static_scratch);
    if (e != cudaSuccess || sz != (size_t)INR_STATIC_SCRATCH_BYTES) {
      set_error("static scratch on device %d has %zu bytes, expected %lld (%s)", dev, sz,
                (long long)INR_STATIC_SCRATCH_BYTES, cudaGetErrorString(e));
      return INR_E_CUDA;
    }
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return INR_OK;
}

}  // namespace inr

using namespace inr;

extern "C" int inr_siren_forward_static(const inr_siren_desc* d, const float* params, const float* coords, int64_t n,
                                        int64_t i0, int64_t T_total, float coord_scale, float* out, void* stream) {
  INR_REQUIRE(d != nullptr, INR_E_BADARG, "desc is NULL");
  INR_REQUIRE(n >= 0, INR_E_BADARG, "negative n");
  const int64_t need = inr_siren_workspace_bytes(d, n, 0);   // validates the descriptor as well
  if (need < 0) {
    // descriptor refused: let the ordinary entry point produce its own message and status (it returns before using ws)
    return inr_siren_forward(d, params, coords, n, i0, T_total, coord_scale, out, nullptr, 0, stream);
  }
  INR_REQUIRE(need <= (int64_t)INR_STATIC_SCRATCH_BYTES, INR_E_WORKSPACE,
              "inr_siren_forward_static: %lld rows of this network need %lld scratch bytes, the built-in array holds %lld "
              "(INR_STATIC_SCRATCH_BYTES); pass fewer rows per call or use inr_siren_forward with your own workspace",
              (long long)n, (long long)need, (long long)INR_STATIC_SCRATCH_BYTES);
  void* ws = nullptr;
  int rc = scratch_for_current_device(&ws);
  if (rc) return rc;
  return inr_siren_forward(d, params, coords, n, i0, T_total, coord_scale, out, ws, (int64_t)INR_STATIC_SCRATCH_BYTES,
                           stream);
}
