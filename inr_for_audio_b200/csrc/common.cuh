// common.cuh -- shared helpers for libinr_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <algorithm>

#include "../../include/inr_b200.h"
#include "../../include/inr_b200_debug.h"

namespace inr {

void set_error(const char* fmt, ...);

#define INR_CUDA_CHECK(expr)                                                                 \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess) {                                                                 \
      inr::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));  \
      return INR_E_CUDA;                                                                     \
    }                                                                                        \
  } while (0)

#define INR_LAUNCH_CHECK()                                                                   \
  do {                                                                                       \
    cudaError_t _e = cudaGetLastError();                                                     \
    if (_e != cudaSuccess) {                                                                 \
      inr::set_error("%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(_e)); \
      return INR_E_CUDA;                                                                     \
    }                                                                                        \
  } while (0)

#define INR_REQUIRE(cond, code, ...)                                                         \
  do {                                                                                       \
    if (!(cond)) {                                                                           \
      inr::set_error(__VA_ARGS__);                                                           \
      return (code);                                                                         \
    }                                                                                        \
  } while (0)

// Every kernel launch of the library goes through a LaunchScope: it counts launches (inr_launch_count) and,
// while profiling is on (inr_profile_begin .. inr_profile_end), brackets the launch with CUDA events on the
// launching stream so bench.py can report per-kernel device time without a profiler attached.
struct LaunchScope {
  const char* tag;
  cudaStream_t st;
  cudaEvent_t e0;
  LaunchScope(const char* tag, cudaStream_t st);
  ~LaunchScope();
};

static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int64_t round_up(int64_t a, int64_t b) { return cdiv(a, b) * b; }

int num_sms();

// utils.py:99-109 -> torch.linspace(-scale, scale, T)[i] in fp32, bit-exact with the CPU kernel:
//   step = (end - start) / float(T-1);  i < T/2 ? fma(step, i, start) : fma(-step, T-1-i, end)
__device__ __forceinline__ float linspace_at(int64_t i, int64_t T, float scale) {
  if (T <= 1) return -scale;
  const float start = -scale, end = scale;
  const float step = (end - start) / (float)(T - 1);
  return (i < T / 2) ? __fmaf_rn(step, (float)i, start) : __fmaf_rn(-step, (float)(T - 1 - i), end);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Layout of the flat parameter vector (state_dict order) and the activation kind of every layer.
enum { ACT_SINE = 0, ACT_SNAKE = 1, ACT_TANH = 2 };

struct SirenLayout {
  int in, H, L, out;          // L = total number of HxH layers = ns + nk + nt
  int ns, nk, nt, first_snake;
  __host__ __device__ int kind(int layer) const {        // layer 0..L
    if (layer == 0) return first_snake ? ACT_SNAKE : ACT_SINE;
    if (layer <= ns) return ACT_SINE;
    return (layer <= ns + nk) ? ACT_SNAKE : ACT_TANH;
  }
  // number of Snake layers with index < layer (each owns an extra H-vector `a` after its bias)
  __host__ __device__ int snakes_before(int layer) const {
    if (layer <= 0) return 0;
    int k = layer - 1 - ns;
    k = k < 0 ? 0 : (k > nk ? nk : k);
    return k + (first_snake ? 1 : 0);
  }
  __host__ __device__ int num_snake_layers() const { return nk + (first_snake ? 1 : 0); }
  __host__ __device__ int64_t w_off(int layer) const {  // layer 0..L+1
    if (layer == 0) return 0;
    int64_t o = (int64_t)H * in + H;
    o += (int64_t)(layer - 1) * ((int64_t)H * H + H);
    o += (int64_t)H * snakes_before(layer);
    return o;
  }
  __host__ __device__ int64_t b_off(int layer) const {
    if (layer == 0) return (int64_t)H * in;
    if (layer <= L) return w_off(layer) + (int64_t)H * H;
    return w_off(layer) + (int64_t)out * H;
  }
  __host__ __device__ int64_t a_off(int layer) const { return b_off(layer) + H; }   // Snake layers only
  __host__ __device__ int64_t total() const { return b_off(L + 1) + out; }
};

static inline SirenLayout layout_of(const inr_siren_desc* d) {
  SirenLayout s;
  s.in = d->in_features; s.H = d->hidden_features; s.out = d->out_features;
  s.ns = d->hidden_layers; s.nk = d->num_snake; s.nt = d->num_tanh; s.first_snake = d->first_linear ? 1 : 0;
  s.L = s.ns + s.nk + s.nt;
  return s;
}

// ---- fp32 (CUDA-core) path, siren_fp32.cu
int siren_fp32_workspace_per_row(const inr_siren_desc* d, int training);
int siren_fp32_forward(const inr_siren_desc* d, const float* params, const float* coords, int64_t n, int64_t i0,
                       int64_t T_total, float scale, float* out, void* ws, int64_t ws_bytes, cudaStream_t st);
int siren_fp32_loss_backward(const inr_siren_desc* d, const float* params, const float* coords, const float* target,
                             int64_t n, int64_t i0, int64_t T_total, float scale, int loss_kind, float grad_scale,
                             const float* extra_gout, float* grads, inr_opt_state* opt, float* out, void* ws,
                             int64_t ws_bytes, cudaStream_t st);

int siren_fp32_train(const inr_siren_desc* d, const float* params, const float* coords, const float* target, int64_t n,
                     int64_t i0, int64_t T_total, float scale, int loss_kind, float grad_scale, const float* extra_gout,
                     float* grads, inr_opt_state* opt, float* out, void* ws, int64_t ws_bytes, int phase, cudaStream_t st);

// ---- bf16 tcgen05 path, siren_tc.cu
int siren_tc_train(const inr_siren_desc* d, const float* params, const float* coords, const float* target, int64_t n,
                   int64_t i0, int64_t T_total, float scale, int loss_kind, float grad_scale, const float* extra_gout,
                   float* grads, inr_opt_state* opt, float* out, void* ws, int64_t ws_bytes, int phase, cudaStream_t st);
bool siren_tc_supported(const inr_siren_desc* d);
int64_t siren_tc_workspace_bytes(const inr_siren_desc* d, int64_t n_rows, int training);
int siren_tc_forward(const inr_siren_desc* d, const float* params, const float* coords, int64_t n, int64_t i0,
                     int64_t T_total, float scale, float* out, void* ws, int64_t ws_bytes, cudaStream_t st);
int siren_tc_loss_backward(const inr_siren_desc* d, const float* params, const float* coords, const float* target,
                           int64_t n, int64_t i0, int64_t T_total, float scale, int loss_kind, float grad_scale,
                           const float* extra_gout, float* grads, inr_opt_state* opt, float* out, void* ws,
                           int64_t ws_bytes, cudaStream_t st);

// ---- shared elementwise / reduction kernels (siren_common.cu)
// first layer: S0 = sin(w0*(W x + b)), C0 = cos(..), written as fp32 or bf16
// (first_linear: S0 = snake(W x + b), C0 = d snake/dz, E0 = d snake/da)
template <typename OutT>
int launch_first_layer(const inr_siren_desc* d, const float* params, const float* coords, int64_t rows, int64_t g0,
                       int64_t T_total, float scale, OutT* S0, OutT* C0, OutT* E0, cudaStream_t st);
// last layer + loss + dZ_L (in place in C_L) + last-layer grads
// (EL: d y_L / d a image when layer L is a Snake layer, else NULL; its `a` gradient is accumulated here)
template <typename ActT>
int launch_last_layer(const inr_siren_desc* d, const float* params, const ActT* SL, ActT* CL_dZ, const ActT* EL,
                      const float* target, const float* extra_gout, int64_t rows, int loss_kind, float grad_scale,
                      float* grads, inr_opt_state* opt, float* out, int training, cudaStream_t st);
// column reductions: gW[h*in + i] += sum_r dZ[r,h]*x[r,i] (x NULL -> generated), gb[h] += sum_r dZ[r,h]
template <typename ActT>
int launch_first_layer_grad(const inr_siren_desc* d, const ActT* dZ0, const float* coords, int64_t rows, int64_t g0,
                            int64_t T_total, float scale, float* gW, float* gb, cudaStream_t st);
template <typename ActT>
int launch_colsum(const ActT* dZ, int64_t rows, int H, float* gb, cudaStream_t st);

// ---- optimiser (optim.cu)
int opt_init(inr_opt_state* st, float lr, float min_lr, int hist_cap, cudaStream_t s);
int adam_step(float* params, float* grads, float* m, float* v, int64_t P, inr_opt_state* opt, float loss_scale,
              float* loss_hist, cudaStream_t s);

}  // namespace inr
