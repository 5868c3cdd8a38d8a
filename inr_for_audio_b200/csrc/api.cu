// api.cu -- the extern "C" boundary of libinr_b200.so (include/inr_b200.h) for the SIREN fit path.
// Argument validation + dispatch on inr_siren_desc.precision; no torch types, no C++ exceptions.
#include <stdarg.h>
#include <string.h>

#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"

namespace inr {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// inr_set_sm_limit: the persistent kernels size their grids from num_sms(); a limit below the device's SM count makes
// every launch occupy only that many SMs, so that independent fits enqueued on different streams (clip-parallel mode)
// run side by side on disjoint SMs instead of one after the other with one short tile loop per CTA.
// Per HOST THREAD (thread_local): it scopes the launches the calling thread enqueues between set(n) and set(0), so two
// threads driving two fitters do not see each other's limit.
static thread_local int g_sm_limit = 0;

int num_sms() {
  if (g_sm_limit > 0) return g_sm_limit;
  static int n = 0;
  if (n == 0) {
    int dev = 0, v = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0)
      n = v;
    else
      n = 148;
  }
  return n;
}

// ---- launch accounting / event profiling ---------------------------------------------------------
static std::atomic<int64_t> g_launches{0};
static bool g_profiling = false;
struct ProfRec { const char* tag; cudaEvent_t e0, e1; };
static std::vector<ProfRec> g_prof;
static std::mutex g_prof_mu;

LaunchScope::LaunchScope(const char* t, cudaStream_t s) : tag(t), st(s), e0(nullptr) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (g_profiling) {
    cudaEventCreate(&e0);
    cudaEventRecord(e0, st);
  }
}
LaunchScope::~LaunchScope() {
  if (e0) {
    cudaEvent_t e1;
    cudaEventCreate(&e1);
    cudaEventRecord(e1, st);
    std::lock_guard<std::mutex> lk(g_prof_mu);
    g_prof.push_back({tag, e0, e1});
  }
}

static int check_desc(const inr_siren_desc* d) {
  INR_REQUIRE(d != nullptr, INR_E_BADARG, "desc is NULL");
  INR_REQUIRE(d->in_features >= 1 && d->in_features <= 4096, INR_E_UNSUPPORTED, "in_features must be in [1,4096], got %d",
              d->in_features);
  INR_REQUIRE(d->out_features == 1, INR_E_UNSUPPORTED, "out_features must be 1, got %d", d->out_features);
  INR_REQUIRE(d->hidden_features >= 8 && d->hidden_features % 8 == 0 && d->hidden_features <= 1024,
              INR_E_UNSUPPORTED, "hidden_features must be a multiple of 8 in [8,1024], got %d", d->hidden_features);
  INR_REQUIRE(d->hidden_layers >= 0 && d->hidden_layers <= 16, INR_E_UNSUPPORTED,
              "hidden_layers must be in [0,16], got %d", d->hidden_layers);
  INR_REQUIRE(d->num_snake >= 0 && d->num_tanh >= 0 && d->hidden_layers + d->num_snake + d->num_tanh <= 16,
              INR_E_UNSUPPORTED, "num_snake/num_tanh must be >= 0 and sine+snake+tanh layers <= 16 (got %d+%d+%d)",
              d->hidden_layers, d->num_snake, d->num_tanh);
  INR_REQUIRE(d->first_linear == 0 || d->first_linear == 1, INR_E_BADARG, "first_linear must be 0 or 1");
  INR_REQUIRE(d->precision != INR_PREC_BF16, INR_E_UNSUPPORTED,
              "SIREN has no bf16 mode (bf16 operands floor the fit at ~24 dB): use INR_PREC_FP16 for the tcgen05 path");
  INR_REQUIRE(d->precision == INR_PREC_FP32 || d->precision == INR_PREC_FP16 || d->precision == INR_PREC_FP16S, INR_E_BADARG, "unknown precision %d",
              d->precision);
  if (d->precision == INR_PREC_FP16 || d->precision == INR_PREC_FP16S)
    INR_REQUIRE(siren_tc_supported(d), INR_E_UNSUPPORTED,
                "fp16 tcgen05 path needs in_features <= hidden_features, hidden_features in {128,256,512} and >= 1 hidden layer "
                "(got in=%d H=%d sine=%d snake=%d tanh=%d first_linear=%d); use INR_PREC_FP32",
                d->in_features, d->hidden_features, d->hidden_layers, d->num_snake, d->num_tanh, d->first_linear);
  return INR_OK;
}

__global__ void k_linspace(float* out, int64_t i0, int64_t n, int64_t T, float scale) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x)
    out[j] = linspace_at(i0 + j, T, scale);
}

__global__ void k_grid2d(float* out, int64_t i0, int64_t n, int64_t height, int64_t width) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    const int64_t idx = i0 + j;
    const int64_t r = idx / width, c = idx - r * width;
    out[2 * j] = linspace_at(r, height, 1.f);
    out[2 * j + 1] = linspace_at(c, width, 1.f);
  }
}

}  // namespace inr

using namespace inr;

extern "C" {

const char* inr_last_error(void) { return g_err; }
int inr_version(void) { return INR_ABI_VERSION; }

int64_t inr_launch_count(void) { return g_launches.load(); }

int inr_set_sm_limit(int n_sms) {
  INR_REQUIRE(n_sms >= 0, INR_E_BADARG, "sm limit must be >= 0 (0 = all SMs)");
  int dev = 0, v = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
  g_sm_limit = (n_sms == 0 || n_sms >= v) ? 0 : std::max(8, n_sms / 2 * 2);   // whole CTA pairs; the launchers round further
  return g_sm_limit;
}

int inr_profile_begin(void) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (auto& r : g_prof) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
  g_prof.clear();
  g_profiling = true;
  return INR_OK;
}

int64_t inr_profile_end(char* buf, int64_t cap) {
  g_profiling = false;
  INR_CUDA_CHECK(cudaDeviceSynchronize());
  std::lock_guard<std::mutex> lk(g_prof_mu);
  std::map<std::string, std::pair<int64_t, double>> agg;
  for (auto& r : g_prof) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.e0, r.e1) == cudaSuccess) {
      auto& a = agg[r.tag];
      a.first += 1;
      a.second += ms;
    }
    cudaEventDestroy(r.e0);
    cudaEventDestroy(r.e1);
  }
  g_prof.clear();
  std::string out;
  char line[256];
  for (auto& kv : agg) {
    snprintf(line, sizeof(line), "%s,%lld,%.6f\n", kv.first.c_str(), (long long)kv.second.first, kv.second.second);
    out += line;
  }
  if (buf && cap > 0) {
    const int64_t n = std::min<int64_t>((int64_t)out.size(), cap - 1);
    memcpy(buf, out.data(), n);
    buf[n] = 0;
  }
  return (int64_t)out.size();
}

int64_t inr_siren_num_params(const inr_siren_desc* d) {
  if (!d) return INR_E_BADARG;
  return layout_of(d).total();
}

int64_t inr_siren_workspace_bytes(const inr_siren_desc* d, int64_t n_rows, int training) {
  if (check_desc(d) != INR_OK) return INR_E_BADARG;
  if (n_rows < 0) return INR_E_BADARG;
  if (d->precision == INR_PREC_FP16 || d->precision == INR_PREC_FP16S) return siren_tc_workspace_bytes(d, n_rows, training);
  // fp32 path: whole-problem activations up to a cap; the host code loops over row chunks beyond it
  const int64_t per_row = siren_fp32_workspace_per_row(d, training);
  const int64_t cap = (int64_t)64 << 30;   // of 180 GB
  int64_t rows = std::max<int64_t>(n_rows, 1);
  if (rows * per_row > cap) rows = std::max<int64_t>(cap / per_row, 1024);
  return rows * per_row;
}

int inr_coords_linspace(float* out, int64_t i0, int64_t n, int64_t T_total, float scale, void* stream) {
  INR_REQUIRE(out && n >= 0 && i0 >= 0 && i0 + n <= T_total, INR_E_BADARG, "bad linspace range i0=%lld n=%lld T=%lld",
              (long long)i0, (long long)n, (long long)T_total);
  if (n == 0) return INR_OK;
  int blocks = (int)std::min<int64_t>(cdiv(n, 256), (int64_t)num_sms() * 8);
  { LaunchScope ls("k_linspace", (cudaStream_t)stream);
  k_linspace<<<blocks, 256, 0, (cudaStream_t)stream>>>(out, i0, n, T_total, scale); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

int inr_coords_grid2d(float* out, int64_t i0, int64_t n, int64_t height, int64_t width, void* stream) {
  INR_REQUIRE(out && n >= 0 && i0 >= 0 && height > 0 && width > 0 && i0 + n <= height * width, INR_E_BADARG,
              "bad grid range i0=%lld n=%lld %lldx%lld", (long long)i0, (long long)n, (long long)height,
              (long long)width);
  if (n == 0) return INR_OK;
  int blocks = (int)std::min<int64_t>(cdiv(n, 256), (int64_t)num_sms() * 8);
  { LaunchScope ls("k_grid2d", (cudaStream_t)stream);
  k_grid2d<<<blocks, 256, 0, (cudaStream_t)stream>>>(out, i0, n, height, width); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

int inr_siren_forward(const inr_siren_desc* d, const float* params, const float* coords, int64_t n, int64_t i0,
                      int64_t T_total, float coord_scale, float* out, void* workspace, int64_t workspace_bytes,
                      void* stream) {
  int rc = check_desc(d);
  if (rc) return rc;
  INR_REQUIRE(params && out && n >= 0, INR_E_BADARG, "NULL params/out or negative n");
  INR_REQUIRE(coords || d->in_features == 1, INR_E_BADARG, "coords == NULL needs in_features == 1");
  INR_REQUIRE(coords || (i0 >= 0 && i0 + n <= T_total), INR_E_BADARG, "generated coords: rows [%lld,%lld) outside T=%lld",
              (long long)i0, (long long)(i0 + n), (long long)T_total);
  if (n == 0) return INR_OK;
  INR_REQUIRE(workspace && workspace_bytes > 0, INR_E_WORKSPACE, "workspace missing");
  if (d->precision == INR_PREC_FP16 || d->precision == INR_PREC_FP16S)
    return siren_tc_forward(d, params, coords, n, i0, T_total, coord_scale, out, workspace, workspace_bytes,
                            (cudaStream_t)stream);
  return siren_fp32_forward(d, params, coords, n, i0, T_total, coord_scale, out, workspace, workspace_bytes,
                            (cudaStream_t)stream);
}

int inr_siren_loss_backward(const inr_siren_desc* d, const float* params, const float* coords, const float* target,
                            int64_t n, int64_t i0, int64_t T_total, float coord_scale, int loss_kind, float grad_scale,
                            const float* extra_gout_or_null, float* grads, inr_opt_state* opt, float* out_or_null,
                            void* workspace, int64_t workspace_bytes, void* stream) {
  int rc = check_desc(d);
  if (rc) return rc;
  INR_REQUIRE(params && grads && opt && n >= 0, INR_E_BADARG, "NULL params/grads/opt or negative n");
  INR_REQUIRE(target || extra_gout_or_null, INR_E_BADARG, "need a target and/or an upstream gradient");
  INR_REQUIRE(loss_kind == INR_LOSS_MSE || loss_kind == INR_LOSS_MAE, INR_E_BADARG, "unknown loss kind %d", loss_kind);
  INR_REQUIRE(coords || d->in_features == 1, INR_E_BADARG, "coords == NULL needs in_features == 1");
  INR_REQUIRE(coords || (i0 >= 0 && i0 + n <= T_total), INR_E_BADARG, "generated coords: rows [%lld,%lld) outside T=%lld",
              (long long)i0, (long long)(i0 + n), (long long)T_total);
  if (n == 0) return INR_OK;
  INR_REQUIRE(workspace && workspace_bytes > 0, INR_E_WORKSPACE, "workspace missing");
  if (d->precision == INR_PREC_FP16 || d->precision == INR_PREC_FP16S)
    return siren_tc_loss_backward(d, params, coords, target, n, i0, T_total, coord_scale, loss_kind, grad_scale,
                                  extra_gout_or_null, grads, opt, out_or_null, workspace, workspace_bytes,
                                  (cudaStream_t)stream);
  return siren_fp32_loss_backward(d, params, coords, target, n, i0, T_total, coord_scale, loss_kind, grad_scale,
                                  extra_gout_or_null, grads, opt, out_or_null, workspace, workspace_bytes,
                                  (cudaStream_t)stream);
}

// shared argument checks of the training entry points
static int check_train_args(const inr_siren_desc* d, const float* params, const float* coords, int64_t n, int64_t i0,
                            int64_t T_total, const void* workspace, int64_t workspace_bytes) {
  int rc = check_desc(d);
  if (rc) return rc;
  INR_REQUIRE(params && n >= 0, INR_E_BADARG, "NULL params or negative n");
  INR_REQUIRE(coords || d->in_features == 1, INR_E_BADARG, "coords == NULL needs in_features == 1");
  INR_REQUIRE(coords || (i0 >= 0 && i0 + n <= T_total), INR_E_BADARG, "generated coords: rows [%lld,%lld) outside T=%lld",
              (long long)i0, (long long)(i0 + n), (long long)T_total);
  INR_REQUIRE(n == 0 || (workspace && workspace_bytes > 0), INR_E_WORKSPACE, "workspace missing");
  return INR_OK;
}

int inr_siren_train_forward(const inr_siren_desc* d, const float* params, const float* coords, int64_t n, int64_t i0,
                            int64_t T_total, float coord_scale, float* out, void* workspace, int64_t workspace_bytes,
                            void* stream) {
  int rc = check_train_args(d, params, coords, n, i0, T_total, workspace, workspace_bytes);
  if (rc) return rc;
  INR_REQUIRE(out, INR_E_BADARG, "train_forward: out is NULL");
  if (n == 0) return INR_OK;
  if (d->precision == INR_PREC_FP16 || d->precision == INR_PREC_FP16S)
    return siren_tc_train(d, params, coords, nullptr, n, i0, T_total, coord_scale, 0, 0.f, nullptr, nullptr, nullptr, out,
                          workspace, workspace_bytes, 1, (cudaStream_t)stream);
  return siren_fp32_train(d, params, coords, nullptr, n, i0, T_total, coord_scale, 0, 0.f, nullptr, nullptr, nullptr, out,
                          workspace, workspace_bytes, 1, (cudaStream_t)stream);
}

int inr_siren_train_backward(const inr_siren_desc* d, const float* params, const float* coords, const float* target,
                             int64_t n, int64_t i0, int64_t T_total, float coord_scale, int loss_kind, float grad_scale,
                             const float* extra_gout_or_null, float* grads, inr_opt_state* opt, void* workspace,
                             int64_t workspace_bytes, void* stream) {
  int rc = check_train_args(d, params, coords, n, i0, T_total, workspace, workspace_bytes);
  if (rc) return rc;
  INR_REQUIRE(grads && opt, INR_E_BADARG, "train_backward: NULL grads/opt");
  INR_REQUIRE(target || extra_gout_or_null, INR_E_BADARG, "need a target and/or an upstream gradient");
  INR_REQUIRE(loss_kind == INR_LOSS_MSE || loss_kind == INR_LOSS_MAE, INR_E_BADARG, "unknown loss kind %d", loss_kind);
  if (n == 0) return INR_OK;
  if (d->precision == INR_PREC_FP16 || d->precision == INR_PREC_FP16S)
    return siren_tc_train(d, params, coords, target, n, i0, T_total, coord_scale, loss_kind, grad_scale, extra_gout_or_null,
                          grads, opt, nullptr, workspace, workspace_bytes, 2, (cudaStream_t)stream);
  return siren_fp32_train(d, params, coords, target, n, i0, T_total, coord_scale, loss_kind, grad_scale, extra_gout_or_null,
                          grads, opt, nullptr, workspace, workspace_bytes, 2, (cudaStream_t)stream);
}

int inr_opt_init(inr_opt_state* opt_dev, float lr, float min_lr, int32_t hist_cap, void* stream) {
  INR_REQUIRE(opt_dev && lr > 0.f && hist_cap >= 0, INR_E_BADARG, "bad optimiser init arguments");
  return opt_init(opt_dev, lr, min_lr, hist_cap, (cudaStream_t)stream);
}

int inr_adam_step(float* params, float* grads, float* m, float* v, int64_t P, inr_opt_state* opt, float loss_scale,
                  float* loss_hist, void* stream) {
  INR_REQUIRE(params && grads && m && v && opt && P > 0, INR_E_BADARG, "bad adam arguments");
  return adam_step(params, grads, m, v, P, opt, loss_scale, loss_hist, (cudaStream_t)stream);
}

int inr_siren_fit_step(const inr_siren_desc* d, float* params, float* grads, float* m, float* v, const float* coords,
                       const float* target, int64_t n, int64_t i0, int64_t T_total, float coord_scale, int loss_kind,
                       float alpha, inr_opt_state* opt, float* loss_hist, void* workspace, int64_t workspace_bytes,
                       void* stream) {
  INR_REQUIRE(T_total > 0 && target, INR_E_BADARG, "fit step needs a target and T_total > 0");
  const float w = 1.f - alpha;
  int rc = inr_siren_loss_backward(d, params, coords, target, n, i0, T_total, coord_scale, loss_kind,
                                   w / (float)T_total, nullptr, grads, opt, nullptr, workspace, workspace_bytes, stream);
  if (rc) return rc;
  return inr_adam_step(params, grads, m, v, layout_of(d).total(), opt, w / (float)T_total, loss_hist, stream);
}

}  // extern "C"
