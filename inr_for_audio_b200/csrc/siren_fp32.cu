// siren_fp32.cu -- the fp32 ("parity mode") SIREN path on CUDA cores.
//
// Same layer-wise structure as the bf16 tcgen05 path (siren_tc.cu) so both share the host logic:
//   fwd  layer l : S_l, C_l = sincos(w * (S_{l-1} W_l^T + b_l))          models.py:114-115
//   bwd  layer l : dZ_{l-1} = (dZ_l W_l) * w_{l-1} * C_{l-1}   (in place over C_{l-1})
//   dW   layer l : gW_l += dZ_l^T S_{l-1}   (split over coordinates, fp32 atomics), gb_l += colsum(dZ_l)
// One generic 128x128x16 register-tiled SGEMM (8x8 per thread, register prefetch) with fused epilogues.
// This path exists for parity (fp32 FMA everywhere, like the reference) and for shapes the tcgen05 path
// does not take; the throughput path is siren_tc.cu.
#include "sgemm.cuh"

namespace inr {

int siren_fp32_workspace_per_row(const inr_siren_desc* d, int training) {
  const SirenLayout lay = layout_of(d);
  const int H = lay.H, L = lay.L;
  // training: Y_0..Y_L (layer outputs) and D_0..D_L (derivative images, overwritten by dZ_l), + one dy/da image per
  // Snake layer;  inference: two ping-pong buffers
  return training ? (2 * (L + 1) + lay.num_snake_layers()) * H * (int)sizeof(float) : 2 * H * (int)sizeof(float);
}

// forward of hidden layer `layer` (1..L): Sout = act(Sin W^T + b), Cout = derivative image, Eout = dy/da (Snake)
static int fwd_layer(const inr_siren_desc* d, const float* params, int layer, const float* Sin, float* Sout,
                     float* Cout, float* Eout, int64_t rows, cudaStream_t st) {
  SirenLayout lay = layout_of(d);
  const int H = d->hidden_features;
  dim3 grid((unsigned)cdiv(rows, BM), (unsigned)cdiv(H, BN), 1);
  const float* W = params + lay.w_off(layer);
  const float* b = params + lay.b_off(layer);
  switch (lay.kind(layer)) {
    case ACT_SINE: {
      LaunchScope ls("k_sgemm_fwd_sincos", st);
      sgemm_launch<true, true, EPI_SINCOS>(grid, st, Sin, H, 1, W, H, 1, rows, H, H, H, b, d->hidden_omega_0, Sout,
                                                            Cout, H);
    } break;
    case ACT_SNAKE: {
      LaunchScope ls("k_sgemm_fwd_snake", st);
      sgemm_launch<true, true, EPI_SNAKE>(grid, st, Sin, H, 1, W, H, 1, rows, H, H, H, b, 1.f, Sout, Cout, H,
                                                           params + lay.a_off(layer), Eout);
    } break;
    default: {
      LaunchScope ls("k_sgemm_fwd_tanh", st);
      sgemm_launch<true, true, EPI_TANH>(grid, st, Sin, H, 1, W, H, 1, rows, H, H, H, b, 1.f, Sout, Cout, H);
    } break;
  }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

// first layer for any in_features: 1 or 2 inputs (raw coordinates) run the elementwise kernel (bit-exact with F.linear at
// in = 1); wider inputs -- encoded coordinates, run.py:81-82,141-144 -- are a GEMM with K = in_features
static int first_layer_any(const inr_siren_desc* d, const float* params, const float* xc, int64_t rows, int64_t g0,
                           int64_t T_total, float scale, float* S0, float* C0, float* E0, cudaStream_t st) {
  if (d->in_features <= 2) return launch_first_layer<float>(d, params, xc, rows, g0, T_total, scale, S0, C0, E0, st);
  const SirenLayout lay = layout_of(d);
  const int H = lay.H, K = lay.in;
  dim3 grid((unsigned)cdiv(rows, BM), (unsigned)cdiv(H, BN), 1);
  if (lay.first_snake) {
    LaunchScope ls("k_sgemm_first_snake", st);
    sgemm_launch<true, true, EPI_SNAKE>(grid, st, xc, K, 1, params + lay.w_off(0), K, 1, rows, H, K, K,
                                                         params + lay.b_off(0), 1.f, S0, C0, H, params + lay.a_off(0), E0);
  } else {
    LaunchScope ls("k_sgemm_first_sincos", st);
    sgemm_launch<true, true, EPI_SINCOS>(grid, st, xc, K, 1, params + lay.w_off(0), K, 1, rows, H, K, K,
                                                          params + lay.b_off(0), d->first_omega_0, S0, C0, H);
  }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

// gW_0 (H x in) += dZ_0^T X, gb_0 += colsum(dZ_0)
static int first_layer_grad_any(const inr_siren_desc* d, const float* dZ0, const float* xc, int64_t rows, int64_t g0,
                                int64_t T_total, float scale, float* gW, float* gb, cudaStream_t st) {
  if (d->in_features <= 2) return launch_first_layer_grad<float>(d, dZ0, xc, rows, g0, T_total, scale, gW, gb, st);
  const int H = d->hidden_features, K = d->in_features;
  const int64_t want = std::max<int64_t>(1, (int64_t)num_sms() * 2 / (cdiv(H, BM) * cdiv(K, BN)));
  const int64_t k_chunk = round_up(std::max<int64_t>(cdiv(rows, want), 256), BK);
  dim3 gdw((unsigned)cdiv(H, BM), (unsigned)cdiv(K, BN), (unsigned)cdiv(rows, k_chunk));
  { LaunchScope ls("k_sgemm_dw0", st);
    sgemm_launch<false, false, EPI_ATOMIC>(gdw, st, dZ0, 1, H, xc, 1, K, H, K, rows, k_chunk, nullptr, 0.f, gW,
                                                           nullptr, K); }
  INR_LAUNCH_CHECK();
  return launch_colsum<float>(dZ0, rows, H, gb, st);
}

int siren_fp32_forward(const inr_siren_desc* d, const float* params, const float* coords, int64_t n, int64_t i0,
                       int64_t T_total, float scale, float* out, void* ws, int64_t ws_bytes, cudaStream_t st) {
  const SirenLayout lay = layout_of(d);
  const int H = lay.H, L = lay.L;
  const int64_t per_row = siren_fp32_workspace_per_row(d, 0);
  int64_t chunk = std::min<int64_t>(n, ws_bytes / per_row);
  INR_REQUIRE(chunk >= 1, INR_E_WORKSPACE, "workspace of %lld bytes holds no row (need %lld/row)", (long long)ws_bytes, (long long)per_row);
  for (int64_t r0 = 0; r0 < n; r0 += chunk) {
    const int64_t rows = std::min(chunk, n - r0);
    float* S[2] = {(float*)ws, (float*)ws + rows * H};
    int rc = first_layer_any(d, params, coords ? coords + r0 * d->in_features : nullptr, rows, i0 + r0, T_total, scale,
                             S[0], nullptr, nullptr, st);
    if (rc) return rc;
    int cur = 0;
    for (int l = 1; l <= L; ++l) {
      rc = fwd_layer(d, params, l, S[cur], S[cur ^ 1], nullptr, nullptr, rows, st);
      if (rc) return rc;
      cur ^= 1;
    }
    rc = launch_last_layer<float>(d, params, S[cur], nullptr, nullptr, nullptr, nullptr, rows, 0, 0.f, nullptr, nullptr,
                                  out + r0, 0, st);
    if (rc) return rc;
  }
  return INR_OK;
}

// phase bits as in siren_tc_train: 1 = forward keeping activations, 2 = backward from them, 3 = fused.
int siren_fp32_train(const inr_siren_desc* d, const float* params, const float* coords, const float* target,
                     int64_t n, int64_t i0, int64_t T_total, float scale, int loss_kind, float grad_scale,
                     const float* extra_gout, float* grads, inr_opt_state* opt, float* out, void* ws,
                     int64_t ws_bytes, int phase, cudaStream_t st) {
  SirenLayout lay = layout_of(d);
  const int H = lay.H, L = lay.L;
  const int64_t per_row = siren_fp32_workspace_per_row(d, 1);
  int64_t chunk = std::min<int64_t>(n, ws_bytes / per_row);
  INR_REQUIRE(chunk >= 1, INR_E_WORKSPACE, "workspace of %lld bytes holds no row (need %lld/row)", (long long)ws_bytes, (long long)per_row);
  INR_REQUIRE(phase == 3 || chunk >= n, INR_E_WORKSPACE,
              "split forward/backward needs the whole row range in one workspace chunk (%lld rows > %lld)", (long long)n,
              (long long)chunk);
  for (int64_t r0 = 0; r0 < n; r0 += chunk) {
    const int64_t rows = std::min(chunk, n - r0);
    float* base = (float*)ws;
    auto S = [&](int l) { return base + (int64_t)l * rows * H; };
    auto C = [&](int l) { return base + (int64_t)(L + 1 + l) * rows * H; };
    // dy/da image of Snake layer l (NULL for the other kinds)
    auto E = [&](int l) -> float* {
      return lay.kind(l) == ACT_SNAKE ? base + (int64_t)(2 * (L + 1) + lay.snakes_before(l)) * rows * H : nullptr;
    };
    const float* xc = coords ? coords + r0 * d->in_features : nullptr;
    int rc;
    if (phase & 1) {
      rc = first_layer_any(d, params, xc, rows, i0 + r0, T_total, scale, S(0), C(0), E(0), st);
      if (rc) return rc;
      for (int l = 1; l <= L; ++l) {
        rc = fwd_layer(d, params, l, S(l - 1), S(l), C(l), E(l), rows, st);
        if (rc) return rc;
      }
      if (phase == 1) {
        rc = launch_last_layer<float>(d, params, S(L), nullptr, nullptr, nullptr, nullptr, rows, 0, 0.f, nullptr, nullptr,
                                      out + r0, 0, st);
        if (rc) return rc;
        continue;
      }
    }
    rc = launch_last_layer<float>(d, params, S(L), C(L), E(L), target ? target + r0 : nullptr,
                                  extra_gout ? extra_gout + r0 : nullptr, rows, loss_kind, grad_scale, grads, opt,
                                  out ? out + r0 : nullptr, 1, st);
    if (rc) return rc;
    for (int l = L; l >= 1; --l) {
      // dW_l += dZ_l^T S_{l-1}: M = H (out), N = H (in), K = rows; both operands are "m/n contiguous"
      const int64_t want = std::max<int64_t>(1, (int64_t)num_sms() * 2 / (cdiv(H, BM) * cdiv(H, BN)));
      int64_t k_chunk = round_up(std::max<int64_t>(cdiv(rows, want), 256), BK);
      dim3 gdw((unsigned)cdiv(H, BM), (unsigned)cdiv(H, BN), (unsigned)cdiv(rows, k_chunk));
      { LaunchScope ls("k_sgemm_dw", st);
      sgemm_launch<false, false, EPI_ATOMIC>(gdw, st, C(l), 1, H, S(l - 1), 1, H, H, H, rows, k_chunk, nullptr,
                                                             0.f, grads + lay.w_off(l), nullptr, H); }
      INR_LAUNCH_CHECK();
      rc = launch_colsum<float>(C(l), rows, H, grads + lay.b_off(l), st);
      if (rc) return rc;
      // dZ_{l-1} = (dZ_l W_l) * (d y_{l-1}/d z_{l-1}): A = dZ_l (k contiguous), B(n,k) = W_l[k][n] (n contiguous).
      // Sine layers keep cos(omega z) in C (times omega here); Snake/Tanh keep the whole derivative; a Snake layer
      // also takes its `a` gradient from the same product (dY = dZ_l W_l) against its dy/da image.
      dim3 gbw((unsigned)cdiv(rows, BM), (unsigned)cdiv(H, BN), 1);
      const int kprev = lay.kind(l - 1);
      if (kprev == ACT_SNAKE) {
        LaunchScope ls("k_sgemm_bwd_snake", st);
        sgemm_launch<true, false, EPI_DMUL_GA>(gbw, st, C(l), H, 1, params + lay.w_off(l), 1, H, rows, H, H, H,
                                                               nullptr, 1.f, C(l - 1), nullptr, H, E(l - 1),
                                                               grads + lay.a_off(l - 1));
      } else {
        const float om = kprev == ACT_TANH ? 1.f : ((l - 1 == 0) ? d->first_omega_0 : d->hidden_omega_0);
        LaunchScope ls("k_sgemm_bwd_cosmul", st);
        sgemm_launch<true, false, EPI_COSMUL>(gbw, st, C(l), H, 1, params + lay.w_off(l), 1, H, rows, H, H, H,
                                                              nullptr, om, C(l - 1), nullptr, H);
      }
      INR_LAUNCH_CHECK();
    }
    rc = first_layer_grad_any(d, C(0), xc, rows, i0 + r0, T_total, scale, grads + lay.w_off(0), grads + lay.b_off(0), st);
    if (rc) return rc;
  }
  return INR_OK;
}

int siren_fp32_loss_backward(const inr_siren_desc* d, const float* params, const float* coords, const float* target,
                             int64_t n, int64_t i0, int64_t T_total, float scale, int loss_kind, float grad_scale,
                             const float* extra_gout, float* grads, inr_opt_state* opt, float* out, void* ws,
                             int64_t ws_bytes, cudaStream_t st) {
  return siren_fp32_train(d, params, coords, target, n, i0, T_total, scale, loss_kind, grad_scale, extra_gout, grads, opt,
                          out, ws, ws_bytes, 3, st);
}

}  // namespace inr
