// siren_fp32.cu -- the fp32 ("parity mode") SIREN path on CUDA cores.
//
// Same layer-wise structure as the bf16 tcgen05 path (siren_tc.cu) so both share the host logic:
//   fwd  layer l : S_l, C_l = sincos(w * (S_{l-1} W_l^T + b_l))          models.py:114-115
//   bwd  layer l : dZ_{l-1} = (dZ_l W_l) * w_{l-1} * C_{l-1}   (in place over C_{l-1})
//   dW   layer l : gW_l += dZ_l^T S_{l-1}   (split over coordinates, fp32 atomics), gb_l += colsum(dZ_l)
// One generic 128x128x16 register-tiled SGEMM (8x8 per thread, register prefetch) with fused epilogues.
// This path exists for parity (fp32 FMA everywhere, like the reference) and for shapes the tcgen05 path
// does not take; the throughput path is siren_tc.cu.
#include "common.cuh"

namespace inr {

enum { EPI_SINCOS = 0, EPI_COSMUL = 1, EPI_ATOMIC = 2 };

constexpr int BM = 128, BN = 128, BK = 16;

template <bool A_KC, bool B_KC, int EPI>
__global__ void __launch_bounds__(256)
k_sgemm(const float* __restrict__ A, int64_t sAm, int64_t sAk, const float* __restrict__ B, int64_t sBn, int64_t sBk,
        int64_t M, int N, int64_t K, int64_t k_chunk, const float* __restrict__ bias, float omega,
        float* out0, float* out1, int64_t ldo) {
  __shared__ __align__(16) float As[BK][BM + 4];
  __shared__ __align__(16) float Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int64_t kbeg = (int64_t)blockIdx.z * k_chunk;
  const int64_t kend = min(K, kbeg + k_chunk);
  const int ty = tid / 16, tx = tid % 16;

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  float ra[8], rb[8];
  auto gload = [&](int64_t k0) {
#pragma unroll
    for (int p = 0; p < 8; ++p) {
      const int e = tid + p * 256;
      int mm, kk;
      if (A_KC) { mm = e / BK; kk = e % BK; } else { kk = e / BM; mm = e % BM; }
      const int64_t gm = m0 + mm, gk = k0 + kk;
      ra[p] = (gm < M && gk < kend) ? A[gm * sAm + gk * sAk] : 0.f;
      int nn;
      if (B_KC) { nn = e / BK; kk = e % BK; } else { kk = e / BN; nn = e % BN; }
      const int gn = n0 + nn;
      const int64_t gk2 = k0 + kk;
      rb[p] = (gn < N && gk2 < kend) ? B[(int64_t)gn * sBn + gk2 * sBk] : 0.f;
    }
  };
  auto sstore = [&]() {
#pragma unroll
    for (int p = 0; p < 8; ++p) {
      const int e = tid + p * 256;
      int mm, kk;
      if (A_KC) { mm = e / BK; kk = e % BK; } else { kk = e / BM; mm = e % BM; }
      As[kk][mm] = ra[p];
      int nn;
      if (B_KC) { nn = e / BK; kk = e % BK; } else { kk = e / BN; nn = e % BN; }
      Bs[kk][nn] = rb[p];
    }
  };

  gload(kbeg);
  for (int64_t k0 = kbeg; k0 < kend; k0 += BK) {
    sstore();
    __syncthreads();
    if (k0 + BK < kend) gload(k0 + BK);
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[kk][64 + ty * 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[kk][64 + tx * 4]);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = __fmaf_rn(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t gm = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (gm >= M) continue;
#pragma unroll
    for (int jh = 0; jh < 2; ++jh) {
      const int gn = n0 + (jh == 0 ? tx * 4 : 64 + tx * 4);
      if (gn >= N) continue;
      const bool full = (gn + 3 < N) && ((ldo & 3) == 0);
      float v[4] = {acc[i][jh * 4 + 0], acc[i][jh * 4 + 1], acc[i][jh * 4 + 2], acc[i][jh * 4 + 3]};
      float* p0 = out0 + gm * ldo + gn;
      if (EPI == EPI_SINCOS) {
        float s[4], c[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float z = v[q] + ((gn + q < N) ? bias[gn + q] : 0.f);
          sincosf(omega * z, &s[q], &c[q]);
        }
        float* p1 = out1 ? out1 + gm * ldo + gn : nullptr;
        if (full) {
          *reinterpret_cast<float4*>(p0) = make_float4(s[0], s[1], s[2], s[3]);
          if (p1) *reinterpret_cast<float4*>(p1) = make_float4(c[0], c[1], c[2], c[3]);
        } else {
          for (int q = 0; q < 4 && gn + q < N; ++q) { p0[q] = s[q]; if (p1) p1[q] = c[q]; }
        }
      } else if (EPI == EPI_COSMUL) {
        if (full) {
          float4 c = *reinterpret_cast<float4*>(p0);
          c.x = v[0] * omega * c.x; c.y = v[1] * omega * c.y; c.z = v[2] * omega * c.z; c.w = v[3] * omega * c.w;
          *reinterpret_cast<float4*>(p0) = c;
        } else {
          for (int q = 0; q < 4 && gn + q < N; ++q) p0[q] = v[q] * omega * p0[q];
        }
      } else {
        for (int q = 0; q < 4 && gn + q < N; ++q) atomicAdd(p0 + q, v[q]);
      }
    }
  }
}

int siren_fp32_workspace_per_row(const inr_siren_desc* d, int training) {
  const int H = d->hidden_features, L = d->hidden_layers;
  // training: S_0..S_L and C_0..C_L (C_l is overwritten by dZ_l);  inference: two ping-pong S buffers
  return training ? 2 * (L + 1) * H * (int)sizeof(float) : 2 * H * (int)sizeof(float);
}

static int fwd_layer(const inr_siren_desc* d, const float* params, int layer, const float* Sin, float* Sout,
                     float* Cout, int64_t rows, cudaStream_t st) {
  SirenLayout lay = layout_of(d);
  const int H = d->hidden_features;
  dim3 grid((unsigned)cdiv(rows, BM), (unsigned)cdiv(H, BN), 1);
  { LaunchScope ls("k_sgemm_fwd_sincos", st);
  k_sgemm<true, true, EPI_SINCOS><<<grid, 256, 0, st>>>(Sin, H, 1, params + lay.w_off(layer), H, 1, rows, H, H, H,
                                                        params + lay.b_off(layer), d->hidden_omega_0, Sout, Cout, H); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

int siren_fp32_forward(const inr_siren_desc* d, const float* params, const float* coords, int64_t n, int64_t i0,
                       int64_t T_total, float scale, float* out, void* ws, int64_t ws_bytes, cudaStream_t st) {
  const int H = d->hidden_features, L = d->hidden_layers;
  const int64_t per_row = siren_fp32_workspace_per_row(d, 0);
  int64_t chunk = std::min<int64_t>(n, ws_bytes / per_row);
  INR_REQUIRE(chunk >= 1, INR_E_WORKSPACE, "workspace of %lld bytes holds no row (need %lld/row)", (long long)ws_bytes, (long long)per_row);
  for (int64_t r0 = 0; r0 < n; r0 += chunk) {
    const int64_t rows = std::min(chunk, n - r0);
    float* S[2] = {(float*)ws, (float*)ws + rows * H};
    int rc = launch_first_layer<float>(d, params, coords ? coords + r0 * d->in_features : nullptr, rows, i0 + r0,
                                       T_total, scale, S[0], nullptr, st);
    if (rc) return rc;
    int cur = 0;
    for (int l = 1; l <= L; ++l) {
      rc = fwd_layer(d, params, l, S[cur], S[cur ^ 1], nullptr, rows, st);
      if (rc) return rc;
      cur ^= 1;
    }
    rc = launch_last_layer<float>(d, params, S[cur], nullptr, nullptr, nullptr, rows, 0, 0.f, nullptr, nullptr,
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
  const int H = d->hidden_features, L = d->hidden_layers;
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
    const float* xc = coords ? coords + r0 * d->in_features : nullptr;
    int rc;
    if (phase & 1) {
      rc = launch_first_layer<float>(d, params, xc, rows, i0 + r0, T_total, scale, S(0), C(0), st);
      if (rc) return rc;
      for (int l = 1; l <= L; ++l) {
        rc = fwd_layer(d, params, l, S(l - 1), S(l), C(l), rows, st);
        if (rc) return rc;
      }
      if (phase == 1) {
        rc = launch_last_layer<float>(d, params, S(L), nullptr, nullptr, nullptr, rows, 0, 0.f, nullptr, nullptr,
                                      out + r0, 0, st);
        if (rc) return rc;
        continue;
      }
    }
    rc = launch_last_layer<float>(d, params, S(L), C(L), target ? target + r0 : nullptr,
                                  extra_gout ? extra_gout + r0 : nullptr, rows, loss_kind, grad_scale, grads, opt,
                                  out ? out + r0 : nullptr, 1, st);
    if (rc) return rc;
    for (int l = L; l >= 1; --l) {
      // dW_l += dZ_l^T S_{l-1}: M = H (out), N = H (in), K = rows; both operands are "m/n contiguous"
      const int64_t want = std::max<int64_t>(1, (int64_t)num_sms() * 2 / (cdiv(H, BM) * cdiv(H, BN)));
      int64_t k_chunk = round_up(std::max<int64_t>(cdiv(rows, want), 256), BK);
      dim3 gdw((unsigned)cdiv(H, BM), (unsigned)cdiv(H, BN), (unsigned)cdiv(rows, k_chunk));
      { LaunchScope ls("k_sgemm_dw", st);
      k_sgemm<false, false, EPI_ATOMIC><<<gdw, 256, 0, st>>>(C(l), 1, H, S(l - 1), 1, H, H, H, rows, k_chunk, nullptr,
                                                             0.f, grads + lay.w_off(l), nullptr, H); }
      INR_LAUNCH_CHECK();
      rc = launch_colsum<float>(C(l), rows, H, grads + lay.b_off(l), st);
      if (rc) return rc;
      // dZ_{l-1} = (dZ_l W_l) * omega_{l-1} * C_{l-1}: A = dZ_l (k contiguous), B(n,k) = W_l[k][n] (n contiguous)
      const float om = (l - 1 == 0) ? d->first_omega_0 : d->hidden_omega_0;
      dim3 gbw((unsigned)cdiv(rows, BM), (unsigned)cdiv(H, BN), 1);
      { LaunchScope ls("k_sgemm_bwd_cosmul", st);
      k_sgemm<true, false, EPI_COSMUL><<<gbw, 256, 0, st>>>(C(l), H, 1, params + lay.w_off(l), 1, H, rows, H, H, H,
                                                            nullptr, om, C(l - 1), nullptr, H); }
      INR_LAUNCH_CHECK();
    }
    rc = launch_first_layer_grad<float>(d, C(0), xc, rows, i0 + r0, T_total, scale, grads + lay.w_off(0),
                                        grads + lay.b_off(0), st);
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
