// kan.cu -- efficient-kan KAN fit path (kan.py:6-285), fp32 parity mode on CUDA cores.
//
// A KANLinear (kan.py:153-166) is  y = SiLU(x) W_b^T + vec(B(x)) (W_s * scaler)^T.  With the EXPANDED features
//     Phi[r, i*(nb+1) + 0] = SiLU(x[r,i]),   Phi[r, i*(nb+1) + 1 + k] = B_k(x[r,i])      (nb = grid_size + order = 8)
// and the combined weight  Wc[o, i*(nb+1) + 0] = W_b[o,i],  Wc[o, i*(nb+1) + 1 + k] = W_s[o,i,k] * scaler[o,i]
// it is ONE dense contraction  y = Phi Wc^T  with K = (nb+1) * in, and so are its two gradients
//     dPhi = g Wc,  dWc = g^T Phi,   dx[r,i] = sum_j dPhi[r, i*(nb+1)+j] * phi_j'(x[r,i]).
// The reference materialises ~10 (rows, in, 8..11) intermediates per layer and keeps them for autograd
// (kan.py:93-104); here Phi is produced chunk-wise into ONE scratch buffer by k_kan_expand, recomputed in the backward
// from the saved layer input, overwritten in place by dPhi, and never kept per layer.
//
// b_splines follows kan.py:78-111 operation for operation (half-open order-0 test, the two divisions and products of
// each Cox-de Boor level in the reference's order, no FMA contraction) on the layer's own `grid` buffer.
#include "sgemm.cuh"

namespace inr {

constexpr int KAN_ORDER = 3;
constexpr int KAN_MAX_KNOTS = 24;   // G + 2K + 1 <= 24  (G <= 17)

struct KanLayout {
  int L, nb;           // layers, bases per input (G + K)
  int w[INR_KAN_MAX_LAYERS + 1];
  int kind[INR_KAN_MAX_LAYERS];   // kernel family per layer (KL_*), filled by kan_layout_k
  __host__ __device__ int64_t layer_params(int l) const { return (int64_t)w[l + 1] * w[l] * (nb + 2); }
  __host__ __device__ int64_t base_off(int l) const {
    int64_t o = 0;
    for (int i = 0; i < l; ++i) o += layer_params(i);
    return o;
  }
  __host__ __device__ int64_t spline_off(int l) const { return base_off(l) + (int64_t)w[l + 1] * w[l]; }
  __host__ __device__ int64_t scaler_off(int l) const { return spline_off(l) + (int64_t)w[l + 1] * w[l] * nb; }
  __host__ __device__ int64_t total() const { return base_off(L); }
  __host__ __device__ int64_t grid_off(int l, int knots) const {
    int64_t o = 0;
    for (int i = 0; i < l; ++i) o += (int64_t)w[i] * knots;
    return o;
  }
};

static KanLayout kan_layout(const inr_kan_desc* d) {
  KanLayout k;
  k.L = d->num_layers;
  k.nb = d->grid_size + d->spline_order;
  for (int i = 0; i <= INR_KAN_MAX_LAYERS; ++i) k.w[i] = (i <= d->num_layers) ? d->widths[i] : 0;
  for (int i = 0; i < INR_KAN_MAX_LAYERS; ++i) k.kind[i] = 0;
  return k;
}

// kan.py:93-104: bases (and, if DERIV, d bases / dx as autograd would form them) for one scalar x on knots g[0..nk).
template <bool DERIV>
__device__ __forceinline__ void bsplines_at(float x, const float* __restrict__ g, int nk, float* b, float* db) {
  const int n0 = nk - 1;
  for (int j = 0; j < n0; ++j) {
    b[j] = (x >= g[j] && x < g[j + 1]) ? 1.f : 0.f;
    if (DERIV) db[j] = 0.f;
  }
  for (int k = 1; k <= KAN_ORDER; ++k) {
    for (int j = 0; j < n0 - k; ++j) {
      const float dl = __fsub_rn(g[j + k], g[j]);
      const float dr = __fsub_rn(g[j + k + 1], g[j + 1]);
      const float l = __fdiv_rn(__fsub_rn(x, g[j]), dl);
      const float r = __fdiv_rn(__fsub_rn(g[j + k + 1], x), dr);
      if (DERIV) {
        const float d0 = __fadd_rn(__fdiv_rn(b[j], dl), __fmul_rn(l, db[j]));
        const float d1 = __fsub_rn(__fmul_rn(r, db[j + 1]), __fdiv_rn(b[j + 1], dr));
        db[j] = __fadd_rn(d0, d1);
      }
      b[j] = __fadd_rn(__fmul_rn(l, b[j]), __fmul_rn(r, b[j + 1]));
    }
  }
}

__device__ __forceinline__ float silu_f(float x) { return x / (1.f + expf(-x)); }
__device__ __forceinline__ float silu_grad_f(float x) {
  const float s = 1.f / (1.f + expf(-x));
  return s * (1.f + x * (1.f - s));
}

// kan.py:78-111 as an API of its own (tests, parity of the basis itself): bases (rows, in, nb)
__global__ void __launch_bounds__(256) k_kan_bsplines(const float* __restrict__ x, const float* __restrict__ grid, int64_t rows,
                                                      int in, int nk, int nb, float* __restrict__ out) {
  const int64_t total = rows * in;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(idx % in);
    float g[KAN_MAX_KNOTS], b[KAN_MAX_KNOTS];
    for (int j = 0; j < nk; ++j) g[j] = grid[(int64_t)i * nk + j];
    bsplines_at<false>(x[idx], g, nk, b, nullptr);
    for (int k = 0; k < nb; ++k) out[idx * nb + k] = b[k];
  }
}

// Phi[r, i*(nb+1) + j]; x == NULL (in == 1): the coordinate of global row g0 + r is generated (utils.py:99-109)
__global__ void __launch_bounds__(256) k_kan_expand(const float* __restrict__ x, const float* __restrict__ grid, int64_t rows,
                                                    int in, int nk, int nb, int64_t g0, int64_t T_total, float scale,
                                                    float* __restrict__ phi) {
  const int64_t total = rows * in;
  const int nf = nb + 1;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(idx % in);
    const float xv = x ? x[idx] : linspace_at(g0 + idx, T_total, scale);
    float g[KAN_MAX_KNOTS], b[KAN_MAX_KNOTS];
    for (int j = 0; j < nk; ++j) g[j] = __ldg(grid + (int64_t)i * nk + j);
    bsplines_at<false>(xv, g, nk, b, nullptr);
    float* p = phi + idx * nf;
    p[0] = silu_f(xv);
    for (int k = 0; k < nb; ++k) p[1 + k] = b[k];
  }
}

// dx[r,i] = sum_j dPhi[r, i*(nb+1)+j] * phi_j'(x[r,i])
__global__ void __launch_bounds__(256) k_kan_dx(const float* __restrict__ x, const float* __restrict__ grid,
                                                const float* __restrict__ dphi, int64_t rows, int in, int nk, int nb,
                                                float* __restrict__ dx) {
  const int64_t total = rows * in;
  const int nf = nb + 1;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(idx % in);
    const float xv = x[idx];
    float g[KAN_MAX_KNOTS], b[KAN_MAX_KNOTS], db[KAN_MAX_KNOTS];
    for (int j = 0; j < nk; ++j) g[j] = __ldg(grid + (int64_t)i * nk + j);
    bsplines_at<true>(xv, g, nk, b, db);
    const float* p = dphi + idx * nf;
    float acc = p[0] * silu_grad_f(xv);
    for (int k = 0; k < nb; ++k) acc = __fmaf_rn(p[1 + k], db[k], acc);
    dx[idx] = acc;
  }
}

// Wc[o, i*(nb+1)+j] from (base, spline, scaler)          kan.py:145-151 scaled_spline_weight
__global__ void __launch_bounds__(256) k_kan_pack(const float* __restrict__ base, const float* __restrict__ spline,
                                                  const float* __restrict__ scaler, int64_t n_oi, int nb,
                                                  float* __restrict__ wc) {
  const int nf = nb + 1;
  for (int64_t oi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; oi < n_oi; oi += (int64_t)gridDim.x * blockDim.x) {
    const float sc = scaler[oi];
    wc[oi * nf] = base[oi];
    for (int k = 0; k < nb; ++k) wc[oi * nf + 1 + k] = spline[oi * nb + k] * sc;
  }
}

// gradient of the combined weight -> gradients of (base, spline, scaler); ACCUMULATES, then clears gWc
__global__ void __launch_bounds__(256) k_kan_unpack_grad(float* __restrict__ gwc, const float* __restrict__ spline,
                                                         const float* __restrict__ scaler, int64_t n_oi, int nb,
                                                         float* __restrict__ g_base, float* __restrict__ g_spline,
                                                         float* __restrict__ g_scaler) {
  const int nf = nb + 1;
  for (int64_t oi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; oi < n_oi; oi += (int64_t)gridDim.x * blockDim.x) {
    const float sc = scaler[oi];
    g_base[oi] += gwc[oi * nf];
    float gs = 0.f;
    for (int k = 0; k < nb; ++k) {
      const float gk = gwc[oi * nf + 1 + k];
      g_spline[oi * nb + k] += gk * sc;
      gs = __fmaf_rn(gk, spline[oi * nb + k], gs);
    }
    g_scaler[oi] += gs;
    for (int j = 0; j < nf; ++j) gwc[oi * nf + j] = 0.f;
  }
}

// run.py:124-125,161-169 on the network output: loss sum, g = d loss / d out (+ upstream)
__global__ void __launch_bounds__(256) k_out_loss(const float* __restrict__ out, const float* __restrict__ target,
                                                  const float* __restrict__ extra_g, int64_t n, int loss_kind,
                                                  float grad_scale, float* __restrict__ g, double* __restrict__ loss_acc) {
  double local = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float gv = 0.f;
    if (target) {
      const float diff = out[i] - target[i];
      if (loss_kind == INR_LOSS_MAE) {
        local += (double)fabsf(diff);
        gv = grad_scale * ((diff > 0.f) ? 1.f : ((diff < 0.f) ? -1.f : 0.f));
      } else {
        local += (double)diff * (double)diff;
        gv = grad_scale * 2.f * diff;
      }
    }
    if (extra_g) gv += extra_g[i];
    g[i] = gv;
  }
  local = warp_sum(local);
  if (target && (threadIdx.x & 31) == 0) atomicAdd(loss_acc, local);
}

}  // namespace inr

#include "kan_tc.cuh"

// ================================================================================================ host side
namespace inr {

// which kernel family runs layer l in the bf16 mode
enum { KL_GENERIC = 0, KL_TC = 1, KL_THIN = 2, KL_IN1 = 3 };
static int kan_layer_kind(const inr_kan_desc* d, int l) {
  if (d->precision != INR_PREC_BF16) return KL_GENERIC;
  const int in = d->widths[l], out = d->widths[l + 1];
  if (in % KT_SLAB == 0 && out == KanFwdCfg::OUT) return KL_TC;
  if (out == 1 && in <= 32 * KT_THIN_PER) return KL_THIN;
  if (in == 1 && out % 4 == 0 && out <= 1024) return KL_IN1;
  return KL_GENERIC;
}
static KanLayout kan_layout_k(const inr_kan_desc* d) {
  KanLayout k = kan_layout(d);
  for (int l = 0; l < k.L; ++l) k.kind[l] = kan_layer_kind(d, l);
  return k;
}

static int check_kan_desc(const inr_kan_desc* d) {
  INR_REQUIRE(d != nullptr, INR_E_BADARG, "kan desc is NULL");
  INR_REQUIRE(d->num_layers >= 1 && d->num_layers <= INR_KAN_MAX_LAYERS, INR_E_UNSUPPORTED, "KAN: num_layers %d not in [1,%d]",
              d->num_layers, INR_KAN_MAX_LAYERS);
  INR_REQUIRE(d->spline_order == KAN_ORDER, INR_E_UNSUPPORTED, "KAN: spline_order must be 3, got %d", d->spline_order);
  INR_REQUIRE(d->grid_size >= 1 && d->grid_size + 2 * d->spline_order + 1 <= KAN_MAX_KNOTS, INR_E_UNSUPPORTED,
              "KAN: grid_size %d unsupported", d->grid_size);
  for (int i = 0; i <= d->num_layers; ++i)
    INR_REQUIRE(d->widths[i] >= 1 && d->widths[i] <= 4096, INR_E_UNSUPPORTED, "KAN: layer width %d unsupported", d->widths[i]);
  INR_REQUIRE(d->precision == INR_PREC_FP32 || d->precision == INR_PREC_BF16, INR_E_BADARG, "KAN: unknown precision %d", d->precision);
  if (d->precision == INR_PREC_BF16)
    INR_REQUIRE(d->grid_size + d->spline_order + 1 == KT_NF, INR_E_UNSUPPORTED,
                "KAN bf16 mode is built for grid_size 5 / spline_order 3 (9 expanded features per input), got grid_size %d", d->grid_size);
  return INR_OK;
}

struct KanWs {
  float* wc[INR_KAN_MAX_LAYERS];   // combined weights per layer
  __nv_bfloat16* wf[INR_KAN_MAX_LAYERS];   // bf16 mode: packed forward / backward operands of the tensor-core layers
  __nv_bfloat16* wb[INR_KAN_MAX_LAYERS];
  float* gwc;                      // combined-weight gradient scratch (largest layer), zero between uses
  float* act[INR_KAN_MAX_LAYERS + 1];  // act[l] = input of layer l (act[0] unused: coords), act[L] = network output
  float* phi;                      // rows x max_l (nf * in_l)
  float* g[2];                     // rows x max width, ping-pong upstream gradients
  __nv_bfloat16* gimg;             // bf16 mode: tile image of the upstream gradient of a tensor-core layer
  int64_t chunk_rows;
};

static int64_t kan_fixed_floats(const KanLayout& k) {
  int64_t tot = 0, mx = 0;
  for (int l = 0; l < k.L; ++l) {
    const int64_t n = (int64_t)k.w[l + 1] * k.w[l] * (k.nb + 1);
    tot += round_up(n, 64) * (k.kind[l] == KL_TC ? 2 : 1);   // + WF and WB (2 x bf16 = one float per element)
    mx = std::max(mx, n);
  }
  return tot + round_up(mx, 64);
}
static int64_t kan_row_floats(const KanLayout& k, int training) {
  int64_t wmax = 0, kmax = 0, acts = 0;
  for (int l = 0; l <= k.L; ++l) wmax = std::max<int64_t>(wmax, k.w[l]);
  for (int l = 0; l < k.L; ++l)
    if (k.kind[l] == KL_GENERIC) kmax = std::max<int64_t>(kmax, (int64_t)k.w[l] * (k.nb + 1));   // Phi scratch: fp32 layers only
  for (int l = 1; l <= k.L; ++l) acts += k.w[l];
  int64_t gimg = 0;
  for (int l = 0; l < k.L; ++l)
    if (k.kind[l] == KL_TC) gimg = KanBwdCfg::OUT / 2;          // bf16 image of g: OUT/2 floats per row
  return training ? acts + kmax + 2 * wmax + gimg : 2 * wmax + kmax;
}

static int kan_carve(const KanLayout& k, int training, int64_t n, void* ws, int64_t ws_bytes, KanWs* w) {
  float* base = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(ws) + 255) & ~(uintptr_t)255);
  const int64_t usable = (ws_bytes - (reinterpret_cast<uint8_t*>(base) - reinterpret_cast<uint8_t*>(ws))) / 4;
  const int64_t fixed = kan_fixed_floats(k);
  const int64_t per_row = kan_row_floats(k, training) + 8;
  int64_t rows = (usable - fixed - 1024 - TILE_ROWS * KanBwdCfg::OUT) / per_row;
  INR_REQUIRE(rows >= 1, INR_E_WORKSPACE, "KAN workspace of %lld bytes is too small", (long long)ws_bytes);
  rows = std::min(rows, n);
  w->chunk_rows = rows;
  float* p = base;
  int64_t mx = 0;
  for (int l = 0; l < k.L; ++l) {
    const int64_t nn = (int64_t)k.w[l + 1] * k.w[l] * (k.nb + 1);
    w->wc[l] = p;
    p += round_up(nn, 64);
    w->wf[l] = w->wb[l] = nullptr;
    if (k.kind[l] == KL_TC) {
      w->wf[l] = reinterpret_cast<__nv_bfloat16*>(p);
      w->wb[l] = w->wf[l] + nn;
      p += round_up(nn, 64);
    }
    mx = std::max(mx, nn);
  }
  w->gwc = p;
  p += round_up(mx, 64);
  int64_t wmax = 0, kmax = 0;
  for (int l = 0; l <= k.L; ++l) wmax = std::max<int64_t>(wmax, k.w[l]);
  for (int l = 0; l < k.L; ++l)
    if (k.kind[l] == KL_GENERIC) kmax = std::max<int64_t>(kmax, (int64_t)k.w[l] * (k.nb + 1));
  auto take = [&](int64_t floats) { float* r = p; p += round_up(floats, 64); return r; };
  w->act[0] = nullptr;
  if (training) {
    for (int l = 1; l <= k.L; ++l) w->act[l] = take(rows * k.w[l]);
  } else {
    // inference: outputs ping-pong through the two g buffers
    for (int l = 1; l <= k.L; ++l) w->act[l] = nullptr;
  }
  w->phi = take(rows * kmax);
  w->g[0] = take(rows * wmax);
  w->g[1] = take(rows * wmax);
  w->gimg = nullptr;
  if (training)
    for (int l = 0; l < k.L; ++l)
      if (k.kind[l] == KL_TC && !w->gimg) w->gimg = reinterpret_cast<__nv_bfloat16*>(take(round_up(rows, TILE_ROWS) * (KanBwdCfg::OUT / 2)));
  INR_REQUIRE((p - base) <= usable, INR_E_WORKSPACE, "KAN workspace carve overflow");
  return INR_OK;
}

static inline int blocks_for(int64_t n) { return (int)std::max<int64_t>(1, std::min<int64_t>(cdiv(n, 256), (int64_t)num_sms() * 16)); }

static int kan_pack_all(const KanLayout& k, const float* params, const KanWs& w, cudaStream_t st) {
  for (int l = 0; l < k.L; ++l) {
    const int64_t n_oi = (int64_t)k.w[l + 1] * k.w[l];
    LaunchScope ls("k_kan_pack", st);
    k_kan_pack<<<blocks_for(n_oi), 256, 0, st>>>(params + k.base_off(l), params + k.spline_off(l), params + k.scaler_off(l),
                                                 n_oi, k.nb, w.wc[l]);
  }
  INR_LAUNCH_CHECK();
  for (int l = 0; l < k.L; ++l)
    if (k.kind[l] == KL_TC) {
      LaunchScope ls("k_kan_tc_pack", st);
      k_kan_tc_pack<<<blocks_for((int64_t)k.w[l + 1] * k.w[l] * KT_NF), 256, 0, st>>>(w.wc[l], k.w[l], k.w[l + 1], w.wf[l], w.wb[l]);
    }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

static int kan_tc_attrs() {
  static bool done = false;
  if (!done) {
    INR_CUDA_CHECK(cudaFuncSetAttribute(k_kan_tc_fwd, cudaFuncAttributeMaxDynamicSharedMemorySize, KanFwdCfg::SMEM_BYTES));
    INR_CUDA_CHECK(cudaFuncSetAttribute(k_kan_tc_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, KanBwdCfg::SMEM_BYTES));
    done = true;
  }
  return INR_OK;
}

// y (rows x out) = Phi(x) Wc^T for layer l
static int kan_layer_fwd(const KanLayout& k, int l, const float* x, const float* grids, int nk, int64_t rows, int64_t g0,
                         int64_t T_total, float scale, const KanWs& w, float* y, cudaStream_t st) {
  const int in = k.w[l], out = k.w[l + 1], Kf = in * (k.nb + 1);
  if (k.kind[l] == KL_TC) {
    int rc = kan_tc_attrs();
    if (rc) return rc;
    const int blocks = (int)std::min<int64_t>(cdiv(rows, 2 * TILE_ROWS), num_sms());
    { LaunchScope ls("k_kan_tc_fwd", st);
      k_kan_tc_fwd<<<blocks, KanFwdCfg::THREADS, KanFwdCfg::SMEM_BYTES, st>>>(x, grids + k.grid_off(l, nk), nk, w.wf[l], rows, in, y); }
    INR_LAUNCH_CHECK();
    return INR_OK;
  }
  if (k.kind[l] == KL_IN1) {
    { LaunchScope ls("k_kan_in1_fwd", st);
      k_kan_in1_fwd<<<blocks_for(rows * (out / 4)), 256, out * KT_NF * sizeof(float), st>>>(x, grids + k.grid_off(l, nk), nk, w.wc[l], rows,
                                                                                           out, g0, T_total, scale, y); }
    INR_LAUNCH_CHECK();
    return INR_OK;
  }
  if (k.kind[l] == KL_THIN) {
    { LaunchScope ls("k_kan_thin_fwd", st);
      k_kan_thin_fwd<<<blocks_for(rows * 32), 256, in * KT_NF * sizeof(float), st>>>(x, grids + k.grid_off(l, nk), nk, w.wc[l], rows, in, y); }
    INR_LAUNCH_CHECK();
    return INR_OK;
  }
  { LaunchScope ls("k_kan_expand", st);
    k_kan_expand<<<blocks_for(rows * in), 256, 0, st>>>(x, grids + k.grid_off(l, nk), rows, in, nk, k.nb, g0, T_total, scale, w.phi); }
  INR_LAUNCH_CHECK();
  dim3 grid((unsigned)cdiv(rows, BM), (unsigned)cdiv(out, BN), 1);
  { LaunchScope ls("k_sgemm_kan_fwd", st);
    sgemm_launch<true, true, EPI_STORE>(grid, st, w.phi, Kf, 1, w.wc[l], Kf, 1, rows, out, Kf, Kf, nullptr, 0.f, y, nullptr, out); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

int kan_forward(const inr_kan_desc* d, const float* params, const float* grids, const float* coords, int64_t n, int64_t i0,
                int64_t T_total, float scale, float* out, void* ws, int64_t ws_bytes, cudaStream_t st) {
  const KanLayout k = kan_layout_k(d);
  const int nk = d->grid_size + 2 * d->spline_order + 1;
  KanWs w;
  int rc = kan_carve(k, 0, n, ws, ws_bytes, &w);
  if (rc) return rc;
  if ((rc = kan_pack_all(k, params, w, st))) return rc;
  const int wl = k.w[k.L];
  for (int64_t r0 = 0; r0 < n; r0 += w.chunk_rows) {
    const int64_t rows = std::min(w.chunk_rows, n - r0);
    const float* x = coords ? coords + r0 * k.w[0] : nullptr;
    for (int l = 0; l < k.L; ++l) {
      float* y = (l == k.L - 1) ? out + r0 * wl : w.g[l & 1];
      if ((rc = kan_layer_fwd(k, l, x, grids, nk, rows, i0 + r0, T_total, scale, w, y, st))) return rc;
      x = y;
    }
  }
  return INR_OK;
}

int kan_loss_backward(const inr_kan_desc* d, const float* params, const float* grids, const float* coords, const float* target,
                      int64_t n, int64_t i0, int64_t T_total, float scale, int loss_kind, float grad_scale,
                      const float* extra_gout, float* grads, inr_opt_state* opt, float* out, void* ws, int64_t ws_bytes,
                      cudaStream_t st) {
  const KanLayout k = kan_layout_k(d);
  const int nk = d->grid_size + 2 * d->spline_order + 1;
  const int nf = k.nb + 1;
  KanWs w;
  int rc = kan_carve(k, 1, n, ws, ws_bytes, &w);
  if (rc) return rc;
  if ((rc = kan_pack_all(k, params, w, st))) return rc;
  {
    int64_t mx = 0;
    for (int l = 0; l < k.L; ++l) mx = std::max(mx, (int64_t)k.w[l + 1] * k.w[l] * nf);
    INR_CUDA_CHECK(cudaMemsetAsync(w.gwc, 0, mx * sizeof(float), st));
  }
  for (int64_t r0 = 0; r0 < n; r0 += w.chunk_rows) {
    const int64_t rows = std::min(w.chunk_rows, n - r0);
    const float* x0 = coords ? coords + r0 * k.w[0] : nullptr;
    auto input_of = [&](int l) -> const float* { return l == 0 ? x0 : w.act[l]; };
    for (int l = 0; l < k.L; ++l)
      if ((rc = kan_layer_fwd(k, l, input_of(l), grids, nk, rows, i0 + r0, T_total, scale, w, w.act[l + 1], st))) return rc;
    if (out) INR_CUDA_CHECK(cudaMemcpyAsync(out + r0, w.act[k.L], rows * sizeof(float), cudaMemcpyDeviceToDevice, st));
    { LaunchScope ls("k_out_loss", st);
      k_out_loss<<<blocks_for(rows), 256, 0, st>>>(w.act[k.L], target ? target + r0 : nullptr, extra_gout ? extra_gout + r0 : nullptr,
                                                   rows, loss_kind, grad_scale, w.g[0], &opt->loss_acc[0]); }
    INR_LAUNCH_CHECK();
    int cur = 0;
    for (int l = k.L - 1; l >= 0; --l) {
      const int in = k.w[l], outw = k.w[l + 1], Kf = in * nf;
      const float* x = input_of(l);
      const float* g = w.g[cur];
      if (k.kind[l] != KL_GENERIC) {
        if (k.kind[l] == KL_TC) {
          if ((rc = kan_tc_attrs())) return rc;
          const int nslab = in / KT_SLAB;
          const int Q = (int)std::max<int64_t>(1, std::min<int64_t>(num_sms() / nslab, cdiv(rows, TILE_ROWS)));
          { LaunchScope ls("k_rows_to_image", st);
            k_rows_to_image<<<(unsigned)cdiv(rows, TILE_ROWS), 256, 0, st>>>(g, rows, outw, w.gimg); }
          INR_LAUNCH_CHECK();
          LaunchScope ls("k_kan_tc_bwd", st);
          k_kan_tc_bwd<<<nslab * Q, KanBwdCfg::THREADS, KanBwdCfg::SMEM_BYTES, st>>>(x, grids + k.grid_off(l, nk), nk, w.wb[l], w.gimg, rows,
                                                                                    in, l > 0 ? w.g[cur ^ 1] : nullptr, w.gwc);
        } else if (k.kind[l] == KL_IN1) {
          const int64_t rpb = std::max<int64_t>(64, cdiv(rows, (int64_t)num_sms() * 8));
          LaunchScope ls("k_kan_in1_dw", st);
          k_kan_in1_dw<<<(unsigned)cdiv(rows, rpb), 256, 0, st>>>(x, grids + k.grid_off(l, nk), nk, g, rows, outw, rpb, i0 + r0, T_total,
                                                                 scale, w.gwc);
        } else {
          LaunchScope ls("k_kan_thin_bwd", st);
          k_kan_thin_bwd<<<std::min(blocks_for(rows * 32), num_sms() * 4), 256, 2 * in * KT_NF * sizeof(float), st>>>(
              x, grids + k.grid_off(l, nk), nk, w.wc[l], g, rows, in, w.g[cur ^ 1], w.gwc);
        }
        INR_LAUNCH_CHECK();
        const int64_t n_oi = (int64_t)outw * in;
        { LaunchScope ls("k_kan_unpack_grad", st);
          k_kan_unpack_grad<<<blocks_for(n_oi), 256, 0, st>>>(w.gwc, params + k.spline_off(l), params + k.scaler_off(l), n_oi, k.nb,
                                                              grads + k.base_off(l), grads + k.spline_off(l), grads + k.scaler_off(l)); }
        INR_LAUNCH_CHECK();
        cur ^= 1;
        continue;
      }
      { LaunchScope ls("k_kan_expand", st);
        k_kan_expand<<<blocks_for(rows * in), 256, 0, st>>>(x, grids + k.grid_off(l, nk), rows, in, nk, k.nb, i0 + r0, T_total, scale, w.phi); }
      INR_LAUNCH_CHECK();
      // dWc (out x Kf) += g^T Phi : split over rows
      {
        const int64_t tiles = cdiv(outw, BM) * cdiv(Kf, BN);
        const int64_t want = std::max<int64_t>(1, (int64_t)num_sms() * 2 / tiles);
        const int64_t k_chunk = round_up(std::max<int64_t>(cdiv(rows, want), 256), BK);
        dim3 gdw((unsigned)cdiv(outw, BM), (unsigned)cdiv(Kf, BN), (unsigned)cdiv(rows, k_chunk));
        LaunchScope ls("k_sgemm_kan_dw", st);
        sgemm_launch<false, false, EPI_ATOMIC>(gdw, st, g, 1, outw, w.phi, 1, Kf, outw, Kf, rows, k_chunk, nullptr, 0.f,
                                                               w.gwc, nullptr, Kf);
      }
      INR_LAUNCH_CHECK();
      if (l > 0) {
        // dPhi (rows x Kf) = g Wc, in place over Phi; then dx
        dim3 gdp((unsigned)cdiv(rows, BM), (unsigned)cdiv(Kf, BN), 1);
        { LaunchScope ls("k_sgemm_kan_dphi", st);
          sgemm_launch<true, false, EPI_STORE>(gdp, st, g, outw, 1, w.wc[l], 1, Kf, rows, Kf, outw, outw, nullptr, 0.f,
                                                               w.phi, nullptr, Kf); }
        INR_LAUNCH_CHECK();
        { LaunchScope ls("k_kan_dx", st);
          k_kan_dx<<<blocks_for(rows * in), 256, 0, st>>>(x, grids + k.grid_off(l, nk), w.phi, rows, in, nk, k.nb, w.g[cur ^ 1]); }
        INR_LAUNCH_CHECK();
      }
      {
        const int64_t n_oi = (int64_t)outw * in;
        LaunchScope ls("k_kan_unpack_grad", st);
        k_kan_unpack_grad<<<blocks_for(n_oi), 256, 0, st>>>(w.gwc, params + k.spline_off(l), params + k.scaler_off(l), n_oi, k.nb,
                                                            grads + k.base_off(l), grads + k.spline_off(l), grads + k.scaler_off(l));
      }
      INR_LAUNCH_CHECK();
      cur ^= 1;
    }
  }
  return INR_OK;
}

}  // namespace inr

using namespace inr;

extern "C" {

int64_t inr_kan_num_params(const inr_kan_desc* d) {
  if (check_kan_desc(d) != INR_OK) return INR_E_BADARG;
  return kan_layout_k(d).total();
}

int64_t inr_kan_num_grid_floats(const inr_kan_desc* d) {
  if (check_kan_desc(d) != INR_OK) return INR_E_BADARG;
  return kan_layout(d).grid_off(d->num_layers, d->grid_size + 2 * d->spline_order + 1);
}

int64_t inr_kan_workspace_bytes(const inr_kan_desc* d, int64_t n_rows, int training) {
  if (check_kan_desc(d) != INR_OK || n_rows < 0) return INR_E_BADARG;
  const KanLayout k = kan_layout_k(d);
  const int64_t per_row = (kan_row_floats(k, training) + 8) * 4;
  const int64_t cap = (int64_t)48 << 30;
  int64_t rows = std::max<int64_t>(n_rows, 1);
  if (rows * per_row > cap) rows = std::max<int64_t>(cap / per_row, 1024);
  return (kan_fixed_floats(k) + 2048 + TILE_ROWS * KanBwdCfg::OUT) * 4 + rows * per_row + 4096;
}

int inr_kan_b_splines(const float* x, const float* grid, int64_t rows, int in_features, int grid_size, int spline_order,
                      float* bases, void* stream) {
  INR_REQUIRE(x && grid && bases && rows >= 0 && in_features >= 1, INR_E_BADARG, "b_splines: bad arguments");
  INR_REQUIRE(spline_order == KAN_ORDER && grid_size >= 1 && grid_size + 2 * spline_order + 1 <= KAN_MAX_KNOTS, INR_E_UNSUPPORTED,
              "b_splines: order must be 3 and grid_size <= %d", KAN_MAX_KNOTS - 2 * KAN_ORDER - 1);
  if (rows == 0) return INR_OK;
  const int nk = grid_size + 2 * spline_order + 1;
  { LaunchScope ls("k_kan_bsplines", (cudaStream_t)stream);
    k_kan_bsplines<<<blocks_for(rows * in_features), 256, 0, (cudaStream_t)stream>>>(x, grid, rows, in_features, nk,
                                                                                     grid_size + spline_order, bases); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

int inr_kan_forward(const inr_kan_desc* d, const float* params, const float* grids, const float* coords, int64_t n,
                    int64_t i0, int64_t T_total, float coord_scale, float* out, void* workspace, int64_t workspace_bytes,
                    void* stream) {
  int rc = check_kan_desc(d);
  if (rc) return rc;
  INR_REQUIRE(params && grids && out && n >= 0, INR_E_BADARG, "kan forward: NULL params/grids/out or negative n");
  INR_REQUIRE(coords || d->widths[0] == 1, INR_E_BADARG, "coords == NULL needs layers_hidden[0] == 1");
  INR_REQUIRE(coords || (i0 >= 0 && i0 + n <= T_total), INR_E_BADARG, "generated coords: rows [%lld,%lld) outside T=%lld",
              (long long)i0, (long long)(i0 + n), (long long)T_total);
  if (n == 0) return INR_OK;
  INR_REQUIRE(workspace && workspace_bytes > 0, INR_E_WORKSPACE, "workspace missing");
  return kan_forward(d, params, grids, coords, n, i0, T_total, coord_scale, out, workspace, workspace_bytes, (cudaStream_t)stream);
}

int inr_kan_loss_backward(const inr_kan_desc* d, const float* params, const float* grids, const float* coords,
                          const float* target, int64_t n, int64_t i0, int64_t T_total, float coord_scale, int loss_kind,
                          float grad_scale, const float* extra_gout_or_null, float* grads, inr_opt_state* opt,
                          float* out_or_null, void* workspace, int64_t workspace_bytes, void* stream) {
  int rc = check_kan_desc(d);
  if (rc) return rc;
  INR_REQUIRE(params && grids && grads && opt && n >= 0, INR_E_BADARG, "kan backward: NULL params/grids/grads/opt or negative n");
  INR_REQUIRE(d->widths[d->num_layers] == 1, INR_E_UNSUPPORTED, "kan fit: the last width must be 1 (run.py:110)");
  INR_REQUIRE(target || extra_gout_or_null, INR_E_BADARG, "need a target and/or an upstream gradient");
  INR_REQUIRE(loss_kind == INR_LOSS_MSE || loss_kind == INR_LOSS_MAE, INR_E_BADARG, "unknown loss kind %d", loss_kind);
  INR_REQUIRE(coords || d->widths[0] == 1, INR_E_BADARG, "coords == NULL needs layers_hidden[0] == 1");
  INR_REQUIRE(coords || (i0 >= 0 && i0 + n <= T_total), INR_E_BADARG, "generated coords: rows [%lld,%lld) outside T=%lld",
              (long long)i0, (long long)(i0 + n), (long long)T_total);
  if (n == 0) return INR_OK;
  INR_REQUIRE(workspace && workspace_bytes > 0, INR_E_WORKSPACE, "workspace missing");
  return kan_loss_backward(d, params, grids, coords, target, n, i0, T_total, coord_scale, loss_kind, grad_scale,
                           extra_gout_or_null, grads, opt, out_or_null, workspace, workspace_bytes, (cudaStream_t)stream);
}

}  // extern "C"
