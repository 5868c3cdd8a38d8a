// siren_common.cu -- the non-GEMM stages of the SIREN fit step, shared by the fp32 and bf16 paths.
//   first layer   models.py:114-115 (SineLayer is_first)      fp32 CUDA cores, accurate sincosf
//   last layer    models.py:258-262 (nn.Linear) + run.py:124-125,161-169 loss + first backward step
//   column sums   the bias / first-layer weight gradients (reduction over coordinates)
// All are HBM-bound streaming kernels: coalesced 16-byte accesses, grid = multiple of the SM count.
#include "common.cuh"

namespace inr {

template <typename T> struct Vec;
template <> struct Vec<float> {
  static constexpr int N = 4;
  using Raw = float4;
  __device__ static void store(float* p, const float* v) { *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]); }
};
template <> struct Vec<__nv_bfloat16> {
  static constexpr int N = 8;
  using Raw = uint4;
  __device__ static void store(__nv_bfloat16* p, const float* v) {
    __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
    __nv_bfloat162 c = __floats2bfloat162_rn(v[4], v[5]), d = __floats2bfloat162_rn(v[6], v[7]);
    uint4 u;
    u.x = *reinterpret_cast<uint32_t*>(&a); u.y = *reinterpret_cast<uint32_t*>(&b);
    u.z = *reinterpret_cast<uint32_t*>(&c); u.w = *reinterpret_cast<uint32_t*>(&d);
    *reinterpret_cast<uint4*>(p) = u;
  }
};

__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ void from_f(float* p, float v) { *p = v; }
__device__ __forceinline__ void from_f(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

// ------------------------------------------------------------------------------------------------
// first layer
// ------------------------------------------------------------------------------------------------
// Snake (models.py:235-241): y = z + (1/a) sin^2(z a);  dy/dz = 1 + 2 sin(za) cos(za);
// dy/da = (2 z sin(za) cos(za) - sin^2(za)/a) / a.   Evaluated in the reference's operation order.
__device__ __forceinline__ void snake_eval(float z, float a, float* y, float* dz, float* da) {
  float sn, cs;
  sincosf(z * a, &sn, &cs);
  const float ia = 1.0f / a;
  const float s2 = sn * sn;
  const float sc2 = 2.f * sn * cs;
  *y = z + ia * s2;
  *dz = 1.f + sc2;
  *da = ia * (z * sc2 - ia * s2);
}

template <typename OutT>
__global__ void __launch_bounds__(256) k_first_layer(const float* __restrict__ W, const float* __restrict__ B,
                                                     const float* __restrict__ Aparam,
                                                     const float* __restrict__ coords, int in, int H, int64_t rows,
                                                     int64_t g0, int64_t T_total, float scale, float omega,
                                                     OutT* __restrict__ S0, OutT* __restrict__ C0,
                                                     OutT* __restrict__ E0) {
  constexpr int V = Vec<OutT>::N;
  const int groups = H / V;
  const int64_t total = rows * groups;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = idx / groups;
    const int h0 = (int)(idx - r * groups) * V;
    float x0, x1 = 0.f;
    if (coords) {
      x0 = coords[r * in];
      if (in == 2) x1 = coords[r * in + 1];
    } else {
      x0 = linspace_at(g0 + r, T_total, scale);
    }
    float s[V], c[V], e[V];
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const int h = h0 + j;
      float z;
      if (in == 1) {
        z = __fmaf_rn(x0, W[h], B[h]);  // bit-exact with F.linear on CPU for in_features == 1
      } else {
        z = __fmaf_rn(x1, W[h * 2 + 1], x0 * W[h * 2]) + B[h];
      }
      if (Aparam) {                     // first_linear: nn.Linear + Snake (models.py:343-347)
        snake_eval(z, Aparam[h], &s[j], &c[j], &e[j]);
      } else {
        sincosf(omega * z, &s[j], &c[j]);
      }
    }
    Vec<OutT>::store(S0 + r * H + h0, s);
    if (C0) Vec<OutT>::store(C0 + r * H + h0, c);
    if (Aparam && E0) Vec<OutT>::store(E0 + r * H + h0, e);
  }
}

template <typename OutT>
int launch_first_layer(const inr_siren_desc* d, const float* params, const float* coords, int64_t rows, int64_t g0,
                       int64_t T_total, float scale, OutT* S0, OutT* C0, OutT* E0, cudaStream_t st) {
  SirenLayout lay = layout_of(d);
  const int64_t total = rows * (d->hidden_features / Vec<OutT>::N);
  int blocks = (int)std::min<int64_t>(cdiv(total, 256), (int64_t)num_sms() * 16);
  if (blocks < 1) blocks = 1;
  { LaunchScope ls("k_first_layer", st);
  k_first_layer<OutT><<<blocks, 256, 0, st>>>(params + lay.w_off(0), params + lay.b_off(0),
                                              lay.first_snake ? params + lay.a_off(0) : nullptr, coords, d->in_features,
                                              d->hidden_features, rows, g0, T_total, scale, d->first_omega_0, S0, C0, E0); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}
template int launch_first_layer<float>(const inr_siren_desc*, const float*, const float*, int64_t, int64_t, int64_t, float, float*, float*, float*, cudaStream_t);

// ------------------------------------------------------------------------------------------------
// last layer + loss + first backward step.  One warp per row (grid-stride); lane j owns h = j, j+32, ...
// ------------------------------------------------------------------------------------------------
constexpr int kMaxHPerLane = 32;  // H <= 1024

// NPER = per-lane feature slots (H <= 32*NPER), a compile-time bound so that the three per-lane arrays take NPER registers
// each: at H = 256 that is ~50 registers instead of ~110, i.e. five resident blocks per SM instead of two -- this kernel
// lives on memory-level parallelism (one row per warp in flight).
template <typename ActT, bool SNAKE, int NPER>
__global__ void __launch_bounds__(256) k_last_layer(const ActT* __restrict__ S, ActT* __restrict__ CdZ,
                                                    const ActT* __restrict__ E, float* __restrict__ gA,
                                                    const float* __restrict__ w, const float* __restrict__ bptr,
                                                    const float* __restrict__ target, const float* __restrict__ extra_g,
                                                    int64_t rows, int H, float omega_prev, float omega_hidden, int outermost_linear,
                                                    int loss_kind, float grad_scale, float* __restrict__ gW,
                                                    float* __restrict__ gb, double* __restrict__ loss_acc,
                                                    float* __restrict__ out, int training) {
  extern __shared__ float s_gw[];  // H floats + 1 (+ H for the Snake `a` gradient)
  const int lane = threadIdx.x & 31;
  const int warp_in_block = threadIdx.x >> 5;
  const int warps_per_block = blockDim.x >> 5;
  const int nper = (H + 31) / 32;
  if (training) {
    for (int i = threadIdx.x; i <= (SNAKE ? 2 * H : H); i += blockDim.x) s_gw[i] = 0.f;
    __syncthreads();
  }
  float wreg[NPER], gacc[NPER], aacc[SNAKE ? NPER : 1];
#pragma unroll
  for (int j = 0; j < NPER; ++j) {
    const int h = j * 32 + lane;
    wreg[j] = (j < nper && h < H) ? w[h] : 0.f;
    gacc[j] = 0.f;
    if (SNAKE) aacc[j] = 0.f;
  }
  const float bias = bptr[0];
  double loss_local = 0.0;
  float gb_local = 0.f;
  const int64_t warp0 = (int64_t)blockIdx.x * warps_per_block + warp_in_block;
  const int64_t nwarps = (int64_t)gridDim.x * warps_per_block;
  for (int64_t r = warp0; r < rows; r += nwarps) {
    const ActT* srow = S + r * H;
    float sreg[NPER];
    float dot = 0.f;
#pragma unroll
    for (int j = 0; j < NPER; ++j) {
      const int h = j * 32 + lane;
      if (j < nper && h < H) {
        sreg[j] = to_f(srow[h]);
        dot = __fmaf_rn(sreg[j], wreg[j], dot);
      } else {
        sreg[j] = 0.f;
      }
    }
    dot = warp_sum(dot);
    float z = dot + bias;
    float o = z, dact = 1.f;
    if (!outermost_linear) {  // final SineLayer (models.py:263-265)
      float sn, cs;
      sincosf(omega_hidden * z, &sn, &cs);
      o = sn;
      dact = omega_hidden * cs;
    }
    if (out && lane == 0) out[r] = o;
    if (!training) continue;
    float g = 0.f;
    if (target) {
      const float diff = o - target[r];
      if (loss_kind == INR_LOSS_MAE) {
        loss_local += (double)fabsf(diff);
        g = grad_scale * ((diff > 0.f) ? 1.f : ((diff < 0.f) ? -1.f : 0.f));
      } else {
        loss_local += (double)diff * (double)diff;
        g = grad_scale * 2.f * diff;
      }
    }
    if (extra_g) g += extra_g[r];
    g *= dact;
    gb_local += g;
    ActT* crow = CdZ + r * H;
    const float gw = g * omega_prev;
#pragma unroll
    for (int j = 0; j < NPER; ++j) {
      const int h = j * 32 + lane;
      if (j < nper && h < H) {
        gacc[j] = __fmaf_rn(g, sreg[j], gacc[j]);
        const float c = to_f(crow[h]);
        from_f(crow + h, gw * wreg[j] * c);  // dZ_L = g * w[h] * omega * cos(omega z_L), in place over C_L
        if (SNAKE) aacc[j] = __fmaf_rn(g * wreg[j], to_f(E[r * H + h]), aacc[j]);   // d loss/d a_L
      }
    }
  }
  if (!training) return;
#pragma unroll
  for (int j = 0; j < NPER; ++j) {
    const int h = j * 32 + lane;
    if (j < nper && h < H) {
      atomicAdd(&s_gw[h], gacc[j]);
      if (SNAKE) atomicAdd(&s_gw[H + 1 + h], aacc[j]);
    }
  }
  if (lane == 0) {
    atomicAdd(&s_gw[H], gb_local);
    if (target) atomicAdd(loss_acc, loss_local);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < H; i += blockDim.x) {
    atomicAdd(&gW[i], s_gw[i]);
    if (SNAKE) atomicAdd(&gA[i], s_gw[H + 1 + i]);
  }
  if (threadIdx.x == 0) atomicAdd(gb, s_gw[H]);
}

template <typename ActT>
int launch_last_layer(const inr_siren_desc* d, const float* params, const ActT* SL, ActT* CL_dZ, const ActT* EL,
                      const float* target, const float* extra_gout, int64_t rows, int loss_kind, float grad_scale,
                      float* grads, inr_opt_state* opt, float* out, int training, cudaStream_t st) {
  SirenLayout lay = layout_of(d);
  const int H = d->hidden_features;
  INR_REQUIRE(H <= 32 * kMaxHPerLane, INR_E_UNSUPPORTED, "hidden_features %d > %d unsupported", H, 32 * kMaxHPerLane);
  const int last = lay.L + 1;
  const int kindL = lay.kind(lay.L);
  // factor between the stored derivative image and d y_L/d z_L: the sine layers keep cos(omega z) (x omega here),
  // Snake/Tanh layers keep the full derivative
  const float dfac = kindL == ACT_SINE ? (lay.L == 0 ? d->first_omega_0 : d->hidden_omega_0) : 1.f;
  const bool snake = training && kindL == ACT_SNAKE && EL != nullptr;
  int blocks = (int)std::min<int64_t>(cdiv(rows, 8), (int64_t)num_sms() * 4);
  if (blocks < 1) blocks = 1;
  const int nper = (H + 31) / 32;
  blocks = (int)std::min<int64_t>(cdiv(rows, 8), (int64_t)num_sms() * (nper <= 8 ? 10 : 4));
  if (blocks < 1) blocks = 1;
  { LaunchScope ls("k_last_layer", st);
#define INR_LAST_LAUNCH(NP)                                                                                              \
  do {                                                                                                                   \
    if (snake)                                                                                                           \
      k_last_layer<ActT, true, NP><<<blocks, 256, (2 * H + 1) * sizeof(float), st>>>(                                    \
          SL, CL_dZ, EL, grads + lay.a_off(lay.L), params + lay.w_off(last), params + lay.b_off(last), target,           \
          extra_gout, rows, H, dfac, d->hidden_omega_0, d->outermost_linear, loss_kind, grad_scale,                      \
          grads + lay.w_off(last), grads + lay.b_off(last), opt ? &opt->loss_acc[0] : nullptr, out, training);           \
    else                                                                                                                 \
      k_last_layer<ActT, false, NP><<<blocks, 256, (H + 1) * sizeof(float), st>>>(                                       \
          SL, CL_dZ, nullptr, nullptr, params + lay.w_off(last), params + lay.b_off(last), target, extra_gout, rows, H,  \
          dfac, d->hidden_omega_0, d->outermost_linear, loss_kind, grad_scale,                                           \
          grads ? grads + lay.w_off(last) : nullptr, grads ? grads + lay.b_off(last) : nullptr,                          \
          opt ? &opt->loss_acc[0] : nullptr, out, training);                                                             \
  } while (0)
    if (nper <= 2) INR_LAST_LAUNCH(2);
    else if (nper <= 4) INR_LAST_LAUNCH(4);
    else if (nper <= 8) INR_LAST_LAUNCH(8);
    else if (nper <= 16) INR_LAST_LAUNCH(16);
    else INR_LAST_LAUNCH(32);
#undef INR_LAST_LAUNCH
  }
  INR_LAUNCH_CHECK();
  return INR_OK;
}
template int launch_last_layer<float>(const inr_siren_desc*, const float*, const float*, float*, const float*, const float*, const float*, int64_t, int, float, float*, inr_opt_state*, float*, int, cudaStream_t);

// ------------------------------------------------------------------------------------------------
// column reductions over coordinates: gb[h] += sum_r dZ[r,h];  gW[h*in+i] += sum_r dZ[r,h] * x[r,i]
// block = 128 threads, thread owns 2 adjacent columns; grid = (col blocks of 256, row chunks)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float2 load2(const float* p) { return *reinterpret_cast<const float2*>(p); }
__device__ __forceinline__ float2 load2(const __nv_bfloat16* p) {
  return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
}

template <typename ActT, bool WITH_X>
__global__ void __launch_bounds__(128) k_colsum(const ActT* __restrict__ dZ, const float* __restrict__ coords, int in,
                                                int64_t rows, int H, int64_t rows_per_block, int64_t g0,
                                                int64_t T_total, float scale, float* __restrict__ gW,
                                                float* __restrict__ gb) {
  const int c = (blockIdx.x * 128 + threadIdx.x) * 2;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_block;
  const int64_t r1 = min(rows, r0 + rows_per_block);
  if (c >= H) return;
  float sb0 = 0.f, sb1 = 0.f, sx0a = 0.f, sx0b = 0.f, sx1a = 0.f, sx1b = 0.f;
#pragma unroll 4
  for (int64_t r = r0; r < r1; ++r) {
    const float2 v = load2(dZ + r * H + c);
    sb0 += v.x; sb1 += v.y;
    if (WITH_X) {
      float x0, x1 = 0.f;
      if (coords) {
        x0 = coords[r * in];
        if (in == 2) x1 = coords[r * in + 1];
      } else {
        x0 = linspace_at(g0 + r, T_total, scale);
      }
      sx0a = __fmaf_rn(v.x, x0, sx0a); sx0b = __fmaf_rn(v.y, x0, sx0b);
      if (in == 2) { sx1a = __fmaf_rn(v.x, x1, sx1a); sx1b = __fmaf_rn(v.y, x1, sx1b); }
    }
  }
  atomicAdd(&gb[c], sb0); atomicAdd(&gb[c + 1], sb1);
  if (WITH_X) {
    atomicAdd(&gW[(int64_t)c * in], sx0a); atomicAdd(&gW[(int64_t)(c + 1) * in], sx0b);
    if (in == 2) { atomicAdd(&gW[(int64_t)c * in + 1], sx1a); atomicAdd(&gW[(int64_t)(c + 1) * in + 1], sx1b); }
  }
}

template <typename ActT, bool WITH_X>
static int launch_colsum_impl(const ActT* dZ, const float* coords, int in, int64_t rows, int H, int64_t g0,
                              int64_t T_total, float scale, float* gW, float* gb, cudaStream_t st) {
  const int col_blocks = (int)cdiv(H, 256);
  int64_t want_row_blocks = std::max<int64_t>(1, (int64_t)num_sms() * 16 / col_blocks);
  int64_t rpb = std::max<int64_t>(64, cdiv(rows, want_row_blocks));
  dim3 grid(col_blocks, (unsigned)cdiv(rows, rpb));
  { LaunchScope ls(WITH_X ? "k_colsum_x" : "k_colsum", st);
  k_colsum<ActT, WITH_X><<<grid, 128, 0, st>>>(dZ, coords, in, rows, H, rpb, g0, T_total, scale, gW, gb); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

template <typename ActT>
int launch_first_layer_grad(const inr_siren_desc* d, const ActT* dZ0, const float* coords, int64_t rows, int64_t g0,
                            int64_t T_total, float scale, float* gW, float* gb, cudaStream_t st) {
  return launch_colsum_impl<ActT, true>(dZ0, coords, d->in_features, rows, d->hidden_features, g0, T_total, scale, gW, gb, st);
}
template <typename ActT>
int launch_colsum(const ActT* dZ, int64_t rows, int H, float* gb, cudaStream_t st) {
  return launch_colsum_impl<ActT, false>(dZ, nullptr, 1, rows, H, 0, 1, 1.f, nullptr, gb, st);
}
template int launch_first_layer_grad<float>(const inr_siren_desc*, const float*, const float*, int64_t, int64_t, int64_t, float, float*, float*, cudaStream_t);
template int launch_colsum<float>(const float*, int64_t, int, float*, cudaStream_t);

}  // namespace inr
