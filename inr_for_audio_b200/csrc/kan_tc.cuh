// kan_tc.cuh -- bf16 tensor-core mode of the KAN fit path (INR_PREC_BF16), included by kan.cu.
//
// The B-spline basis evaluation and its spline-weight contraction are ONE kernel per direction: the expanded features
//   Phi[r, (i, j)]  (j = 0: SiLU(x[r,i]); j = 1..8: B_{j-1}(x[r,i]))                                  kan.py:153-166
// are generated on-chip by producer warps straight into the UMMA-canonical bf16 operand layout in shared memory and are
// never written to HBM (the reference materialises ~10 (rows, in, 8..11) fp32 tensors per layer, kan.py:93-104).
//
//   k_kan_tc_fwd   y (rows x 256) = Phi(x) Wc^T.  One CTA = 256 rows (two 128-row accumulators filling TMEM), the
//                  packed weight is STREAMED from L2 through a bulk-copy ring (1.18 MB per layer does not fit in shared
//                  memory; two row tiles per weight stage keep the L2->SM stream at ~43 GB/s per SM).
//   k_kan_tc_bwd   weight-stationary over a 16-input slab (144 expanded features): per 128-row tile it converts the
//                  upstream gradient g to a bf16 operand image, regenerates the slab's Phi, and issues
//                     dWc[:, slab] += g^T Phi        (both operands MN-major; accumulator resident in TMEM for the kernel)
//                     dPhi = g Wc[:, slab]           (K-major) -> epilogue contracts with phi_j'(x) -> dx[:, slab inputs]
//   k_kan_thin_*   the 256 -> 1 output layer on CUDA cores (a GEMV: no tensor-core shape), warp per row.
// Layers that do not have the tensor shape (in % 16 != 0 or out != 256, e.g. the 1 -> 256 input layer) run the fp32
// kernels of kan.cu; activations between layers stay fp32 row-major, so the two kinds mix freely.
//
// This mode uses the closed form of the cubic B-spline on a UNIFORM knot vector (the reference's construction,
// kan.py:26-35): it reads t0 and h from the first row of the layer's `grid` and requires all rows to be that same
// uniform vector (checked on the host side by KanFitter; grids refined by update_grid need the fp32 mode).
#pragma once
#include "tc_common.cuh"

namespace inr {
using namespace tc;

constexpr int KT_NF = 9;          // expanded features per input (SiLU + 8 bases): grid_size 5, order 3
constexpr int KT_SLAB = 16;       // inputs per slab
constexpr int KT_SLABF = KT_SLAB * KT_NF;   // 144 expanded features per slab

struct UGrid { float t0, inv_h; int nk; };

__device__ __forceinline__ UGrid load_ugrid(const float* __restrict__ grid, int nk) {
  UGrid u;
  u.t0 = __ldg(grid);
  u.inv_h = (float)(nk - 1) / (__ldg(grid + nk - 1) - u.t0);
  u.nk = nk;
  return u;
}

// cubic B-spline on the uniform knots t_j = t0 + j h: interval m = floor((x - t0)/h); the four bases that are non-zero
// on it are b = m-3 .. m with values w[0..3] (and d/dx dw[0..3]); every basis is zero outside [t_0, t_{nk-1}).
template <bool DERIV>
__device__ __forceinline__ int basis4(float x, const UGrid& ug, float* w, float* dw) {
  const float s = (x - ug.t0) * ug.inv_h;
  const float mf = floorf(s);
  const float u = s - mf;
  const float u2 = u * u, u3 = u2 * u;
  const float omu = 1.f - u;
  const bool in = (s >= 0.f) && (mf <= (float)(ug.nk - 2));
  const float k = in ? (1.f / 6.f) : 0.f;
  w[3] = k * u3;
  w[2] = k * (-3.f * u3 + 3.f * u2 + 3.f * u + 1.f);
  w[1] = k * (3.f * u3 - 6.f * u2 + 4.f);
  w[0] = k * omu * omu * omu;
  if (DERIV) {
    const float kd = in ? 0.5f * ug.inv_h : 0.f;
    dw[3] = kd * u2;
    dw[2] = kd * (-3.f * u2 + 2.f * u + 1.f);
    dw[1] = kd * (3.f * u2 - 4.f * u);
    dw[0] = -kd * omu * omu;
  }
  return in ? (int)mf : -100;
}
// value of expanded feature j (0 = SiLU, 1.. = basis j-1) given the interval and its four weights
__device__ __forceinline__ float feat_of(int j, float silu, int m, const float* w) {
  if (j == 0) return silu;
  const int q = (j - 1) - (m - 3);
  float v = 0.f;
#pragma unroll
  for (int t = 0; t < 4; ++t) v = (q == t) ? w[t] : v;
  return v;
}

// ------------------------------------------------------------------------------------------------ weight packing
// Wc (fp32, [o][i*9 + j], kan.cu k_kan_pack)  ->
//   WF: forward B operand, stage (slab, jg) of 48 K:  [stage][kg = jj*2 + hf][n = o (OUT)][e]  = Wc[o][i = slab*16+hf*8+e][j = 3jg+jj]
//   WB: backward B operand per slab (K-major, K = o):  [slab][kg = o/8][n = j*16 + ii (144)][e = o%8] = Wc[o][i = slab*16+ii][j]
__global__ void __launch_bounds__(256) k_kan_tc_pack(const float* __restrict__ wc, int in, int out, __nv_bfloat16* __restrict__ WF,
                                                     __nv_bfloat16* __restrict__ WB) {
  const int64_t total = (int64_t)out * in * KT_NF;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int o = (int)(idx / (in * KT_NF));
    const int rem = (int)(idx - (int64_t)o * in * KT_NF);
    const int i = rem / KT_NF, j = rem - i * KT_NF;
    const __nv_bfloat16 v = __float2bfloat16_rn(wc[idx]);
    const int slab = i / KT_SLAB, ii = i % KT_SLAB, hf = ii >> 3, e = ii & 7;
    const int jg = j / 3, jj = j - 3 * jg;
    WF[((((int64_t)(slab * 3 + jg) * 6 + (jj * 2 + hf)) * out + o) << 3) + e] = v;
    WB[(((int64_t)slab * (out >> 3) + (o >> 3)) * KT_SLABF + (j * KT_SLAB + ii)) * 8 + (o & 7)] = v;
  }
}

// ------------------------------------------------------------------------------------------------ thin output layer
// y[r] = sum_i sum_j phi_j(x[r,i]) wc[i*9+j]      (out_features == 1): warp per row, lane owns inputs lane, lane+32, ...
constexpr int KT_THIN_PER = 8;   // inputs per lane: in <= 256
__global__ void __launch_bounds__(256) k_kan_thin_fwd(const float* __restrict__ x, const float* __restrict__ grid, int nk,
                                                      const float* __restrict__ wc, int64_t rows, int in, float* __restrict__ y) {
  extern __shared__ float s_wc[];
  for (int i = threadIdx.x; i < in * KT_NF; i += blockDim.x) s_wc[i] = wc[i];
  __syncthreads();
  const UGrid ug = load_ugrid(grid, nk);
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp0; r < rows; r += nwarps) {
    float acc = 0.f;
    for (int i = lane; i < in; i += 32) {
      const float xv = x[r * in + i];
      float w[4];
      const int m = basis4<false>(xv, ug, w, nullptr);
      const float* wp = s_wc + i * KT_NF;
      acc = __fmaf_rn(xv / (1.f + __expf(-xv)), wp[0], acc);
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int b = m - 3 + t;
        if (b >= 0 && b < KT_NF - 1) acc = __fmaf_rn(w[t], wp[1 + b], acc);
      }
    }
    acc = warp_sum(acc);
    if (lane == 0) y[r] = acc;
  }
}

// backward of the thin layer: dx[r,i] = g[r] * sum_j phi_j'(x) wc[i*9+j];  gwc[i*9+j] += sum_r g[r] phi_j(x[r,i]).
// Only the 5 non-zero features of an input (SiLU + the 4 bases of its knot interval) are touched: their weights are
// read from, and their gradient sums accumulated into, shared memory by index (lanes own different inputs, so the
// stride-9 addresses of a warp fall into different banks); one global atomic per weight and block at the end.
__global__ void __launch_bounds__(256) k_kan_thin_bwd(const float* __restrict__ x, const float* __restrict__ grid, int nk,
                                                      const float* __restrict__ wc, const float* __restrict__ g, int64_t rows,
                                                      int in, float* __restrict__ dx, float* __restrict__ gwc) {
  extern __shared__ float s_mem[];
  float* s_wc = s_mem;                 // in*9
  float* s_gw = s_mem + in * KT_NF;    // in*9 block-level partial sums
  for (int i = threadIdx.x; i < in * KT_NF; i += blockDim.x) { s_wc[i] = wc[i]; s_gw[i] = 0.f; }
  __syncthreads();
  const UGrid ug = load_ugrid(grid, nk);
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp0; r < rows; r += nwarps) {
    const float gv = g[r];
    for (int i = lane; i < in; i += 32) {
      const float xv = x[r * in + i];
      float w[4], dw[4];
      const int m = basis4<true>(xv, ug, w, dw);
      const float* wp = s_wc + i * KT_NF;
      float* gp = s_gw + i * KT_NF;
      const float sg = 1.f / (1.f + __expf(-xv));
      float d = sg * (1.f + xv * (1.f - sg)) * wp[0];
      atomicAdd(gp, gv * xv * sg);
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int b = m - 3 + t;
        if (b >= 0 && b < KT_NF - 1) {
          d = __fmaf_rn(dw[t], wp[1 + b], d);
          atomicAdd(gp + 1 + b, gv * w[t]);
        }
      }
      dx[r * in + i] = gv * d;
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < in * KT_NF; i += blockDim.x) atomicAdd(&gwc[i], s_gw[i]);
}

// ------------------------------------------------------------------------------------------------ thin input layer
// all 9 expanded features of one input value (closed form on the uniform grid)
__device__ __forceinline__ void feats9(float xv, const UGrid& ug, float* f) {
  float w[4];
  const int m = basis4<false>(xv, ug, w, nullptr);
  f[0] = xv / (1.f + __expf(-xv));
#pragma unroll
  for (int b = 0; b < KT_NF - 1; ++b) {
    const int q = b - (m - 3);
    float v = 0.f;
#pragma unroll
    for (int t = 0; t < 4; ++t) v = (q == t) ? w[t] : v;
    f[1 + b] = v;
  }
}

// in_features == 1 (the coordinate layer): y[r,o] = sum_j phi_j(x_r) wc[o*9+j];  gwc[o*9+j] += sum_r g[r,o] phi_j(x_r).
// x == NULL: the coordinate of global row g0 + r is generated (utils.py:99-109).
__global__ void __launch_bounds__(256) k_kan_in1_fwd(const float* __restrict__ x, const float* __restrict__ grid, int nk,
                                                     const float* __restrict__ wc, int64_t rows, int out, int64_t g0,
                                                     int64_t T_total, float scale, float* __restrict__ y) {
  extern __shared__ float s_wc1[];
  for (int i = threadIdx.x; i < out * KT_NF; i += blockDim.x) s_wc1[i] = wc[i];
  __syncthreads();
  const UGrid ug = load_ugrid(grid, nk);
  const int og = out >> 2;                                       // out % 4 == 0
  const int64_t total = rows * og;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = idx / og;
    const int o0 = (int)(idx - r * og) * 4;
    const float xv = x ? x[r] : linspace_at(g0 + r, T_total, scale);
    float f[KT_NF];
    feats9(xv, ug, f);
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int j = 0; j < KT_NF; ++j) acc[u] = __fmaf_rn(f[j], s_wc1[(o0 + u) * KT_NF + j], acc[u]);
    *reinterpret_cast<float4*>(y + r * out + o0) = make_float4(acc[0], acc[1], acc[2], acc[3]);
  }
}

__global__ void __launch_bounds__(256) k_kan_in1_dw(const float* __restrict__ x, const float* __restrict__ grid, int nk,
                                                    const float* __restrict__ g, int64_t rows, int out, int64_t rows_per_block,
                                                    int64_t g0, int64_t T_total, float scale, float* __restrict__ gwc) {
  // the features of a row are shared by every output: the block evaluates them once for 256 rows into shared memory,
  // then thread o sweeps those rows with coalesced reads of g[r, o] and broadcast reads of the features
  __shared__ float s_f[256][KT_NF + 1];
  const UGrid ug = load_ugrid(grid, nk);
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = min(rows, r0 + rows_per_block);
  float acc[4][KT_NF];                                           // out <= 1024: up to 4 outputs per thread
#pragma unroll
  for (int u = 0; u < 4; ++u)
#pragma unroll
    for (int j = 0; j < KT_NF; ++j) acc[u][j] = 0.f;
  for (int64_t rb = r0; rb < r1; rb += 256) {
    const int nr = (r1 - rb) < 256 ? (int)(r1 - rb) : 256;
    __syncthreads();
    if ((int)threadIdx.x < nr) {
      const int64_t r = rb + threadIdx.x;
      const float xv = x ? __ldg(x + r) : linspace_at(g0 + r, T_total, scale);
      float f[KT_NF];
      feats9(xv, ug, f);
#pragma unroll
      for (int j = 0; j < KT_NF; ++j) s_f[threadIdx.x][j] = f[j];
    }
    __syncthreads();
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int o = threadIdx.x + u * 256;
      if (o < out) {
        for (int rr = 0; rr < nr; ++rr) {
          const float gv = g[(rb + rr) * out + o];
#pragma unroll
          for (int j = 0; j < KT_NF; ++j) acc[u][j] = __fmaf_rn(gv, s_f[rr][j], acc[u][j]);
        }
      }
    }
  }
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int o = threadIdx.x + u * 256;
    if (o < out)
#pragma unroll
      for (int j = 0; j < KT_NF; ++j) atomicAdd(&gwc[o * KT_NF + j], acc[u][j]);
  }
}

// ------------------------------------------------------------------------------------------------ forward GEMM
struct KanFwdCfg {
  static constexpr int OUT = 256;
  static constexpr int KG = 6;                                  // K groups (8 K each) per stage: 3 features x 16 inputs
  static constexpr int A_TILE_BYTES = KG * TILE_ROWS * 16;      // 12 KB
  static constexpr int A_BYTES = 2 * A_TILE_BYTES;              // two row tiles per stage
  static constexpr int B_BYTES = KG * OUT * 16;                 // 24 KB
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;         // 48 KB
  static constexpr int STAGES = 4;
  static constexpr int OFF_BAR = STAGES * STAGE_BYTES;
  static constexpr int SMEM_BYTES = OFF_BAR + 128;
  static constexpr int PROD_WARPS = 16;
  static constexpr int THREADS = 32 * (2 + PROD_WARPS + 4);     // B producer, MMA, 16 Phi producers, 4 epilogue
};

// x (rows x in) fp32 row-major, WF packed by k_kan_tc_pack, y (rows x 256) fp32 row-major.  in % 16 == 0.
__global__ void __launch_bounds__(KanFwdCfg::THREADS, 1) k_kan_tc_fwd(const float* __restrict__ x, const float* __restrict__ grid,
                                                                       int nk, const __nv_bfloat16* __restrict__ WF, int64_t rows,
                                                                       int in, float* __restrict__ y) {
  using Cfg = KanFwdCfg;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t bar0 = smem_base + Cfg::OFF_BAR;
  auto bar_full = [&](int s) { return bar0 + 8u * s; };
  auto bar_empty = [&](int s) { return bar0 + 8u * (Cfg::STAGES + s); };
  const uint32_t bar_tfull = bar0 + 8u * (2 * Cfg::STAGES);
  const uint32_t bar_tempty = bar0 + 8u * (2 * Cfg::STAGES + 1);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + Cfg::OFF_BAR + 8 * (2 * Cfg::STAGES + 2));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nslab = in / KT_SLAB;
  const int nstage = nslab * 3;
  const int64_t num_dt = (rows + 2 * TILE_ROWS - 1) / (2 * TILE_ROWS);

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::STAGES; ++s) { mbar_init(bar_full(s), 1 + Cfg::PROD_WARPS); mbar_init(bar_empty(s), 1); }
    mbar_init(bar_tfull, 1);
    mbar_init(bar_tempty, 4);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== weight producer: the packed stage (24 KB, contiguous) from L2 through the ring ==========
    if (lane == 0) {
      int s = 0;
      uint32_t phase = 0;
      for (int64_t dt = blockIdx.x; dt < num_dt; dt += gridDim.x) {
        for (int st = 0; st < nstage; ++st) {
          mbar_wait(bar_empty(s), phase ^ 1);
          mbar_arrive_expect_tx(bar_full(s), Cfg::B_BYTES);
          bulk_g2s(smem_base + s * Cfg::STAGE_BYTES + Cfg::A_BYTES, reinterpret_cast<const uint8_t*>(WF) + (size_t)st * Cfg::B_BYTES,
                   Cfg::B_BYTES, bar_full(s));
          if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, Cfg::OUT, 0, 0);
      int s = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int64_t dt = blockIdx.x; dt < num_dt; dt += gridDim.x, ++it) {
        mbar_wait(bar_tempty, (it & 1) ^ 1);                    // epilogue drained the previous double tile
        tc_fence_after();
        for (int st = 0; st < nstage; ++st) {
          mbar_wait(bar_full(s), phase);
          tc_fence_after();
          const uint32_t a0 = smem_base + s * Cfg::STAGE_BYTES;
          const uint32_t b0 = a0 + Cfg::A_BYTES;
#pragma unroll
          for (int t = 0; t < 2; ++t)
#pragma unroll
            for (int jj = 0; jj < 3; ++jj) {
              const uint64_t ad = make_smem_desc(a0 + t * Cfg::A_TILE_BYTES + jj * 2 * (TILE_ROWS * 16), TILE_ROWS * 16, 128);
              const uint64_t bd = make_smem_desc(b0 + jj * 2 * (Cfg::OUT * 16), Cfg::OUT * 16, 128);
              umma_f16(tmem_base + t * Cfg::OUT, ad, bd, idesc, (st | jj) != 0 ? 1u : 0u);
            }
          umma_commit(bar_empty(s));
          if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
        }
        umma_commit(bar_tfull);
      }
    }
  } else if (warp < 2 + Cfg::PROD_WARPS) {
    // ===================== Phi producers: thread = (row of the double tile, 8 inputs of the slab) ==================
    // All 9 expanded features of the thread's 8 inputs are evaluated ONCE per slab and kept packed in registers
    // (9 x 16 bytes); the slab's three ring stages (feature triples jg) then only wait for their slot and store.
    const UGrid ug = load_ugrid(grid, nk);
    const int p = threadIdx.x - 64;                             // 0..511
    const int hf = p >> 8, rr = p & 255;
    const int t = rr >> 7, r = rr & 127;
    int s = 0;
    uint32_t phase = 0;
    for (int64_t dt = blockIdx.x; dt < num_dt; dt += gridDim.x) {
      int64_t grow = dt * (2 * TILE_ROWS) + rr;
      if (grow >= rows) grow = rows - 1;                        // pad rows replicate the last row; never stored
      const float* xrow = x + grow * in + hf * 8;
      float4 f0 = *reinterpret_cast<const float4*>(xrow), f1 = *reinterpret_cast<const float4*>(xrow + 4);
      for (int slab = 0; slab < nslab; ++slab) {
        const float xv[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
        if (slab + 1 < nslab) {                                 // next slab's inputs: in flight under this slab's math
          f0 = *reinterpret_cast<const float4*>(xrow + (slab + 1) * KT_SLAB);
          f1 = *reinterpret_cast<const float4*>(xrow + (slab + 1) * KT_SLAB + 4);
        }
        uint32_t o[KT_NF][4];
#pragma unroll
        for (int e = 0; e < 8; e += 2) {
          float f[2][KT_NF];
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const float xx = xv[e + u];
            float w[4];
            const int m = basis4<false>(xx, ug, w, nullptr);
            f[u][0] = xx / (1.f + __expf(-xx));
#pragma unroll
            for (int b = 0; b < KT_NF - 1; ++b) {
              const int q = b - (m - 3);
              float v = 0.f;
#pragma unroll
              for (int tt = 0; tt < 4; ++tt) v = (q == tt) ? w[tt] : v;
              f[u][1 + b] = v;
            }
          }
#pragma unroll
          for (int j = 0; j < KT_NF; ++j) o[j][e >> 1] = pack_bf16x2(f[0][j], f[1][j]);
        }
#pragma unroll
        for (int jg = 0; jg < 3; ++jg) {
          mbar_wait(bar_empty(s), phase ^ 1);
          uint8_t* abase = smem + s * Cfg::STAGE_BYTES + t * Cfg::A_TILE_BYTES;
#pragma unroll
          for (int jj = 0; jj < 3; ++jj)
            *reinterpret_cast<uint4*>(abase + ((jj * 2 + hf) * TILE_ROWS + r) * 16) =
                make_uint4(o[3 * jg + jj][0], o[3 * jg + jj][1], o[3 * jg + jj][2], o[3 * jg + jj][3]);
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_full(s));
          if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ===================== epilogue: TMEM -> fp32 row-major y =====================
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    int it = 0;
    for (int64_t dt = blockIdx.x; dt < num_dt; dt += gridDim.x, ++it) {
      mbar_wait(bar_tfull, it & 1);
      tc_fence_after();
#pragma unroll 1
      for (int t = 0; t < 2; ++t) {
        const int64_t grow = dt * (2 * TILE_ROWS) + t * TILE_ROWS + row;
#pragma unroll 1
        for (int c = 0; c < Cfg::OUT / 32; ++c) {
          float v[32];
          tmem_ld32(tmem_base + t * Cfg::OUT + c * 32 + ((uint32_t)(quad * 32) << 16), v);
          tmem_ld_wait();
          if (grow < rows) {
            float4* dst = reinterpret_cast<float4*>(y + grow * Cfg::OUT + c * 32);
#pragma unroll
            for (int q = 0; q < 8; ++q) dst[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_tempty);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------------ backward GEMMs
// 32 lanes x 8 consecutive fp32 columns (the x16 load lives in tc_common.cuh)
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&v)[8]) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}

// fp32 row-major (rows x H) -> bf16 tile image (tc_common.cuh), pad rows zero.  One block per 128-row tile; the tile
// goes through shared memory in 64-feature chunks so that both the global reads (along features) and the image
// writes (along rows) are coalesced.
__global__ void __launch_bounds__(256) k_rows_to_image(const float* __restrict__ src, int64_t rows, int H,
                                                       __nv_bfloat16* __restrict__ img) {
  __shared__ float s[TILE_ROWS][65];
  const int64_t tile = blockIdx.x;
  const int tid = threadIdx.x;
  for (int c0 = 0; c0 < H; c0 += 64) {
    for (int e = tid; e < TILE_ROWS * 16; e += 256) {           // 16 float4 per row chunk
      const int r = e >> 4, v4 = e & 15;
      const int64_t grow = tile * TILE_ROWS + r;
      float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
      if (grow < rows) f = *reinterpret_cast<const float4*>(src + grow * H + c0 + v4 * 4);
      s[r][v4 * 4] = f.x; s[r][v4 * 4 + 1] = f.y; s[r][v4 * 4 + 2] = f.z; s[r][v4 * 4 + 3] = f.w;
    }
    __syncthreads();
    for (int e = tid; e < TILE_ROWS * 8; e += 256) {            // 8 feature groups per chunk
      const int r = e & 127, fg = e >> 7;
      const float* q = &s[r][fg * 8];
      const uint4 o = make_uint4(pack_bf16x2(q[0], q[1]), pack_bf16x2(q[2], q[3]), pack_bf16x2(q[4], q[5]), pack_bf16x2(q[6], q[7]));
      *reinterpret_cast<uint4*>(img + tile * (TILE_ROWS * (int64_t)H) + ((int64_t)((c0 >> 3) + fg) * TILE_ROWS + r) * 8) = o;
    }
    __syncthreads();
  }
}

struct KanBwdCfg {
  static constexpr int OUT = 256;
  static constexpr int WB_BYTES = (OUT / 8) * KT_SLABF * 16;            // 72 KB: Wc[:, slab]^T, K-major (K = out)
  static constexpr int G_BYTES = (OUT / 8) * TILE_ROWS * 16;            // 64 KB: g tile image, two halves of 128 features
  static constexpr int G_HALF = G_BYTES / 2;
  static constexpr int PHI_BYTES = (KT_SLABF / 8) * TILE_ROWS * 16;     // 36 KB: Phi slab tile image
  static constexpr int OFF_G = WB_BYTES;
  static constexpr int OFF_PHI = OFF_G + G_BYTES;
  static constexpr int OFF_BAR = OFF_PHI + PHI_BYTES;
  static constexpr int SMEM_BYTES = OFF_BAR + 128;
  static constexpr int PROD_WARPS = 8;
  static constexpr int EPI_WARPS = 8;
  static constexpr int THREADS = 32 * (2 + PROD_WARPS + EPI_WARPS);
  static constexpr int COL_DW = 0, COL_DPHI = 2 * KT_SLABF;             // TMEM columns: dW mb0 | dW mb1 | dPhi
};

// x (rows x in) fp32 row-major; gimg = the upstream gradient g (rows x 256) as a bf16 tile image (k_rows_to_image);
// dx (rows x in) fp32 (may be NULL: first layer); gwc fp32 [256][in*9] (the i-major layout of kan.cu), accumulated with
// atomics.  grid = nslab * Q CTAs.  The g image sits in two half buffers (features 0..127 | 128..255): the MMAs are
// ordered [dPhi k 0..7, dW mb0 | dPhi k 8..15, dW mb1] so half 0 is refilled by the bulk-copy engine for the next tile
// while the tensor pipe still works on half 1.
__global__ void __launch_bounds__(KanBwdCfg::THREADS, 1) k_kan_tc_bwd(const float* __restrict__ x, const float* __restrict__ grid,
                                                                       int nk, const __nv_bfloat16* __restrict__ WB,
                                                                       const __nv_bfloat16* __restrict__ gimg, int64_t rows, int in,
                                                                       float* __restrict__ dx, float* __restrict__ gwc) {
  using Cfg = KanBwdCfg;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t sWB = smem_base, sG = smem_base + Cfg::OFF_G, sPhi = smem_base + Cfg::OFF_PHI;
  const uint32_t bar0 = smem_base + Cfg::OFF_BAR;
  const uint32_t bar_bfull = bar0, bar_phi_full = bar0 + 8, bar_phi_empty = bar0 + 16, bar_dphi_full = bar0 + 24,
                 bar_dphi_empty = bar0 + 32, bar_done = bar0 + 40;
  auto bar_g_full = [&](int h) { return bar0 + 48u + 8u * h; };
  auto bar_g_empty = [&](int h) { return bar0 + 64u + 8u * h; };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + Cfg::OFF_BAR + 80);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nslab = in / KT_SLAB;
  const int slab = blockIdx.x % nslab;
  const int q0 = blockIdx.x / nslab, Q = gridDim.x / nslab;
  const int64_t num_tiles = (rows + TILE_ROWS - 1) / TILE_ROWS;
  const bool has_work = q0 < num_tiles;

  if (threadIdx.x == 0) {
    mbar_init(bar_bfull, 1);
    mbar_init(bar_phi_full, Cfg::PROD_WARPS);
    mbar_init(bar_phi_empty, 1);
    mbar_init(bar_dphi_full, 1);
    mbar_init(bar_dphi_empty, Cfg::EPI_WARPS);
    mbar_init(bar_done, 1);
    for (int h = 0; h < 2; ++h) { mbar_init(bar_g_full(h), 1); mbar_init(bar_g_empty(h), 1); }
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== bulk-copy producer: the slab's weights once, then the g image halves per tile ===========
    if (lane == 0 && has_work) {
      mbar_arrive_expect_tx(bar_bfull, Cfg::WB_BYTES);
      const uint8_t* src = reinterpret_cast<const uint8_t*>(WB) + (size_t)slab * Cfg::WB_BYTES;
      for (int off = 0; off < Cfg::WB_BYTES; off += 24576) bulk_g2s(sWB + off, src + off, 24576, bar_bfull);
      int it = 0;
      for (int64_t tile = q0; tile < num_tiles; tile += Q, ++it) {
        const uint8_t* gsrc = reinterpret_cast<const uint8_t*>(gimg) + (size_t)tile * Cfg::G_BYTES;
        for (int h = 0; h < 2; ++h) {
          mbar_wait(bar_g_empty(h), (it & 1) ^ 1);
          mbar_arrive_expect_tx(bar_g_full(h), Cfg::G_HALF);
          bulk_g2s(sG + h * Cfg::G_HALF, gsrc + h * Cfg::G_HALF, Cfg::G_HALF, bar_g_full(h));
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && has_work) {
      constexpr uint32_t idesc_dphi = make_idesc_bf16(128, KT_SLABF, 0, 0);
      constexpr uint32_t idesc_dw = make_idesc_bf16(128, KT_SLABF, 1, 1);
      mbar_wait(bar_bfull, 0);
      int it = 0;
      for (int64_t tile = q0; tile < num_tiles; tile += Q, ++it) {
        mbar_wait(bar_phi_full, it & 1);
        mbar_wait(bar_dphi_empty, (it & 1) ^ 1);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          mbar_wait(bar_g_full(h), it & 1);
          tc_fence_after();
          // dPhi (128 rows x 144) += g[:, 128h .. 128h+128) (K-major) x Wc[128h .. , slab]
#pragma unroll
          for (int k8 = 0; k8 < 8; ++k8) {
            const int kk = h * 8 + k8;
            const uint64_t ad = make_smem_desc(sG + kk * 2 * (TILE_ROWS * 16), TILE_ROWS * 16, 128);
            const uint64_t bd = make_smem_desc(sWB + kk * 2 * (KT_SLABF * 16), KT_SLABF * 16, 128);
            umma_f16(tmem_base + Cfg::COL_DPHI, ad, bd, idesc_dphi, kk != 0 ? 1u : 0u);
          }
          if (h == 1) umma_commit(bar_dphi_full);
          // dWc[128h .. +128, slab] += g[:, 128h ..]^T Phi  (both MN-major, K = 16 rows per step)
#pragma unroll
          for (int kk = 0; kk < TILE_ROWS / 16; ++kk) {
            const uint64_t ad = make_smem_desc(sG + h * Cfg::G_HALF + kk * 256, 128, TILE_ROWS * 16);
            const uint64_t bd = make_smem_desc(sPhi + kk * 256, 128, TILE_ROWS * 16);
            umma_f16(tmem_base + Cfg::COL_DW + h * KT_SLABF, ad, bd, idesc_dw, (it | kk) != 0 ? 1u : 0u);
          }
          umma_commit(bar_g_empty(h));
        }
        umma_commit(bar_phi_empty);
      }
      umma_commit(bar_done);
    }
  } else if (warp < 2 + Cfg::PROD_WARPS) {
    // ===================== Phi producers: thread = (row, 8 inputs of the slab) =====================
    const UGrid ug = load_ugrid(grid, nk);
    const int p = threadIdx.x - 64;                             // 0..255
    const int r = p & 127, hf = p >> 7;
    int it = 0;
    for (int64_t tile = q0; tile < num_tiles; tile += Q, ++it) {
      const int64_t grow = tile * TILE_ROWS + r;
      const int64_t crow = grow < rows ? grow : rows - 1;       // pad rows: finite Phi against a zero g row
      const float4 f0 = *reinterpret_cast<const float4*>(x + crow * in + slab * KT_SLAB + hf * 8);
      const float4 f1 = *reinterpret_cast<const float4*>(x + crow * in + slab * KT_SLAB + hf * 8 + 4);
      const float xv[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
      uint32_t ph[KT_NF][4];
#pragma unroll
      for (int e = 0; e < 8; e += 2) {
        float fa[KT_NF], fb[KT_NF];
        feats9(xv[e], ug, fa);
        feats9(xv[e + 1], ug, fb);
#pragma unroll
        for (int j = 0; j < KT_NF; ++j) ph[j][e >> 1] = pack_bf16x2(fa[j], fb[j]);
      }
      mbar_wait(bar_phi_empty, (it & 1) ^ 1);
#pragma unroll
      for (int j = 0; j < KT_NF; ++j)
        *reinterpret_cast<uint4*>(smem + Cfg::OFF_PHI + ((j * 2 + hf) * TILE_ROWS + r) * 16) = make_uint4(ph[j][0], ph[j][1], ph[j][2], ph[j][3]);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_phi_full);
    }
  } else {
    // ===================== epilogue: dPhi x phi' -> dx ; at the end dW accumulators -> gwc =====================
    // 8 warps: two per TMEM lane quadrant, one for each half (8 inputs) of the slab / each out-feature block of dW.
    const UGrid ug = load_ugrid(grid, nk);
    const int quad = warp & 3;
    const int hf = (warp - (2 + Cfg::PROD_WARPS)) >> 2;
    const int row = quad * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    int it = 0;
    for (int64_t tile = q0; tile < num_tiles; tile += Q, ++it) {
      const int64_t grow = tile * TILE_ROWS + row;
      const int64_t crow = grow < rows ? grow : rows - 1;
      const float4 f0 = *reinterpret_cast<const float4*>(x + crow * in + slab * KT_SLAB + hf * 8);
      const float4 f1 = *reinterpret_cast<const float4*>(x + crow * in + slab * KT_SLAB + hf * 8 + 4);
      const float xv[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
      float dwq[8][4], acc[8];
      int mm[8];
      if (dx) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          float w[4];
          mm[e] = basis4<true>(xv[e], ug, w, dwq[e]);
          const float sg = 1.f / (1.f + __expf(-xv[e]));
          acc[e] = sg * (1.f + xv[e] * (1.f - sg));             // SiLU'(x); multiplied by dPhi_0 below
        }
      }
      mbar_wait(bar_dphi_full, it & 1);
      tc_fence_after();
      if (dx) {
#pragma unroll
        for (int grp = 0; grp < 3; ++grp) {                     // features j = 3 grp .. 3 grp + 2: three loads, one wait
          float v[3][8];
#pragma unroll
          for (int jj = 0; jj < 3; ++jj) tmem_ld8(tmem_base + Cfg::COL_DPHI + (3 * grp + jj) * KT_SLAB + hf * 8 + lane_addr, v[jj]);
          tmem_ld_wait();
#pragma unroll
          for (int jj = 0; jj < 3; ++jj) {
            const int j = 3 * grp + jj;
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              if (j == 0) {
                acc[e] *= v[jj][e];
              } else {
                const int qq = (j - 1) - (mm[e] - 3);
                float dv = 0.f;
#pragma unroll
                for (int t = 0; t < 4; ++t) dv = (qq == t) ? dwq[e][t] : dv;
                acc[e] = __fmaf_rn(v[jj][e], dv, acc[e]);
              }
            }
          }
        }
        if (grow < rows) {
          float4* dst = reinterpret_cast<float4*>(dx + grow * in + slab * KT_SLAB + hf * 8);
          dst[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
          dst[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_dphi_empty);
    }
    if (has_work) {
      mbar_wait(bar_done, 0);
      tc_fence_after();
      const int Kf = in * KT_NF;
      const int mb = hf;
      float* gro = gwc + (size_t)(mb * 128 + row) * Kf + (size_t)slab * KT_SLAB * KT_NF;
#pragma unroll 1
      for (int j = 0; j < KT_NF; ++j) {
        float v[16];
        tmem_ld16(tmem_base + Cfg::COL_DW + mb * KT_SLABF + j * KT_SLAB + lane_addr, v);
        tmem_ld_wait();
#pragma unroll
        for (int ii = 0; ii < KT_SLAB; ++ii) atomicAdd(gro + ii * KT_NF + j, v[ii]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace inr
