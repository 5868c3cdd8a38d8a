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

// backward of the thin layer: dx[r,i] = g[r] * sum_j phi_j'(x) wc[i*9+j];  gwc[i*9+j] += sum_r g[r] phi_j(x[r,i])
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
  float gacc[KT_THIN_PER][KT_NF];
#pragma unroll
  for (int p = 0; p < KT_THIN_PER; ++p)
#pragma unroll
    for (int j = 0; j < KT_NF; ++j) gacc[p][j] = 0.f;
  for (int64_t r = warp0; r < rows; r += nwarps) {
    const float gv = g[r];
#pragma unroll
    for (int p = 0; p < KT_THIN_PER; ++p) {
      const int i = lane + 32 * p;
      if (i < in) {
        const float xv = x[r * in + i];
        float w[4], dw[4];
        const int m = basis4<true>(xv, ug, w, dw);
        const float* wp = s_wc + i * KT_NF;
        const float sg = 1.f / (1.f + __expf(-xv));
        float d = sg * (1.f + xv * (1.f - sg)) * wp[0];
        gacc[p][0] = __fmaf_rn(gv, xv * sg, gacc[p][0]);
#pragma unroll
        for (int b = 0; b < KT_NF - 1; ++b) {
          const int q = b - (m - 3);
          float wv = 0.f, dv = 0.f;
#pragma unroll
          for (int t = 0; t < 4; ++t) { wv = (q == t) ? w[t] : wv; dv = (q == t) ? dw[t] : dv; }
          d = __fmaf_rn(dv, wp[1 + b], d);
          gacc[p][1 + b] = __fmaf_rn(gv, wv, gacc[p][1 + b]);
        }
        dx[r * in + i] = gv * d;
      }
    }
  }
#pragma unroll
  for (int p = 0; p < KT_THIN_PER; ++p) {
    const int i = lane + 32 * p;
    if (i < in)
#pragma unroll
      for (int j = 0; j < KT_NF; ++j) atomicAdd(&s_gw[i * KT_NF + j], gacc[p][j]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < in * KT_NF; i += blockDim.x) atomicAdd(&gwc[i], s_gw[i]);
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
  static constexpr int PROD_WARPS = 8;
  static constexpr int THREADS = 32 * (2 + PROD_WARPS + 4);     // B producer, MMA, 8 Phi producers, 4 epilogue
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
              umma_bf16(tmem_base + t * Cfg::OUT, ad, bd, idesc, (st | jj) != 0 ? 1u : 0u);
            }
          umma_commit(bar_empty(s));
          if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
        }
        umma_commit(bar_tfull);
      }
    }
  } else if (warp < 2 + Cfg::PROD_WARPS) {
    // ===================== Phi producers: thread = one row of the double tile, all 16 inputs of the slab ==========
    const UGrid ug = load_ugrid(grid, nk);
    const int p = threadIdx.x - 64;                             // 0..255
    const int t = p >> 7, r = p & 127;
    int s = 0;
    uint32_t phase = 0;
    for (int64_t dt = blockIdx.x; dt < num_dt; dt += gridDim.x) {
      int64_t grow = dt * (2 * TILE_ROWS) + p;
      if (grow >= rows) grow = rows - 1;                        // pad rows replicate the last row; never stored
      const float* xrow = x + grow * in;
      for (int slab = 0; slab < nslab; ++slab) {
        float xv[KT_SLAB];
#pragma unroll
        for (int v4 = 0; v4 < KT_SLAB / 4; ++v4) {
          const float4 f = *reinterpret_cast<const float4*>(xrow + slab * KT_SLAB + v4 * 4);
          xv[v4 * 4] = f.x; xv[v4 * 4 + 1] = f.y; xv[v4 * 4 + 2] = f.z; xv[v4 * 4 + 3] = f.w;
        }
#pragma unroll 1
        for (int jg = 0; jg < 3; ++jg) {
          uint32_t o[3][2][4];
#pragma unroll
          for (int ii = 0; ii < KT_SLAB; ii += 2) {
            float f[2][3];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              const float xx = xv[ii + u];
              float w[4];
              const int m = basis4<false>(xx, ug, w, nullptr);
              const float silu = xx / (1.f + __expf(-xx));
#pragma unroll
              for (int jj = 0; jj < 3; ++jj) f[u][jj] = feat_of(3 * jg + jj, silu, m, w);
            }
#pragma unroll
            for (int jj = 0; jj < 3; ++jj) o[jj][ii >> 3][(ii & 7) >> 1] = pack_bf16x2(f[0][jj], f[1][jj]);
          }
          mbar_wait(bar_empty(s), phase ^ 1);
          uint8_t* abase = smem + s * Cfg::STAGE_BYTES + t * Cfg::A_TILE_BYTES;
#pragma unroll
          for (int jj = 0; jj < 3; ++jj)
#pragma unroll
            for (int hf = 0; hf < 2; ++hf)
              *reinterpret_cast<uint4*>(abase + ((jj * 2 + hf) * TILE_ROWS + r) * 16) =
                  make_uint4(o[jj][hf][0], o[jj][hf][1], o[jj][hf][2], o[jj][hf][3]);
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
// 32 lanes x 16 / 8 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&v)[8]) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}

struct KanBwdCfg {
  static constexpr int OUT = 256;
  static constexpr int WB_BYTES = (OUT / 8) * KT_SLABF * 16;            // 72 KB: Wc[:, slab]^T, K-major (K = out)
  static constexpr int G_BYTES = (OUT / 8) * TILE_ROWS * 16;            // 64 KB: g tile image
  static constexpr int PHI_BYTES = (KT_SLABF / 8) * TILE_ROWS * 16;     // 36 KB: Phi slab tile image
  static constexpr int OFF_G = WB_BYTES;
  static constexpr int OFF_PHI = OFF_G + G_BYTES;
  static constexpr int OFF_BAR = OFF_PHI + PHI_BYTES;
  static constexpr int SMEM_BYTES = OFF_BAR + 128;
  static constexpr int PROD_WARPS = 8;
  static constexpr int THREADS = 32 * (2 + PROD_WARPS + 4);
  static constexpr int COL_DW = 0, COL_DPHI = 2 * KT_SLABF;             // TMEM columns: dW mb0 | dW mb1 | dPhi
};

// x (rows x in), g (rows x 256) fp32 row-major; dx (rows x in) fp32 (may be NULL: first layer); gwc fp32 [256][in*9]
// (the i-major layout of kan.cu), accumulated with atomics.  grid = nslab * Q CTAs.
__global__ void __launch_bounds__(KanBwdCfg::THREADS, 1) k_kan_tc_bwd(const float* __restrict__ x, const float* __restrict__ grid,
                                                                       int nk, const __nv_bfloat16* __restrict__ WB,
                                                                       const float* __restrict__ g, int64_t rows, int in,
                                                                       float* __restrict__ dx, float* __restrict__ gwc) {
  using Cfg = KanBwdCfg;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t sWB = smem_base, sG = smem_base + Cfg::OFF_G, sPhi = smem_base + Cfg::OFF_PHI;
  const uint32_t bar0 = smem_base + Cfg::OFF_BAR;
  const uint32_t bar_bfull = bar0, bar_img_full = bar0 + 8, bar_img_empty = bar0 + 16, bar_dphi_full = bar0 + 24,
                 bar_dphi_empty = bar0 + 32, bar_done = bar0 + 40;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + Cfg::OFF_BAR + 48);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nslab = in / KT_SLAB;
  const int slab = blockIdx.x % nslab;
  const int q0 = blockIdx.x / nslab, Q = gridDim.x / nslab;
  const int64_t num_tiles = (rows + TILE_ROWS - 1) / TILE_ROWS;
  const bool has_work = q0 < num_tiles;

  if (threadIdx.x == 0) {
    mbar_init(bar_bfull, 1);
    mbar_init(bar_img_full, Cfg::PROD_WARPS);
    mbar_init(bar_img_empty, 1);
    mbar_init(bar_dphi_full, 1);
    mbar_init(bar_dphi_empty, 4);
    mbar_init(bar_done, 1);
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
    if (lane == 0 && has_work) {
      mbar_arrive_expect_tx(bar_bfull, Cfg::WB_BYTES);
      const uint8_t* src = reinterpret_cast<const uint8_t*>(WB) + (size_t)slab * Cfg::WB_BYTES;
      for (int off = 0; off < Cfg::WB_BYTES; off += 24576) bulk_g2s(sWB + off, src + off, 24576, bar_bfull);
    }
  } else if (warp == 1) {
    if (lane == 0 && has_work) {
      constexpr uint32_t idesc_dphi = make_idesc_bf16(128, KT_SLABF, 0, 0);
      constexpr uint32_t idesc_dw = make_idesc_bf16(128, KT_SLABF, 1, 1);
      mbar_wait(bar_bfull, 0);
      int it = 0;
      for (int64_t tile = q0; tile < num_tiles; tile += Q, ++it) {
        mbar_wait(bar_img_full, it & 1);
        mbar_wait(bar_dphi_empty, (it & 1) ^ 1);
        tc_fence_after();
        // dPhi (128 rows x 144) = g (K-major, K = out) x Wc[:, slab]
#pragma unroll
        for (int kk = 0; kk < Cfg::OUT / 16; ++kk) {
          const uint64_t ad = make_smem_desc(sG + kk * 2 * (TILE_ROWS * 16), TILE_ROWS * 16, 128);
          const uint64_t bd = make_smem_desc(sWB + kk * 2 * (KT_SLABF * 16), KT_SLABF * 16, 128);
          umma_bf16(tmem_base + Cfg::COL_DPHI, ad, bd, idesc_dphi, kk != 0 ? 1u : 0u);
        }
        umma_commit(bar_dphi_full);
        // dWc[mb*128 .. +128, slab] += g^T Phi  (both MN-major, K = 16 rows per step)
#pragma unroll
        for (int mb = 0; mb < Cfg::OUT / 128; ++mb)
#pragma unroll
          for (int kk = 0; kk < TILE_ROWS / 16; ++kk) {
            const uint64_t ad = make_smem_desc(sG + mb * 16 * (TILE_ROWS * 16) + kk * 256, 128, TILE_ROWS * 16);
            const uint64_t bd = make_smem_desc(sPhi + kk * 256, 128, TILE_ROWS * 16);
            umma_bf16(tmem_base + Cfg::COL_DW + mb * KT_SLABF, ad, bd, idesc_dw, (it | kk) != 0 ? 1u : 0u);
          }
        umma_commit(bar_img_empty);
      }
      umma_commit(bar_done);
    }
  } else if (warp < 2 + Cfg::PROD_WARPS) {
    // ===================== producers: g tile -> bf16 image, Phi slab tile -> bf16 image =====================
    const UGrid ug = load_ugrid(grid, nk);
    const int p = threadIdx.x - 64;
    const int r = p & 127, hf = p >> 7;
    int it = 0;
    for (int64_t tile = q0; tile < num_tiles; tile += Q, ++it) {
      const int64_t grow = tile * TILE_ROWS + r;
      const bool valid = grow < rows;
      const int64_t crow = valid ? grow : rows - 1;
      // Phi features of this thread's 8 inputs
      uint32_t ph[KT_NF][4];
      {
        const float4 f0 = *reinterpret_cast<const float4*>(x + crow * in + slab * KT_SLAB + hf * 8);
        const float4 f1 = *reinterpret_cast<const float4*>(x + crow * in + slab * KT_SLAB + hf * 8 + 4);
        const float xv[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
#pragma unroll
        for (int e = 0; e < 8; e += 2) {
          float f[2][KT_NF];
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            float w[4];
            const int m = basis4<false>(xv[e + u], ug, w, nullptr);
            const float silu = xv[e + u] / (1.f + __expf(-xv[e + u]));
#pragma unroll
            for (int j = 0; j < KT_NF; ++j) f[u][j] = feat_of(j, silu, m, w);
          }
#pragma unroll
          for (int j = 0; j < KT_NF; ++j) ph[j][e >> 1] = pack_bf16x2(f[0][j], f[1][j]);
        }
      }
      mbar_wait(bar_img_empty, (it & 1) ^ 1);
#pragma unroll
      for (int j = 0; j < KT_NF; ++j)
        *reinterpret_cast<uint4*>(smem + Cfg::OFF_PHI + ((j * 2 + hf) * TILE_ROWS + r) * 16) = make_uint4(ph[j][0], ph[j][1], ph[j][2], ph[j][3]);
      // g: this thread converts feature groups hf*16 .. hf*16+15 of its row (zeros for pad rows: they must not reach dW)
      const float* grow_p = g + crow * Cfg::OUT + hf * 128;
#pragma unroll 4
      for (int fg = 0; fg < 16; ++fg) {
        uint4 o = make_uint4(0u, 0u, 0u, 0u);
        if (valid) {
          const float4 a = *reinterpret_cast<const float4*>(grow_p + fg * 8);
          const float4 b = *reinterpret_cast<const float4*>(grow_p + fg * 8 + 4);
          o = make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
        }
        *reinterpret_cast<uint4*>(smem + Cfg::OFF_G + ((hf * 16 + fg) * TILE_ROWS + r) * 16) = o;
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_img_full);
    }
  } else {
    // ===================== epilogue: dPhi x phi' -> dx ; at the end dW accumulators -> gwc =====================
    const UGrid ug = load_ugrid(grid, nk);
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    int it = 0;
    for (int64_t tile = q0; tile < num_tiles; tile += Q, ++it) {
      const int64_t grow = tile * TILE_ROWS + row;
      const int64_t crow = grow < rows ? grow : rows - 1;
      mbar_wait(bar_dphi_full, it & 1);
      tc_fence_after();
      if (dx) {
#pragma unroll 1
        for (int hf = 0; hf < 2; ++hf) {
          const float4 f0 = *reinterpret_cast<const float4*>(x + crow * in + slab * KT_SLAB + hf * 8);
          const float4 f1 = *reinterpret_cast<const float4*>(x + crow * in + slab * KT_SLAB + hf * 8 + 4);
          const float xv[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
          float dwq[8][4], dsl[8], acc[8];
          int mm[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            float w[4];
            mm[e] = basis4<true>(xv[e], ug, w, dwq[e]);
            const float sg = 1.f / (1.f + __expf(-xv[e]));
            dsl[e] = sg * (1.f + xv[e] * (1.f - sg));
          }
          {
            float v[8];
            tmem_ld8(tmem_base + Cfg::COL_DPHI + hf * 8 + lane_addr, v);
            tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] = v[e] * dsl[e];
          }
#pragma unroll
          for (int b = 0; b < KT_NF - 1; ++b) {
            float v[8];
            tmem_ld8(tmem_base + Cfg::COL_DPHI + (b + 1) * KT_SLAB + hf * 8 + lane_addr, v);
            tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const int qq = b - (mm[e] - 3);
              float dv = 0.f;
#pragma unroll
              for (int t = 0; t < 4; ++t) dv = (qq == t) ? dwq[e][t] : dv;
              acc[e] = __fmaf_rn(v[e], dv, acc[e]);
            }
          }
          if (grow < rows) {
            float4* dst = reinterpret_cast<float4*>(dx + grow * in + slab * KT_SLAB + hf * 8);
            dst[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
            dst[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
          }
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
#pragma unroll 1
      for (int mb = 0; mb < Cfg::OUT / 128; ++mb) {
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
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace inr
