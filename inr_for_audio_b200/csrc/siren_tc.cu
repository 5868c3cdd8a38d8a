// siren_tc.cu -- the fp16 tensor-core SIREN fit path (INR_PREC_FP16): host orchestration + the streaming kernels
// around the tcgen05 GEMMs of tc_gemm.cuh.  All activations live in HBM as fp16 "tile images" (tc_common.cuh).
//
//   per step and row chunk:
//     k_tc_pack        fp32 master weights -> fp16 UMMA operand images W_l (fwd: [hi | lo] x wscale in the split mode) and
//                      W_l^T (dX); sets the step's loss scale when there is no upstream gradient
//     H <= 256, sine stacks:   k_tc_fused_fwd / k_tc_fused_fwd_s   first layer + all hidden layers + head + loss + dZ_L
//     otherwise:
//       k_tc_first       S_0 [+ lo], C_0 = sincos(omega_0 (W_0 x + b_0))      fp32 argument, Cody-Waite + MUFU (models.py:114-115)
//       k_tc_layer2<FWD> S_l [+ lo], C_l = sincos(omega (S_{l-1} W_l^T + b_l))                          l = 1..L
//       k_tc_last2       out = S_L w + b; loss; g = dloss/dout; dZ_L = gs g w omega C_L (over C_L); last-layer gradients
//     k_tc_dw / dw2    gW_l += dZ_l^T S_{l-1} / gs, gb_l += colsum(dZ_l) / gs                            l = L..1
//     k_tc_layer2<BWD> dZ_{l-1} = (dZ_l W_l) omega C_{l-1}  (over C_{l-1})                               l = L..1
//     k_tc_colsum      gW_0 += dZ_0^T x / gs, gb_0 += colsum(dZ_0) / gs; `a` gradients of Snake layers
#include <stdlib.h>

#include "tc_gemm2.cuh"
#include "tc_fused.cuh"
#include "tc_fused_split.cuh"

namespace inr {
using namespace tc;

// ------------------------------------------------------------------------------------------------ loss scale
// fp16 dZ images need a loss scale: g = d loss / d out is O(1/T).  gsc = {gs, 1/gs} with gs the power of two that puts
// max |g| of the PREVIOUS step (opt->gmax_prev, measured by the last-layer kernels) at 2^6.  Round 2 aimed at 2^10, sized
// for the INITIAL weights (|w_head| omega ~ 0.15): once the head weights have trained to O(1), |dZ_L| = |g w omega C| of
// the largest-|g| rows passed 65504 and cvt.rn.satfinite clipped them silently -- SirenFitter.dz_saturated() counted
// 1e4..1e5 clipped elements per step from step ~2000 on in the C1 and 5x512 fits (profiles/r03_notes.md).  2^6 leaves a
// factor 1000 for |w| omega and the per-layer backward gains; typical dZ elements are then ~0.1..1, four decades above
// the fp16 normal minimum (6e-5; subnormals degrade gracefully below it).  Without history (first step, or no optimiser state: the autograd path) the estimate is
// 4 * grad_scale (|2 diff| <= 4 for peak-normalised audio) plus max |extra_g|.  Consumers of dZ (dW, column sums)
// multiply by 1/gs: exact, it is a power of two.
__device__ __forceinline__ void gscale_compute(const inr_opt_state* opt, float extra_max, float grad_scale, int has_target,
                                               float* gsc) {
  float est = opt ? opt->gmax_prev : 0.f;
  if (!(est > 0.f)) est = (has_target ? 4.f * grad_scale : 0.f) + extra_max;
  est = fmaxf(est, extra_max);
  int e = 0;
  if (est > 0.f && est < 3e38f) e = 6 - (int)ceilf(log2f(est));
  e = max(-100, min(100, e));
  gsc[0] = exp2f((float)e);
  gsc[1] = exp2f((float)-e);
}

// ------------------------------------------------------------------------------------------------ pack
// Forward operand slices are [slice][hi | lo (split)][K/8][NSF][8] (NSF columns per CTA: tc_ns_fwd); the dX operand is
// never split: [slice][K/8][NSB][8].
// split: the forward operand carries W * wscale (a power of two, undone by the epilogue's acc_scale) so that the lo halves
// |lo| <= 2^-12 |W wscale| stay in fp16's NORMAL range: trained SIREN weights are ~0.02, their unscaled lo halves ~5e-6 would
// be subnormal (absolute spacing 6e-8 = only 18 bits of W, a -64 dB forward-noise floor at C1; with the scale < -75 dB).
__global__ void __launch_bounds__(256) k_tc_pack(const float* __restrict__ params, SirenLayout lay, int NSF, int NSB, int split,
                                                 float wscale, op16* __restrict__ Wf, op16* __restrict__ Wb,
                                                 const inr_opt_state* __restrict__ opt, float grad_scale, int has_target,
                                                 float* __restrict__ gsc) {
  // gsc != NULL: the fused step without an upstream gradient -- this launch also sets the step's loss scale
  if (gsc && blockIdx.x == 0 && threadIdx.x == 0) gscale_compute(opt, 0.f, grad_scale, has_target, gsc);
  const int H = lay.H;
  const int64_t per = (int64_t)H * H;
  const int64_t total = per * lay.L;
  const int64_t fper = per * (split ? 2 : 1);         // forward-operand elements per layer
  const int64_t fslice = (int64_t)H * NSF * (split ? 2 : 1);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int l = (int)(i / per);
    const int rem = (int)(i - (int64_t)l * per);
    const int o = rem / H, in = rem - o * H;
    const float wv = params[lay.w_off(l + 1) + rem];
    const op16 w = to_op(wv);
    // forward operand: B[n = out][k = in]
    const int64_t fo = (int64_t)l * fper + (int64_t)(o / NSF) * fslice + ((int64_t)(in >> 3) * NSF + (o % NSF)) * 8 + (in & 7);
    const op16 wf = to_op(wv * wscale);
    Wf[fo] = wf;
    if (split) Wf[fo + (int64_t)H * NSF] = to_op(wv * wscale - __half2float(wf));
    // dX operand: B[n = in][k = out]
    Wb[(int64_t)l * per + (int64_t)(in / NSB) * ((int64_t)H * NSB) + ((int64_t)(o >> 3) * NSB + (in % NSB)) * 8 + (o & 7)] = w;
  }
}

// ------------------------------------------------------------------------------------------------ wide inputs
// Encoded coordinates (in_features = 2*num_freq > 2, run.py:81-82,141-144) on the tcgen05 path: the input rows become one
// more tile image X (fp16, zero-padded to H columns) and layer 0 one more H x H layer GEMM with W_0 zero-padded to H
// input columns; its weight gradient is the dW GEMM against X, cut back to (H x in).  Needs in_features <= H.
__global__ void __launch_bounds__(256) k_tc_pack0(const float* __restrict__ W0, int in, int H, int NS, int split, float wscale,
                                                  op16* __restrict__ Wf0) {
  const int total = H * H;
  const int64_t fslice = (int64_t)H * NS * (split ? 2 : 1);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int o = i / H, k = i - o * H;
    const float v = (k < in ? W0[(int64_t)o * in + k] : 0.f) * wscale;
    const op16 w = to_op(v);
    const int64_t fo = (int64_t)(o / NS) * fslice + ((int64_t)(k >> 3) * NS + (o % NS)) * 8 + (k & 7);
    Wf0[fo] = w;
    if (split) Wf0[fo + (int64_t)H * NS] = to_op(v - __half2float(w));
  }
}

// X[tile][group][row][8] <- coords (rows x in) fp32, zero beyond `in` columns and `rows` rows
__global__ void __launch_bounds__(256) k_tc_rows_to_image(const float* __restrict__ coords, int in, int H, int64_t rows,
                                                          int64_t rows_pad, op16* __restrict__ X, op16* __restrict__ Xlo) {
  const int lg = 31 - __clz(H >> 3);
  const int64_t total = rows_pad << lg;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(idx & (TILE_ROWS - 1));
    const int64_t tg = idx >> 7;
    const int kg = (int)(tg & ((1 << lg) - 1));
    const int64_t row = (tg >> lg) * TILE_ROWS + r;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int f = kg * 8 + j;
      v[j] = (row < rows && f < in) ? coords[row * in + f] : 0.f;
    }
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int h = 0; h < 4; ++h) pack_op2_split(v[2 * h], v[2 * h + 1], &hi[h], &lo[h]);
    *reinterpret_cast<uint4*>(X + idx * 8) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    if (Xlo) *reinterpret_cast<uint4*>(Xlo + idx * 8) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
  }
}

// gW_0 (H x in) += the first `in` columns of the padded H x H gradient
__global__ void __launch_bounds__(256) k_tc_unpad_add(const float* __restrict__ gpad, int in, int H, float* __restrict__ gW) {
  const int total = H * in;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int o = i / in, k = i - o * in;
    gW[i] += gpad[(int64_t)o * H + k];
  }
}

// ------------------------------------------------------------------------------------------------ first layer
// one thread = one row x one feature group (8 features); consecutive threads = consecutive rows of a tile.
// sincos_reduced (tc_fused.cuh): fp32 argument omega*z as the reference forms it, Cody-Waite reduction, MUFU sincos.
__global__ void __launch_bounds__(256) k_tc_first(const float* __restrict__ W, const float* __restrict__ B,
                                                  const float* __restrict__ coords, int in, int H, int64_t rows,
                                                  int64_t rows_pad, int64_t g0, int64_t T_total, float scale,
                                                  float omega, op16* __restrict__ S0,
                                                  op16* __restrict__ C0, const float* __restrict__ Aparam,
                                                  op16* __restrict__ E0, op16* __restrict__ S0lo) {
  // Aparam != NULL: first_linear (models.py:343-347), layer 0 = nn.Linear + Snake: S0 = y, C0 = dy/dz, E0 = dy/da
  const int lg = 31 - __clz(H >> 3);                      // H/8 feature groups: a power of two on this path
  const int64_t total = rows_pad << lg;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(idx & (TILE_ROWS - 1));
    const int64_t tg = idx >> 7;
    const int kg = (int)(tg & ((1 << lg) - 1));
    const int64_t tile = tg >> lg;
    int64_t row = tile * TILE_ROWS + r;
    if (row >= rows) row = rows - 1;   // pad rows replicate the last row: finite values, zero gradient later
    float x0, x1 = 0.f;
    if (coords) {
      x0 = coords[row * in];
      if (in == 2) x1 = coords[row * in + 1];
    } else {
      x0 = linspace_at(g0 + row, T_total, scale);
    }
    uint32_t so[4], co[4], eo[4], sl[4];
#pragma unroll
    for (int h = 0; h < 4; ++h) {
      float s[2], c[2], e[2] = {0.f, 0.f};
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int f = kg * 8 + h * 2 + u;
        const float z = (in == 1) ? __fmaf_rn(x0, W[f], B[f]) : (__fmaf_rn(x1, W[f * 2 + 1], x0 * W[f * 2]) + B[f]);
        if (Aparam) {
          const float a = Aparam[f], ia = 1.0f / a;
          float sn, cs;
          sincos_reduced(z * a, &sn, &cs);
          const float s2 = sn * sn, sc2 = 2.f * sn * cs;
          s[u] = __fmaf_rn(ia, s2, z);
          c[u] = 1.f + sc2;
          e[u] = ia * (z * sc2 - ia * s2);
        } else {
          sincos_reduced(omega * z, &s[u], &c[u]);
        }
      }
      pack_op2_split(s[0], s[1], &so[h], &sl[h]);
      co[h] = pack_op2(c[0], c[1]);
      eo[h] = pack_op2(e[0], e[1]);
    }
    const int64_t off = idx * 8;   // == ((tile*groups + kg)*128 + r)*8
    *reinterpret_cast<uint4*>(S0 + off) = make_uint4(so[0], so[1], so[2], so[3]);
    if (S0lo) *reinterpret_cast<uint4*>(S0lo + off) = make_uint4(sl[0], sl[1], sl[2], sl[3]);
    if (C0) *reinterpret_cast<uint4*>(C0 + off) = make_uint4(co[0], co[1], co[2], co[3]);
    if (Aparam && E0) *reinterpret_cast<uint4*>(E0 + off) = make_uint4(eo[0], eo[1], eo[2], eo[3]);
  }
}

// ------------------------------------------------------------------------------------------------ last layer, one pass
// k_tc_last2<H, SNAKE>: the head, the loss, dZ_L AND the last-layer weight gradient (+ the `a` gradient of a Snake layer
// L) in ONE pass over S_L / C_L: 3 units of HBM traffic instead of the 4 (5) of k_tc_last + k_tc_colsum_w (+ k_tc_colsum).
//
// Work unit = 32 rows of a tile (one lane per row).  512 threads; a 16-byte cp.async ring (NSTAGE stages of [S | C] slabs,
// 2 ahead in flight) decouples the HBM latency from the two block barriers a unit needs:
//   phase A   warp w owns feature groups w, w+16, ..: partial dots of its 32 rows from the S slab (kept in registers)
//   barrier   every warp sums the 16 partials (fixed order), forms g = dloss/dout redundantly; warp 0 writes out / loss
//   phase C   dZ_L = g w omega C_L straight to global, acc_w[f] += g S[f] in registers (one butterfly at the very end)
// LO (operand split): the lo image of S_L rides the same ring as a third slab per stage.  (Round 2 fetched it with plain
// loads at the top of each unit and consumed it at once: ~1 us of exposed latency per 5.5 us unit at H = 512, the kernel
// ran at 0.53 of the HBM peak.)  At H = 512 three 32 KB slabs leave room for 2 stages = one unit in flight while one is
// consumed; 96 KB in flight per SM is twice what the bandwidth-latency product asks for (6.5 TB/s x ~1 us / 148 SMs).
template <int H, bool LO = false>
struct Last2Cfg {
  static constexpr int G = H / 8;                  // feature groups
  static constexpr int KPW = G / 16;               // groups per warp
  static constexpr int SLAB = G * 32 * 16;         // bytes of a 32-row slab of one image
  static constexpr int NSLAB = LO ? 3 : 2;         // [S | C | S_lo]
  static constexpr int NSTAGE = (H >= 512) ? (LO ? 2 : 3) : 4;
  static constexpr int STAGE_BYTES = NSLAB * SLAB;
  static constexpr int OFF_W = NSTAGE * STAGE_BYTES;
  static constexpr int OFF_DOT = OFF_W + H * 4;
  static constexpr int SMEM_BYTES = OFF_DOT + 16 * 32 * 4;
  static_assert(H == 128 || H == 256 || H == 512, "hidden width");
  static_assert(SMEM_BYTES <= 232448, "shared-memory budget");
};

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int H, bool SNAKE, bool LO>
__global__ void __launch_bounds__(512, 1) k_tc_last2(const op16* __restrict__ S, op16* __restrict__ CdZ,
                                                     const float* __restrict__ w, const float* __restrict__ bptr,
                                                     const float* __restrict__ target, const float* __restrict__ extra_g,
                                                     int64_t rows, int num_tiles, float omega_prev, float omega_hidden,
                                                     int outermost_linear, int loss_kind, float grad_scale,
                                                     float* __restrict__ gw, float* __restrict__ gb,
                                                     double* __restrict__ loss_acc, float* __restrict__ out, int training,
                                                     const op16* __restrict__ Eimg, float* __restrict__ ga,
                                                     const float* __restrict__ gsc, unsigned* __restrict__ gmax_acc,
                                                     const op16* __restrict__ Slo) {
  // LO (operand split, Slo != NULL): S_L = S + Slo, the head's dot product takes both (third slab of the ring)
  using Cfg = Last2Cfg<H, LO>;
  constexpr int KPW = Cfg::KPW;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t smem_base = smem_u32(smem);
  float* s_w = reinterpret_cast<float*>(smem + Cfg::OFF_W);
  float* s_dot = reinterpret_cast<float*>(smem + Cfg::OFF_DOT);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < H; i += 512) s_w[i] = w[i];
  __syncthreads();
  const float bias = bptr[0];
  const int num_units = num_tiles * 4;
  const float gs = (training && gsc) ? gsc[0] : 1.f;   // power-of-two loss scale of the fp16 dZ images (set_gscale)
  float gmax_local = 0.f;

  auto issue = [&](int unit, int stage) {
    if (unit < num_units) {
      const int tile = unit >> 2, rb = unit & 3;
      const size_t tile_off = (size_t)tile * (TILE_ROWS * (size_t)H);
      const uint32_t dstS = smem_base + stage * Cfg::STAGE_BYTES;
#pragma unroll
      for (int c = tid; c < Cfg::G * 32; c += 512) {
        const int kg = c >> 5, r = c & 31;
        const size_t off = tile_off + ((size_t)kg * TILE_ROWS + rb * 32 + r) * 8;
        cp_async16(dstS + c * 16, S + off);
        if (training) cp_async16(dstS + Cfg::SLAB + c * 16, CdZ + off);
        if (LO) cp_async16(dstS + 2 * Cfg::SLAB + c * 16, Slo + off);
      }
    }
    cp_async_commit();
  };

  float acc_w[KPW][8], acc_a[SNAKE ? KPW : 1][8];
#pragma unroll
  for (int i = 0; i < KPW; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) { acc_w[i][j] = 0.f; if (SNAKE) acc_a[i][j] = 0.f; }
  double loss_local = 0.0;
  float gb_local = 0.f;

#pragma unroll
  for (int s = 0; s < Cfg::NSTAGE - 1; ++s) issue(blockIdx.x + s * gridDim.x, s);
  int it = 0;
  for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x, ++it) {
    const int stage = it % Cfg::NSTAGE;
    const int tile = unit >> 2, rb = unit & 3;
    const int r = rb * 32 + lane;
    const int64_t row = (int64_t)tile * TILE_ROWS + r;
    const bool valid = row < rows;
    // g needs the target: fetch it before the ring wait (every warp forms g itself; same addresses -> L1 hits)
    float tgt = 0.f, eg = 0.f;
    if (training && valid) {
      if (target) tgt = target[row];
      if (extra_g) eg = extra_g[row];
    }
    cp_async_wait<Cfg::NSTAGE - 2>();
    __syncthreads();   // this unit's slabs have landed for every thread; the stage refilled below is no longer read
    issue(unit + (Cfg::NSTAGE - 1) * gridDim.x, (it + Cfg::NSTAGE - 1) % Cfg::NSTAGE);
    const uint8_t* sS = smem + stage * Cfg::STAGE_BYTES;
    const uint8_t* sC = sS + Cfg::SLAB;
    uint4 sv[KPW];
    float pd = 0.f;
#pragma unroll
    for (int i = 0; i < KPW; ++i) {
      const int kg = warp + 16 * i;
      sv[i] = *reinterpret_cast<const uint4*>(sS + (kg * 32 + lane) * 16);
      const uint32_t uw[4] = {sv[i].x, sv[i].y, sv[i].z, sv[i].w};
#pragma unroll
      for (int h = 0; h < 4; ++h) {
        const float2 f = unpack_op2(uw[h]);
        pd = __fmaf_rn(f.x, s_w[kg * 8 + 2 * h], pd);
        pd = __fmaf_rn(f.y, s_w[kg * 8 + 2 * h + 1], pd);
      }
      if (LO) {
        const uint4 lo4 = *reinterpret_cast<const uint4*>(sS + 2 * Cfg::SLAB + (kg * 32 + lane) * 16);
        const uint32_t lw[4] = {lo4.x, lo4.y, lo4.z, lo4.w};
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          const float2 f = unpack_op2(lw[h]);
          pd = __fmaf_rn(f.x, s_w[kg * 8 + 2 * h], pd);
          pd = __fmaf_rn(f.y, s_w[kg * 8 + 2 * h + 1], pd);
        }
      }
    }
    s_dot[warp * 32 + lane] = pd;
    __syncthreads();
    float dot = 0.f;
#pragma unroll
    for (int q = 0; q < 16; ++q) dot += s_dot[q * 32 + lane];
    const float z = dot + bias;
    float o = z, dact = 1.f;
    if (!outermost_linear) {
      float sn, cs;
      sincosf(omega_hidden * z, &sn, &cs);
      o = sn;
      dact = omega_hidden * cs;
    }
    if (warp == 0 && out && valid) out[row] = o;
    if (!training) continue;
    float g = 0.f;
    if (valid) {
      if (target) {
        const float diff = o - tgt;
        if (loss_kind == INR_LOSS_MAE) {
          if (warp == 0) loss_local += (double)fabsf(diff);
          g = grad_scale * ((diff > 0.f) ? 1.f : ((diff < 0.f) ? -1.f : 0.f));
        } else {
          if (warp == 0) loss_local += (double)diff * (double)diff;
          g = grad_scale * 2.f * diff;
        }
      }
      g = (g + eg) * dact;
    }
    if (warp == 0) { gb_local += g; gmax_local = fmaxf(gmax_local, fabsf(g)); }
    const float gwm = g * gs * omega_prev;
    const size_t tile_off = (size_t)tile * (TILE_ROWS * (size_t)H);
#pragma unroll
    for (int i = 0; i < KPW; ++i) {
      const int kg = warp + 16 * i;
      const uint4 cu = *reinterpret_cast<const uint4*>(sC + (kg * 32 + lane) * 16);
      const uint32_t cw[4] = {cu.x, cu.y, cu.z, cu.w};
      const uint32_t uw[4] = {sv[i].x, sv[i].y, sv[i].z, sv[i].w};
      const size_t off = tile_off + ((size_t)kg * TILE_ROWS + r) * 8;
      uint32_t ow[4];
#pragma unroll
      for (int h = 0; h < 4; ++h) {
        const float2 c = unpack_op2(cw[h]);
        const float2 f = unpack_op2(uw[h]);
        ow[h] = pack_op2(gwm * s_w[kg * 8 + 2 * h] * c.x, gwm * s_w[kg * 8 + 2 * h + 1] * c.y);
        acc_w[i][2 * h] = __fmaf_rn(g, f.x, acc_w[i][2 * h]);
        acc_w[i][2 * h + 1] = __fmaf_rn(g, f.y, acc_w[i][2 * h + 1]);
      }
      *reinterpret_cast<uint4*>(CdZ + off) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
      if (SNAKE) {   // layer L is a Snake layer: grad a[f] = w[f] * sum_r g[r] (dy/da)[r][f]
        const uint4 eu = *reinterpret_cast<const uint4*>(Eimg + off);
        const uint32_t ew[4] = {eu.x, eu.y, eu.z, eu.w};
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          const float2 e = unpack_op2(ew[h]);
          acc_a[i][2 * h] = __fmaf_rn(g, e.x, acc_a[i][2 * h]);
          acc_a[i][2 * h + 1] = __fmaf_rn(g, e.y, acc_a[i][2 * h + 1]);
        }
      }
    }
  }
  cp_async_wait<0>();
  if (!training) return;
#pragma unroll
  for (int i = 0; i < KPW; ++i) {
    const int kg = warp + 16 * i;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float sw = warp_sum(acc_w[i][j]);
      if (lane == j) atomicAdd(&gw[kg * 8 + j], sw);
      if (SNAKE) {
        const float sa = warp_sum(acc_a[i][j]);
        if (lane == j) atomicAdd(&ga[kg * 8 + j], s_w[kg * 8 + j] * sa);
      }
    }
  }
  if (warp == 0) {
    loss_local = warp_sum(loss_local);
    gb_local = warp_sum(gb_local);
    for (int o = 16; o > 0; o >>= 1) gmax_local = fmaxf(gmax_local, __shfl_xor_sync(0xffffffffu, gmax_local, o));
    if (lane == 0) {
      if (target) atomicAdd(loss_acc, loss_local);
      atomicAdd(gb, gb_local);
      if (gmax_acc) atomicMax(gmax_acc, __float_as_uint(gmax_local));   // non-negative floats order as unsigned ints
    }
  }
}

// ------------------------------------------------------------------------------------------------ loss scale (stand-alone launches)
__global__ void __launch_bounds__(256) k_tc_absmax(const float* __restrict__ x, int64_t n, unsigned* __restrict__ out) {
  float m = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    m = fmaxf(m, fabsf(x[i]));
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(out, __float_as_uint(m));
}
__global__ void k_tc_gscale(const inr_opt_state* __restrict__ opt, const unsigned* __restrict__ extra_max, float grad_scale,
                            int has_target, float* __restrict__ gsc) {
  gscale_compute(opt, extra_max ? __uint_as_float(*extra_max) : 0.f, grad_scale, has_target, gsc);
}

// ------------------------------------------------------------------------------------------------ column sums
// out_plain[f] += sum_r X[r][f]          (bias gradients)
// out_w[f*wstride + j] += sum_r wrow_j[r] X[r][f]   (last-layer weight: wrow = gvec; first layer: wrow = coordinates)
// A warp owns one feature group for a span of tiles: lane = row % 32, 16-byte loads (512 B per warp instruction),
// 8 register accumulators per output, ONE butterfly reduction at the end, then atomics.
template <int NW>   // number of weighted outputs per feature: 0, 1 or 2
__global__ void __launch_bounds__(256) k_tc_colsum(const op16* __restrict__ X, int H, int num_tiles,
                                                   int tiles_per_task, const float* __restrict__ wrow,
                                                   const float* __restrict__ coords, int in, int64_t rows, int64_t g0,
                                                   int64_t T_total, float scale, float* __restrict__ out_plain,
                                                   float* __restrict__ out_w, int wstride,
                                                   const float* __restrict__ gsc) {
  const float inv_gs = gsc ? gsc[1] : 1.f;   // X is a loss-scaled dZ-like image
  const int groups = H >> 3;
  const int lane = threadIdx.x & 31;
  const int64_t warp_id = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int tasks_per_group = (num_tiles + tiles_per_task - 1) / tiles_per_task;
  const int64_t ntasks = (int64_t)groups * tasks_per_group;
  for (int64_t task = warp_id; task < ntasks; task += nwarps) {
    const int kg = (int)(task % groups);
    const int t0 = (int)(task / groups) * tiles_per_task;
    const int t1 = min(num_tiles, t0 + tiles_per_task);
    float acc[8], accw[NW > 0 ? NW : 1][8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      acc[j] = 0.f;
#pragma unroll
      for (int q = 0; q < (NW > 0 ? NW : 1); ++q) accw[q][j] = 0.f;
    }
    for (int tile = t0; tile < t1; ++tile) {
      const op16* base = X + (size_t)tile * (TILE_ROWS * (size_t)H) + (size_t)kg * TILE_ROWS * 8;
#pragma unroll
      for (int rr = 0; rr < TILE_ROWS; rr += 32) {
        const int r = rr + lane;
        const int64_t row = (int64_t)tile * TILE_ROWS + r;
        const uint4 u = *reinterpret_cast<const uint4*>(base + (size_t)r * 8);
        const uint32_t uw[4] = {u.x, u.y, u.z, u.w};
        float wv[NW > 0 ? NW : 1];
        if (NW > 0) {
          if (wrow) {
            wv[0] = wrow[row];
          } else {   // first layer: the weights are the (generated or given) coordinates; pad rows carry dZ = 0
            const int64_t rc = row < rows ? row : rows - 1;
            if (coords) {
              wv[0] = coords[rc * in];
              if (NW > 1) wv[NW - 1] = coords[rc * in + 1];
            } else {
              wv[0] = linspace_at(g0 + rc, T_total, scale);
            }
          }
        }
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          const float2 f = unpack_op2(uw[h]);
          acc[2 * h] += f.x;
          acc[2 * h + 1] += f.y;
#pragma unroll
          for (int q = 0; q < NW; ++q) {
            accw[q][2 * h] = __fmaf_rn(wv[q], f.x, accw[q][2 * h]);
            accw[q][2 * h + 1] = __fmaf_rn(wv[q], f.y, accw[q][2 * h + 1]);
          }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float s = warp_sum(acc[j]) * inv_gs;
      if (out_plain && lane == j) atomicAdd(&out_plain[kg * 8 + j], s);
#pragma unroll
      for (int q = 0; q < NW; ++q) {
        const float sw = warp_sum(accw[q][j]) * inv_gs;
        if (lane == j) atomicAdd(&out_w[(int64_t)(kg * 8 + j) * wstride + q], sw);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------ host side
bool siren_tc_supported(const inr_siren_desc* d) {
  const int H = d->hidden_features;
  const int Ltot = d->hidden_layers + d->num_snake + d->num_tanh;
  return (H == 128 || H == 256 || H == 512) && Ltot >= 1 && d->out_features == 1 &&
         (d->in_features <= 2 || d->in_features <= H);   // encoded inputs: zero-padded to H (k_tc_rows_to_image)
}

// columns of the packed weight slices ([slice][K/8][ns][8]): one slice per CTA of a pair
int tc_ns(int H) { return H >= 256 ? 128 : H / 2; }
static int ns_of(int H) { return tc_ns(H); }
// INR_PREC_FP16S: forward GEMM operands as fp16 hi + lo pairs (k_tc_layer2<.., SPLIT>)
static bool is_split(const inr_siren_desc* d) { return d->precision == INR_PREC_FP16S; }
// columns per CTA of the FORWARD operand slices
static int ns_fwd(int H, bool split) { return split ? (H >= 512 ? 64 : H / 2) : ns_of(H); }
// power-of-two scale of the packed FORWARD weights (k_tc_pack): keeps the lo halves of the split mode out of fp16's subnormals
static float wscale_fwd(bool split) { return split ? 64.f : 1.f; }

struct TcPlan {
  int64_t pack_bytes, img_bytes_per_row, fixed_bytes;
  int n_images;
  int x_image;     // wide inputs: index of the input image X, else -1
  int lo_image;    // operand split: first of the two ping-pong lo images, else -1
  int xlo_image;   // operand split + wide inputs: lo image of X, else -1
};

static TcPlan plan_of(const inr_siren_desc* d, int training) {
  TcPlan pl;
  const SirenLayout lay = layout_of(d);
  const int64_t H = lay.H, L = lay.L;
  const bool wide = lay.in > 2, split = is_split(d);
  const int64_t fw = split ? 2 : 1;                    // forward operands: hi + lo
  // forward + dX operands of the L layers, + padded W_0 operand and padded fp32 gW_0 scratch (wide inputs)
  pl.pack_bytes = (fw + 1) * L * H * H * 2 + (wide ? fw * H * H * 2 + H * H * 4 : 0);
  int n = training ? 2 * (int)(L + 1) + lay.num_snake_layers() : 2;   // S_l, C_l, dy/da images | two ping-pong images
  pl.x_image = wide ? n++ : -1;
  pl.lo_image = split ? n : -1;
  if (split) n += 2;
  pl.xlo_image = (split && wide) ? n++ : -1;
  pl.n_images = n;
  pl.img_bytes_per_row = pl.n_images * H * 2;
  pl.fixed_bytes = round_up(pl.pack_bytes, 1024) + 4096;
  return pl;
}

int64_t siren_tc_workspace_bytes(const inr_siren_desc* d, int64_t n_rows, int training) {
  const TcPlan pl = plan_of(d, training);
  const int64_t cap = (int64_t)48 << 30;   // row chunks beyond this (180 GB parts hold a 60 s clip at 5x512 in one chunk)
  int64_t rows = round_up(std::max<int64_t>(n_rows, 1), TILE_ROWS);
  if (rows * pl.img_bytes_per_row > cap) rows = std::max<int64_t>(cap / pl.img_bytes_per_row / TILE_ROWS, 8) * TILE_ROWS;
  return pl.fixed_bytes + rows * pl.img_bytes_per_row + 1024;
}

template <int H>
static int launch_dw(const DwParams& p, cudaStream_t st) {
  using Cfg = DwCfg<H>;
  static bool attr_set = false;
  if (!attr_set) {
    INR_CUDA_CHECK(cudaFuncSetAttribute(k_tc_dw<H>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    attr_set = true;
  }
  int grid = num_sms() / Cfg::NBLOCKS * Cfg::NBLOCKS;
  const int max_useful = (int)std::min<int64_t>((int64_t)p.num_tiles * Cfg::NBLOCKS, 1 << 30);
  if (grid > max_useful) grid = max_useful;
  { LaunchScope ls("k_tc_dw", st);
    k_tc_dw<H><<<grid, Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(p); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

template <int H, int MODE, bool SPLIT = false>
static int launch_layer2(const LayerParams& p, cudaStream_t st) {
  using Cfg = Layer2Cfg<H, SPLIT>;
  static bool attr_set = false;
  if (!attr_set) {
    INR_CUDA_CHECK(cudaFuncSetAttribute(k_tc_layer2<H, MODE, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    attr_set = true;
  }
  int pairs = (num_sms() / 2) / Cfg::NQ * Cfg::NQ;
  const int max_useful = ((p.num_tiles + 1) / 2) * Cfg::NQ;
  if (pairs > max_useful) pairs = max_useful;
  { LaunchScope ls(mode_is_fwd(MODE) ? "k_tc_layer_fwd" : "k_tc_layer_bwd", st);
    k_tc_layer2<H, MODE, SPLIT><<<2 * pairs, Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(p); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

template <int MODE, bool SPLIT = false>
static int launch_layer2_h(int H, const LayerParams& p, cudaStream_t st) {
  switch (H) {
    case 128: return launch_layer2<128, MODE, SPLIT>(p, st);
    case 256: return launch_layer2<256, MODE, SPLIT>(p, st);
    case 512: return launch_layer2<512, MODE, SPLIT>(p, st);
  }
  set_error("tcgen05 layer kernel: unsupported H=%d", H);
  return INR_E_UNSUPPORTED;
}

int tc_layer(int H, int mode, const LayerParams& p, cudaStream_t st) {
  if (p.A_lo) {   // operand split (forward modes): A = A + A_lo, Bpack = [hi | lo] slices, out0_lo written
    INR_REQUIRE(mode_is_fwd(mode) && p.out0_lo, INR_E_BADARG, "operand split: forward layer modes with an out0_lo image only");
    if (mode == MODE_FWD_SNAKE) return launch_layer2_h<MODE_FWD_SNAKE, true>(H, p, st);
    if (mode == MODE_FWD_TANH) return launch_layer2_h<MODE_FWD_TANH, true>(H, p, st);
    return launch_layer2_h<MODE_FWD, true>(H, p, st);
  }
  if (mode == MODE_FWD_SNAKE) return launch_layer2_h<MODE_FWD_SNAKE>(H, p, st);
  if (mode == MODE_FWD_TANH) return launch_layer2_h<MODE_FWD_TANH>(H, p, st);
  if (mode == MODE_BWD_SNAKE) return launch_layer2_h<MODE_BWD_SNAKE>(H, p, st);
  if (mode == MODE_FWD) return launch_layer2_h<MODE_FWD>(H, p, st);
  return launch_layer2_h<MODE_BWD>(H, p, st);
}

template <int H>
static int launch_dw2(const DwParams& p, cudaStream_t st) {
  using Cfg = Dw2Cfg<H>;
  static bool attr_set = false;
  if (!attr_set) {
    INR_CUDA_CHECK(cudaFuncSetAttribute(k_tc_dw2<H>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    attr_set = true;
  }
  int pairs = (num_sms() / 2) / Cfg::NBLOCKS * Cfg::NBLOCKS;
  const int max_useful = (int)std::min<int64_t>((int64_t)p.num_tiles * Cfg::NBLOCKS, 1 << 30);
  if (pairs > max_useful) pairs = max_useful;
  { LaunchScope ls("k_tc_dw", st);
    k_tc_dw2<H><<<2 * pairs, Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(p); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}


// H = 512: the CTA-pair kernel k_tc_dw2 (64 KB instead of 96 KB of operands per tile and CTA; -15 % on the five dW
// launches of a C2 step); H <= 256: the single-CTA kernel is faster (C1 0.203 vs 0.223 ms, profiles/r01c_notes.md).
int tc_dw(int H, const DwParams& p, cudaStream_t st) {
  switch (H) {
    case 128: return launch_dw<128>(p, st);
    case 256: return launch_dw<256>(p, st);
    case 512: return launch_dw2<512>(p, st);
  }
  set_error("tcgen05 dW kernel: unsupported H=%d", H);
  return INR_E_UNSUPPORTED;
}

static bool fused_ok(const inr_siren_desc* d) {
  // fp16: k_tc_fused_fwd (two ping-pong slots); fp16s: k_tc_fused_fwd_s (one hi + lo slot, layers pipelined by K chunk)
  return (d->hidden_features == 128 || d->hidden_features == 256) &&
         d->hidden_layers >= 1 && d->hidden_layers <= FUSED_MAX_L && d->num_snake == 0 && d->num_tanh == 0 &&
         d->first_linear == 0 && d->in_features <= 2;
}

template <int H, int EW>
static int launch_fused(const FusedParams& p, cudaStream_t st) {
  using Cfg = FusedCfg<H>;
  static bool attr_set = false;
  if (!attr_set) {
    INR_CUDA_CHECK(cudaFuncSetAttribute(k_tc_fused_fwd<H, EW>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    attr_set = true;
  }
  int pairs = num_sms() / 2;
  const int num_dt = (p.num_tiles + 1) / 2;
  if (pairs > num_dt) pairs = num_dt;
  { LaunchScope ls("k_tc_fused_fwd", st);
    k_tc_fused_fwd<H, EW><<<2 * pairs, 64 + 32 * EW, Cfg::SMEM_BYTES, st>>>(p); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

// first layer + all hidden layers of one row chunk in ONE kernel; S[l] / C[l] (l = 0..L) are the stash images to
// write (NULL entries are skipped: inference keeps only S_L)
// hand-off trace buffer of k_tc_fused_fwd (allocated by inr_dbg_fused_trace_enable; NULL = tracing off)
static long long* g_fused_trace = nullptr;
constexpr int kFusedTraceJobs = 96;

struct FusedLast {   // arguments of the fused head + loss + dZ_L epilogue (FusedParams::fuse_last)
  const float* target; const float* extra_g; int loss_kind; float grad_scale; float* grads; double* loss_acc; float* out;
  const float* gsc; unsigned* gmax_acc;
};

template <int H>
static int launch_fused_s(const FusedSParams& p, cudaStream_t st) {
  using Cfg = FusedSCfg<H>;
  static bool attr_set = false;
  if (!attr_set) {
    INR_CUDA_CHECK(cudaFuncSetAttribute(k_tc_fused_fwd_s<H>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    attr_set = true;
  }
  int pairs = num_sms() / 2;
  const int num_dt = (p.f.num_tiles + 1) / 2;
  if (pairs > num_dt) pairs = num_dt;
  { LaunchScope ls("k_tc_fused_fwd", st);
    k_tc_fused_fwd_s<H><<<2 * pairs, Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(p); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

static int fused_forward(const inr_siren_desc* d, const float* params, const op16* Wf, const float* coords,
                         int64_t rows, int tiles, int64_t g0, int64_t T_total, float scale, op16* const* S,
                         op16* const* C, const FusedLast* fl, cudaStream_t st, op16* S_lo_last = nullptr) {
  SirenLayout lay = layout_of(d);
  FusedParams p;
  p.Wf = Wf; p.W0 = params + lay.w_off(0); p.B0 = params + lay.b_off(0);
  for (int l = 0; l <= FUSED_MAX_L; ++l) { p.S[l] = nullptr; p.C[l] = nullptr; }
  for (int l = 1; l <= lay.L; ++l) p.bias[l - 1] = params + lay.b_off(l);
  for (int l = lay.L; l < FUSED_MAX_L; ++l) p.bias[l] = nullptr;
  for (int l = 0; l <= lay.L; ++l) { p.S[l] = S[l]; p.C[l] = C ? C[l] : nullptr; }
  p.coords = coords; p.in = lay.in; p.L = lay.L; p.num_tiles = tiles; p.rows = rows; p.g0 = g0; p.T_total = T_total;
  p.scale = scale; p.omega0 = d->first_omega_0; p.omega_h = d->hidden_omega_0;
  p.fuse_last = fl ? 1 : 0; p.loss_kind = fl ? fl->loss_kind : 0; p.grad_scale = fl ? fl->grad_scale : 0.f;
  p.w_last = params + lay.w_off(lay.L + 1); p.b_last = params + lay.b_off(lay.L + 1);
  p.target = fl ? fl->target : nullptr; p.extra_g = fl ? fl->extra_g : nullptr; p.out = fl ? fl->out : nullptr;
  p.g_wlast = fl ? fl->grads + lay.w_off(lay.L + 1) : nullptr; p.g_blast = fl ? fl->grads + lay.b_off(lay.L + 1) : nullptr;
  p.loss_acc = fl ? fl->loss_acc : nullptr;
  p.gsc = fl ? fl->gsc : nullptr; p.gmax_acc = fl ? fl->gmax_acc : nullptr;
  p.trace = g_fused_trace; p.trace_jobs = g_fused_trace ? kFusedTraceJobs : 0;
  if (is_split(d)) {
    FusedSParams ps;
    ps.f = p; ps.f.trace = nullptr; ps.f.trace_jobs = 0;
    ps.S_lo_last = fl ? nullptr : S_lo_last;
    ps.acc_scale = 1.f / wscale_fwd(true);
    return lay.H == 256 ? launch_fused_s<256>(ps, st) : launch_fused_s<128>(ps, st);
  }
  return lay.H == 256 ? launch_fused<256, 16>(p, st) : launch_fused<128, 16>(p, st);
}

static LayerParams layer_params(int H, const op16* A, const op16* Bp, const float* bias, float omega,
                                op16* out0, op16* out1, const op16* Cin, int tiles, int ns = 0) {
  LayerParams p;
  p.A = A; p.Bpack = Bp; p.bias = bias; p.omega = omega; p.out0 = out0; p.out1 = out1; p.Cin = Cin; p.num_tiles = tiles;
  p.avec = nullptr; p.out2 = nullptr;
  p.a_lbo = TILE_ROWS * 16; p.a_sbo = 128;            // K-major, no swizzle: LBO = next 8-feature group, SBO = next 8 rows
  p.b_lbo = (uint32_t)(ns ? ns : ns_of(H)) * 16; p.b_sbo = 128;
  return p;
}

static int colsum(const op16* X, int H, int tiles, const float* wrow, const float* coords, int in, int64_t rows,
                  int64_t g0, int64_t T_total, float scale, float* out_plain, float* out_w, int wstride, int nw,
                  const float* gsc, cudaStream_t st) {
  const int groups = H / 8;
  int tiles_per_task = std::max(1, (int)cdiv((int64_t)tiles * groups, (int64_t)num_sms() * 32));
  const int64_t ntasks = (int64_t)groups * cdiv(tiles, tiles_per_task);
  int blocks = (int)std::min<int64_t>(cdiv(ntasks, 8), (int64_t)num_sms() * 8);
  if (blocks < 1) blocks = 1;
  { LaunchScope ls(nw == 0 ? "k_tc_colsum" : "k_tc_colsum_w", st);
    if (nw == 0)
      k_tc_colsum<0><<<blocks, 256, 0, st>>>(X, H, tiles, tiles_per_task, wrow, coords, in, rows, g0, T_total, scale, out_plain, out_w, wstride, gsc);
    else if (nw == 1)
      k_tc_colsum<1><<<blocks, 256, 0, st>>>(X, H, tiles, tiles_per_task, wrow, coords, in, rows, g0, T_total, scale, out_plain, out_w, wstride, gsc);
    else
      k_tc_colsum<2><<<blocks, 256, 0, st>>>(X, H, tiles, tiles_per_task, wrow, coords, in, rows, g0, T_total, scale, out_plain, out_w, wstride, gsc); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

struct TcWorkspace {
  op16 *Wf, *Wb;
  op16* Wf0;   // wide inputs: W_0 zero-padded to H x H (forward operand)
  float* gW0;           // wide inputs: H x H fp32 scratch of the padded first-layer weight gradient
  int x_image;          // wide inputs: index of the input image X, else -1
  int lo_image, xlo_image;   // operand split: see TcPlan
  bool split;
  float* gsc;           // {gs, 1/gs}: loss scale of the fp16 dZ images of this call (k_tc_gscale); [2] = max |extra_g| bits
  op16* img;
  int64_t chunk_rows;   // multiple of 128
  int64_t img_elems;    // elements per image
};

static int carve(const inr_siren_desc* d, int training, int64_t n, void* ws, int64_t ws_bytes, TcWorkspace* w) {
  const TcPlan pl = plan_of(d, training);
  const int64_t H = d->hidden_features, L = layout_of(d).L;
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(ws) + 255) & ~(uintptr_t)255);
  const int64_t usable = ws_bytes - (base - reinterpret_cast<uint8_t*>(ws));
  w->split = is_split(d);
  const int64_t fw = w->split ? 2 : 1;
  w->Wf = reinterpret_cast<op16*>(base);
  w->Wb = w->Wf + fw * L * H * H;
  w->Wf0 = w->Wb + L * H * H;
  w->gW0 = reinterpret_cast<float*>(w->Wf0 + fw * H * H);
  w->x_image = pl.x_image; w->lo_image = pl.lo_image; w->xlo_image = pl.xlo_image;
  int64_t avail = usable - pl.fixed_bytes;
  int64_t rows = avail / pl.img_bytes_per_row / TILE_ROWS * TILE_ROWS;
  INR_REQUIRE(rows >= TILE_ROWS, INR_E_WORKSPACE, "workspace of %lld bytes is too small for one 128-row tile", (long long)ws_bytes);
  rows = std::min<int64_t>(rows, round_up(n, TILE_ROWS));
  w->chunk_rows = rows;
  w->img_elems = rows * H;
  uint8_t* p = base + pl.fixed_bytes;
  w->gsc = reinterpret_cast<float*>(p - 256);   // inside the 4 KB pad that follows the packed weights
  w->img = reinterpret_cast<op16*>(p);
  return INR_OK;
}

struct GscaleFold { const inr_opt_state* opt; float grad_scale; int has_target; };

static int pack_weights(const inr_siren_desc* d, const float* params, const TcWorkspace& w, cudaStream_t st,
                        const GscaleFold* gf = nullptr) {
  SirenLayout lay = layout_of(d);
  const int64_t total = (int64_t)lay.H * lay.H * lay.L;
  int blocks = (int)std::min<int64_t>(cdiv(total, 256), (int64_t)num_sms() * 8);
  { LaunchScope ls("k_tc_pack", st);
    k_tc_pack<<<blocks, 256, 0, st>>>(params, lay, ns_fwd(lay.H, w.split), ns_of(lay.H), w.split ? 1 : 0, wscale_fwd(w.split), w.Wf, w.Wb,
                                      gf ? gf->opt : nullptr, gf ? gf->grad_scale : 0.f, gf ? gf->has_target : 0,
                                      gf ? w.gsc : nullptr); }
  INR_LAUNCH_CHECK();
  if (lay.in > 2) {
    LaunchScope ls("k_tc_pack", st);
    k_tc_pack0<<<(lay.H * lay.H + 255) / 256, 256, 0, st>>>(params + lay.w_off(0), lay.in, lay.H, ns_fwd(lay.H, w.split),
                                                            w.split ? 1 : 0, wscale_fwd(w.split), w.Wf0);
    INR_LAUNCH_CHECK();
  }
  return INR_OK;
}

static int first_layer(const inr_siren_desc* d, const float* params, const float* coords, int64_t rows, int64_t rows_pad,
                       int64_t g0, int64_t T_total, float scale, op16* S0, op16* C0, cudaStream_t st,
                       op16* E0 = nullptr, op16* S0lo = nullptr) {
  SirenLayout lay = layout_of(d);
  const int64_t total = rows_pad * (lay.H / 8);
  int blocks = (int)std::min<int64_t>(cdiv(total, 256), (int64_t)num_sms() * 16);
  { LaunchScope ls("k_tc_first", st);
    k_tc_first<<<blocks, 256, 0, st>>>(params + lay.w_off(0), params + lay.b_off(0), coords, lay.in, lay.H, rows, rows_pad,
                                       g0, T_total, scale, d->first_omega_0, S0, C0,
                                       lay.first_snake ? params + lay.a_off(0) : nullptr, E0, S0lo); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

template <int H, bool SNAKE>
static int launch_last2(const op16* SL, op16* CL, const float* w, const float* b, const float* target,
                        const float* extra_g, int64_t rows, int tiles, float dfac, float omega_h, int outermost_linear,
                        int loss_kind, float grad_scale, float* gw, float* gb, double* loss_acc, float* out, int training,
                        const op16* EL, float* ga, const float* gsc, unsigned* gmax_acc, const op16* SLlo,
                        cudaStream_t st) {
  const int blocks = std::max(1, std::min(tiles * 4, num_sms()));
  if (SLlo) {
    using Cfg = Last2Cfg<H, true>;
    static bool attr_done = false;
    if (!attr_done) {
      INR_CUDA_CHECK(cudaFuncSetAttribute(k_tc_last2<H, SNAKE, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
      attr_done = true;
    }
    LaunchScope ls("k_tc_last", st);
    k_tc_last2<H, SNAKE, true><<<blocks, 512, Cfg::SMEM_BYTES, st>>>(SL, CL, w, b, target, extra_g, rows, tiles, dfac, omega_h,
                                                                     outermost_linear, loss_kind, grad_scale, gw, gb, loss_acc,
                                                                     out, training, EL, ga, gsc, gmax_acc, SLlo);
  } else {
    using Cfg = Last2Cfg<H, false>;
    static bool attr_done = false;
    if (!attr_done) {
      INR_CUDA_CHECK(cudaFuncSetAttribute(k_tc_last2<H, SNAKE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
      attr_done = true;
    }
    LaunchScope ls("k_tc_last", st);
    k_tc_last2<H, SNAKE, false><<<blocks, 512, Cfg::SMEM_BYTES, st>>>(SL, CL, w, b, target, extra_g, rows, tiles, dfac, omega_h,
                                                                      outermost_linear, loss_kind, grad_scale, gw, gb, loss_acc,
                                                                      out, training, EL, ga, gsc, gmax_acc, nullptr);
  }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

// out = act(S_L w + b); training: loss, g = dloss/dout, dZ_L over C_L, last-layer gradients (and the `a` gradient of a
// Snake layer L from its dy/da image EL)
static int last_layer(const inr_siren_desc* d, const float* params, const op16* SL, op16* CL,
                      const float* target, const float* extra_g, int64_t rows, int tiles, int loss_kind, float grad_scale,
                      const float* gsc, float* grads, inr_opt_state* opt, float* out, int training, cudaStream_t st,
                      op16* EL = nullptr, const op16* SLlo = nullptr) {
  SirenLayout lay = layout_of(d);
  const int last = lay.L + 1;
  const int H = lay.H;
  // sine layers keep cos(omega z) in C_L (times omega here); Snake / Tanh layers keep the whole derivative
  const float dfac = lay.kind(lay.L) == ACT_SINE ? d->hidden_omega_0 : 1.f;
  const float* w = params + lay.w_off(last);
  const float* b = params + lay.b_off(last);
  float* gw = grads ? grads + lay.w_off(last) : nullptr;
  float* gb = grads ? grads + lay.b_off(last) : nullptr;
  double* loss_acc = opt ? &opt->loss_acc[0] : nullptr;
  unsigned* gmax = (opt && training) ? reinterpret_cast<unsigned*>(&opt->gmax_acc) : nullptr;
  if (!training) EL = nullptr;
  float* ga = (EL && grads) ? grads + lay.a_off(lay.L) : nullptr;
#define INR_LAST2(HH)                                                                                                   \
  return EL ? launch_last2<HH, true>(SL, CL, w, b, target, extra_g, rows, tiles, dfac, d->hidden_omega_0,              \
                                     d->outermost_linear, loss_kind, grad_scale, gw, gb, loss_acc, out, training, EL,   \
                                     ga, gsc, gmax, SLlo, st)                                                           \
            : launch_last2<HH, false>(SL, CL, w, b, target, extra_g, rows, tiles, dfac, d->hidden_omega_0,             \
                                      d->outermost_linear, loss_kind, grad_scale, gw, gb, loss_acc, out, training,      \
                                      nullptr, nullptr, gsc, gmax, SLlo, st)
  if (H == 512) { INR_LAST2(512); }
  if (H == 256) { INR_LAST2(256); }
  INR_LAST2(128);
#undef INR_LAST2
}

// forward of hidden layer l (1..L) by its activation kind: Sout = act(Sin W^T + b), Cout = derivative image,
// Eout = dy/da image (Snake layers)
static int fwd_layer_tc(const inr_siren_desc* d, const SirenLayout& lay, const float* params, const TcWorkspace& w, int l,
                        const op16* Sin, op16* Sout, op16* Cout, op16* Eout, int tiles,
                        cudaStream_t st, const op16* Sin_lo = nullptr, op16* Sout_lo = nullptr) {
  const int H = lay.H;
  const int kind = lay.kind(l);
  const int64_t fw = w.split ? 2 : 1;
  LayerParams p = layer_params(H, Sin, w.Wf + (int64_t)(l - 1) * fw * H * H, params + lay.b_off(l),
                               kind == ACT_SINE ? d->hidden_omega_0 : 1.f, Sout, Cout, nullptr, tiles, ns_fwd(H, w.split));
  p.A_lo = Sin_lo; p.out0_lo = Sout_lo; p.acc_scale = 1.f / wscale_fwd(w.split);
  if (kind == ACT_SNAKE) { p.avec = params + lay.a_off(l); p.out2 = Eout; }
  return tc_layer(H, kind == ACT_SINE ? MODE_FWD : (kind == ACT_SNAKE ? MODE_FWD_SNAKE : MODE_FWD_TANH), p, st);
}

// layer 0 on an encoded coordinate: X image from the rows, then the layer GEMM with the padded W_0
static int first_layer_wide(const inr_siren_desc* d, const SirenLayout& lay, const float* params, const TcWorkspace& w,
                            const float* xc, int64_t rows, int64_t rows_pad, op16* X, op16* S0,
                            op16* C0, op16* E0, int tiles, cudaStream_t st, op16* Xlo = nullptr, op16* S0lo = nullptr) {
  const int H = lay.H;
  const int64_t total = rows_pad * (H / 8);
  const int blocks = (int)std::min<int64_t>(cdiv(total, 256), (int64_t)num_sms() * 16);
  { LaunchScope ls("k_tc_rows_to_image", st);
    k_tc_rows_to_image<<<blocks, 256, 0, st>>>(xc, lay.in, H, rows, rows_pad, X, Xlo); }
  INR_LAUNCH_CHECK();
  const int kind = lay.kind(0);
  LayerParams p = layer_params(H, X, w.Wf0, params + lay.b_off(0), kind == ACT_SINE ? d->first_omega_0 : 1.f, S0, C0,
                               nullptr, tiles, ns_fwd(H, w.split));
  p.A_lo = Xlo; p.out0_lo = S0lo; p.acc_scale = 1.f / wscale_fwd(w.split);
  if (kind == ACT_SNAKE) { p.avec = params + lay.a_off(0); p.out2 = E0; }
  return tc_layer(H, kind == ACT_SINE ? MODE_FWD : MODE_FWD_SNAKE, p, st);
}

int siren_tc_forward(const inr_siren_desc* d, const float* params, const float* coords, int64_t n, int64_t i0,
                     int64_t T_total, float scale, float* out, void* ws, int64_t ws_bytes, cudaStream_t st) {
  SirenLayout lay = layout_of(d);
  const int H = lay.H, L = lay.L;
  TcWorkspace w;
  int rc = carve(d, 0, n, ws, ws_bytes, &w);
  if (rc) return rc;
  if ((rc = pack_weights(d, params, w, st))) return rc;
  for (int64_t r0 = 0; r0 < n; r0 += w.chunk_rows) {
    const int64_t rows = std::min(w.chunk_rows, n - r0);
    const int64_t rows_pad = round_up(rows, TILE_ROWS);
    const int tiles = (int)(rows_pad / TILE_ROWS);
    op16* S[2] = {w.img, w.img + w.img_elems};
    op16* Slo[2] = {w.split ? w.img + (int64_t)w.lo_image * w.img_elems : nullptr,
                    w.split ? w.img + (int64_t)(w.lo_image + 1) * w.img_elems : nullptr};
    op16* Xlo = w.xlo_image >= 0 ? w.img + (int64_t)w.xlo_image * w.img_elems : nullptr;
    if (fused_ok(d)) {   // all layers in one kernel; only S_L goes to HBM, for the output dot below
      op16* Sl[FUSED_MAX_L + 1] = {nullptr};
      Sl[L] = S[0];
      if ((rc = fused_forward(d, params, w.Wf, coords ? coords + r0 * lay.in : nullptr, rows, tiles, i0 + r0, T_total, scale,
                              Sl, nullptr, nullptr, st, Slo[0])))
        return rc;
      if ((rc = last_layer(d, params, S[0], nullptr, nullptr, nullptr, rows, tiles, 0, 0.f, nullptr, nullptr, nullptr,
                           out + r0, 0, st, nullptr, Slo[0])))
        return rc;
      continue;
    }
    if (w.x_image >= 0) {
      rc = first_layer_wide(d, lay, params, w, coords + r0 * lay.in, rows, rows_pad, w.img + (int64_t)w.x_image * w.img_elems,
                            S[0], nullptr, nullptr, tiles, st, Xlo, Slo[0]);
    } else {
      rc = first_layer(d, params, coords ? coords + r0 * lay.in : nullptr, rows, rows_pad, i0 + r0, T_total, scale, S[0],
                       nullptr, st, nullptr, Slo[0]);
    }
    if (rc) return rc;
    int cur = 0;
    for (int l = 1; l <= L; ++l) {
      if ((rc = fwd_layer_tc(d, lay, params, w, l, S[cur], S[cur ^ 1], nullptr, nullptr, tiles, st, Slo[cur], Slo[cur ^ 1])))
        return rc;
      cur ^= 1;
    }
    if ((rc = last_layer(d, params, S[cur], nullptr, nullptr, nullptr, rows, tiles, 0, 0.f, nullptr, nullptr, nullptr,
                         out + r0, 0, st, nullptr, Slo[cur])))
      return rc;
  }
  return INR_OK;
}

// loss scale of this call's dZ images -> w.gsc (see k_tc_gscale)
static int set_gscale(const TcWorkspace& w, const inr_opt_state* opt, const float* extra_g, int64_t n, const float* target,
                      float grad_scale, cudaStream_t st) {
  unsigned* exmax = nullptr;
  if (extra_g) {
    exmax = reinterpret_cast<unsigned*>(w.gsc + 2);
    INR_CUDA_CHECK(cudaMemsetAsync(exmax, 0, sizeof(unsigned), st));
    const int blocks = (int)std::min<int64_t>(cdiv(n, 256 * 8), (int64_t)num_sms() * 4);
    LaunchScope ls("k_tc_gscale", st);
    k_tc_absmax<<<std::max(blocks, 1), 256, 0, st>>>(extra_g, n, exmax);
    INR_LAUNCH_CHECK();
  }
  { LaunchScope ls("k_tc_gscale", st);
    k_tc_gscale<<<1, 1, 0, st>>>(opt, exmax, grad_scale, target != nullptr, w.gsc); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

// phase bits: 1 = forward (activations kept in the workspace), 2 = backward.  phase 3 is the fused step (the last-layer
// kernel computes the output, the loss gradient and dZ_L in one pass); phase 1 writes the model output and stops so a
// spectral loss can be evaluated on it; phase 2 resumes from the SAME workspace with that loss's gradient in extra_gout.
int siren_tc_train(const inr_siren_desc* d, const float* params, const float* coords, const float* target,
                   int64_t n, int64_t i0, int64_t T_total, float scale, int loss_kind, float grad_scale,
                   const float* extra_gout, float* grads, inr_opt_state* opt, float* out, void* ws,
                   int64_t ws_bytes, int phase, cudaStream_t st) {
  SirenLayout lay = layout_of(d);
  const int H = lay.H, L = lay.L;
  TcWorkspace w;
  int rc = carve(d, 1, n, ws, ws_bytes, &w);
  if (rc) return rc;
  INR_REQUIRE(phase == 3 || w.chunk_rows >= n, INR_E_WORKSPACE,
              "split forward/backward needs the whole row range in one workspace chunk (%lld rows > %lld)", (long long)n,
              (long long)w.chunk_rows);
  // the fused step without an upstream gradient sets its loss scale inside the pack launch
  const bool fold_gs = phase == 3 && !extra_gout;
  const GscaleFold gf = {opt, grad_scale, target != nullptr};
  if ((phase & 1) && (rc = pack_weights(d, params, w, st, fold_gs ? &gf : nullptr))) return rc;
  if ((phase & 2) && !fold_gs && (rc = set_gscale(w, opt, extra_gout, n, target, grad_scale, st))) return rc;
  for (int64_t r0 = 0; r0 < n; r0 += w.chunk_rows) {
    const int64_t rows = std::min(w.chunk_rows, n - r0);
    const int64_t rows_pad = round_up(rows, TILE_ROWS);
    const int tiles = (int)(rows_pad / TILE_ROWS);
    auto S = [&](int l) { return w.img + (int64_t)l * w.img_elems; };
    auto C = [&](int l) { return w.img + (int64_t)(L + 1 + l) * w.img_elems; };
    // dy/da image of Snake layer l (NULL for the other kinds)
    auto E = [&](int l) -> op16* {
      return lay.kind(l) == ACT_SNAKE ? w.img + (int64_t)(2 * (L + 1) + lay.snakes_before(l)) * w.img_elems : nullptr;
    };
    // operand split: S_l's lo image is read by the forward of layer l+1 (and by the head for l = L) only: two ping-pong
    auto Slo = [&](int l) -> op16* { return w.split ? w.img + (int64_t)(w.lo_image + (l & 1)) * w.img_elems : nullptr; };
    op16* Xlo = w.xlo_image >= 0 ? w.img + (int64_t)w.xlo_image * w.img_elems : nullptr;
    const float* xc = coords ? coords + r0 * lay.in : nullptr;
    bool fused_tail = false;
    if (phase & 1) {
      if (fused_ok(d)) {   // one kernel: S_l never re-read from HBM, only the stash the backward needs is written
        op16 *Sl[FUSED_MAX_L + 1], *Cl[FUSED_MAX_L + 1];
        for (int l = 0; l <= L; ++l) { Sl[l] = S(l); Cl[l] = C(l); }
        // the fused step with the nn.Linear head also runs the head, the loss and dZ_L in the same kernel:
        // S_L / C_L never reach HBM (dZ_L is written where C_L would go)
        fused_tail = (phase == 3) && d->outermost_linear;
        FusedLast fl = {target ? target + r0 : nullptr, extra_gout ? extra_gout + r0 : nullptr, loss_kind, grad_scale, grads,
                        opt ? &opt->loss_acc[0] : nullptr, out ? out + r0 : nullptr, w.gsc,
                        opt ? reinterpret_cast<unsigned*>(&opt->gmax_acc) : nullptr};
        if (fused_tail) Sl[L] = nullptr;
        if ((rc = fused_forward(d, params, w.Wf, xc, rows, tiles, i0 + r0, T_total, scale, Sl, Cl, fused_tail ? &fl : nullptr, st,
                                Slo(L))))
          return rc;
      } else {
        if (w.x_image >= 0) {
          rc = first_layer_wide(d, lay, params, w, xc, rows, rows_pad, w.img + (int64_t)w.x_image * w.img_elems, S(0), C(0),
                                E(0), tiles, st, Xlo, Slo(0));
        } else {
          rc = first_layer(d, params, xc, rows, rows_pad, i0 + r0, T_total, scale, S(0), C(0), st, E(0), Slo(0));
        }
        if (rc) return rc;
        for (int l = 1; l <= L; ++l)
          if ((rc = fwd_layer_tc(d, lay, params, w, l, S(l - 1), S(l), C(l), E(l), tiles, st, Slo(l - 1), Slo(l)))) return rc;
      }
      if (phase == 1) {
        if ((rc = last_layer(d, params, S(L), nullptr, nullptr, nullptr, rows, tiles, 0, 0.f, nullptr, nullptr, nullptr,
                             out + r0, 0, st, nullptr, Slo(L))))
          return rc;
        continue;
      }
    }
    if (!fused_tail &&
        (rc = last_layer(d, params, S(L), C(L), target ? target + r0 : nullptr, extra_gout ? extra_gout + r0 : nullptr, rows,
                         tiles, loss_kind, grad_scale, w.gsc, grads, opt, out ? out + r0 : nullptr, 1, st, E(L), Slo(L))))
      return rc;
    for (int l = L; l >= 1; --l) {
      DwParams dp;
      dp.dZ = C(l); dp.S = S(l - 1); dp.gW = grads + lay.w_off(l); dp.gb = grads + lay.b_off(l); dp.num_tiles = tiles;
      dp.lbo = 128; dp.sbo = TILE_ROWS * 16;   // MN-major, no swizzle: LBO = next 8 rows (K), SBO = next 8-feature group
      dp.gsc = w.gsc;
      if ((rc = tc_dw(H, dp, st))) return rc;
      // dZ_{l-1} = (dZ_l W_l) * d y_{l-1}/d z_{l-1}: sine layers keep cos(omega z) (times omega here), Snake / Tanh the
      // whole derivative; a Snake layer also takes its `a` gradient from the same product against its dy/da image
      const int kprev = lay.kind(l - 1);
      const float om_prev = kprev != ACT_SINE ? 1.f : ((l - 1 == 0) ? d->first_omega_0 : d->hidden_omega_0);
      LayerParams p = layer_params(H, C(l), w.Wb + (int64_t)(l - 1) * H * H, nullptr, om_prev, C(l - 1), nullptr, C(l - 1), tiles);
      if (kprev == ACT_SNAKE) {
        p.out2 = E(l - 1);
        if ((rc = tc_layer(H, MODE_BWD_SNAKE, p, st))) return rc;
        if ((rc = colsum(E(l - 1), H, tiles, nullptr, nullptr, 1, rows, 0, 1, 1.f, grads + lay.a_off(l - 1), nullptr, 1, 0,
                         w.gsc, st)))
          return rc;
      } else if ((rc = tc_layer(H, MODE_BWD, p, st))) {
        return rc;
      }
    }
    if (w.x_image >= 0) {
      // encoded input: gW_0 = dZ_0^T X as one more dW GEMM into the padded scratch, cut back to (H x in); gb_0 with it
      INR_CUDA_CHECK(cudaMemsetAsync(w.gW0, 0, (size_t)H * H * sizeof(float), st));
      DwParams dp;
      dp.dZ = C(0); dp.S = w.img + (int64_t)w.x_image * w.img_elems; dp.gW = w.gW0; dp.gb = grads + lay.b_off(0);
      dp.num_tiles = tiles; dp.lbo = 128; dp.sbo = TILE_ROWS * 16; dp.gsc = w.gsc;
      if ((rc = tc_dw(H, dp, st))) return rc;
      { LaunchScope ls("k_tc_unpad_add", st);
        k_tc_unpad_add<<<(H * lay.in + 255) / 256, 256, 0, st>>>(w.gW0, lay.in, H, grads + lay.w_off(0)); }
      INR_LAUNCH_CHECK();
      continue;
    }
    // first layer: gW_0[f][i] += sum_r dZ_0[r][f] x[r][i], gb_0 += colsum(dZ_0)
    if ((rc = colsum(C(0), H, tiles, nullptr, xc, lay.in, rows, i0 + r0, T_total, scale, grads + lay.b_off(0),
                     grads + lay.w_off(0), lay.in, lay.in, w.gsc, st)))
      return rc;
  }
  return INR_OK;
}

int siren_tc_loss_backward(const inr_siren_desc* d, const float* params, const float* coords, const float* target,
                           int64_t n, int64_t i0, int64_t T_total, float scale, int loss_kind, float grad_scale,
                           const float* extra_gout, float* grads, inr_opt_state* opt, float* out, void* ws,
                           int64_t ws_bytes, cudaStream_t st) {
  return siren_tc_train(d, params, coords, target, n, i0, T_total, scale, loss_kind, grad_scale, extra_gout, grads, opt,
                        out, ws, ws_bytes, 3, st);
}

}  // namespace inr

// ---- debug: clock64 stamps of the slot hand-offs in pair 0 of k_tc_fused_fwd (profiles/r01c_notes.md)
extern "C" int inr_dbg_fused_trace_enable(int on) {
  using namespace inr;
  if (on && !g_fused_trace) {
    INR_CUDA_CHECK(cudaMalloc(&g_fused_trace, kFusedTraceJobs * 4 * sizeof(long long)));
    INR_CUDA_CHECK(cudaMemset(g_fused_trace, 0, kFusedTraceJobs * 4 * sizeof(long long)));
  } else if (!on && g_fused_trace) {
    cudaFree(g_fused_trace);
    g_fused_trace = nullptr;
  }
  return INR_OK;
}
extern "C" int inr_dbg_fused_trace_read(long long* host_out, int n) {
  using namespace inr;
  INR_REQUIRE(g_fused_trace && host_out && n > 0, INR_E_BADARG, "trace not enabled");
  INR_CUDA_CHECK(cudaDeviceSynchronize());
  INR_CUDA_CHECK(cudaMemcpy(host_out, g_fused_trace, std::min(n, kFusedTraceJobs * 4) * sizeof(long long), cudaMemcpyDeviceToHost));
  return INR_OK;
}
