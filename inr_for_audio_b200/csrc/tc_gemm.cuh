// tc_gemm.cuh -- the two tcgen05 GEMM kernels of the fp16 SIREN path (sm_100a).
//
//  k_tc_layer<H, MODE>   weight-stationary layer GEMM over 128-row activation tiles
//      MODE_FWD :  S_l, C_l  = sincos( omega * (S_{l-1} W_l^T + b_l) )          models.py:114-115
//      MODE_BWD :  dZ_{l-1}  = (dZ_l W_l) * omega_{l-1} * C_{l-1}               (autograd of the above)
//    Each CTA keeps its N-slice of the (pre-packed, fp16) weight matrix resident in shared memory for the whole
//    kernel, streams A tiles (tile images, tc_common.cuh) through a 4-stage cp.async.bulk ring, accumulates in
//    TMEM (two accumulator buffers) and drains them with 8 epilogue warps that apply the activation math in
//    fp32 and write the next tile images with coalesced 16-byte stores.
//    warp 0 = bulk-copy producer, warp 1 = MMA issuer (+ TMEM owner), warps 2..9 = epilogue.
//
//  k_tc_dw<H>            dW_l += dZ_l^T S_{l-1}, both operands MN-major views of the SAME tile images, split over
//    row tiles across CTAs (split-K), accumulated in TMEM across all of a CTA's tiles, reduced into the fp32
//    gradient with red.global.add.v4.f32 once at the end.
#pragma once
#include "common.cuh"
#include "tc_common.cuh"

namespace inr {
namespace tc {

// MODE_FWD / MODE_BWD: SineLayer.  The Snake / Tanh variants (Linear+Snake, Linear+Tanh layers of SirenWithSnakeTanh,
// models.py:185-241,356-372) exist in the CTA-pair kernel (tc_gemm2.cuh) only:
//   MODE_FWD_SNAKE : y = z + sin^2(a z)/a -> out0, dy/dz -> out1, dy/da -> out2      (z = acc + b, a = avec[col])
//   MODE_FWD_TANH  : y = tanh(z) -> out0, 1 - y^2 -> out1
//   MODE_BWD_SNAKE : dZ_{l-1} = v * Cin (Cin = dy/dz) and, in place over E (= dy/da), the products v * E whose column
//                    sums are the gradient of `a` (k_tc_colsum afterwards)
enum { MODE_FWD = 0, MODE_BWD = 1, MODE_FWD_SNAKE = 2, MODE_FWD_TANH = 3, MODE_BWD_SNAKE = 4 };
__host__ __device__ constexpr bool mode_is_fwd(int m) { return m == MODE_FWD || m == MODE_FWD_SNAKE || m == MODE_FWD_TANH; }

struct LayerParams {
  const op16* A;      // input tile images: S_{l-1} (fwd) or dZ_l (bwd)
  const op16* Bpack;  // packed weights [NQ][H/8][NS][8]: W_l (fwd) or W_l^T (bwd)
  const float* bias;           // fwd: b_l (H floats)
  float omega;                 // fwd: omega_l;  bwd: omega_{l-1}
  op16* out0;         // fwd: S_l;  bwd: dZ_{l-1}
  op16* out1;         // fwd: C_l or NULL (inference)
  const op16* Cin;    // bwd: C_{l-1} (may alias out0: read before written, same thread)
  const float* avec;           // MODE_FWD_SNAKE: the layer's `a` (H floats)
  op16* out2;         // MODE_FWD_SNAKE: dy/da image;  MODE_BWD_SNAKE: dy/da image, overwritten with v * dy/da
  const op16* A_lo = nullptr;  // operand split (k_tc_layer2<.., SPLIT>): lo image of A; Bpack then holds [hi | lo] per slice
  op16* out0_lo = nullptr;     // operand split: lo image of out0 (the next layer's A_lo)
  float acc_scale = 1.f;       // forward modes: multiplies the accumulator (1 / the power-of-two scale of the packed weights)
  int num_tiles;
  uint32_t a_lbo, a_sbo, b_lbo, b_sbo;  // descriptor strides (bytes); passed in so a test can probe the convention
};

// ------------------------------------------------------------------------------------------------------------
template <int H>
struct DwCfg {
  static constexpr int NB = (H >= 256) ? 256 : 128;   // in-feature block (MMA N)
  static constexpr int MBLK = H / 128;                // out-feature blocks (MMA M = 128)
  static constexpr int NBLK = H / NB;
  static constexpr int NBLOCKS = MBLK * NBLK;
  static constexpr int A_BYTES = 128 * TILE_ROWS * 2;  // 16 feature groups of dZ: 32 KB
  static constexpr int B_BYTES = NB * TILE_ROWS * 2;   // 64 KB / 32 KB of S
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = 2;
  static constexpr int TMEM_COLS = 2 * NB;             // [0,NB) dW block, [NB,NB+16) bias column sums (power of two)
  static constexpr int OFF_ONES = STAGES * STAGE_BYTES; // 512 B of fp16 1.0: the B operand of the bias-gradient MMA
  static constexpr int OFF_BAR = OFF_ONES + 512;
  static constexpr int SMEM_BYTES = OFF_BAR + 64;
  static constexpr int THREADS = 192;                  // producer, MMA, 4 epilogue warps
};

struct DwParams {
  const op16* dZ;   // tile images of dZ_l   (M operand: out features)
  const op16* S;    // tile images of S_{l-1} (N operand: in features)
  float* gW;                 // fp32 gradient of W_l, row-major [H out][H in], accumulated with atomics
  float* gb;                 // fp32 gradient of b_l (H), or NULL: gb[o] += sum_r dZ[r][o] = dZ^T 1, an N=16 MMA against ones
  int num_tiles;
  uint32_t lbo, sbo;         // MN-major strides: LBO = 128 (8-row k groups), SBO = 2048 (feature groups)
  const float* gsc = nullptr; // {gs, 1/gs}: dZ carries the loss scale gs; the reduction into gW / gb multiplies by 1/gs
};

template <int H>
__global__ void __launch_bounds__(DwCfg<H>::THREADS, 1) k_tc_dw(const DwParams p) {
  using Cfg = DwCfg<H>;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t bar0 = smem_base + Cfg::OFF_BAR;
  auto bar_full = [&](int s) { return bar0 + 8u * s; };
  auto bar_empty = [&](int s) { return bar0 + 8u * (Cfg::STAGES + s); };
  const uint32_t bar_tfull = bar0 + 8u * (2 * Cfg::STAGES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + Cfg::OFF_BAR + 8 * (2 * Cfg::STAGES + 1));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int blk = blockIdx.x % Cfg::NBLOCKS;
  const int mb = blk / Cfg::NBLK, nb = blk % Cfg::NBLK;
  const int tile0 = blockIdx.x / Cfg::NBLOCKS;
  const int tile_stride = gridDim.x / Cfg::NBLOCKS;
  const bool has_work = tile0 < p.num_tiles;
  const bool do_bias = (p.gb != nullptr);   // the NBLK CTAs that share an A slice take alternate tiles of the bias sums

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::STAGES; ++s) { mbar_init(bar_full(s), 1); mbar_init(bar_empty(s), 1); }
    mbar_init(bar_tfull, 1);
    mbar_fence_init();
  }
  for (int i = threadIdx.x; i < 128; i += blockDim.x) reinterpret_cast<uint32_t*>(smem + Cfg::OFF_ONES)[i] = OP16_ONE_X2;
  fence_proxy_async_smem();
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), Cfg::TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      constexpr int PF = 3;   // tiles pulled into L2 ahead of the 2-stage ring (see k_tc_layer)
      auto a_of = [&](int tile) {
        return reinterpret_cast<const uint8_t*>(p.dZ) + (size_t)tile * (TILE_ROWS * H * 2) + (size_t)(mb * 16) * (TILE_ROWS * 16);
      };
      auto b_of = [&](int tile) {
        return reinterpret_cast<const uint8_t*>(p.S) + (size_t)tile * (TILE_ROWS * H * 2) + (size_t)(nb * (Cfg::NB / 8)) * (TILE_ROWS * 16);
      };
      // the NBLK CTAs sharing an A slice and the MBLK CTAs sharing a B slice split the prefetch between them
      auto prefetch_tile = [&](int tile) {
        if (tile >= p.num_tiles) return;
        if ((tile / tile_stride) % Cfg::NBLK == nb) bulk_prefetch_l2(a_of(tile), Cfg::A_BYTES);
        if ((tile / tile_stride) % Cfg::MBLK == mb) bulk_prefetch_l2(b_of(tile), Cfg::B_BYTES);
      };
      for (int d = 0; d < PF; ++d) prefetch_tile(tile0 + d * tile_stride);
      int s = 0;
      uint32_t phase = 0;
      for (int tile = tile0; tile < p.num_tiles; tile += tile_stride) {
        prefetch_tile(tile + PF * tile_stride);
        const uint8_t* a_src = a_of(tile);
        const uint8_t* b_src = b_of(tile);
        mbar_wait(bar_empty(s), phase ^ 1);
        mbar_arrive_expect_tx(bar_full(s), Cfg::STAGE_BYTES);
        const uint32_t dst = smem_base + s * Cfg::STAGE_BYTES;
        bulk_g2s(dst, a_src, Cfg::A_BYTES, bar_full(s));
        for (int off = 0; off < Cfg::B_BYTES; off += 32768)
          bulk_g2s(dst + Cfg::A_BYTES + off, b_src + off, 32768, bar_full(s));
        if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && has_work) {
      constexpr uint32_t idesc = make_idesc_op(128, Cfg::NB, 1, 1);   // both operands MN-major
      int s = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = tile0; tile < p.num_tiles; tile += tile_stride, ++it) {
        mbar_wait(bar_full(s), phase);
        tc_fence_after();
        const uint32_t a0 = smem_base + s * Cfg::STAGE_BYTES;
        const uint32_t b0 = a0 + Cfg::A_BYTES;
#pragma unroll
        for (int kk = 0; kk < TILE_ROWS / 16; ++kk) {   // K = 16 rows per MMA: 16 rows x 16 B inside each feature group
          const uint64_t ad = make_smem_desc(a0 + kk * 256, p.lbo, p.sbo);
          const uint64_t bd = make_smem_desc(b0 + kk * 256, p.lbo, p.sbo);
          umma_f16(tmem_base, ad, bd, idesc, (it | kk) != 0 ? 1u : 0u);
        }
        if (do_bias && (it % Cfg::NBLK) == nb) {   // every element of the ones block is 1.0, so its strides only have to stay inside the block
          constexpr uint32_t idesc1 = make_idesc_op(128, 16, 1, 1);
          const uint64_t od = make_smem_desc(smem_base + Cfg::OFF_ONES, 128, 256);
#pragma unroll
          for (int kk = 0; kk < TILE_ROWS / 16; ++kk)
            umma_f16(tmem_base + Cfg::NB, make_smem_desc(a0 + kk * 256, p.lbo, p.sbo), od, idesc1, (it >= Cfg::NBLK || kk != 0) ? 1u : 0u);
        }
        umma_commit(bar_empty(s));
        if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
      }
      umma_commit(bar_tfull);
    }
  } else if (has_work) {
    const int quad = warp & 3;
    const int row = quad * 32 + lane;                 // out-feature index inside the M block
    mbar_wait(bar_tfull, 0);
    tc_fence_after();
    float* grow = p.gW + (size_t)(mb * 128 + row) * H + nb * Cfg::NB;
    const float inv = p.gsc ? p.gsc[1] : 1.f;
    for (int c = 0; c < Cfg::NB / 32; ++c) {
      float v[32];
      tmem_ld32(tmem_base + c * 32 + ((uint32_t)(quad * 32) << 16), v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 8; ++j)
        red_add_v4(grow + c * 32 + j * 4, v[4 * j] * inv, v[4 * j + 1] * inv, v[4 * j + 2] * inv, v[4 * j + 3] * inv);
    }
    if (do_bias && (p.num_tiles - tile0 + tile_stride - 1) / tile_stride > nb) {   // this CTA issued at least one bias MMA
      float v[32];
      tmem_ld32(tmem_base + Cfg::NB + ((uint32_t)(quad * 32) << 16), v);   // columns 0..15 all hold the column sum
      tmem_ld_wait();
      atomicAdd(p.gb + mb * 128 + row, v[0] * inv);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

}  // namespace tc
}  // namespace inr
