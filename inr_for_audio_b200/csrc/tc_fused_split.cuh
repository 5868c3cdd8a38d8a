// tc_fused_split.cuh -- fused all-layer SIREN forward for the operand-split mode (INR_PREC_FP16S), H <= 256.
//
// Same job as k_tc_fused_fwd (tc_fused.cuh): first layer + ALL hidden layers (+ head, loss, dZ_L) of a row tile in one
// kernel, activations never leave the SM between layers (models.py:114-115, 271-275; run.py:161-169).  In the split mode a
// layer input is an fp16 hi + lo image pair (128 KB per 128-row tile at H = 256), so a CTA holds ONE tile, not two
// ping-pong slots.  What keeps the tensor pipe busy instead is a pipeline INSIDE the tile:
//
//   * two TMEM accumulators alternate between consecutive layers of the tile;
//   * the epilogue of layer l (omega*acc/wscale + omega*b -> sincos -> hi, lo) overwrites the slot IN PLACE, 64 input
//     columns (one K chunk of layer l+1) at a time, ALL 16 warps on the same chunk (32 rows x 16 columns each) -- layer l's
//     MMAs have all completed by then, so the old columns are dead -- and signals a per-chunk barrier;
//   * the MMA thread starts layer l+1 on K chunk c as soon as chunk c is in place, accumulating into the OTHER
//     accumulator while the epilogue warps are still draining the first: the second half of epilogue(l) overlaps the
//     first half of MMA(l+1).
//   * weights hi | lo stream through a bulk-copy ring in 32-column K chunks (two lanes issue the two copies).
//
// HBM traffic: only the stash the backward needs (S_l hi for dW, C_l) -- the lo images never reach HBM.
//
// Barriers (leader = cluster rank 0; s = ring stage, a = accumulator j & 1 of job j, c = 64-column input chunk):
//   wfull[s]    leader 2 (own producer tx + peer relay), peer 1      wempty[s]  both 1 (multicast commit)
//   accfull[a]  both 1 (multicast commit after the job's last MMA)   accfree[a] leader 2*EW (every epilogue warp of the pair)
//   inready[c]  leader 2*EW: every epilogue warp of the pair writes a 16-column strip of chunk c
#pragma once
#include "tc_fused.cuh"

namespace inr {
namespace tc {

template <int H>
struct FusedSCfg {
  static constexpr int EW = 16;                       // epilogue warps
  static constexpr int NP = H;                        // MMA N: the whole layer
  static constexpr int NC = NP / 2;                   // weight columns staged by each CTA
  static constexpr int KCW = 32;                      // K columns per weight ring stage
  static constexpr int NKW = H / KCW;                 // ring stages per layer
  static constexpr int W_HALF = (KCW / 8) * NC * 16;  // hi (or lo) chunk: 8 KB (H = 256), 4 KB (128)
  static constexpr int W_STAGE_BYTES = 2 * W_HALF;
  static constexpr int STAGES = 4;
  static constexpr int A_BYTES = (H / 8) * TILE_ROWS * 16;   // one image (hi or lo): 64 KB (H = 256)
  static constexpr int NIN = H / 64;                  // 64-column input chunks = inready barriers
  static constexpr int TMEM_COLS = 2 * NP;
  static constexpr int OFF_AHI = 0;
  static constexpr int OFF_ALO = A_BYTES;
  static constexpr int OFF_W = 2 * A_BYTES;
  static constexpr int OFF_BIAS = OFF_W + STAGES * W_STAGE_BYTES;   // omega * b_l, l = 1..L
  static constexpr int OFF_W0 = OFF_BIAS + FUSED_MAX_L * H * 4;
  static constexpr int OFF_B0 = OFF_W0 + 2 * H * 4;
  static constexpr int OFF_WL = OFF_B0 + H * 4;
  static constexpr int OFF_DOT = OFF_WL + H * 4;                    // partial output dots [job parity][4 column parts][128 rows]
  static constexpr int OFF_G = OFF_DOT + 2 * 4 * TILE_ROWS * 4;
  static constexpr int OFF_BAR = OFF_G + TILE_ROWS * 4;
  static constexpr int SMEM_BYTES = OFF_BAR + 256;
  static constexpr int THREADS = 64 + 32 * EW;
  static_assert(H == 128 || H == 256, "fused split forward: hidden width 128 or 256");
  static_assert(SMEM_BYTES <= 232448, "shared-memory budget");
};

struct FusedSParams {
  FusedParams f;        // same fields as the non-split kernel; f.Wf holds [layer][rank][hi | lo][K/8][NC][8] (k_tc_pack, split)
  op16* S_lo_last;      // !fuse_last: lo image of S_L for the head kernel (k_tc_last2), else NULL
  float acc_scale;      // 1 / wscale of the packed forward weights
};

template <int H>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(FusedSCfg<H>::THREADS, 1) k_tc_fused_fwd_s(const FusedSParams ps) {
  using Cfg = FusedSCfg<H>;
  constexpr int EW = Cfg::EW;
  constexpr int NP = Cfg::NP;
  constexpr int NC = Cfg::NC;
  const FusedParams& p = ps.f;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t sAhi = smem_base + Cfg::OFF_AHI, sAlo = smem_base + Cfg::OFF_ALO;
  const uint32_t sW = smem_base + Cfg::OFF_W;
  float* s_bias = reinterpret_cast<float*>(smem + Cfg::OFF_BIAS);
  float* s_w0 = reinterpret_cast<float*>(smem + Cfg::OFF_W0);
  float* s_b0 = reinterpret_cast<float*>(smem + Cfg::OFF_B0);
  float* s_wl = reinterpret_cast<float*>(smem + Cfg::OFF_WL);
  float* s_dot = reinterpret_cast<float*>(smem + Cfg::OFF_DOT);
  float* s_g = reinterpret_cast<float*>(smem + Cfg::OFF_G);
  const uint32_t bar0 = smem_base + Cfg::OFF_BAR;
  auto bar_wfull = [&](int s) { return bar0 + 8u * s; };
  auto bar_wempty = [&](int s) { return bar0 + 8u * (Cfg::STAGES + s); };
  auto bar_accfull = [&](int a) { return bar0 + 8u * (2 * Cfg::STAGES + a); };
  auto bar_accfree = [&](int a) { return bar0 + 8u * (2 * Cfg::STAGES + 2 + a); };
  auto bar_inready = [&](int c) { return bar0 + 8u * (2 * Cfg::STAGES + 4 + c); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + Cfg::OFF_BAR + 8 * (2 * Cfg::STAGES + 4 + Cfg::NIN));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1;
  const int npairs = gridDim.x >> 1;
  const int num_dt = (p.num_tiles + 1) >> 1;
  const int n_k = pair < num_dt ? (num_dt - pair + npairs - 1) / npairs : 0;   // double tiles of this pair
  const int L = p.L;
  auto dt_of = [&](int k) { return pair + k * npairs; };
  constexpr int NPART = EW / 4;                          // column parts (warps per TMEM lane quadrant): 4 x 16 columns = one chunk
  // every epilogue warp of both CTAs writes its 32 rows x 16 columns of each 64-column input chunk
  constexpr int INREADY_COUNT = 2 * EW;

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::STAGES; ++s) { mbar_init(bar_wfull(s), leader ? 2 : 1); mbar_init(bar_wempty(s), 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(bar_accfull(a), 1); mbar_init(bar_accfree(a), 2 * EW); }
    for (int c = 0; c < Cfg::NIN; ++c) mbar_init(bar_inready(c), INREADY_COUNT);
    mbar_fence_init();
  }
  for (int l = 0; l < L; ++l)
    for (int i = threadIdx.x; i < H; i += blockDim.x) s_bias[l * H + i] = p.omega_h * p.bias[l][i];
  for (int i = threadIdx.x; i < H * p.in; i += blockDim.x) s_w0[i] = p.W0[i];
  for (int i = threadIdx.x; i < H; i += blockDim.x) s_b0[i] = p.B0[i];
  if (p.fuse_last)
    for (int i = threadIdx.x; i < H; i += blockDim.x) s_wl[i] = p.w_last[i];
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tmem_alloc2(smem_u32(tmem_slot), Cfg::TMEM_COLS);
    tmem_relinquish2();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int n_jobs = n_k * L;

  if (warp == 0) {
    // ===================== producer: this CTA's half of every job's weights, hi (lane 0) and lo (lane 1) ==========
    if (lane < 2) {
      int s = 0;
      uint32_t phase = 0;
      for (int k = 0; k < n_k; ++k) {
        for (int l = 1; l <= L; ++l) {
          const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(p.Wf) + (size_t)(l - 1) * (2 * H * H * 2) +
                                (size_t)rank * (2 * H * NC * 2) + (size_t)lane * (H * NC * 2);
          for (int c = 0; c < Cfg::NKW; ++c) {
            mbar_wait(bar_wempty(s), phase ^ 1);
            if (lane == 0) mbar_arrive_expect_tx(bar_wfull(s), Cfg::W_STAGE_BYTES);
            bulk_g2s(sW + s * Cfg::W_STAGE_BYTES + lane * Cfg::W_HALF, wsrc + (size_t)c * Cfg::W_HALF, Cfg::W_HALF, bar_wfull(s));
            if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      if (leader) {
        // ===================== MMA issuer =====================
        constexpr uint32_t idesc = make_idesc_op(256, NP, 0, 0);
        int s = 0;
        uint32_t phase = 0;
        for (int j = 0; j < n_jobs; ++j) {
          const int a = j & 1;
          mbar_wait(bar_accfree(a), (uint32_t)((j >> 1) & 1) ^ 1u);      // epilogue of job j-2 has drained this accumulator
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + a * NP;
          for (int c = 0; c < Cfg::NKW; ++c) {
            if ((c & 1) == 0) {                                          // entering a new 64-column input chunk
              mbar_wait(bar_inready(c >> 1), (uint32_t)(j & 1));
              tc_fence_after();
            }
            mbar_wait(bar_wfull(s), phase);
            tc_fence_after();
#pragma unroll
            for (int kk = 0; kk < Cfg::KCW / 16; ++kk) {
              const uint32_t aoff = (c * (Cfg::KCW / 8) + kk * 2) * (TILE_ROWS * 16);
              const uint32_t boff = sW + s * Cfg::W_STAGE_BYTES + kk * 2 * (NC * 16);
              const uint64_t ah = make_smem_desc(sAhi + aoff, TILE_ROWS * 16, 128);
              const uint64_t al = make_smem_desc(sAlo + aoff, TILE_ROWS * 16, 128);
              const uint64_t bh = make_smem_desc(boff, NC * 16, 128);
              const uint64_t bl = make_smem_desc(boff + Cfg::W_HALF, NC * 16, 128);
              umma2_f16(d_tmem, al, bh, idesc, (c | kk) != 0 ? 1u : 0u);
              umma2_f16(d_tmem, ah, bl, idesc, 1u);
              umma2_f16(d_tmem, ah, bh, idesc, 1u);
            }
            umma2_commit_mc(bar_wempty(s), 3);
            if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
          }
          umma2_commit_mc(bar_accfull(a), 3);
        }
      } else {
        // ===================== peer: forward "my weight stage has landed" to the leader =====================
        int s = 0;
        uint32_t phase = 0;
        for (int j = 0; j < n_jobs * Cfg::NKW; ++j) {
          mbar_wait(bar_wfull(s), phase);
          mbar_arrive_remote(bar_wfull(s), 0);
          if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
        }
      }
    }
  } else {
#include "tc_fused_split_epi.inc"
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc2(tmem_base, Cfg::TMEM_COLS);
  }
}

}  // namespace tc
}  // namespace inr
