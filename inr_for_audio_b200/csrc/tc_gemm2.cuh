// tc_gemm2.cuh -- CTA-pair (tcgen05 cta_group::2) version of the weight-stationary layer GEMM of tc_gemm.cuh.
//
// Why pairs (profiles/r01_layer_ablation.md): with one CTA per 128x128 output slice the A stream needs 64 B per MMA
// cycle per SM; the 64 KB ring left beside the 128 KB weight slice sustains ~20 B/cycle at the measured bulk-copy
// latency, so the tensor pipe idled 68 % of the time and every A tile crossed the L2 fabric four times.  A pair
// runs ONE 256-row x NP-column MMA per K step: each CTA stages only its own 128 rows of A and HALF of the
// weight slice (NP/2 columns), and receives its own 128 rows x NP columns of D in its own TMEM.  Per SM that is
// half the A bytes per MAC, half the fabric traffic, the same shared-memory budget.
//
//   k_tc_layer2<H, MODE>   grid = 2 * pairs, __cluster_dims__(2); pair g handles N-slice q = g % NQ of double
//                          tiles dt = g / NQ, + pairs/NQ, ...; CTA rank r of the pair owns row tile 2*dt + r.
//   warp 0 = bulk-copy producer (own A rows, own B half), warp 1 = MMA issuer in the leader (rank 0) / barrier
//   relay in the peer, warps 2..9 = epilogue (as in tc_gemm.cuh, over NP columns).
//
// Barriers (s = ring stage, b = accumulator buffer; "both" = same offset in both CTAs):
//   full[s]    leader: 2 arrivals = own producer (with tx bytes) + peer relay;  peer: 1 = own producer (tx)
//   empty[s]   both: 1 arrival, multicast by the leader's tcgen05.commit once the MMAs reading stage s are done
//   tfull[b]   both: 1 arrival, multicast commit once the tile's last MMA is done
//   tempty[b]  leader: 16 arrivals = 8 local epilogue warps + 8 remote (peer's epilogue warps)
//   bfull      leader: 2 (own producer tx + peer relay); peer: 1 (own producer tx)
#pragma once
#include "tc_gemm.cuh"

namespace inr {
namespace tc {

// SPLIT (INR_PREC_FP16S, forward modes only): both operands are fp16 hi + lo pairs (x = hi + lo, lo = fp16(x - hi):
// 22 significant bits) and every K step issues THREE MMAs into the same accumulator, hi*hi + lo*hi + hi*lo -- fp32-grade
// products (the lo*lo term is below 2^-22).  Only the hi half of the weight slice stays resident (a pair covers half as
// many output features at H = 512); a ring stage carries the hi and the lo chunk of the A rows AND the matching K chunk of
// the lo weights (L2-resident, re-streamed per tile: +25 % fabric bytes) -- three copies issued by three lanes (one
// thread sustains only ~1 bulk copy per 0.32 us).  That leaves 160 KB (H = 512) instead of 96 KB for bytes in flight:
// with both weight halves resident the kernel sat at 58 % tensor-pipe activity, waiting on the ring (profiles/r02_notes.md).
template <int H, bool SPLIT = false>
struct Layer2Cfg {
  static constexpr int NP = SPLIT ? ((H >= 512) ? 128 : H) : ((H >= 256) ? 256 : H);   // output features per pair (MMA N)
  static constexpr int NC = NP / 2;                 // B columns staged by each CTA
  static constexpr int NQ = H / NP;                 // N slices
  // One thread sustains ~1 cp.async.bulk per 0.32 us whatever its size (profiles/r01_bulk_probe_*.csv): 16 KB copies
  // cap a CTA at ~51 GB/s, below the ~60 GB/s the MMA consumes; 32 KB copies (128 K columns) lift the cap to ~100 GB/s.
  static constexpr int KC = SPLIT ? 64 : 128;       // K columns per ring stage
  static constexpr int NCHUNK = H / KC;
  static constexpr int B_HALF = H * NC * 2;         // one (hi or lo) weight slice half
  static constexpr int B_SLICE = B_HALF * (SPLIT ? 2 : 1);   // bytes of one packed slice in global memory: [hi | lo]
  static constexpr int B_BYTES = B_HALF;            // resident: 128 KB (H=512), 64 KB (256; split 512), 16 KB (128)
  static constexpr int A_HALF = (KC / 8) * TILE_ROWS * 16;   // one (hi or lo) A chunk: 32 KB, split 16 KB
  static constexpr int WLO_CHUNK = SPLIT ? (KC / 8) * NC * 16 : 0;   // split: K chunk of the lo weights, 8 / 16 KB
  static constexpr int A_STAGE_BYTES = A_HALF * (SPLIT ? 2 : 1) + WLO_CHUNK;  // 32 KB; split 40 / 48 KB
  static constexpr int STAGES_FIT = (232448 - B_BYTES - 2 * NP * 4 - 256) / A_STAGE_BYTES;
  static constexpr int STAGES = STAGES_FIT > 4 ? 4 : STAGES_FIT;   // 96 .. 160 KB in flight
  static constexpr int TMEM_COLS = 2 * NP;          // two accumulator buffers
  static constexpr int OFF_A = B_BYTES;
  static constexpr int OFF_BIAS = OFF_A + STAGES * A_STAGE_BYTES;
  static constexpr int OFF_AVEC = OFF_BIAS + NP * 4;   // Snake `a` of this N slice (forward)
  static constexpr int OFF_BAR = OFF_AVEC + NP * 4;
  static constexpr int SMEM_BYTES = OFF_BAR + 256;
  static constexpr int THREADS = 320;
  static_assert(H == 128 || H == 256 || H == 512, "hidden width");
  static_assert(SMEM_BYTES <= 232448, "shared-memory budget");
};

template <int H, int MODE, bool SPLIT = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(Layer2Cfg<H, SPLIT>::THREADS, 1) k_tc_layer2(const LayerParams p) {
  using Cfg = Layer2Cfg<H, SPLIT>;
  static_assert(!SPLIT || mode_is_fwd(MODE), "operand split: forward modes only");
  constexpr int NP = Cfg::NP;
  constexpr int NC = Cfg::NC;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t sB = smem_base;
  const uint32_t sA = smem_base + Cfg::OFF_A;
  float* s_bias = reinterpret_cast<float*>(smem + Cfg::OFF_BIAS);
  float* s_a = reinterpret_cast<float*>(smem + Cfg::OFF_AVEC);
  constexpr bool IS_FWD = mode_is_fwd(MODE);
  const uint32_t bar0 = smem_base + Cfg::OFF_BAR;
  auto bar_full = [&](int s) { return bar0 + 8u * s; };
  auto bar_empty = [&](int s) { return bar0 + 8u * (Cfg::STAGES + s); };
  auto bar_tfull = [&](int b) { return bar0 + 8u * (2 * Cfg::STAGES + b); };
  auto bar_tempty = [&](int b) { return bar0 + 8u * (2 * Cfg::STAGES + 2 + b); };
  const uint32_t bar_bfull = bar0 + 8u * (2 * Cfg::STAGES + 4);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + Cfg::OFF_BAR + 8 * (2 * Cfg::STAGES + 5));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1;
  const int npairs = gridDim.x >> 1;
  const int q = pair % Cfg::NQ;
  const int dt0 = pair / Cfg::NQ;
  const int dt_stride = npairs / Cfg::NQ;
  const int num_dt = (p.num_tiles + 1) >> 1;          // double tiles (the last may have only its rank-0 half)

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::STAGES; ++s) { mbar_init(bar_full(s), leader ? 2 : 1); mbar_init(bar_empty(s), 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(bar_tfull(b), 1); mbar_init(bar_tempty(b), 16); }
    mbar_init(bar_bfull, leader ? 2 : 1);
    mbar_fence_init();
  }
  if (IS_FWD) {   // sine: omega * b (the argument is omega*acc + omega*b); Snake / Tanh: b
    for (int i = threadIdx.x; i < NP; i += blockDim.x) s_bias[i] = (MODE == MODE_FWD ? p.omega : 1.f) * p.bias[q * NP + i];
  }
  if (MODE == MODE_FWD_SNAKE) {
    for (int i = threadIdx.x; i < NP; i += blockDim.x) s_a[i] = p.avec[q * NP + i];
  }
  __syncthreads();
  cluster_sync_all();                                 // both CTAs' barriers exist before anything can signal them
  if (warp == 1) {
    tmem_alloc2(smem_u32(tmem_slot), Cfg::TMEM_COLS);
    tmem_relinquish2();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== producer: own half of the weight slice once, then own A rows through the ring =========
    if (lane == 0) {
      // packed weights: [slice = q*2 + rank][hi | lo (split)][H/8][NC][8]
      mbar_arrive_expect_tx(bar_bfull, Cfg::B_BYTES);
      const uint8_t* bsrc = reinterpret_cast<const uint8_t*>(p.Bpack) + (size_t)(q * 2 + rank) * Cfg::B_SLICE;
      constexpr int BPIECE = Cfg::B_BYTES < 32768 ? Cfg::B_BYTES : 32768;
      for (int off = 0; off < Cfg::B_BYTES; off += BPIECE) bulk_g2s(sB + off, bsrc + off, BPIECE, bar_bfull);
    }
    if (lane < (SPLIT ? 3 : 1)) {                     // split: lane 0 copies the hi chunk, lane 1 the lo chunk, lane 2 the lo weights
      int s = 0;
      uint32_t phase = 0;
      const op16* abase = (SPLIT && lane == 1) ? p.A_lo : p.A;
      const uint8_t* wlo = reinterpret_cast<const uint8_t*>(p.Bpack) + (size_t)(q * 2 + rank) * Cfg::B_SLICE + Cfg::B_HALF;
      for (int dt = dt0; dt < num_dt; dt += dt_stride) {
        const int tile = 2 * dt + (int)rank;
        const bool valid = tile < p.num_tiles;
        const uint8_t* asrc = reinterpret_cast<const uint8_t*>(abase) + (size_t)tile * (TILE_ROWS * H * 2);
        for (int c = 0; c < Cfg::NCHUNK; ++c) {
          mbar_wait(bar_empty(s), phase ^ 1);
          if (lane == 0) {
            // odd tile count: the MMA multiplies stale A rows of this CTA, nobody stores them -- but this CTA's half of
            // the lo weights feeds the PEER's rows too, so it is staged regardless
            if (valid) mbar_arrive_expect_tx(bar_full(s), Cfg::A_STAGE_BYTES);
            else if (SPLIT) mbar_arrive_expect_tx(bar_full(s), Cfg::WLO_CHUNK);
            else mbar_arrive(bar_full(s));
          }
          if (SPLIT && lane == 2) {
            bulk_g2s(sA + s * Cfg::A_STAGE_BYTES + 2 * Cfg::A_HALF, wlo + (size_t)c * Cfg::WLO_CHUNK, Cfg::WLO_CHUNK, bar_full(s));
          } else if (valid) {
            bulk_g2s(sA + s * Cfg::A_STAGE_BYTES + lane * Cfg::A_HALF, asrc + (size_t)c * Cfg::A_HALF, Cfg::A_HALF, bar_full(s));
          }
          if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      if (leader) {
        // ===================== MMA issuer (one thread of the pair) =====================
        constexpr uint32_t idesc = make_idesc_op(256, NP, 0, 0);
        mbar_wait(bar_bfull, 0);
        int s = 0;
        uint32_t phase = 0;
        int it = 0;
        for (int dt = dt0; dt < num_dt; dt += dt_stride, ++it) {
          const int buf = it & 1;
          mbar_wait(bar_tempty(buf), ((it >> 1) & 1) ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + buf * NP;
          for (int c = 0; c < Cfg::NCHUNK; ++c) {
            mbar_wait(bar_full(s), phase);
            tc_fence_after();
#pragma unroll
            for (int kk = 0; kk < Cfg::KC / 16; ++kk) {
              const uint32_t a_addr = sA + s * Cfg::A_STAGE_BYTES + kk * 2 * (TILE_ROWS * 16);
              const uint32_t b_addr = sB + (c * (Cfg::KC / 8) + kk * 2) * (NC * 16);
              const uint64_t ad = make_smem_desc(a_addr, p.a_lbo, p.a_sbo);
              const uint64_t bd = make_smem_desc(b_addr, p.b_lbo, p.b_sbo);
              if (SPLIT) {   // small terms first: lo*hi + hi*lo, then hi*hi
                const uint64_t adl = make_smem_desc(a_addr + Cfg::A_HALF, p.a_lbo, p.a_sbo);
                const uint64_t bdl = make_smem_desc(sA + s * Cfg::A_STAGE_BYTES + 2 * Cfg::A_HALF + kk * 2 * (NC * 16), p.b_lbo, p.b_sbo);
                umma2_f16(d_tmem, adl, bd, idesc, (c | kk) != 0 ? 1u : 0u);
                umma2_f16(d_tmem, ad, bdl, idesc, 1u);
                umma2_f16(d_tmem, ad, bd, idesc, 1u);
              } else {
                umma2_f16(d_tmem, ad, bd, idesc, (c | kk) != 0 ? 1u : 0u);
              }
            }
            umma2_commit_mc(bar_empty(s), 3);
            if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
          }
          umma2_commit_mc(bar_tfull(buf), 3);
        }
      } else {
        // ===================== peer: forward "my operands have landed" to the leader's barriers ==============
        mbar_wait(bar_bfull, 0);
        mbar_arrive_remote(bar_bfull, 0);
        int s = 0;
        uint32_t phase = 0;
        for (int dt = dt0; dt < num_dt; dt += dt_stride) {
          for (int c = 0; c < Cfg::NCHUNK; ++c) {
            mbar_wait(bar_full(s), phase);
            mbar_arrive_remote(bar_full(s), 0);
            if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
          }
        }
      }
    }
  } else {
    // ===================== epilogue: 8 warps, warp%4 = TMEM lane quadrant, two column-chunk parities =====
    const int e = warp - 2;
    const int quad = warp & 3;
    const int half = e >> 2;
    const int row = quad * 32 + lane;
    const float om_acc = p.omega * p.acc_scale;          // forward: the packed weights carry 1 / acc_scale (a power of two)
    constexpr int NCH = NP / 64;                          // 32-column chunks per warp
    uint4 cin[NCH][4];
    auto cin_ptr = [&](int tile, int ci, int j) {
      const int g0 = (q * NP + (half + 2 * ci) * 32) >> 3;
      return reinterpret_cast<const uint4*>(p.Cin + (size_t)tile * (TILE_ROWS * H) + ((size_t)(g0 + j) * TILE_ROWS + row) * 8);
    };
    if (!IS_FWD && 2 * dt0 + (int)rank < p.num_tiles) {
#pragma unroll
      for (int ci = 0; ci < NCH; ++ci)
#pragma unroll
        for (int j = 0; j < 4; ++j) cin[ci][j] = *cin_ptr(2 * dt0 + (int)rank, ci, j);
    }
    int it = 0;
    for (int dt = dt0; dt < num_dt; dt += dt_stride, ++it) {
      const int buf = it & 1;
      const int tile = 2 * dt + (int)rank;
      const bool valid = tile < p.num_tiles;
      mbar_wait(bar_tfull(buf), (it >> 1) & 1);
      tc_fence_after();
      const size_t tile_off = (size_t)tile * (TILE_ROWS * H);
      const int next_tile = tile + 2 * dt_stride;
      if (valid) {
#pragma unroll
        for (int ci = 0; ci < NCH; ++ci) {
          const int c = half + 2 * ci;
          const int g0 = (q * NP + c * 32) >> 3;
          float v[32];
          tmem_ld32(tmem_base + buf * NP + c * 32 + ((uint32_t)(quad * 32) << 16), v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint32_t o0[4], o1[4], ol[4];
            const size_t off = tile_off + ((size_t)(g0 + j) * TILE_ROWS + row) * 8;
            if (MODE == MODE_FWD_SNAKE || MODE == MODE_FWD_TANH) {
              uint32_t o2[4];
#pragma unroll
              for (int h = 0; h < 4; ++h) {
                const int cc = j * 8 + h * 2;
                float y[2], dz[2], da[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                  const float z = __fmaf_rn(v[cc + u], p.acc_scale, s_bias[c * 32 + cc + u]);
                  if (MODE == MODE_FWD_SNAKE) {
                    const float a = s_a[c * 32 + cc + u], ia = __fdividef(1.0f, a);
                    float sn, cs;
                    __sincosf(z * a, &sn, &cs);
                    const float s2 = sn * sn, sc2 = 2.f * sn * cs;
                    y[u] = __fmaf_rn(ia, s2, z);
                    dz[u] = 1.f + sc2;
                    da[u] = ia * (z * sc2 - ia * s2);
                  } else {
                    float th;
                    asm("tanh.approx.f32 %0, %1;" : "=f"(th) : "f"(z));
                    y[u] = th;
                    dz[u] = __fmaf_rn(-th, th, 1.f);
                    da[u] = 0.f;
                  }
                }
                if (SPLIT) pack_op2_split(y[0], y[1], &o0[h], &ol[h]);
                else o0[h] = pack_op2(y[0], y[1]);
                o1[h] = pack_op2(dz[0], dz[1]);
                o2[h] = pack_op2(da[0], da[1]);
              }
              *reinterpret_cast<uint4*>(p.out0 + off) = make_uint4(o0[0], o0[1], o0[2], o0[3]);
              if (SPLIT) *reinterpret_cast<uint4*>(p.out0_lo + off) = make_uint4(ol[0], ol[1], ol[2], ol[3]);
              if (p.out1) *reinterpret_cast<uint4*>(p.out1 + off) = make_uint4(o1[0], o1[1], o1[2], o1[3]);
              if (MODE == MODE_FWD_SNAKE && p.out2) *reinterpret_cast<uint4*>(p.out2 + off) = make_uint4(o2[0], o2[1], o2[2], o2[3]);
            } else if (MODE == MODE_FWD) {
#pragma unroll
              for (int h = 0; h < 4; ++h) {
                const int cc = j * 8 + h * 2;
                const float t0 = __fmaf_rn(v[cc], om_acc, s_bias[c * 32 + cc]);
                const float t1 = __fmaf_rn(v[cc + 1], om_acc, s_bias[c * 32 + cc + 1]);
                float s0, c0, s1, c1;
                __sincosf(t0, &s0, &c0);
                __sincosf(t1, &s1, &c1);
                if (SPLIT) pack_op2_split(s0, s1, &o0[h], &ol[h]);
                else o0[h] = pack_op2(s0, s1);
                o1[h] = pack_op2(c0, c1);
              }
              *reinterpret_cast<uint4*>(p.out0 + off) = make_uint4(o0[0], o0[1], o0[2], o0[3]);
              if (SPLIT) *reinterpret_cast<uint4*>(p.out0_lo + off) = make_uint4(ol[0], ol[1], ol[2], ol[3]);
              if (p.out1) *reinterpret_cast<uint4*>(p.out1 + off) = make_uint4(o1[0], o1[1], o1[2], o1[3]);
            } else {
              const uint32_t cw[4] = {cin[ci][j].x, cin[ci][j].y, cin[ci][j].z, cin[ci][j].w};
#pragma unroll
              for (int h = 0; h < 4; ++h) {
                const int cc = j * 8 + h * 2;
                const float2 cf = unpack_op2(cw[h]);
                o0[h] = pack_op2(v[cc] * p.omega * cf.x, v[cc + 1] * p.omega * cf.y);
              }
              *reinterpret_cast<uint4*>(p.out0 + off) = make_uint4(o0[0], o0[1], o0[2], o0[3]);
              if (MODE == MODE_BWD_SNAKE) {   // v * dy/da in place over the dy/da image (column-summed afterwards)
                const uint4 eu = *reinterpret_cast<const uint4*>(p.out2 + off);
                const uint32_t ew[4] = {eu.x, eu.y, eu.z, eu.w};
#pragma unroll
                for (int h = 0; h < 4; ++h) {
                  const int cc = j * 8 + h * 2;
                  const float2 ef = unpack_op2(ew[h]);
                  o1[h] = pack_op2(v[cc] * ef.x, v[cc + 1] * ef.y);
                }
                *reinterpret_cast<uint4*>(p.out2 + off) = make_uint4(o1[0], o1[1], o1[2], o1[3]);
              }
              if (next_tile < p.num_tiles) cin[ci][j] = *cin_ptr(next_tile, ci, j);
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (leader) mbar_arrive(bar_tempty(buf));
        else mbar_arrive_remote(bar_tempty(buf), 0);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                                 // the peer's TMEM reads and remote arrivals are done
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc2(tmem_base, Cfg::TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------------------
// k_tc_dw2<H> -- CTA-pair version of k_tc_dw (tc_gemm.cuh) for H >= 256.
//
// Why (profiles/r01b_ncu_full_summary.md): the single-CTA dW kernel moves 96 KB of operands per 128x256x128 MMA
// block through the L2->SM fabric -- 16.3 GB per launch at H = 512, i.e. 13.5 TB/s, the fabric ceiling -- while DRAM
// (55 %) and the tensor pipe (69 %) idle.  A pair runs ONE M = 256 MMA per K step: each CTA stages its own 128
// out-features of dZ (32 KB) and only HALF of the S block's in-features (32 KB), and receives its own 128 x NB block of
// dW in its own TMEM: 64 KB per CTA per row tile instead of 96 KB, 3 ring stages instead of 2.
//
//   grid = 2 * pairs; pair g handles block (mp, nb) = g % (MPAIRS*NBLK) of row tiles g / (MPAIRS*NBLK), + stride, ...
//   CTA rank r of the pair owns out-feature block mb = 2*mp + r.
// Barriers: full[s] leader 2 (own producer tx + peer relay) / peer 1; empty[s] 1 (multicast commit); tfull 1 (multicast).
template <int H>
struct Dw2Cfg {
  static constexpr int NB = 256;                       // in-feature block (MMA N)
  static constexpr int MPAIRS = H / 256;               // out-feature blocks of 256 (MMA M over the pair)
  static constexpr int NBLK = H / NB;
  static constexpr int NBLOCKS = MPAIRS * NBLK;        // pair-blocks per row tile
  static constexpr int A_BYTES = 128 * TILE_ROWS * 2;  // own 16 feature groups of dZ: 32 KB
  static constexpr int B_BYTES = (NB / 2) * TILE_ROWS * 2;   // own half of the S block: 32 KB
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = 3;
  static constexpr int TMEM_COLS = 2 * NB;             // [0,NB) dW block, [NB,NB+16) bias column sums
  static constexpr int OFF_ONES = STAGES * STAGE_BYTES;
  static constexpr int OFF_BAR = OFF_ONES + 512;
  static constexpr int SMEM_BYTES = OFF_BAR + 128;
  static constexpr int THREADS = 192;                  // producer, MMA / relay, 4 epilogue warps
  static_assert(H == 256 || H == 512, "pair dW kernel: hidden width 256 or 512");
};

template <int H>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(Dw2Cfg<H>::THREADS, 1) k_tc_dw2(const DwParams p) {
  using Cfg = Dw2Cfg<H>;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t bar0 = smem_base + Cfg::OFF_BAR;
  auto bar_full = [&](int s) { return bar0 + 8u * s; };
  auto bar_empty = [&](int s) { return bar0 + 8u * (Cfg::STAGES + s); };
  const uint32_t bar_tfull = bar0 + 8u * (2 * Cfg::STAGES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + Cfg::OFF_BAR + 8 * (2 * Cfg::STAGES + 1));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1;
  const int npairs = gridDim.x >> 1;
  const int blk = pair % Cfg::NBLOCKS;
  const int mp = blk / Cfg::NBLK, nb = blk % Cfg::NBLK;
  const int mb = 2 * mp + (int)rank;                   // this CTA's 128 out-features
  const int tile0 = pair / Cfg::NBLOCKS;
  const int tile_stride = npairs / Cfg::NBLOCKS;
  const bool has_work = tile0 < p.num_tiles;
  const bool do_bias = (p.gb != nullptr);

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::STAGES; ++s) { mbar_init(bar_full(s), leader ? 2 : 1); mbar_init(bar_empty(s), 1); }
    mbar_init(bar_tfull, 1);
    mbar_fence_init();
  }
  for (int i = threadIdx.x; i < 128; i += blockDim.x) reinterpret_cast<uint32_t*>(smem + Cfg::OFF_ONES)[i] = OP16_ONE_X2;
  fence_proxy_async_smem();
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

  if (warp == 0) {
    if (lane == 0) {
      int s = 0;
      uint32_t phase = 0;
      for (int tile = tile0; tile < p.num_tiles; tile += tile_stride) {
        const uint8_t* a_src = reinterpret_cast<const uint8_t*>(p.dZ) + (size_t)tile * (TILE_ROWS * H * 2) +
                               (size_t)(mb * 16) * (TILE_ROWS * 16);
        const uint8_t* b_src = reinterpret_cast<const uint8_t*>(p.S) + (size_t)tile * (TILE_ROWS * H * 2) +
                               (size_t)(nb * (Cfg::NB / 8) + (int)rank * (Cfg::NB / 16)) * (TILE_ROWS * 16);
        mbar_wait(bar_empty(s), phase ^ 1);
        mbar_arrive_expect_tx(bar_full(s), Cfg::STAGE_BYTES);
        const uint32_t dst = smem_base + s * Cfg::STAGE_BYTES;
        bulk_g2s(dst, a_src, Cfg::A_BYTES, bar_full(s));
        bulk_g2s(dst + Cfg::A_BYTES, b_src, Cfg::B_BYTES, bar_full(s));
        if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && has_work) {
      if (leader) {
        constexpr uint32_t idesc = make_idesc_op(256, Cfg::NB, 1, 1);   // both operands MN-major
        int s = 0;
        uint32_t phase = 0;
        int it = 0;
        for (int tile = tile0; tile < p.num_tiles; tile += tile_stride, ++it) {
          mbar_wait(bar_full(s), phase);
          tc_fence_after();
          const uint32_t a0 = smem_base + s * Cfg::STAGE_BYTES;
          const uint32_t b0 = a0 + Cfg::A_BYTES;
#pragma unroll
          for (int kk = 0; kk < TILE_ROWS / 16; ++kk) {   // K = 16 rows per MMA
            const uint64_t ad = make_smem_desc(a0 + kk * 256, p.lbo, p.sbo);
            const uint64_t bd = make_smem_desc(b0 + kk * 256, p.lbo, p.sbo);
            umma2_f16(tmem_base, ad, bd, idesc, (it | kk) != 0 ? 1u : 0u);
          }
          if (do_bias && (it % Cfg::NBLK) == nb) {   // gb = dZ^T 1: N = 16 ones-MMA (8 ones columns from each CTA)
            constexpr uint32_t idesc1 = make_idesc_op(256, 16, 1, 1);
            const uint64_t od = make_smem_desc(smem_base + Cfg::OFF_ONES, 128, 256);
#pragma unroll
            for (int kk = 0; kk < TILE_ROWS / 16; ++kk)
              umma2_f16(tmem_base + Cfg::NB, make_smem_desc(a0 + kk * 256, p.lbo, p.sbo), od, idesc1,
                         (it >= Cfg::NBLK || kk != 0) ? 1u : 0u);
          }
          umma2_commit_mc(bar_empty(s), 3);
          if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
        }
        umma2_commit_mc(bar_tfull, 3);
      } else {
        int s = 0;
        uint32_t phase = 0;
        for (int tile = tile0; tile < p.num_tiles; tile += tile_stride) {
          mbar_wait(bar_full(s), phase);
          mbar_arrive_remote(bar_full(s), 0);
          if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
        }
      }
    }
  } else if (has_work) {
    const int quad = warp & 3;
    const int row = quad * 32 + lane;                 // out-feature index inside this CTA's M block
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
    if (do_bias && (p.num_tiles - tile0 + tile_stride - 1) / tile_stride > nb) {   // this pair issued at least one bias MMA
      float v[32];
      tmem_ld32(tmem_base + Cfg::NB + ((uint32_t)(quad * 32) << 16), v);   // columns 0..15 all hold the column sum
      tmem_ld_wait();
      atomicAdd(p.gb + mb * 128 + row, v[0] * inv);
    }
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
