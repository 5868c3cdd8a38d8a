// tc_fused.cuh -- fused all-layer SIREN forward for H <= 256 (sm_100a, tcgen05 cta_group::2).
//
// One kernel runs the FIRST layer and ALL hidden layers of a row tile (models.py:114-115 per layer, 271-275 for the
// stack) without the activations leaving the SM between layers:
//
//   * a CTA pair owns two 256-row "slots" (X, Y), each CTA 128 rows of each; the current layer input of a slot lives
//     in shared memory as a fp16 tile image (the UMMA K-major operand layout of tc_common.cuh) and is OVERWRITTEN by
//     the epilogue with the next layer's input -- S_l is never read back from HBM;
//   * the weights are the streamed operand: every layer's packed fp16 matrix (L2-resident, 128 KB at H = 256) is
//     pulled through a bulk-copy ring, each CTA staging its half of the N columns (cta_group::2 B split);
//   * accumulators: one TMEM buffer per slot; while the 8 epilogue warps drain slot X (omega*acc + omega*b -> sincos ->
//     fp16 -> shared-memory image + the S_l / C_l stash the backward needs), the tensor pipe runs slot Y;
//   * the first layer (fp32 FMA, Cody-Waite reduction, MUFU sincos) is computed by the epilogue warps straight into the
//     slot's image, from coordinates generated in-kernel (0 HBM bytes) or given.
//
// HBM traffic of the forward: only the stash writes (S_l, C_l: 2 units per layer) -- the layer-major path
// additionally READS S_{l-1} once per layer.  What the backward needs is written in the same tile-image layout as the
// layer-major forward, so k_tc_layer2<BWD>, k_tc_dw and k_tc_last run unchanged.
//
// Job order of a pair (k-th double tile of the pair -> slot k&1, round r = k>>1):
//   MMA / weight ring:  for r: for l = 1..L: (X, l), (Y, l)
//   epilogue:           gen X(0), gen Y(0); for r: for l = 1..L: epi(X, l) [l == L: gen X(r+1)], epi(Y, l) [gen Y(r+1)]
// Barriers (s = ring stage, b = slot):
//   wfull[s]   leader: 2 arrivals (own producer with tx bytes + peer relay); peer: 1 (own producer, tx)
//   wempty[s]  both: 1, multicast tcgen05.commit once the MMAs reading stage s are done
//   tfull[b]   both: 1, multicast commit after the job's last MMA
//   aready[b]  leader: 16 arrivals (8 local + 8 remote epilogue warps): the slot's image holds the next layer's input
//              in BOTH CTAs and its accumulator has been drained
#pragma once
#include "tc_gemm.cuh"

namespace inr {
namespace tc {

// The argument omega*z of the first layer is formed in fp32 exactly as the reference does (Linear -> * omega_0 -> sin,
// models.py:115); it can reach +-4.4e4 rad (run.py's default omega = 22000), so it is reduced with a two-constant
// Cody-Waite step (k = rint(t/2pi); r = t - k*C1 - k*C2 with FMAs, error < 2e-7 rad for |t| < 1e6) and the reduced
// angle goes to the SFU (MUFU.SIN/COS, abs error ~5e-7 on [-pi, pi]) -- two orders below the fp16 rounding of the
// stored result.
__device__ __forceinline__ void sincos_reduced(float t, float* s, float* c) {
  // k = rint(t / 2pi) by the 1.5 * 2^23 magic constant: two FMA-pipe instructions instead of an FRND on the XU pipe,
  // which the two MUFUs below already saturate (|t / 2pi| < 2^22 holds for any omega_0 * z of interest)
  const float k = __fadd_rn(__fmaf_rn(t, 0.15915494309189535f, 12582912.0f), -12582912.0f);
  float r = __fmaf_rn(-k, 6.2831854820251465f, t);      // C1 = fl32(2 pi)
  r = __fmaf_rn(-k, -1.7484555e-7f, r);                 // C2 = 2 pi - C1
  *s = __sinf(r);
  *c = __cosf(r);
}

constexpr int FUSED_MAX_L = 8;

template <int H>
struct FusedCfg {
  static constexpr int NP = H;                      // output features per pair (MMA N): the whole layer
  static constexpr int NC = NP / 2;                 // weight columns staged by each CTA
  static constexpr int KC = 64;                     // K columns per weight ring stage
  static constexpr int NCHUNK = H / KC;
  static constexpr int W_STAGE_BYTES = (KC / 8) * NC * 16;          // 16 KB (H = 256), 8 KB (128)
  static constexpr int STAGES = 5;
  static constexpr int A_BYTES = (H / 8) * TILE_ROWS * 16;          // one slot image: 64 KB (H = 256)
  static constexpr int TMEM_COLS = 2 * NP;          // one accumulator per slot
  static constexpr int OFF_A = 0;
  static constexpr int OFF_W = 2 * A_BYTES;
  static constexpr int OFF_BIAS = OFF_W + STAGES * W_STAGE_BYTES;   // omega * b_l, l = 1..L (FUSED_MAX_L x H floats)
  static constexpr int OFF_W0 = OFF_BIAS + FUSED_MAX_L * H * 4;     // first-layer weight (H x in, in <= 2)
  static constexpr int OFF_B0 = OFF_W0 + 2 * H * 4;
  static constexpr int OFF_WL = OFF_B0 + H * 4;                     // last nn.Linear weight (H floats)
  static constexpr int OFF_DOT = OFF_WL + H * 4;                    // partial output dots [job parity][column part <= 4][128 rows]
  static constexpr int OFF_G = OFF_DOT + 2 * 4 * TILE_ROWS * 4;     // d loss / d out per row of the current tile
  static constexpr int OFF_BAR = OFF_G + TILE_ROWS * 4;
  static constexpr int SMEM_BYTES = OFF_BAR + 256;
  static_assert(H == 128 || H == 256, "fused forward: hidden width 128 or 256");
  static_assert(SMEM_BYTES <= 232448, "shared-memory budget");
};

struct FusedParams {
  const op16* Wf;     // packed hidden weights, layer l-1 at Wf + (l-1)*H*H: [2 slices][H/8][NC][8]
  const float* W0;             // first layer weight (H x in), bias (H)
  const float* B0;
  const float* bias[FUSED_MAX_L];        // hidden biases b_1..b_L
  op16* S[FUSED_MAX_L + 1];     // stash of S_0..S_L (tile images) or NULL
  op16* C[FUSED_MAX_L + 1];     // stash of C_0..C_L or NULL
  const float* coords;         // (rows, in) or NULL: generated from the global index
  int in, L, num_tiles;
  int64_t rows, g0, T_total;
  float scale, omega0, omega_h;
  // fuse_last: the final nn.Linear (models.py:258-262), the time-domain loss (run.py:124-125,161-169) and the first
  // backward step run in the epilogue of layer L: out = S_L w + b, g = d loss/d out (+ extra_g),
  // dZ_L = g w omega C_L is written where C_L would go (C[L]); S_L / C_L never reach HBM; the last layer's weight
  // and bias gradients and the loss numerator are reduced in the kernel.
  int fuse_last, loss_kind;
  float grad_scale;
  const float* w_last;         // (H), bias (1)
  const float* b_last;
  const float* target;         // (rows) or NULL
  const float* extra_g;        // (rows) upstream gradient or NULL
  float* out;                  // (rows) or NULL
  float* g_wlast;              // gradient of w_last (H), of b_last (1): accumulated with atomics
  float* g_blast;
  double* loss_acc;            // += sum (out - target)^2 or |out - target|
  const float* gsc;            // {gs, 1/gs}: power-of-two loss scale of the fp16 dZ images (siren_tc.cu k_tc_gscale)
  unsigned* gmax_acc;          // max |g| of this step (float bits, atomicMax) or NULL: the next step's scale
  // hand-off trace (inr_dbg_fused_trace): pair 0's leader CTA writes clock64() stamps, 4 per job in MMA order:
  // [MMA thread passed aready, MMA thread committed tfull, epilogue warp 0 passed tfull, epilogue warp 0 signalled]
  long long* trace;
  int trace_jobs;
};

// EW = number of epilogue warps (8 or 16): warp%4 picks the TMEM lane quadrant, (warp-2)/4 the column part.
template <int H, int EW>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(64 + 32 * EW, 1) k_tc_fused_fwd(const FusedParams p) {
  using Cfg = FusedCfg<H>;
  constexpr int NP = Cfg::NP;
  constexpr int NC = Cfg::NC;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t smem_base = smem_u32(smem);
  auto sA = [&](int slot) { return smem_base + Cfg::OFF_A + slot * Cfg::A_BYTES; };
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
  auto bar_tfull = [&](int b) { return bar0 + 8u * (2 * Cfg::STAGES + b); };
  auto bar_aready = [&](int b) { return bar0 + 8u * (2 * Cfg::STAGES + 2 + b); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + Cfg::OFF_BAR + 8 * (2 * Cfg::STAGES + 4));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1;
  const int npairs = gridDim.x >> 1;
  const int num_dt = (p.num_tiles + 1) >> 1;          // double tiles (the last may have only its rank-0 half)
  const int n_k = pair < num_dt ? (num_dt - pair + npairs - 1) / npairs : 0;   // double tiles of this pair
  const int L = p.L;
  auto dt_of = [&](int k) { return pair + k * npairs; };

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::STAGES; ++s) { mbar_init(bar_wfull(s), leader ? 2 : 1); mbar_init(bar_wempty(s), 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(bar_tfull(b), 1); mbar_init(bar_aready(b), 2 * EW); }
    mbar_fence_init();
  }
  for (int l = 0; l < L; ++l)
    for (int i = threadIdx.x; i < H; i += blockDim.x) s_bias[l * H + i] = p.omega_h * p.bias[l][i];
  for (int i = threadIdx.x; i < H * p.in; i += blockDim.x) s_w0[i] = p.W0[i];
  for (int i = threadIdx.x; i < H; i += blockDim.x) s_b0[i] = p.B0[i];
  if (p.fuse_last)
    for (int i = threadIdx.x; i < H; i += blockDim.x) s_wl[i] = p.w_last[i];
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
    // ===================== producer: this CTA's half of every job's weight matrix through the ring ===============
    if (lane == 0) {
      int s = 0;
      uint32_t phase = 0;
      for (int r = 0; 2 * r < n_k; ++r) {
        const int nslots = (2 * r + 1 < n_k) ? 2 : 1;
        for (int l = 1; l <= L; ++l) {
          const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(p.Wf) + (size_t)(l - 1) * H * H * 2 +
                                (size_t)rank * (H * NC * 2);
          for (int b = 0; b < nslots; ++b) {
            for (int c = 0; c < Cfg::NCHUNK; ++c) {
              mbar_wait(bar_wempty(s), phase ^ 1);
              mbar_arrive_expect_tx(bar_wfull(s), Cfg::W_STAGE_BYTES);
              bulk_g2s(sW + s * Cfg::W_STAGE_BYTES, wsrc + (size_t)c * Cfg::W_STAGE_BYTES, Cfg::W_STAGE_BYTES, bar_wfull(s));
              if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      if (leader) {
        // ===================== MMA issuer (one thread of the pair) =====================
        constexpr uint32_t idesc = make_idesc_op(256, NP, 0, 0);
        int s = 0;
        uint32_t phase = 0;
        for (int r = 0; 2 * r < n_k; ++r) {
          const int nslots = (2 * r + 1 < n_k) ? 2 : 1;
          for (int l = 1; l <= L; ++l) {
            const uint32_t par = (uint32_t)(r * L + l - 1) & 1u;
            for (int b = 0; b < nslots; ++b) {
              mbar_wait(bar_aready(b), par);          // image written in both CTAs, accumulator drained
              tc_fence_after();
              const int tj = (r * L + l - 1) * 2 + b;
              if (p.trace && blockIdx.x == 0 && tj < p.trace_jobs) p.trace[tj * 4 + 0] = clock64();
              const uint32_t d_tmem = tmem_base + b * NP;
              for (int c = 0; c < Cfg::NCHUNK; ++c) {
                mbar_wait(bar_wfull(s), phase);
                tc_fence_after();
#pragma unroll
                for (int kk = 0; kk < Cfg::KC / 16; ++kk) {
                  const uint64_t ad = make_smem_desc(sA(b) + (c * (Cfg::KC / 8) + kk * 2) * (TILE_ROWS * 16), TILE_ROWS * 16, 128);
                  const uint64_t bd = make_smem_desc(sW + s * Cfg::W_STAGE_BYTES + kk * 2 * (NC * 16), NC * 16, 128);
                  umma2_f16(d_tmem, ad, bd, idesc, (c | kk) != 0 ? 1u : 0u);
                }
                umma2_commit_mc(bar_wempty(s), 3);
                if (++s == Cfg::STAGES) { s = 0; phase ^= 1; }
              }
              umma2_commit_mc(bar_tfull(b), 3);
              if (p.trace && blockIdx.x == 0 && tj < p.trace_jobs) p.trace[tj * 4 + 1] = clock64();
            }
          }
        }
      } else {
        // ===================== peer: forward "my weight stage has landed" to the leader's barriers ==============
        int s = 0;
        uint32_t phase = 0;
        for (int r = 0; 2 * r < n_k; ++r) {
          const int nslots = (2 * r + 1 < n_k) ? 2 : 1;
          for (int j = 0; j < L * nslots * Cfg::NCHUNK; ++j) {
            mbar_wait(bar_wfull(s), phase);
            mbar_arrive_remote(bar_wfull(s), 0);
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
    constexpr int NPART = EW / 4;                         // column parts (warps per TMEM lane quadrant)
    constexpr int NCH = NP / (32 * NPART);                // 32-column chunks per warp
    uint8_t* smem_gen = smem;                             // generic-proxy view of the slot images

    auto signal_ready = [&](int b) {
      fence_proxy_async_smem();                           // generic-proxy image writes -> visible to the UMMA reads
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (leader) mbar_arrive(bar_aready(b));
        else mbar_arrive_remote(bar_aready(b), 0);
      }
    };
    // first layer of double tile k into slot b (this CTA's 128 rows), + the S_0 / C_0 stash
    auto gen_first = [&](int b, int k) {
      const int tile = 2 * dt_of(k) + (int)rank;
      if (tile < p.num_tiles) {
        int64_t grow = (int64_t)tile * TILE_ROWS + row;
        if (grow >= p.rows) grow = p.rows - 1;            // pad rows replicate the last row: finite, zero gradient later
        float x0, x1 = 0.f;
        if (p.coords) {
          x0 = p.coords[grow * p.in];
          if (p.in == 2) x1 = p.coords[grow * p.in + 1];
        } else {
          x0 = linspace_at(p.g0 + grow, p.T_total, p.scale);
        }
        const size_t tile_off = (size_t)tile * (TILE_ROWS * H);
        uint8_t* img = smem_gen + Cfg::OFF_A + b * Cfg::A_BYTES;
#pragma unroll 1
        for (int ci = 0; ci < NCH; ++ci) {
          const int g0 = ((half + NPART * ci) * 32) >> 3;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint32_t so[4], co[4];
#pragma unroll
            for (int h = 0; h < 4; ++h) {
              float s[2], c[2];
#pragma unroll
              for (int u = 0; u < 2; ++u) {
                const int f = (g0 + j) * 8 + h * 2 + u;
                const float z = (p.in == 1) ? __fmaf_rn(x0, s_w0[f], s_b0[f])
                                            : (__fmaf_rn(x1, s_w0[f * 2 + 1], x0 * s_w0[f * 2]) + s_b0[f]);
                sincos_reduced(p.omega0 * z, &s[u], &c[u]);
              }
              so[h] = pack_op2(s[0], s[1]);
              co[h] = pack_op2(c[0], c[1]);
            }
            const size_t loc = ((size_t)(g0 + j) * TILE_ROWS + row) * 8;   // elements inside the tile image
            *reinterpret_cast<uint4*>(img + loc * 2) = make_uint4(so[0], so[1], so[2], so[3]);
            if (p.S[0]) *reinterpret_cast<uint4*>(p.S[0] + tile_off + loc) = make_uint4(so[0], so[1], so[2], so[3]);
            if (p.C[0]) *reinterpret_cast<uint4*>(p.C[0] + tile_off + loc) = make_uint4(co[0], co[1], co[2], co[3]);
          }
        }
      }
      signal_ready(b);
    };

    // fuse_last accumulators: loss numerator, bias gradient (half 0 threads), weight gradient (lane = column % 32)
    double loss_local = 0.0;
    float gb_local = 0.f, gmax_local = 0.f;
    const float gs = (p.fuse_last && p.gsc) ? p.gsc[0] : 1.f;
    float gw_acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) gw_acc[i] = 0.f;
    int last_jobs = 0;

    if (n_k > 0) gen_first(0, 0);
    if (n_k > 1) gen_first(1, 1);
    for (int r = 0; 2 * r < n_k; ++r) {
      const int nslots = (2 * r + 1 < n_k) ? 2 : 1;
      for (int l = 1; l <= L; ++l) {
        const uint32_t par = (uint32_t)(r * L + l - 1) & 1u;
        const float* bias_l = s_bias + (l - 1) * H;
        for (int b = 0; b < nslots; ++b) {
          const int k = 2 * r + b;
          const int tile = 2 * dt_of(k) + (int)rank;
          const bool valid = tile < p.num_tiles;
          mbar_wait(bar_tfull(b), par);
          tc_fence_after();
          const int tj = (r * L + l - 1) * 2 + b;
          const bool tr = p.trace && blockIdx.x == 0 && e == 0 && lane == 0 && tj < p.trace_jobs;
          if (tr) p.trace[tj * 4 + 2] = clock64();
          if (l == L && p.fuse_last) {
            // ---------- last hidden layer + nn.Linear head + loss + dZ_L ----------
            // pass A: S_L = sin(.) (fp16) -> the slot's image (free: no layer L+1) + this thread's part of the output dot
            // pass B: cos(.) from the accumulator again -> dZ_L = g w omega C_L -> HBM
            // pass C: g_w[f] += sum_r g[r] S_L[r][f] from the image (one thread = 8 features x 128/TPG rows)
            const size_t tile_off = (size_t)tile * (TILE_ROWS * H);
            const int64_t grow = (int64_t)tile * TILE_ROWS + row;
            const bool live = valid && grow < p.rows;
            uint8_t* img = smem_gen + Cfg::OFF_A + b * Cfg::A_BYTES;
            float dot = 0.f;
            if (valid) {
#pragma unroll
              for (int ci = 0; ci < NCH; ++ci) {
                const int c = half + NPART * ci;
                const int g0 = (c * 32) >> 3;
                float v[32];
                tmem_ld32(tmem_base + b * NP + c * 32 + ((uint32_t)(quad * 32) << 16), v);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  uint32_t o0[4];
#pragma unroll
                  for (int h = 0; h < 4; ++h) {
                    const int cc = j * 8 + h * 2;
                    const float s0 = __sinf(__fmaf_rn(v[cc], p.omega_h, bias_l[c * 32 + cc]));
                    const float s1 = __sinf(__fmaf_rn(v[cc + 1], p.omega_h, bias_l[c * 32 + cc + 1]));
                    o0[h] = pack_op2(s0, s1);
                    const float2 sb = unpack_op2(o0[h]);   // the head sees the fp16 S_L, as on the layer-major path
                    dot = __fmaf_rn(sb.x, s_wl[c * 32 + cc], dot);
                    dot = __fmaf_rn(sb.y, s_wl[c * 32 + cc + 1], dot);
                  }
                  *reinterpret_cast<uint4*>(img + ((size_t)(g0 + j) * TILE_ROWS + row) * 16) = make_uint4(o0[0], o0[1], o0[2], o0[3]);
                }
              }
            }
            float* sd = s_dot + (last_jobs & 1) * (NPART * TILE_ROWS);
            sd[half * TILE_ROWS + row] = dot;
            asm volatile("bar.sync 1, %0;" ::"n"(32 * EW) : "memory");      // the 8 epilogue warps: both column halves of every row
            float o = p.b_last[0];
#pragma unroll
            for (int q = NPART - 1; q >= 0; --q) o += sd[q * TILE_ROWS + row];
            ++last_jobs;
            float g = 0.f;
            if (live) {
              if (p.out && half == 0) p.out[grow] = o;
              if (p.target) {
                const float diff = o - p.target[grow];
                if (p.loss_kind == INR_LOSS_MAE) {
                  if (half == 0) loss_local += (double)fabsf(diff);
                  g = p.grad_scale * ((diff > 0.f) ? 1.f : ((diff < 0.f) ? -1.f : 0.f));
                } else {
                  if (half == 0) loss_local += (double)diff * (double)diff;
                  g = p.grad_scale * 2.f * diff;
                }
              }
              if (p.extra_g) g += p.extra_g[grow];
              if (half == 0) { gb_local += g; gmax_local = fmaxf(gmax_local, fabsf(g)); }
            }
            if (half == 0) s_g[row] = g;
            if (valid) {
              const float gw = g * gs * p.omega_h;
#pragma unroll
              for (int ci = 0; ci < NCH; ++ci) {
                const int c = half + NPART * ci;
                const int g0 = (c * 32) >> 3;
                float v[32];
                tmem_ld32(tmem_base + b * NP + c * 32 + ((uint32_t)(quad * 32) << 16), v);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  uint32_t o1[4];
#pragma unroll
                  for (int h = 0; h < 4; ++h) {
                    const int cc = j * 8 + h * 2;
                    // same roundings as k_tc_last on the stored fp16 C_L image
                    const float2 cb = unpack_op2(pack_op2(__cosf(__fmaf_rn(v[cc], p.omega_h, bias_l[c * 32 + cc])),
                                                                __cosf(__fmaf_rn(v[cc + 1], p.omega_h, bias_l[c * 32 + cc + 1]))));
                    o1[h] = pack_op2(gw * s_wl[c * 32 + cc] * cb.x, gw * s_wl[c * 32 + cc + 1] * cb.y);
                  }
                  const size_t loc = ((size_t)(g0 + j) * TILE_ROWS + row) * 8;
                  *reinterpret_cast<uint4*>(p.C[L] + tile_off + loc) = make_uint4(o1[0], o1[1], o1[2], o1[3]);
                }
              }
            }
            asm volatile("bar.sync 1, %0;" ::"n"(32 * EW) : "memory");      // S_L image and s_g complete
            if (valid) {
              constexpr int TPG = (32 * EW) / (H / 8);                 // threads per feature group: 8 (H = 256), 16 (H = 128)
              constexpr int RPT = TILE_ROWS / TPG;               // rows per thread
              const int et = threadIdx.x - 64;                   // 0..255 over the epilogue warps
              const int fg = et / TPG, r0 = (et % TPG) * RPT;
#pragma unroll 4
              for (int rr = 0; rr < RPT; ++rr) {
                const uint4 u = *reinterpret_cast<const uint4*>(img + ((size_t)fg * TILE_ROWS + r0 + rr) * 16);
                const float gr = s_g[r0 + rr];
                const uint32_t uw[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                for (int h = 0; h < 4; ++h) {
                  const float2 f = unpack_op2(uw[h]);
                  gw_acc[2 * h] = __fmaf_rn(gr, f.x, gw_acc[2 * h]);
                  gw_acc[2 * h + 1] = __fmaf_rn(gr, f.y, gw_acc[2 * h + 1]);
                }
              }
            }
            // the image is overwritten by the next gen_first (of ANY warp): all 8 warps leave it first
            asm volatile("bar.sync 1, %0;" ::"n"(32 * EW) : "memory");
          } else if (valid) {
            const size_t tile_off = (size_t)tile * (TILE_ROWS * H);
            uint8_t* img = smem_gen + Cfg::OFF_A + b * Cfg::A_BYTES;
            op16* Sl = p.S[l];
            op16* Cl = p.C[l];
#pragma unroll
            for (int ci = 0; ci < NCH; ++ci) {
              const int c = half + NPART * ci;
              const int g0 = (c * 32) >> 3;
              float v[32];
              tmem_ld32(tmem_base + b * NP + c * 32 + ((uint32_t)(quad * 32) << 16), v);
              tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                uint32_t o0[4], o1[4];
#pragma unroll
                for (int h = 0; h < 4; ++h) {
                  const int cc = j * 8 + h * 2;
                  const float t0 = __fmaf_rn(v[cc], p.omega_h, bias_l[c * 32 + cc]);
                  const float t1 = __fmaf_rn(v[cc + 1], p.omega_h, bias_l[c * 32 + cc + 1]);
                  float s0, c0, s1, c1;
                  __sincosf(t0, &s0, &c0);
                  __sincosf(t1, &s1, &c1);
                  o0[h] = pack_op2(s0, s1);
                  o1[h] = pack_op2(c0, c1);
                }
                const size_t loc = ((size_t)(g0 + j) * TILE_ROWS + row) * 8;
                if (l < L) *reinterpret_cast<uint4*>(img + loc * 2) = make_uint4(o0[0], o0[1], o0[2], o0[3]);
                if (Sl) *reinterpret_cast<uint4*>(Sl + tile_off + loc) = make_uint4(o0[0], o0[1], o0[2], o0[3]);
                if (Cl) *reinterpret_cast<uint4*>(Cl + tile_off + loc) = make_uint4(o1[0], o1[1], o1[2], o1[3]);
              }
            }
          }
          if (l < L) {
            signal_ready(b);                              // next layer's input is in place, accumulator drained
          } else if (k + 2 < n_k) {
            gen_first(b, k + 2);                          // the slot's next double tile (also releases the accumulator)
          }
          if (tr) p.trace[tj * 4 + 3] = clock64();
        }
      }
    }
    if (p.fuse_last) {
      constexpr int TPG = (32 * EW) / (H / 8);
      const int et = threadIdx.x - 64;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float a = gw_acc[i];
#pragma unroll
        for (int o = TPG / 2; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        if ((et % TPG) == 0) atomicAdd(p.g_wlast + (et / TPG) * 8 + i, a);
      }
      loss_local = warp_sum(loss_local);
      gb_local = warp_sum(gb_local);
      for (int o = 16; o > 0; o >>= 1) gmax_local = fmaxf(gmax_local, __shfl_xor_sync(0xffffffffu, gmax_local, o));
      if (lane == 0 && half == 0) {
        if (p.target) atomicAdd(p.loss_acc, loss_local);
        atomicAdd(p.g_blast, gb_local);
        if (p.gmax_acc) atomicMax(p.gmax_acc, __float_as_uint(gmax_local));
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

}  // namespace tc
}  // namespace inr
