// sgemm.cuh -- the generic fp32 CUDA-core GEMM of the parity paths (SIREN fp32 mode, KAN fp32 mode).
// One 128x128x16 register-tiled SGEMM (8x8 per thread, register prefetch) with fused epilogues; operands are addressed
// through (row, k) strides so the same kernel serves K-contiguous and M/N-contiguous views, and gridDim.z splits K.
#pragma once
#include <stdlib.h>

#include "common.cuh"

namespace inr {

// EPI_SNAKE / EPI_TANH: forward epilogues of the Linear+Snake / Linear+Tanh layers (models.py:185-241,356-372):
//   out0 = y, out1 = dy/dz, (Snake) aux1 = dy/da image, aux0 = the layer's `a` vector.
// EPI_DMUL_GA: backward epilogue towards a Snake layer: v = dY; out0 (holding dy/dz) <- v * dy/dz;
//   aux1[n] += sum_m v * aux0[m][n]  (aux0 = the dy/da image, aux1 = the gradient of `a`).
enum { EPI_SINCOS = 0, EPI_COSMUL = 1, EPI_ATOMIC = 2, EPI_STORE = 3, EPI_SNAKE = 4, EPI_TANH = 5, EPI_DMUL_GA = 6 };

constexpr int BM = 128, BN = 128, BK = 16;

// VEC: both operands are read with 16-byte loads along their contiguous axis (the launcher checks strides and
// alignment, sgemm_launch below) and shared memory is double-buffered: one block barrier per K tile.  The FMA order per
// output element is the same in both variants (k ascending), so they give identical results.
// F2: the inner product runs on packed fma.rn.f32x2 (two IEEE fp32 FMAs per instruction, same rounding as FFMA): half the
// issue slots for the same FMA-pipe work.
__device__ __forceinline__ unsigned long long pack_f32x2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void ffma2(unsigned long long& d, unsigned long long a, unsigned long long b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}

template <bool A_KC, bool B_KC, int EPI, bool VEC = false, int KU = 16, bool F2 = false>
__global__ void __launch_bounds__(256, VEC ? 2 : 0)
k_sgemm(const float* __restrict__ A, int64_t sAm, int64_t sAk, const float* __restrict__ B, int64_t sBn, int64_t sBk,
        int64_t M, int N, int64_t K, int64_t k_chunk, const float* __restrict__ bias, float omega,
        float* out0, float* out1, int64_t ldo, const float* __restrict__ aux0 = nullptr, float* aux1 = nullptr) {
  constexpr int NBUF = VEC ? 2 : 1;
  __shared__ __align__(16) float As[NBUF][BK][BM + 4];
  __shared__ __align__(16) float Bs[NBUF][BK][BN + 4];
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
    if (VEC) {
#pragma unroll
      for (int p = 0; p < 2; ++p) {
        const int e4 = tid + p * 256;
        float4 va = make_float4(0.f, 0.f, 0.f, 0.f), vb = va;
        if (A_KC) {   // 4 consecutive k of row mm
          const int64_t gm = m0 + (e4 >> 2), gk = k0 + (e4 & 3) * 4;
          if (gm < M && gk < kend) va = *reinterpret_cast<const float4*>(A + gm * sAm + gk);
        } else {      // 4 consecutive m of k-row kk
          const int64_t gk = k0 + (e4 >> 5), gm = m0 + (e4 & 31) * 4;
          if (gm < M && gk < kend) va = *reinterpret_cast<const float4*>(A + gk * sAk + gm);
        }
        if (B_KC) {
          const int64_t gn = n0 + (e4 >> 2), gk = k0 + (e4 & 3) * 4;
          if (gn < N && gk < kend) vb = *reinterpret_cast<const float4*>(B + gn * sBn + gk);
        } else {
          const int64_t gk = k0 + (e4 >> 5), gn = n0 + (e4 & 31) * 4;
          if (gn < N && gk < kend) vb = *reinterpret_cast<const float4*>(B + gk * sBk + gn);
        }
        ra[p * 4 + 0] = va.x; ra[p * 4 + 1] = va.y; ra[p * 4 + 2] = va.z; ra[p * 4 + 3] = va.w;
        rb[p * 4 + 0] = vb.x; rb[p * 4 + 1] = vb.y; rb[p * 4 + 2] = vb.z; rb[p * 4 + 3] = vb.w;
      }
      return;
    }
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
  auto sstore = [&](int buf) {
    if (VEC) {
#pragma unroll
      for (int p = 0; p < 2; ++p) {
        const int e4 = tid + p * 256;
        if (A_KC) {
#pragma unroll
          for (int i = 0; i < 4; ++i) As[buf][(e4 & 3) * 4 + i][e4 >> 2] = ra[p * 4 + i];
        } else {
          *reinterpret_cast<float4*>(&As[buf][e4 >> 5][(e4 & 31) * 4]) = make_float4(ra[p * 4], ra[p * 4 + 1], ra[p * 4 + 2], ra[p * 4 + 3]);
        }
        if (B_KC) {
#pragma unroll
          for (int i = 0; i < 4; ++i) Bs[buf][(e4 & 3) * 4 + i][e4 >> 2] = rb[p * 4 + i];
        } else {
          *reinterpret_cast<float4*>(&Bs[buf][e4 >> 5][(e4 & 31) * 4]) = make_float4(rb[p * 4], rb[p * 4 + 1], rb[p * 4 + 2], rb[p * 4 + 3]);
        }
      }
      return;
    }
#pragma unroll
    for (int p = 0; p < 8; ++p) {
      const int e = tid + p * 256;
      int mm, kk;
      if (A_KC) { mm = e / BK; kk = e % BK; } else { kk = e / BM; mm = e % BM; }
      As[buf][kk][mm] = ra[p];
      int nn;
      if (B_KC) { nn = e / BK; kk = e % BK; } else { kk = e / BN; nn = e % BN; }
      Bs[buf][kk][nn] = rb[p];
    }
  };
  unsigned long long acc2[F2 ? 8 : 1][F2 ? 4 : 1];
  if (F2) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc2[i][j] = 0ull;
  }
  auto compute = [&](int buf) {
#pragma unroll KU
    for (int kk = 0; kk < BK; ++kk) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][64 + ty * 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][kk][64 + tx * 4]);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
      if (F2) {
        const unsigned long long bp[4] = {pack_f32x2(b[0], b[1]), pack_f32x2(b[2], b[3]), pack_f32x2(b[4], b[5]),
                                          pack_f32x2(b[6], b[7])};
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const unsigned long long ap = pack_f32x2(a[i], a[i]);
#pragma unroll
          for (int j = 0; j < 4; ++j) ffma2(acc2[i][j], ap, bp[j]);
        }
        continue;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = __fmaf_rn(a[i], b[j], acc[i][j]);
    }
  };

  gload(kbeg);
  if (VEC) {
    sstore(0);
    __syncthreads();
    int cur = 0;
    for (int64_t k0 = kbeg; k0 < kend; k0 += BK) {
      const bool more = k0 + BK < kend;
      if (more) gload(k0 + BK);
      compute(cur);
      if (more) sstore(cur ^ 1);   // the other buffer was last read before the previous barrier
      __syncthreads();
      cur ^= 1;
    }
  } else {
    for (int64_t k0 = kbeg; k0 < kend; k0 += BK) {
      sstore(0);
      __syncthreads();
      if (k0 + BK < kend) gload(k0 + BK);
      compute(0);
      __syncthreads();
    }
  }

  if (F2) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
        asm("mov.b64 {%0, %1}, %2;" : "=f"(acc[i][2 * j]), "=f"(acc[i][2 * j + 1]) : "l"(acc2[i][j]));
  }

  float ga_part[8];
  if (EPI == EPI_DMUL_GA) {
#pragma unroll
    for (int j = 0; j < 8; ++j) ga_part[j] = 0.f;
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
      } else if (EPI == EPI_SNAKE || EPI == EPI_TANH) {
        float y[4], dz[4], da[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const bool in = gn + q < N;
          const float z = v[q] + (in ? bias[gn + q] : 0.f);
          if (EPI == EPI_SNAKE) {
            const float a = in ? aux0[gn + q] : 1.f;
            float sn, cs;
            sincosf(z * a, &sn, &cs);
            const float ia = 1.0f / a, s2 = sn * sn, sc2 = 2.f * sn * cs;
            y[q] = z + ia * s2;
            dz[q] = 1.f + sc2;
            da[q] = ia * (z * sc2 - ia * s2);
          } else {
            y[q] = tanhf(z);
            dz[q] = 1.f - y[q] * y[q];
            da[q] = 0.f;
          }
        }
        float* p1 = out1 ? out1 + gm * ldo + gn : nullptr;
        float* p2 = (EPI == EPI_SNAKE && aux1) ? aux1 + gm * ldo + gn : nullptr;
        if (full) {
          *reinterpret_cast<float4*>(p0) = make_float4(y[0], y[1], y[2], y[3]);
          if (p1) *reinterpret_cast<float4*>(p1) = make_float4(dz[0], dz[1], dz[2], dz[3]);
          if (p2) *reinterpret_cast<float4*>(p2) = make_float4(da[0], da[1], da[2], da[3]);
        } else {
          for (int q = 0; q < 4 && gn + q < N; ++q) { p0[q] = y[q]; if (p1) p1[q] = dz[q]; if (p2) p2[q] = da[q]; }
        }
      } else if (EPI == EPI_DMUL_GA) {
        const float* pe = aux0 + gm * ldo + gn;
        for (int q = 0; q < 4 && gn + q < N; ++q) {
          ga_part[jh * 4 + q] = __fmaf_rn(v[q], pe[q], ga_part[jh * 4 + q]);
          p0[q] = v[q] * p0[q];
        }
      } else if (EPI == EPI_STORE) {
        if (full) *reinterpret_cast<float4*>(p0) = make_float4(v[0], v[1], v[2], v[3]);
        else for (int q = 0; q < 4 && gn + q < N; ++q) p0[q] = v[q];
      } else {
        for (int q = 0; q < 4 && gn + q < N; ++q) atomicAdd(p0 + q, v[q]);
      }
    }
  }
  if (EPI == EPI_DMUL_GA) {
    // the main loop ended with __syncthreads(): As is free -> block-level column sums, one global atomic per column
    float* s_ga = &As[0][0][0];
    if (tid < BN) s_ga[tid] = 0.f;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 8; ++j) atomicAdd(&s_ga[(j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4))], ga_part[j]);
    __syncthreads();
    if (tid < BN && n0 + tid < N) atomicAdd(&aux1[n0 + tid], s_ga[tid]);
  }
}

// Launch helper: picks the 16-byte-load variant when both operands allow it (contiguous axis with stride 1, the other
// stride and the extent along the contiguous axis multiples of 4 floats, 16-byte-aligned base pointers).
template <bool A_KC, bool B_KC, int EPI>
static inline void sgemm_launch(dim3 grid, cudaStream_t st, const float* A, int64_t sAm, int64_t sAk, const float* B,
                                int64_t sBn, int64_t sBk, int64_t M, int N, int64_t K, int64_t k_chunk, const float* bias,
                                float omega, float* out0, float* out1, int64_t ldo, const float* aux0 = nullptr,
                                float* aux1 = nullptr) {
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  const bool va = A_KC ? (sAk == 1 && sAm % 4 == 0 && K % 4 == 0) : (sAm == 1 && sAk % 4 == 0 && M % 4 == 0);
  const bool vb = B_KC ? (sBk == 1 && sBn % 4 == 0 && K % 4 == 0) : (sBn == 1 && sBk % 4 == 0 && N % 4 == 0);
  // Vector path: K loop unrolled by 2 (the body stays inside the instruction cache: the fully unrolled 17 KB body measured
  // 10 % slower, fp32 C1 9.23 vs 8.32 ms/step, ncu no_instruction stalls) and the inner product on packed fma.rn.f32x2
  // (fp32 C1 step 8.31 -> 7.89 ms); both A/Bs in profiles/r01e_sgemm_ncu_summary.md.
  if (va && vb && al16(A) && al16(B) && k_chunk % 4 == 0) {
    k_sgemm<A_KC, B_KC, EPI, true, 2, true><<<grid, 256, 0, st>>>(A, sAm, sAk, B, sBn, sBk, M, N, K, k_chunk, bias, omega,
                                                                  out0, out1, ldo, aux0, aux1);
  } else
    k_sgemm<A_KC, B_KC, EPI, false><<<grid, 256, 0, st>>>(A, sAm, sAk, B, sBn, sBk, M, N, K, k_chunk, bias, omega, out0,
                                                          out1, ldo, aux0, aux1);
}

}  // namespace inr
