// encode.cu -- input encodings in front of the first layer.
//
//   k_encode_gaussian   random Fourier features: the third-party `rff.layers.GaussianEncoding(sigma, input_size,
//                       encoded_size)` the reference applies to the coordinates once before the loop (run.py:141-144):
//                       vp = (2*pi*v) @ b^T, out = [cos(vp) | sin(vp)]           -> (n, 2*encoded_size)
//   k_encode_nerf       PosEncodingNeRF.forward (models.py:48-59): [x, sin(s^0 pi x), cos(s^0 pi x), sin(s^1 pi x), ...]
//                       (frequency-major, then input feature, then sin/cos)    -> (n, in + 2*in*num_frequencies)
//
// The encoded coordinates are constant over the fit, so they are produced once and stay resident (as the reference does);
// coords == NULL generates coordinate i0+row of linspace(-scale, scale, T_total) in the kernel (in_features == 1), so a
// coordinate shard never needs the raw (n,1) tensor either.  fp32 with full-range sincosf (arguments reach 2*pi*sigma*3).
#include <math.h>

#include "common.cuh"

namespace inr {

constexpr int kMaxNerfFreq = 32;
struct NerfMul { float m[kMaxNerfFreq]; };   // float(scale**i * pi), computed on the host in double like Python does

__global__ void __launch_bounds__(256) k_encode_gaussian(const float* __restrict__ coords, int in,
                                                         const float* __restrict__ b, int F, int64_t n, int64_t g0,
                                                         int64_t T_total, float scale, float* __restrict__ out) {
  const int64_t total = n * F;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = idx / F;
    const int j = (int)(idx - r * F);
    float vp = 0.f;
    if (coords) {
      for (int i = 0; i < in; ++i) vp = __fmaf_rn(__fmul_rn(6.2831855f, coords[r * in + i]), b[(int64_t)j * in + i], vp);
    } else {
      vp = __fmul_rn(__fmul_rn(6.2831855f, linspace_at(g0 + r, T_total, scale)), b[j]);
    }
    float sn, cs;
    sincosf(vp, &sn, &cs);
    out[r * (2 * (int64_t)F) + j] = cs;
    out[r * (2 * (int64_t)F) + F + j] = sn;
  }
}

__global__ void __launch_bounds__(256) k_encode_nerf(const float* __restrict__ coords, int in, int nf, NerfMul mul,
                                                     int64_t n, int64_t g0, int64_t T_total, float scale,
                                                     float* __restrict__ out) {
  const int per_row = in * (nf + 1);   // one task per (row, block, input feature): block 0 = identity, block i+1 = frequency i
  const int out_dim = in + 2 * in * nf;
  const int64_t total = n * per_row;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = idx / per_row;
    const int t = (int)(idx - r * per_row);
    const int blk = t / in, j = t - blk * in;
    const float x = coords ? coords[r * in + j] : linspace_at(g0 + r, T_total, scale);
    float* o = out + r * out_dim;
    if (blk == 0) {
      o[j] = x;
    } else {
      float sn, cs;
      sincosf(__fmul_rn(mul.m[blk - 1], x), &sn, &cs);
      o[in + ((blk - 1) * in + j) * 2] = sn;
      o[in + ((blk - 1) * in + j) * 2 + 1] = cs;
    }
  }
}

}  // namespace inr

using namespace inr;

extern "C" int inr_encode_gaussian(const float* coords, int in_features, int64_t n, int64_t i0, int64_t T_total,
                                   float scale, const float* b, int encoded_size, float* out, void* stream) {
  INR_REQUIRE(out && b && n >= 0 && in_features >= 1 && encoded_size >= 1, INR_E_BADARG,
              "bad gaussian encoding arguments (n=%lld in=%d encoded_size=%d)", (long long)n, in_features, encoded_size);
  INR_REQUIRE(coords || (in_features == 1 && i0 >= 0 && i0 + n <= T_total), INR_E_BADARG,
              "coords == NULL needs in_features == 1 and i0+n <= T_total");
  if (n == 0) return INR_OK;
  const int blocks = (int)std::min<int64_t>(cdiv(n * encoded_size, 256), (int64_t)num_sms() * 16);
  { LaunchScope ls("k_encode_gaussian", (cudaStream_t)stream);
    k_encode_gaussian<<<blocks, 256, 0, (cudaStream_t)stream>>>(coords, in_features, b, encoded_size, n, i0, T_total, scale, out); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

extern "C" int inr_encode_nerf(const float* coords, int in_features, int64_t n, int64_t i0, int64_t T_total, float scale,
                               int num_frequencies, float freq_scale, float* out, void* stream) {
  INR_REQUIRE(out && n >= 0 && in_features >= 1 && num_frequencies >= 0 && num_frequencies <= kMaxNerfFreq, INR_E_BADARG,
              "bad NeRF encoding arguments (n=%lld in=%d num_frequencies=%d, at most %d)", (long long)n, in_features,
              num_frequencies, kMaxNerfFreq);
  INR_REQUIRE(coords || (in_features == 1 && i0 >= 0 && i0 + n <= T_total), INR_E_BADARG,
              "coords == NULL needs in_features == 1 and i0+n <= T_total");
  if (n == 0) return INR_OK;
  NerfMul mul;
  for (int i = 0; i < kMaxNerfFreq; ++i)   // (self.scale ** i) * np.pi in Python doubles, cast when it meets the fp32 tensor
    mul.m[i] = (float)(pow((double)freq_scale, (double)i) * 3.141592653589793);
  const int64_t tasks = n * in_features * (num_frequencies + 1);
  const int blocks = (int)std::min<int64_t>(cdiv(tasks, 256), (int64_t)num_sms() * 16);
  { LaunchScope ls("k_encode_nerf", (cudaStream_t)stream);
    k_encode_nerf<<<blocks, 256, 0, (cudaStream_t)stream>>>(coords, in_features, num_frequencies, mul, n, i0, T_total, scale, out); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}
