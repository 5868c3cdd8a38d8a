// tc_debug.cu -- test hooks for the tcgen05 kernels (exported as inr_dbg_*): run ONE layer GEMM / ONE dW GEMM on
// plain row-major fp32 inputs (converted to tile images on the device) so tests can check them against a matmul
// and probe the UMMA descriptor stride convention (`variant` bits swap LBO/SBO) in a single GPU session.
#include "tc_gemm.cuh"

namespace inr {
using namespace tc;
int tc_layer(int H, int mode, const LayerParams& p, cudaStream_t st);
int tc_dw(int H, const DwParams& p, cudaStream_t st);
int tc_ns(int H);

__global__ void k_img_from_rows(const float* __restrict__ X, int64_t rows, int64_t rows_pad, int H, op16* img) {
  const int64_t total = rows_pad * H;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / H;
    const int f = (int)(i - r * H);
    img[img_offset(H, r, f)] = to_op(r < rows ? X[i] : 0.f);
  }
}
__global__ void k_rows_from_img(const op16* __restrict__ img, int64_t rows, int H, float* X) {
  const int64_t total = rows * H;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / H;
    const int f = (int)(i - r * H);
    X[i] = __half2float(img[img_offset(H, r, f)]);
  }
}
__global__ void k_pack_one(const float* __restrict__ W, int H, int NS, int transpose, op16* out) {
  const int64_t total = (int64_t)H * H;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int o = (int)(i / H), in = (int)(i % H);
    const int n = transpose ? in : o, k = transpose ? o : in;   // B[n][k]
    out[(int64_t)(n / NS) * ((int64_t)H * NS) + ((int64_t)(k >> 3) * NS + (n % NS)) * 8 + (k & 7)] = to_op(W[i]);
  }
}
}  // namespace inr

using namespace inr;
using namespace inr::tc;

extern "C" {

int64_t inr_dbg_tc_scratch_bytes(int H, int64_t rows) {
  const int64_t rp = round_up(std::max<int64_t>(rows, 1), TILE_ROWS);
  return 4 * rp * H * 2 + (int64_t)H * H * 2 + 4096;
}

// mode 0: out0, out1 = sincos(omega * (A W^T + bias));   mode 1: out0 = (A W) * omega * Cin
int inr_dbg_tc_layer(int H, int mode, const float* A, const float* W, const float* bias, float omega, const float* Cin,
                     float* out0, float* out1, int64_t rows, int variant, void* scratch, int64_t scratch_bytes,
                     void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  INR_REQUIRE(H == 128 || H == 256 || H == 512, INR_E_UNSUPPORTED, "H=%d", H);
  INR_REQUIRE(scratch_bytes >= inr_dbg_tc_scratch_bytes(H, rows), INR_E_WORKSPACE, "scratch too small");
  const int64_t rp = round_up(rows, TILE_ROWS);
  const int NS = tc_ns(H);
  op16* iA = (op16*)scratch;
  op16* iO0 = iA + rp * H;
  op16* iO1 = iO0 + rp * H;
  op16* iC = iO1 + rp * H;
  op16* Wp = iC + rp * H;
  k_img_from_rows<<<592, 256, 0, st>>>(A, rows, rp, H, iA);
  if (mode == 1) k_img_from_rows<<<592, 256, 0, st>>>(Cin, rows, rp, H, iC);
  k_pack_one<<<592, 256, 0, st>>>(W, H, NS, mode == 1, Wp);
  LayerParams p;
  p.A = iA; p.Bpack = Wp; p.bias = bias; p.omega = omega; p.out0 = iO0; p.out1 = (mode == 0) ? iO1 : nullptr; p.Cin = iC;
  p.num_tiles = (int)(rp / TILE_ROWS);
  p.a_lbo = TILE_ROWS * 16; p.a_sbo = 128; p.b_lbo = NS * 16; p.b_sbo = 128;
  if (variant & 1) { uint32_t t = p.a_lbo; p.a_lbo = p.a_sbo; p.a_sbo = t; }
  if (variant & 2) { uint32_t t = p.b_lbo; p.b_lbo = p.b_sbo; p.b_sbo = t; }
  int rc = tc_layer(H, mode, p, st);
  if (rc) return rc;
  k_rows_from_img<<<592, 256, 0, st>>>(iO0, rows, H, out0);
  if (mode == 0 && out1) k_rows_from_img<<<592, 256, 0, st>>>(iO1, rows, H, out1);
  INR_LAUNCH_CHECK();
  return INR_OK;
}

// gW[H][H] += dZ^T S
int inr_dbg_tc_dw(int H, const float* dZ, const float* S, float* gW, int64_t rows, int variant, void* scratch,
                  int64_t scratch_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  INR_REQUIRE(H == 128 || H == 256 || H == 512, INR_E_UNSUPPORTED, "H=%d", H);
  INR_REQUIRE(scratch_bytes >= inr_dbg_tc_scratch_bytes(H, rows), INR_E_WORKSPACE, "scratch too small");
  const int64_t rp = round_up(rows, TILE_ROWS);
  op16* iZ = (op16*)scratch;
  op16* iS = iZ + rp * H;
  k_img_from_rows<<<592, 256, 0, st>>>(dZ, rows, rp, H, iZ);
  k_img_from_rows<<<592, 256, 0, st>>>(S, rows, rp, H, iS);
  DwParams p;
  p.dZ = iZ; p.S = iS; p.gW = gW; p.gb = nullptr; p.num_tiles = (int)(rp / TILE_ROWS);
  p.lbo = 128; p.sbo = TILE_ROWS * 16;
  if (variant & 1) { uint32_t t = p.lbo; p.lbo = p.sbo; p.sbo = t; }
  return tc_dw(H, p, st);
}

}  // extern "C"
