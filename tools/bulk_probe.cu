// bulk_probe.cu -- microbenchmark: per-SM throughput of cp.async.bulk global->shared as a function of copy size,
// copies in flight and source residency (L2-resident vs DRAM-sized), with all 148 SMs running.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bulk_probe tools/bulk_probe.cu
// Output: one line per configuration: bytes/copy, in flight, footprint MB, GB/s per SM, TB/s chip, us per copy.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect(uint32_t bar, uint32_t bytes) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// each CTA: `depth` slots of `bytes`; thread 0 keeps all slots in flight, walking its own region of the source
__global__ void __launch_bounds__(128, 1) k_probe(const uint8_t* src, size_t footprint, int bytes, int depth, int iters,
                                                  int split, int nprod, unsigned long long* cycles) {
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t base = smem_u32(smem);
  const uint32_t bars = base + (uint32_t)bytes * depth;
  if (threadIdx.x == 0) {
    for (int i = 0; i < depth; ++i) mbar_init(bars + 8 * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  // producers are lane 0 of warps 0..nprod-1; producer w owns slots w, w+nprod, ... (depth must be a multiple of nprod)
  if ((threadIdx.x & 31) == 0 && (threadIdx.x >> 5) < nprod) {
    const int w = threadIdx.x >> 5;
    const size_t per_cta = footprint / gridDim.x / nprod / bytes * bytes;
    const uint8_t* mine = src + per_cta * (blockIdx.x * nprod + w);
    size_t off = 0;
    const int my_depth = depth / nprod;
    const unsigned long long t0 = clock64();
    iters /= nprod;
    for (int it = 0; it < iters + my_depth; ++it) {
      const int s = (it % my_depth) * nprod + w;
      const int depth = my_depth;
      const uint32_t par = ((it / depth) & 1) ^ 1;
      if (it >= depth) while (!mbar_try(bars + 8 * s, par)) {}
      if (it < iters) {
        mbar_expect(bars + 8 * s, bytes);
        const int piece = bytes / split;
        for (int k = 0; k < split; ++k) bulk_g2s(base + s * bytes + k * piece, mine + off + k * piece, piece, bars + 8 * s);
        off += bytes;
        if (off + bytes > per_cta) off = 0;
      }
    }
    if (w == 0) cycles[blockIdx.x] = clock64() - t0;
  }
}

int main() {
  const size_t big = (size_t)8 << 30;
  uint8_t* src;
  cudaMalloc(&src, big);
  cudaMemset(src, 1, big);
  unsigned long long* cyc;
  cudaMalloc(&cyc, 148 * 8);
  cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  int clk_khz = 0;
  cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
  printf("bytes,depth,split,nprod,footprint_MB,GBps_per_SM,TBps_chip,us_per_copy_latency_est\n");
  const int sizes[] = {4096, 16384, 32768, 65536};
  const size_t foots[] = {(size_t)32 << 20, (size_t)8 << 30};
  for (size_t fp : foots)
    for (int bytes : sizes)
      for (int depth : {2, 4, 8})
        for (int cfg = 0; cfg < 4; ++cfg) {
          const int split = (cfg == 1) ? 4 : 1;
          const int nprod = (cfg == 2) ? 2 : (cfg == 3 ? 4 : 1);
          if ((size_t)bytes * depth > 200 * 1024) continue;
          if (depth % nprod) continue;
          if (fp == foots[1] && cfg != 0) continue;
          if (bytes == 4096 && cfg == 1) continue;
          const int iters = (int)((fp == foots[0] ? 64 : 24) * (size_t)(1 << 20) / bytes);
          cudaEvent_t e0, e1;
          cudaEventCreate(&e0); cudaEventCreate(&e1);
          k_probe<<<148, 128, bytes * depth + 256>>>(src, fp, bytes, depth, iters / 8, split, nprod, cyc);  // warm (fills L2 for the small footprint)
          cudaEventRecord(e0);
          k_probe<<<148, 128, bytes * depth + 256>>>(src, fp, bytes, depth, iters, split, nprod, cyc);
          cudaEventRecord(e1);
          cudaError_t err = cudaDeviceSynchronize();
          if (err != cudaSuccess) { printf("error %s\n", cudaGetErrorString(err)); return 1; }
          float ms = 0;
          cudaEventElapsedTime(&ms, e0, e1);
          const double total = (double)bytes * iters * 148;
          const double per_sm = (double)bytes * iters / (ms * 1e-3) / 1e9;
          printf("%d,%d,%d,%d,%zu,%.1f,%.2f,%.2f\n", bytes, depth, split, nprod, fp >> 20, per_sm, total / (ms * 1e-3) / 1e12,
                 (double)bytes * depth / (per_sm * 1e3));
        }
  return 0;
}
