// Micro-benchmark: how many bytes per clock can ONE SM pull through TMA tensor-map loads, by box shape?
// One CTA per SM; one thread issues boxes into a ring of `stages` buffers and waits for each to land before reusing it
// (no consumer at all).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scratch/tma_rate profiles/tma_rate.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(128, 1)
rate_kernel(const __grid_constant__ CUtensorMap map, int mode, int stages, uint32_t box_bytes, int iters, int rowsH, int nB,
            long long* clocks, int issuers, int lane_mode) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (s32(smem_raw) & 1023u)) & 1023u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)stages * box_bytes);
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bars[s])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  // issuers: how many threads issue boxes (box `it` belongs to issuer it % issuers); lane_mode 0: one thread per warp
  // (threads 0, 32, 64, 96), 1: lanes 0..issuers-1 of warp 0
  const int me = lane_mode ? (int)threadIdx.x : ((threadIdx.x & 31) == 0 ? (int)(threadIdx.x >> 5) : -1);
  if (me >= 0 && me < issuers) {
    const long long t0 = clock64();
    for (int it = me; it < iters; it += issuers) {
      const int s = it % stages;
      if (it >= stages) {
        const uint32_t parity = ((it / stages) - 1) & 1;
        uint32_t done = 0;
        while (!done)
          asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                       : "=r"(done) : "r"(s32(&bars[s])), "r"(parity) : "memory");
      }
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bars[s])), "r"(box_bytes) : "memory");
      // walk distinct rows so that successive boxes are different memory: h over rowsH, b over nB, per-CTA offset
      const int lin = it + blockIdx.x * 977;
      const int h = lin % rowsH, b = (lin / rowsH) % nB;
      unsigned char* dst = smem + (size_t)s * box_bytes;
      if (mode == 2)
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     ::"r"(s32(dst)), "l"(&map), "r"(s32(&bars[s])), "r"(0), "r"((h + b * rowsH) * 128) : "memory");
      else
        asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                     ::"r"(s32(dst)), "l"(&map), "r"(s32(&bars[s])), "r"(0), "r"(0), "r"(h), "r"(b) : "memory");
    }
    for (int s = 0; me == 0 && s < stages && s < iters; ++s) {     // drain (issuer 0 only: an upper bound within one box)
      const int last = ((iters - 1 - s) / stages);
      const int st = (iters - 1 - s) % stages;
      uint32_t done = 0;
      while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(s32(&bars[st])), "r"((uint32_t)(last & 1)) : "memory");
    }
    if (me == 0) clocks[blockIdx.x] = clock64() - t0;
  }
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  void* fnp = nullptr;
  cudaDriverEntryPointQueryResult qr;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fnp, cudaEnableDefault, &qr);
  EncodeFn enc = (EncodeFn)fnp;
  const int B = 64, H = 64, W = 512, C = 64;            // bf16 [B][H][W][C]: 268 MB
  void* src;
  cudaMalloc(&src, (size_t)B * H * W * C * 2);
  cudaMemset(src, 0, (size_t)B * H * W * C * 2);
  long long* clk;
  cudaMalloc(&clk, 148 * sizeof(long long));
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  struct Case { const char* name; int mode, rows, strideW; } cases[] = {
      {"2-D  [128 rows x 128 B] contiguous", 2, 128, 1}, {"2-D  [64 rows x 128 B] contiguous", 2, 64, 1},
      {"4-D  [128 px x 64 ch] unit stride", 4, 128, 1},  {"4-D  [64 px x 64 ch] unit stride", 4, 64, 1},
      {"4-D  [128 px x 64 ch] stride 2 (span 256)", 4, 128, 2}, {"4-D  [64 px x 64 ch] stride 2 (span 128)", 4, 64, 2}};
  for (auto& cs : cases) {
    CUtensorMap m;
    CUresult r;
    if (cs.mode == 2) {
      cuuint64_t gd[2] = {(cuuint64_t)C, (cuuint64_t)B * H * W};
      cuuint64_t gs[1] = {(cuuint64_t)C * 2};
      cuuint32_t bx[2] = {64u, (cuuint32_t)cs.rows}, es[2] = {1u, 1u};
      r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, src, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else {
      cuuint64_t gd[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
      cuuint64_t gs[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
      cuuint32_t bx[4] = {64u, (cuuint32_t)(cs.rows * cs.strideW), 1u, 1u}, es[4] = {1u, (cuuint32_t)cs.strideW, 1u, 1u};
      r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, src, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    }
    if (r != CUDA_SUCCESS) { printf("%s: encode failed %d\n", cs.name, (int)r); continue; }
    const uint32_t box_bytes = (uint32_t)cs.rows * 128u;
    for (int cfg = 0; cfg < 6; ++cfg) {
      const int stages = 8;
      const int issuers = cfg == 0 ? 1 : (cfg == 1 ? 2 : (cfg == 2 ? 4 : (cfg == 3 ? 2 : (cfg == 4 ? 4 : 8))));
      const int lane_mode = cfg >= 3;
      const int iters = 4000;
      const size_t smem = (size_t)stages * box_bytes + stages * 8 + 1024 + 64;
      rate_kernel<<<148, 128, smem>>>(m, cs.mode, stages, box_bytes, iters, H, B, clk, issuers, lane_mode);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("%s: %s\n", cs.name, cudaGetErrorString(e)); return 1; }
      long long h[148];
      cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
      double avg = 0;
      for (int i = 0; i < 148; ++i) avg += (double)h[i];
      avg /= 148;
      printf("%-46s %d %s: %7.1f clk/box  %6.1f B/clk/SM  (%.2f TB/s chip at 1.965 GHz)\n", cs.name, issuers, lane_mode ? "lanes of one warp" : "warps            ", avg / iters,
             box_bytes * (double)iters / avg, box_bytes * (double)iters / avg * 148 * 1.965e9 / 1e12);
    }
  }
  return 0;
}
