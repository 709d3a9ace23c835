// Shared helpers for libvaeclust.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/vaeclust.h"

namespace vc {

void set_error(const char* fmt, ...);

#define VC_REQUIRE(cond, code, ...)                 \
  do {                                              \
    if (!(cond)) {                                  \
      ::vc::set_error(__VA_ARGS__);                 \
      return (code);                                \
    }                                               \
  } while (0)

#define VC_LAUNCH_CHECK()                                                        \
  do {                                                                           \
    cudaError_t e__ = cudaGetLastError();                                        \
    if (e__ != cudaSuccess) {                                                    \
      ::vc::set_error("%s:%d CUDA launch failed: %s", __FILE__, __LINE__,        \
                      cudaGetErrorString(e__));                                  \
      return VC_ERR_CUDA;                                                        \
    }                                                                            \
  } while (0)

#define VC_CUDA(call)                                                            \
  do {                                                                           \
    cudaError_t e__ = (call);                                                    \
    if (e__ != cudaSuccess) {                                                    \
      ::vc::set_error("%s:%d %s: %s", __FILE__, __LINE__, #call,                 \
                      cudaGetErrorString(e__));                                  \
      return VC_ERR_CUDA;                                                        \
    }                                                                            \
  } while (0)

int sm_count();

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Predicated read-only load without a branch region: 0.f when pred == 0 (the address is then never dereferenced).
__device__ __forceinline__ float ldg_pred(const float* p, uint32_t pred) {
  float v;
  asm("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %2, 0;\n\tmov.f32 %0, 0f00000000;\n\t@q ld.global.nc.f32 %0, [%1];\n\t}"
      : "=f"(v)
      : "l"(p), "r"(pred));
  return v;
}

// Sum of p[s * stride] for s = first, first + step, ... < n, added in that order (the result does not depend on the batching).
// Eight loads are in flight at a time: the fold kernels that finish every split-K / per-block reduction were chains of
// dependent L2 round trips (10-30 us for a few hundred partials) when written as a plain loop.
template <typename T>
__device__ __forceinline__ T strided_sum_ordered(const T* __restrict__ p, int64_t stride, int first, int n, int step) {
  T acc = 0;
  int s = first;
  for (; s + 7 * step < n; s += 8 * step) {
    T v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = __ldcs(p + (int64_t)(s + u * step) * stride);
#pragma unroll
    for (int u = 0; u < 8; ++u) acc += v[u];
  }
  for (; s < n; s += step) acc += __ldcs(p + (int64_t)s * stride);
  return acc;
}

// Block-wide sum, result valid in thread 0. `scratch` holds >= 32 elements.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  if (warp == 0) {
    v = lane < nw ? scratch[lane] : T(0);
    v = warp_sum(v);
  }
  return v;
}

}  // namespace vc
