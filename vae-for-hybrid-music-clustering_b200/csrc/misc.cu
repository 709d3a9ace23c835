// Library core (error string, device info) and the small HBM-bound helpers: ReLU backward, bias gradient,
// gate+concat, multi-tensor Adam, column statistics / affine (KMeans centering, StandardScaler).
#include <stdarg.h>
#include "common.cuh"

namespace vc {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count() {
  static int cached = 0;
  if (cached) return cached;
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
  cached = n;
  return n;
}

__global__ void relu_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ y, float* __restrict__ dx,
                                int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    dx[i] = y[i] > 0.f ? dy[i] : 0.f;
}

__global__ void relu_bwd_vec_kernel(const float4* __restrict__ dy, const float4* __restrict__ y,
                                    float4* __restrict__ dx, int64_t n4) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const float4 g = dy[i], v = y[i];
    dx[i] = make_float4(v.x > 0.f ? g.x : 0.f, v.y > 0.f ? g.y : 0.f, v.z > 0.f ? g.z : 0.f,
                        v.w > 0.f ? g.w : 0.f);
  }
}

// db[c] = sum_{o,i} dy[o,C,inner].  Two deterministic stages: block (c, s) sums slice s of the (outer x inner) range of
// channel c into part[c*S + s] (double), then one thread per channel folds the S partials in order.
constexpr int kBiasSplits = 64;

__global__ void __launch_bounds__(256)
bias_grad_stage1(const float* __restrict__ dy, int64_t outer, int C, int64_t inner, int S, double* __restrict__ part) {
  __shared__ double scratch[32];
  const int c = blockIdx.x, sp = blockIdx.y;
  double acc = 0.0;
  if (inner == 1) {
    for (int64_t o = sp * blockDim.x + threadIdx.x; o < outer; o += (int64_t)S * blockDim.x) acc += (double)dy[o * C + c];
  } else {
    // slices of the flattened (outer, inner/4) index space, float4 loads when rows are 16-byte aligned
    const bool vec = (inner % 4 == 0) && ((reinterpret_cast<uintptr_t>(dy) & 15) == 0);
    if (vec && C == 1) {
      // single channel (the reconstruction's bias): dy is one contiguous vector, no index arithmetic
      const int64_t total = outer * (inner / 4);
      const float4* p4 = reinterpret_cast<const float4*>(dy);
      for (int64_t e = sp * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)S * blockDim.x) {
        const float4 v = __ldcs(p4 + e);
        acc += (double)((v.x + v.y) + (v.z + v.w));
      }
    } else if (vec) {
      const int64_t i4 = inner / 4, total = outer * i4;
      for (int64_t e = sp * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)S * blockDim.x) {
        const int64_t o = e / i4, i = e % i4;
        const float4 v = *reinterpret_cast<const float4*>(dy + (o * C + c) * inner + 4 * i);
        acc += (double)((v.x + v.y) + (v.z + v.w));
      }
    } else {
      const int64_t total = outer * inner;
      for (int64_t e = sp * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)S * blockDim.x) {
        const int64_t o = e / inner, i = e % inner;
        acc += (double)dy[(o * C + c) * inner + i];
      }
    }
  }
  acc = block_sum<double>(acc, scratch);
  if (threadIdx.x == 0) part[(int64_t)c * S + sp] = acc;
}

// inner == 1 (Linear layers): thread = column, so every row is read as full coalesced lines; rows sp, sp+S, ...
__global__ void bias_grad_cols_stage1(const float* __restrict__ dy, int64_t outer, int C, int S,
                                      double* __restrict__ part) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x, sp = blockIdx.y;
  if (c >= C) return;
  double acc = 0.0;
  for (int64_t o = sp; o < outer; o += S) acc += (double)dy[o * C + c];
  part[(int64_t)c * S + sp] = acc;
}

// ReLU backward and bias gradient in ONE pass: g = dy * (y > 0) is written and summed per channel while it is in
// registers.  Channel c's elements, flattened over (o, i), are cut into S equal slices and block (c, sp) streams slice
// sp; the host sizes S so that C*S blocks are exactly the blocks the GPU holds at once (one balanced wave: no tail of
// half-empty SMs, which cost the plane-per-block version 13 % at 2048 or 4096 equal blocks on 1184 slots).  Inside a
// slice the loop walks plane by plane, so there is no division per element.  Partials land in the layout
// bias_grad_stage2 folds.
template <typename V>   // float4 (inner % 4 == 0, 16-byte aligned) or float
__device__ __forceinline__ double relu_bwd_slice(const float* __restrict__ dy, const float* __restrict__ y,
                                                 float* __restrict__ g, int64_t outer, int C, int64_t inner, int S,
                                                 int c, int sp) {
  constexpr int W = sizeof(V) / 4;
  const int64_t iw = inner / W, total = outer * iw;
  const int64_t len = (total + S - 1) / S;
  const int64_t lo = (int64_t)sp * len, hi = lo + len < total ? lo + len : total;
  double acc = 0.0;
  for (int64_t o = lo / iw; o * iw < hi; ++o) {
    const int64_t first = o * iw;
    const int64_t s0 = (lo > first ? lo : first) - first, s1 = (hi < first + iw ? hi : first + iw) - first;
    const int64_t base = (o * C + c) * inner;
    const V* dv = reinterpret_cast<const V*>(dy + base);
    const V* yv = reinterpret_cast<const V*>(y + base);
    V* gv = reinterpret_cast<V*>(g + base);
    float a = 0.f;
    if constexpr (W == 4) {
      int64_t i = s0 + threadIdx.x;
      for (; i + blockDim.x < s1; i += 2 * blockDim.x) {   // two independent 16-byte load pairs in flight per thread
        const float4 d0 = __ldcs(dv + i), v0 = __ldcs(yv + i);
        const float4 d1 = __ldcs(dv + i + blockDim.x), v1 = __ldcs(yv + i + blockDim.x);
        const float4 r0 = make_float4(v0.x > 0.f ? d0.x : 0.f, v0.y > 0.f ? d0.y : 0.f, v0.z > 0.f ? d0.z : 0.f,
                                      v0.w > 0.f ? d0.w : 0.f);
        const float4 r1 = make_float4(v1.x > 0.f ? d1.x : 0.f, v1.y > 0.f ? d1.y : 0.f, v1.z > 0.f ? d1.z : 0.f,
                                      v1.w > 0.f ? d1.w : 0.f);
        gv[i] = r0;
        gv[i + blockDim.x] = r1;
        a += ((r0.x + r0.y) + (r0.z + r0.w)) + ((r1.x + r1.y) + (r1.z + r1.w));
      }
      if (i < s1) {
        const float4 d0 = __ldcs(dv + i), v0 = __ldcs(yv + i);
        const float4 r0 = make_float4(v0.x > 0.f ? d0.x : 0.f, v0.y > 0.f ? d0.y : 0.f, v0.z > 0.f ? d0.z : 0.f,
                                      v0.w > 0.f ? d0.w : 0.f);
        gv[i] = r0;
        a += (r0.x + r0.y) + (r0.z + r0.w);
      }
    } else {
      for (int64_t i = s0 + threadIdx.x; i < s1; i += blockDim.x) {
        const float r = yv[i] > 0.f ? dv[i] : 0.f;
        gv[i] = r;
        a += r;
      }
    }
    acc += (double)a;
  }
  return acc;
}

__global__ void __launch_bounds__(256)
relu_bwd_bias_stage1(const float* __restrict__ dy, const float* __restrict__ y, float* __restrict__ g, int64_t outer,
                     int C, int64_t inner, int S, double* __restrict__ part) {
  __shared__ double scratch[32];
  const int c = blockIdx.x, sp = blockIdx.y;
  const bool vec = (inner % 4 == 0) && (((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(y) |
                                          reinterpret_cast<uintptr_t>(g)) & 15) == 0);
  double acc = vec ? relu_bwd_slice<float4>(dy, y, g, outer, C, inner, S, c, sp)
                   : relu_bwd_slice<float>(dy, y, g, outer, C, inner, S, c, sp);
  acc = block_sum<double>(acc, scratch);
  if (threadIdx.x == 0) part[(int64_t)c * S + sp] = acc;
}
// inner == 1 (Linear): thread = column
__global__ void relu_bwd_bias_cols_stage1(const float* __restrict__ dy, const float* __restrict__ y,
                                          float* __restrict__ g, int64_t outer, int C, int S, double* __restrict__ part) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x, sp = blockIdx.y;
  if (c >= C) return;
  double acc = 0.0;
  for (int64_t o = sp; o < outer; o += S) {
    const float r = y[o * C + c] > 0.f ? dy[o * C + c] : 0.f;
    g[o * C + c] = r;
    acc += (double)r;
  }
  part[(int64_t)c * S + sp] = acc;
}

// inner == 1, C % 4 == 0, 16-byte aligned: thread = four columns as one float4, rows sp, sp+S, ... (four row steps in
// flight); partials in [S][C] order so that the four doubles a thread writes are contiguous.  The scalar kernel above
// had one row split (S = 1) at C = 262144, i.e. 64 dependent 4-byte steps per thread: 2.6 TB/s.
__global__ void __launch_bounds__(256)
relu_bwd_bias_cols4_stage1(const float4* __restrict__ dy, const float4* __restrict__ y, float4* __restrict__ g,
                           int64_t outer, int C4, int S, double* __restrict__ part) {
  const int c4 = blockIdx.x * blockDim.x + threadIdx.x, sp = blockIdx.y;
  if (c4 >= C4) return;
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll 4
  for (int64_t o = sp; o < outer; o += S) {
    const float4 d = __ldcs(dy + o * C4 + c4), v = __ldcs(y + o * C4 + c4);
    const float4 r = make_float4(v.x > 0.f ? d.x : 0.f, v.y > 0.f ? d.y : 0.f, v.z > 0.f ? d.z : 0.f,
                                 v.w > 0.f ? d.w : 0.f);
    g[o * C4 + c4] = r;
    a0 += (double)r.x; a1 += (double)r.y; a2 += (double)r.z; a3 += (double)r.w;
  }
  double* p = part + ((int64_t)sp * C4 + c4) * 4;
  reinterpret_cast<double2*>(p)[0] = make_double2(a0, a1);
  reinterpret_cast<double2*>(p)[1] = make_double2(a2, a3);
}
__global__ void bias_grad_stage2_rows(const double* __restrict__ part, int C, int S, float* __restrict__ db) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double acc = 0.0;
  for (int s = 0; s < S; ++s) acc += part[(int64_t)s * C + c];
  db[c] = (float)acc;
}

__global__ void bias_grad_stage2(const double* __restrict__ part, int C, int S, float* __restrict__ db, int accumulate) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double acc = 0.0;
  for (int s = 0; s < S; ++s) acc += part[(int64_t)c * S + s];
  db[c] = (accumulate ? db[c] : 0.f) + (float)acc;
}

// the same with one warp per channel (many partials, few channels: the single-channel reconstruction bias has 512+):
// lane l adds partials l, l+32, ... and the 32 sub-sums meet in a fixed butterfly, so the result does not vary by run
__global__ void bias_grad_stage2_warp(const double* __restrict__ part, int C, int S, float* __restrict__ db,
                                      int accumulate) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (c >= C) return;
  double acc = 0.0;
  for (int s = lane; s < S; s += 32) acc += part[(int64_t)c * S + s];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) db[c] = (accumulate ? db[c] : 0.f) + (float)acc;
}

static void launch_bias_stage2(const double* part, int C, int S, float* db, int accumulate, cudaStream_t st) {
  if (S >= 32 && C <= 65536)
    bias_grad_stage2_warp<<<(C * 32 + 127) / 128, 128, 0, st>>>(part, C, S, db, accumulate);
  else
    bias_grad_stage2<<<(C + 127) / 128, 128, 0, st>>>(part, C, S, db, accumulate);
}

__global__ void cat_gate_fwd_kernel(const float* __restrict__ a, int na, const float* __restrict__ l, int nl,
                                    const float* __restrict__ gate, int append_gate, int B,
                                    float* __restrict__ out) {
  const int wout = na + nl + (append_gate ? 1 : 0);
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * wout) return;
  const int r = i / wout, c = i % wout;
  const float g = gate ? gate[r] : 1.f;
  float v;
  if (c < na) v = a[r * na + c];
  else if (c < na + nl) v = l[r * nl + (c - na)] * g;
  else v = g;
  out[i] = v;
}

__global__ void cat_gate_bwd_kernel(const float* __restrict__ dout, int na, int nl, const float* __restrict__ gate,
                                    int append_gate, int B, float* __restrict__ da, float* __restrict__ dl) {
  const int wout = na + nl + (append_gate ? 1 : 0);
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * (na + nl)) return;
  const int r = i / (na + nl), c = i % (na + nl);
  const float g = gate ? gate[r] : 1.f;
  const float v = dout[r * wout + c];
  if (c < na) {
    if (da) da[r * na + c] = v;
  } else {
    if (dl) dl[r * nl + (c - na)] = v * g;
  }
}

constexpr int kAdamThreads = 256;

__global__ void __launch_bounds__(kAdamThreads)
adam_multi_kernel(const VcAdamChunk* __restrict__ table, float lr, float b1, float b2, float eps, float bc1,
                  float bc2, const VcAdamDeviceStep* __restrict__ dev_step) {
  const VcAdamChunk ch = table[blockIdx.x];
  if (dev_step) {   // graph-capturable form: the bias corrections of this step live in device memory
    bc1 = dev_step->bias_corr1;
    bc2 = dev_step->bias_corr2;
  }
  // torch.optim.Adam (single-tensor, non-capturable) arithmetic order:
  //   m = b1*m + (1-b1)*g ; v = b2*v + (1-b2)*g*g ; p -= (lr/bc1) * m / (sqrt(v)/sqrt(bc2) + eps)
  const float step_size = lr / bc1;
  const float inv_sqrt_bc2 = 1.f / sqrtf(bc2);
  const bool vec = ((reinterpret_cast<uintptr_t>(ch.p) | reinterpret_cast<uintptr_t>(ch.g) |
                     reinterpret_cast<uintptr_t>(ch.m) | reinterpret_cast<uintptr_t>(ch.v)) & 15) == 0;
  const int64_t n = ch.n;
  if (vec) {
    const int64_t n4 = n >> 2;
    float4* p4 = reinterpret_cast<float4*>(ch.p);
    const float4* g4 = reinterpret_cast<const float4*>(ch.g);
    float4* m4 = reinterpret_cast<float4*>(ch.m);
    float4* v4 = reinterpret_cast<float4*>(ch.v);
    for (int64_t i = threadIdx.x; i < n4; i += kAdamThreads) {
      float4 p = p4[i], m = m4[i], v = v4[i];
      const float4 g = __ldcs(g4 + i);
#define VC_ADAM1(c)                                           \
  m.c = b1 * m.c + (1.f - b1) * g.c;                          \
  v.c = b2 * v.c + (1.f - b2) * g.c * g.c;                    \
  p.c -= step_size * (m.c / (sqrtf(v.c) * inv_sqrt_bc2 + eps));
      VC_ADAM1(x) VC_ADAM1(y) VC_ADAM1(z) VC_ADAM1(w)
#undef VC_ADAM1
      p4[i] = p;
      m4[i] = m;
      v4[i] = v;
    }
    for (int64_t i = (n4 << 2) + threadIdx.x; i < n; i += kAdamThreads) {
      const float g = ch.g[i];
      const float m = b1 * ch.m[i] + (1.f - b1) * g;
      const float v = b2 * ch.v[i] + (1.f - b2) * g * g;
      ch.m[i] = m;
      ch.v[i] = v;
      ch.p[i] -= step_size * (m / (sqrtf(v) * inv_sqrt_bc2 + eps));
    }
  } else {
    for (int64_t i = threadIdx.x; i < n; i += kAdamThreads) {
      const float g = ch.g[i];
      const float m = b1 * ch.m[i] + (1.f - b1) * g;
      const float v = b2 * ch.v[i] + (1.f - b2) * g * g;
      ch.m[i] = m;
      ch.v[i] = v;
      ch.p[i] -= step_size * (m / (sqrtf(v) * inv_sqrt_bc2 + eps));
    }
  }
}

// ---- column statistics: two-stage, fixed order.  Stage 1: each block owns a contiguous row range and writes
// per-column (sum, sumsq about a pivot row) in double.  d <= 1024.
constexpr int kColBlocks = 148 * 4;

__global__ void col_stats_stage1(const float* __restrict__ X, int64_t N, int d, double* __restrict__ part) {
  // threads: x = column (<= d rounded), y = row lane
  extern __shared__ double sh[];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int64_t rows_per = (N + gridDim.x - 1) / gridDim.x;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per;
  const int64_t r1 = r0 + rows_per < N ? r0 + rows_per : N;
  for (int c = tx; c < d; c += blockDim.x) {
    const double pivot = (double)X[c];  // row 0 as pivot keeps sumsq well conditioned
    double s = 0.0, q = 0.0;
    for (int64_t r = r0 + ty; r < r1; r += blockDim.y) {
      const double v = (double)X[r * d + c] - pivot;
      s += v;
      q += v * v;
    }
    sh[(ty * d + c) * 2 + 0] = s;
    sh[(ty * d + c) * 2 + 1] = q;
  }
  __syncthreads();
  if (ty == 0) {
    for (int c = tx; c < d; c += blockDim.x) {
      double s = 0.0, q = 0.0;
      for (int y = 0; y < blockDim.y; ++y) {
        s += sh[(y * d + c) * 2 + 0];
        q += sh[(y * d + c) * 2 + 1];
      }
      part[((int64_t)blockIdx.x * d + c) * 2 + 0] = s;
      part[((int64_t)blockIdx.x * d + c) * 2 + 1] = q;
    }
  }
}

__global__ void col_stats_stage2(const float* __restrict__ X, const double* __restrict__ part, int nblocks,
                                 int64_t N, int d, double* __restrict__ mean, double* __restrict__ var) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= d) return;
  double s = 0.0, q = 0.0;
  for (int b = 0; b < nblocks; ++b) {
    s += part[((int64_t)b * d + c) * 2 + 0];
    q += part[((int64_t)b * d + c) * 2 + 1];
  }
  const double pivot = (double)X[c];
  const double ms = s / (double)N;
  mean[c] = pivot + ms;
  if (var) {
    double v = q / (double)N - ms * ms;
    var[c] = v > 0.0 ? v : 0.0;
  }
}

__global__ void col_affine_kernel(const float* __restrict__ X, int64_t n, int d, const double* __restrict__ shift,
                                  const double* __restrict__ scale, float* __restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int c = (int)(i % d);
    // numpy's in-place `X -= mean; X /= scale` on a float32 array with float64 operands: each op is
    // evaluated in double and rounded back to float32.
    float v = (float)((double)X[i] - shift[c]);
    if (scale) v = (float)((double)v / scale[c]);
    out[i] = v;
  }
}

}  // namespace vc

using namespace vc;

extern "C" const char* vc_last_error_string(void) { return g_err; }
extern "C" int vc_version(void) { return 100; }
extern "C" int vc_sm_count(void) { return sm_count(); }

static inline int grid_for(int64_t n, int threads, int per_thread) {
  int64_t b = (n + (int64_t)threads * per_thread - 1) / ((int64_t)threads * per_thread);
  const int64_t cap = (int64_t)sm_count() * 16;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

extern "C" int vc_relu_bwd(const float* dy, const float* y, float* dx, int64_t n, void* stream) {
  VC_REQUIRE(dy && y && dx && n >= 0, VC_ERR_BAD_ARG, "vc_relu_bwd: bad arguments");
  if (n == 0) return VC_OK;
  const bool vec = n % 4 == 0 && ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(y) |
                                   reinterpret_cast<uintptr_t>(dx)) & 15) == 0;
  if (vec)
    relu_bwd_vec_kernel<<<grid_for(n / 4, 256, 2), 256, 0, as_stream(stream)>>>(
        reinterpret_cast<const float4*>(dy), reinterpret_cast<const float4*>(y), reinterpret_cast<float4*>(dx),
        n / 4);
  else
    relu_bwd_kernel<<<grid_for(n, 256, 4), 256, 0, as_stream(stream)>>>(dy, y, dx, n);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" int vc_relu_bwd_bias(const float* dy, const float* y, int64_t outer, int C, int64_t inner, float* g, float* db,
                                void* workspace, size_t workspace_bytes, void* stream) {
  VC_REQUIRE(dy && y && g && db && outer > 0 && C > 0 && inner > 0, VC_ERR_BAD_ARG, "vc_relu_bwd_bias: bad arguments");
  VC_REQUIRE(workspace && workspace_bytes >= vc_bias_grad_workspace_bytes(C), VC_ERR_WORKSPACE,
             "vc_relu_bwd_bias: workspace too small");
  int S = kBiasSplits;
  double* part = static_cast<double*>(workspace);
  if (inner == 1 && C % 4 == 0 && C >= 4096 &&
      ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(g)) & 15) == 0) {
    const int C4 = C / 4;
    S = 1;   // enough row splits for ~2 threads per resident-thread slot, at least 8 rows each
    while (S < kBiasSplits && (int64_t)C4 * S < 2 * 2048 * (int64_t)sm_count() && outer / (2 * S) >= 8) S <<= 1;
    relu_bwd_bias_cols4_stage1<<<dim3((C4 + 255) / 256, S), 256, 0, as_stream(stream)>>>(
        reinterpret_cast<const float4*>(dy), reinterpret_cast<const float4*>(y), reinterpret_cast<float4*>(g), outer, C4,
        S, part);
    VC_LAUNCH_CHECK();
    bias_grad_stage2_rows<<<(C + 127) / 128, 128, 0, as_stream(stream)>>>(part, C, S, db);
    VC_LAUNCH_CHECK();
    return VC_OK;
  } else if (inner == 1) {
    while (S > 1 && (S > outer || (int64_t)C * S > 262144)) S >>= 1;
    relu_bwd_bias_cols_stage1<<<dim3((C + 127) / 128, S), 128, 0, as_stream(stream)>>>(dy, y, g, outer, C, S, part);
  } else {
    // one balanced wave: as many equal slices per channel as there are resident-block slots (never more than the
    // workspace's kBiasSplits per channel, never slices shorter than ~one pass of the block)
    static int slots = 0;
    if (slots == 0) {
      int per_sm = 0;
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, relu_bwd_bias_stage1, 256, 0);
      slots = (per_sm > 0 ? per_sm : 1) * sm_count();
    }
    S = slots / C;
    if (S > kBiasSplits) S = kBiasSplits;
    const int64_t units = outer * (inner % 4 == 0 ? inner / 4 : inner);
    while (S > 1 && units / S < 256) --S;
    if (S < 1) S = 1;
    relu_bwd_bias_stage1<<<dim3(C, S), 256, 0, as_stream(stream)>>>(dy, y, g, outer, C, inner, S, part);
  }
  VC_LAUNCH_CHECK();
  launch_bias_stage2(part, C, S, db, 0, as_stream(stream));
  VC_LAUNCH_CHECK();
  return VC_OK;
}

// kBiasSplits partials per channel, at least 1024 in total so that a few channels still fill the GPU
extern "C" size_t vc_bias_grad_workspace_bytes(int C) {
  const size_t n = (size_t)kBiasSplits * (C > 0 ? C : 1);
  return sizeof(double) * (n > 1024 ? n : 1024);
}

extern "C" int vc_bias_grad(const float* dy, int64_t outer, int C, int64_t inner, float* db, int accumulate,
                            void* workspace, size_t workspace_bytes, void* stream) {
  VC_REQUIRE(dy && db && outer > 0 && C > 0 && inner > 0, VC_ERR_BAD_ARG, "vc_bias_grad: bad arguments");
  VC_REQUIRE(workspace && workspace_bytes >= vc_bias_grad_workspace_bytes(C), VC_ERR_WORKSPACE,
             "vc_bias_grad: workspace too small");
  int S = kBiasSplits;
  double* part = static_cast<double*>(workspace);
  if (inner == 1) {
    while (S > 1 && (S > outer || (int64_t)C * S > 262144)) S >>= 1;
    bias_grad_cols_stage1<<<dim3((C + 127) / 128, S), 128, 0, as_stream(stream)>>>(dy, outer, C, S, part);
  } else {
    while ((int64_t)C * S < 512) S <<= 1;   // few channels: split further so the grid still fills the GPU
    while (S > 1 && outer * inner / S < 1024) S >>= 1;
    bias_grad_stage1<<<dim3(C, S), 256, 0, as_stream(stream)>>>(dy, outer, C, inner, S, part);
  }
  VC_LAUNCH_CHECK();
  launch_bias_stage2(part, C, S, db, accumulate, as_stream(stream));
  VC_LAUNCH_CHECK();
  return VC_OK;
}

// ---- vc_bias_grad without any caller-provided or allocated scratch ---------------------------------------------------
// For callers that cannot allocate (CUDA graph capture, serving loops): the per-channel double partials live in a
// file-scope device array.  Every GPU that loads the library holds these 2 MiB until the process exits.  In host code
// the array's NAME is only a host-side stand-in; the device address is looked up once per device with
// cudaGetSymbolAddress and that pointer is what the kernels receive.  All calls on a device share the array: two calls
// that overlap on the GPU (different streams, no dependency) race.  Nothing in it has to start at zero -- stage 1 writes
// every partial that stage 2 reads -- so leftovers of the previous call are harmless.
constexpr int kNoAllocMaxC = 4096;                                   // channels per call
static __device__ __align__(256) double g_bias_scratch[(size_t)kBiasSplits * kNoAllocMaxC];
constexpr int kNoAllocMaxDevices = 64;
static double* g_bias_scratch_ptr[kNoAllocMaxDevices] = {nullptr};   // device address of the array, per device ordinal

static int bias_scratch_address(double** out) {
  int dev = -1;
  VC_CUDA(cudaGetDevice(&dev));
  VC_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, VC_ERR_UNSUPPORTED, "vc_bias_grad_noalloc: device ordinal %d >= %d",
             dev, kNoAllocMaxDevices);
  if (!g_bias_scratch_ptr[dev]) {
    double* p = nullptr;
    VC_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_bias_scratch));
    VC_REQUIRE(p != nullptr, VC_ERR_CUDA, "vc_bias_grad_noalloc: cudaGetSymbolAddress returned a null address");
    g_bias_scratch_ptr[dev] = p;
  }
  *out = g_bias_scratch_ptr[dev];
  return VC_OK;
}

extern "C" int vc_bias_grad_noalloc_max_channels(void) { return kNoAllocMaxC; }

extern "C" int vc_bias_grad_noalloc(const float* dy, int64_t outer, int C, int64_t inner, float* db, int accumulate,
                                    void* stream) {
  VC_REQUIRE(dy && db && outer > 0 && C > 0 && inner > 0, VC_ERR_BAD_ARG, "vc_bias_grad_noalloc: bad arguments");
  VC_REQUIRE(C <= kNoAllocMaxC, VC_ERR_UNSUPPORTED,
             "vc_bias_grad_noalloc: C = %d exceeds the built-in scratch (at most %d channels); use vc_bias_grad with a "
             "caller workspace", C, kNoAllocMaxC);
  double* part = nullptr;
  const int rc = bias_scratch_address(&part);
  if (rc != VC_OK) return rc;
  int S = kBiasSplits;   // same split rules as vc_bias_grad: S <= kBiasSplits or C * S <= 1024 <= the array
  if (inner == 1) {
    while (S > 1 && (S > outer || (int64_t)C * S > 262144)) S >>= 1;
    bias_grad_cols_stage1<<<dim3((C + 127) / 128, S), 128, 0, as_stream(stream)>>>(dy, outer, C, S, part);
  } else {
    while ((int64_t)C * S < 512) S <<= 1;
    while (S > 1 && outer * inner / S < 1024) S >>= 1;
    bias_grad_stage1<<<dim3(C, S), 256, 0, as_stream(stream)>>>(dy, outer, C, inner, S, part);
  }
  VC_LAUNCH_CHECK();
  launch_bias_stage2(part, C, S, db, accumulate, as_stream(stream));
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" int vc_cat_gate_fwd(const float* a, int na, const float* l, int nl, const float* gate, int append_gate,
                               int B, float* out, void* stream) {
  VC_REQUIRE(out && B > 0 && na >= 0 && nl >= 0 && (na == 0 || a) && (nl == 0 || l), VC_ERR_BAD_ARG,
             "vc_cat_gate_fwd: bad arguments");
  VC_REQUIRE(!append_gate || gate, VC_ERR_BAD_ARG, "vc_cat_gate_fwd: append_gate without gate");
  const int n = B * (na + nl + (append_gate ? 1 : 0));
  cat_gate_fwd_kernel<<<(n + 255) / 256, 256, 0, as_stream(stream)>>>(a, na, l, nl, gate, append_gate, B, out);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" int vc_cat_gate_bwd(const float* dout, int na, int nl, const float* gate, int append_gate, int B,
                               float* da, float* dl, void* stream) {
  VC_REQUIRE(dout && B > 0 && na >= 0 && nl >= 0, VC_ERR_BAD_ARG, "vc_cat_gate_bwd: bad arguments");
  const int n = B * (na + nl);
  if (n == 0) return VC_OK;
  cat_gate_bwd_kernel<<<(n + 255) / 256, 256, 0, as_stream(stream)>>>(dout, na, nl, gate, append_gate, B, da, dl);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" int vc_adam_multi(const VcAdamChunk* table, int n_chunks, float lr, float beta1, float beta2, float eps,
                             float bias_corr1, float bias_corr2, void* stream) {
  VC_REQUIRE(table && n_chunks >= 0, VC_ERR_BAD_ARG, "vc_adam_multi: bad arguments");
  if (n_chunks == 0) return VC_OK;
  adam_multi_kernel<<<n_chunks, kAdamThreads, 0, as_stream(stream)>>>(table, lr, beta1, beta2, eps, bias_corr1,
                                                                     bias_corr2, nullptr);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

// step += 1 and the bias corrections of that step, in double like the host path (1.0 - beta ** t in Python)
__global__ void adam_tick_kernel(VcAdamDeviceStep* st, float b1, float b2) {
  const int t = st->step + 1;
  st->step = t;
  st->bias_corr1 = (float)(1.0 - pow((double)b1, (double)t));
  st->bias_corr2 = (float)(1.0 - pow((double)b2, (double)t));
}

extern "C" int vc_adam_multi_graph(const VcAdamChunk* table, int n_chunks, float lr, float beta1, float beta2, float eps,
                                   VcAdamDeviceStep* step_state, void* stream) {
  VC_REQUIRE(table && n_chunks >= 0 && step_state, VC_ERR_BAD_ARG, "vc_adam_multi_graph: bad arguments");
  adam_tick_kernel<<<1, 1, 0, as_stream(stream)>>>(step_state, beta1, beta2);
  VC_LAUNCH_CHECK();
  if (n_chunks == 0) return VC_OK;
  adam_multi_kernel<<<n_chunks, kAdamThreads, 0, as_stream(stream)>>>(table, lr, beta1, beta2, eps, 1.f, 1.f, step_state);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

// The two halves of vc_adam_multi_graph as separate entry points: a step that updates its parameters in several
// launches (large gradients as soon as backward has produced them, on a side stream, under the rest of backward)
// advances the device step counter once, then every launch reads the bias corrections of that step.
extern "C" int vc_adam_tick(VcAdamDeviceStep* step_state, float beta1, float beta2, void* stream) {
  VC_REQUIRE(step_state, VC_ERR_BAD_ARG, "vc_adam_tick: bad arguments");
  adam_tick_kernel<<<1, 1, 0, as_stream(stream)>>>(step_state, beta1, beta2);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" int vc_adam_multi_at(const VcAdamChunk* table, int n_chunks, float lr, float beta1, float beta2, float eps,
                                const VcAdamDeviceStep* step_state, void* stream) {
  VC_REQUIRE(table && n_chunks >= 0 && step_state, VC_ERR_BAD_ARG, "vc_adam_multi_at: bad arguments");
  if (n_chunks == 0) return VC_OK;
  adam_multi_kernel<<<n_chunks, kAdamThreads, 0, as_stream(stream)>>>(table, lr, beta1, beta2, eps, 1.f, 1.f, step_state);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" size_t vc_col_stats_workspace_bytes(int d) { return sizeof(double) * 2 * (size_t)kColBlocks * d; }

extern "C" int vc_col_stats(const float* X, int64_t N, int d, double* mean, double* var, void* workspace,
                            size_t workspace_bytes, void* stream) {
  VC_REQUIRE(X && mean && N > 0 && d > 0 && d <= 1024, VC_ERR_BAD_ARG, "vc_col_stats: bad arguments (N=%lld d=%d)",
             (long long)N, d);
  VC_REQUIRE(workspace && workspace_bytes >= vc_col_stats_workspace_bytes(d), VC_ERR_WORKSPACE,
             "vc_col_stats: workspace too small");
  int tx = 32;
  while (tx < d && tx < 256) tx <<= 1;
  int ty = 512 / tx;
  // shared memory: ty*d*2 doubles
  while ((size_t)ty * d * 2 * sizeof(double) > 40000 && ty > 1) ty >>= 1;
  int blocks = kColBlocks;
  if ((int64_t)blocks * ty > N) blocks = (int)((N + ty - 1) / ty);
  if (blocks < 1) blocks = 1;
  double* part = static_cast<double*>(workspace);
  col_stats_stage1<<<blocks, dim3(tx, ty), (size_t)ty * d * 2 * sizeof(double), as_stream(stream)>>>(X, N, d, part);
  VC_LAUNCH_CHECK();
  col_stats_stage2<<<(d + 127) / 128, 128, 0, as_stream(stream)>>>(X, part, blocks, N, d, mean, var);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" int vc_col_affine(const float* X, int64_t N, int d, const double* shift, const double* scale, float* out,
                             void* stream) {
  VC_REQUIRE(X && out && shift && N > 0 && d > 0, VC_ERR_BAD_ARG, "vc_col_affine: bad arguments");
  const int64_t n = N * d;
  col_affine_kernel<<<grid_for(n, 256, 4), 256, 0, as_stream(stream)>>>(X, n, d, shift, scale, out);
  VC_LAUNCH_CHECK();
  return VC_OK;
}
