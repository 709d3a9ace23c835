// Silhouette: tiled, row-blocked pairwise-distance reduction (fp32 FFMA path).
// Rows are pre-grouped by cluster (vc_silhouette_gather), so every 128-column tile belongs to exactly one cluster
// and the per-cluster bucketed sum of sklearn's _silhouette_reduce (np.bincount per row,
// sklearn:metrics/cluster/_unsupervised.py:184-189) becomes a plain row sum in registers.
// d^2 = ||x||^2 - 2 x.y + ||y||^2 (pairwise.py:376-427), clamped at 0, sqrt, self-distance forced to 0 (:2285).
// The tcgen05 variant lives in silhouette_tc.cu; this one is the fp32 parity path.
#include "common.cuh"

namespace vc {

constexpr int SB = 128;   // rows and columns per tile
constexpr int SK = 16;    // k-chunk

__global__ void silhouette_gather_kernel(const float* __restrict__ X, int64_t N, int d,
                                         const int64_t* __restrict__ order, const double* __restrict__ shift,
                                         float* __restrict__ Xs, float* __restrict__ sqn) {
  const int lane = threadIdx.x & 31;
  const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t i = w; i < N; i += nw) {
    const int64_t src = order ? order[i] : i;
    float s = 0.f;
    for (int j = lane; j < d; j += 32) {
      float v = X[src * d + j];
      if (shift) v = (float)((double)v - shift[j]);
      Xs[i * d + j] = v;
      s = fmaf(v, v, s);
    }
    s = warp_sum(s);
    if (lane == 0) sqn[i] = s;
  }
}

// Same close-pair rule as the tensor-core kernel (silhouette_tc.cu): below 1/256 of |x|^2 + |y|^2 the norm formula
// has cancelled most of its bits, so the pair is recomputed as sum (x_k - y_k)^2 (duplicates come out exactly 0).
constexpr float kCloseRatioF = 1.f / 256.f;
__device__ __noinline__ float exact_dist_f32(const float* __restrict__ Xs, int d, int64_t i, int64_t j) {
  const float4* a = reinterpret_cast<const float4*>(Xs + i * d);
  const float4* b = reinterpret_cast<const float4*>(Xs + j * d);
  float acc = 0.f;
  for (int t = 0; t < (d >> 2); ++t) {
    const float4 x = a[t], y = b[t];
    const float d0 = x.x - y.x, d1 = x.y - y.y, d2 = x.z - y.z, d3 = x.w - y.w;
    acc = fmaf(d0, d0, acc); acc = fmaf(d1, d1, acc); acc = fmaf(d2, d2, acc); acc = fmaf(d3, d3, acc);
  }
  return sqrtf(acc);
}

__global__ void __launch_bounds__(256)
silhouette_rows_kernel(const float* __restrict__ Xs, const float* __restrict__ sqn, int64_t N, int d,
                       const int64_t* __restrict__ off, int k, int64_t row_begin, int64_t row_end, int nsplit,
                       float* __restrict__ out) {
  __shared__ float As[SK][SB + 4];
  __shared__ float Bs[SK][SB + 4];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t r0 = row_begin + (int64_t)blockIdx.x * SB;
  const int64_t nrows = row_end - row_begin;
  const int split = blockIdx.y;
  float* o = out + (int64_t)split * nrows * k;

  float rn[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t r = r0 + ty + 16 * i;
    rn[i] = r < row_end ? sqn[r] : 0.f;
  }
  int64_t tile_counter = 0;
  for (int c = 0; c < k; ++c) {
    const int64_t c0 = off[c], c1 = off[c + 1];
    float rsum[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) rsum[i] = 0.f;
    for (int64_t j0 = c0; j0 < c1; j0 += SB, ++tile_counter) {
      if (tile_counter % nsplit != split) continue;
      float acc[8][8];
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
      for (int k0 = 0; k0 < d; k0 += SK) {
        __syncthreads();
        // 128 rows x 16 dims = 512 float4 per operand, 2 per thread
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int e = tid + 256 * i;
          const int rr = e & (SB - 1), c4 = e >> 7;
          const int kk = k0 + 4 * c4;
          float4 va = make_float4(0.f, 0.f, 0.f, 0.f), vb = va;
          const int64_t ra = r0 + rr, rb = j0 + rr;
          if (kk < d) {
            if (ra < row_end) va = *reinterpret_cast<const float4*>(Xs + ra * d + kk);
            if (rb < c1) vb = *reinterpret_cast<const float4*>(Xs + rb * d + kk);
          }
          As[4 * c4 + 0][rr] = va.x; As[4 * c4 + 1][rr] = va.y; As[4 * c4 + 2][rr] = va.z; As[4 * c4 + 3][rr] = va.w;
          Bs[4 * c4 + 0][rr] = vb.x; Bs[4 * c4 + 1][rr] = vb.y; Bs[4 * c4 + 2][rr] = vb.z; Bs[4 * c4 + 3][rr] = vb.w;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < SK; ++kk) {
          float a[8], b[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) a[i] = As[kk][ty + 16 * i];
#pragma unroll
          for (int j = 0; j < 8; ++j) b[j] = Bs[kk][tx + 16 * j];
#pragma unroll
          for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
      }
      // epilogue: distances and per-row partial sums
      float cn[8];
      bool cvalid[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int64_t col = j0 + tx + 16 * j;
        cvalid[j] = col < c1;
        cn[j] = cvalid[j] ? sqn[col] : 0.f;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int64_t r = r0 + ty + 16 * i;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int64_t col = j0 + tx + 16 * j;
          const float sn = rn[i] + cn[j];
          float d2 = fmaf(-2.f, acc[i][j], sn);
          const bool close = d2 < kCloseRatioF * sn;
          d2 = fmaxf(d2, 0.f);
          float dist = d2 > 0.f ? d2 * rsqrtf(d2) : 0.f;
          if (!cvalid[j] || col == r || r >= row_end) dist = 0.f;
          else if (close) dist = exact_dist_f32(Xs, d, r, col);
          rsum[i] += dist;
        }
      }
    }
    // fold the 16 column-threads of each row (lanes differing in the low 4 bits), fixed order
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float v = rsum[i];
      v += __shfl_xor_sync(0xffffffffu, v, 8);
      v += __shfl_xor_sync(0xffffffffu, v, 4);
      v += __shfl_xor_sync(0xffffffffu, v, 2);
      v += __shfl_xor_sync(0xffffffffu, v, 1);
      const int64_t r = r0 + ty + 16 * i;
      if (tx == 0 && r < row_end) o[(r - row_begin) * k + c] = v;
    }
  }
}

constexpr int kSilFinBlocks = 148 * 4;

__global__ void silhouette_finalize_kernel(const float* __restrict__ sums, int nsplit,
                                           const int32_t* __restrict__ labels_sorted,
                                           const int64_t* __restrict__ counts, int k, int64_t nrows,
                                           float* __restrict__ s_rows, double* __restrict__ part) {
  __shared__ double scratch[32];
  double acc = 0.0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nrows; i += stride) {
    const int own = labels_sorted[i];
    double a = 0.0, b = INFINITY;
    for (int c = 0; c < k; ++c) {
      double s = 0.0;
      for (int sp = 0; sp < nsplit; ++sp) s += (double)sums[((int64_t)sp * nrows + i) * k + c];
      if (c == own) a = s;
      else if (counts[c] > 0) b = fmin(b, s / (double)counts[c]);
    }
    const int64_t n_own = counts[own];
    double sv = 0.0;
    if (n_own > 1) {
      a /= (double)(n_own - 1);
      const double mx = fmax(a, b);
      if (mx > 0.0 && isfinite(b)) sv = (b - a) / mx;
    }
    if (s_rows) s_rows[i] = (float)sv;
    acc += sv;
  }
  acc = block_sum<double>(acc, scratch);
  if (threadIdx.x == 0) part[blockIdx.x] = acc;
}

__global__ void sum_doubles_kernel(const double* __restrict__ part, int n, double* __restrict__ out) {
  __shared__ double scratch[32];
  double acc = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) acc += part[i];
  acc = block_sum<double>(acc, scratch);
  if (threadIdx.x == 0) *out = acc;
}

}  // namespace vc

using namespace vc;

extern "C" int vc_silhouette_gather(const float* X, int64_t N, int d, const int64_t* order, const double* shift,
                                    float* Xs, float* sqn, void* stream) {
  VC_REQUIRE(X && Xs && sqn && N > 0 && d > 0, VC_ERR_BAD_ARG, "vc_silhouette_gather: bad arguments");
  int64_t blocks = (N + 7) / 8;
  if (blocks > (int64_t)sm_count() * 16) blocks = (int64_t)sm_count() * 16;
  silhouette_gather_kernel<<<(int)blocks, 256, 0, as_stream(stream)>>>(X, N, d, order, shift, Xs, sqn);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" int vc_silhouette_rows(const float* Xs, const float* sqn, int64_t N, int d, const int64_t* off_dev, int k,
                                  int64_t row_begin, int64_t row_end, int nsplit, float* cluster_sums, void* stream) {
  VC_REQUIRE(Xs && sqn && off_dev && cluster_sums && N > 0 && d > 0 && k > 0, VC_ERR_BAD_ARG,
             "vc_silhouette_rows: bad arguments");
  VC_REQUIRE(row_begin >= 0 && row_end <= N && row_begin < row_end, VC_ERR_BAD_ARG,
             "vc_silhouette_rows: bad row range [%lld,%lld) of %lld", (long long)row_begin, (long long)row_end,
             (long long)N);
  VC_REQUIRE(d % 4 == 0 && (reinterpret_cast<uintptr_t>(Xs) & 15) == 0, VC_ERR_UNSUPPORTED,
             "vc_silhouette_rows: need d %% 4 == 0 and 16-byte aligned rows (d=%d)", d);
  VC_REQUIRE(nsplit >= 1 && nsplit <= 65535, VC_ERR_BAD_ARG, "vc_silhouette_rows: bad nsplit %d", nsplit);
  const int64_t tiles = (row_end - row_begin + SB - 1) / SB;
  VC_REQUIRE(tiles < 2147483647LL, VC_ERR_UNSUPPORTED, "vc_silhouette_rows: too many row tiles");
  dim3 grid((unsigned)tiles, (unsigned)nsplit);
  silhouette_rows_kernel<<<grid, 256, 0, as_stream(stream)>>>(Xs, sqn, N, d, off_dev, k, row_begin, row_end, nsplit,
                                                              cluster_sums);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" size_t vc_silhouette_finalize_workspace_bytes(void) { return sizeof(double) * kSilFinBlocks; }

extern "C" int vc_silhouette_finalize(const float* cluster_sums, int nsplit, const int32_t* labels_sorted,
                                      const int64_t* counts_dev, int k, int64_t nrows, float* s_rows, double* s_sum,
                                      void* workspace, size_t workspace_bytes, void* stream) {
  VC_REQUIRE(cluster_sums && labels_sorted && counts_dev && s_sum && k > 0 && nrows > 0 && nsplit >= 1,
             VC_ERR_BAD_ARG, "vc_silhouette_finalize: bad arguments");
  VC_REQUIRE(workspace && workspace_bytes >= vc_silhouette_finalize_workspace_bytes(), VC_ERR_WORKSPACE,
             "vc_silhouette_finalize: workspace too small");
  int blocks = (int)((nrows + 255) / 256);
  if (blocks > kSilFinBlocks) blocks = kSilFinBlocks;
  double* part = static_cast<double*>(workspace);
  silhouette_finalize_kernel<<<blocks, 256, 0, as_stream(stream)>>>(cluster_sums, nsplit, labels_sorted, counts_dev, k,
                                                                    nrows, s_rows, part);
  VC_LAUNCH_CHECK();
  sum_doubles_kernel<<<1, 256, 0, as_stream(stream)>>>(part, blocks, s_sum);
  VC_LAUNCH_CHECK();
  return VC_OK;
}
