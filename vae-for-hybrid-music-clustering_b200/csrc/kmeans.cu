// KMeans Lloyd iteration, fused assign + update in ONE pass over X (132 B/point of HBM traffic at d=32).
//
// Persistent CTAs (2 per SM).  Each CTA streams tiles of TP points: the tile is one contiguous slab of X, moved
// global->shared by a single 1-D bulk async copy (cp.async.bulk, the TMA engine) into a 2-deep mbarrier ring, so
// the copy of tile t+2 overlaps the arithmetic of tile t+1.
//   phase A (thread = point): x row in registers, centroids broadcast from shared memory,
//            pd_j = ||c_j||^2 - 2 x.c_j, first-minimum argmin with strict '<' (sklearn:_k_means_lloyd.pyx:196-213);
//   phase B (warp = the 32 points it just labelled, lane = its own point, walking the dimensions diagonally): the
//            row is added to a warp-private shared-memory accumulator [k,d] with plain, bank-conflict-free,
//            race-free read-modify-writes (32 points per instruction).
// Warp accumulators are folded into a per-CTA double accumulator every kFlush tiles, CTAs write double partials,
// and a second kernel reduces them in a fixed order: bit-reproducible run to run, no floating-point atomics.
#include <string.h>
#include "common.cuh"
#include "tc05.cuh"

namespace vc {

constexpr int kKmThreads = 256;
constexpr int kKmWarps = kKmThreads / 32;
constexpr int kFlush = 16;

using tma::mbar_init;
using tma::mbar_expect_tx;
using tma::mbar_wait;
using tma::mbar_arrive;

__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   tma::s32(dst)),
               "l"(src), "r"(bytes), "r"(tma::s32(bar))
               : "memory");
}

struct KmSmem {
  int tp;             // points per tile
  int ncopies;        // accumulator copies (warps * copies per warp)
  int swz;            // 1: tiles arrive through a 2-D tensor map with the 128-byte swizzle (d % 32 == 0)
  size_t off_acc, off_dsum, off_c, off_c2, off_cnt, off_bar, total;
};

static KmSmem km_layout(int k, int d, int update) {
  KmSmem L;
  int tp = 256;
  while (tp > 32 && (size_t)2 * tp * d * 4 > 128 * 1024) tp >>= 1;
  L.tp = tp;
  L.swz = (d % 32 == 0) ? 1 : 0;
  const int groups = d < 32 ? (31 / d + 1) : 1;  // accumulator copies per warp (see the diagonal update)
  L.ncopies = update ? kKmWarps * groups : 0;
  size_t o = (size_t)2 * tp * d * 4;
  L.off_acc = o;  o += (size_t)L.ncopies * k * d * 4;
  o = (o + 7) & ~(size_t)7;
  L.off_dsum = o; o += update ? (size_t)k * d * 8 : 0;
  L.off_c = o;    o += (size_t)k * d * 4;
  L.off_c2 = o;   o += (size_t)k * 4;
  L.off_cnt = o;  o += (size_t)k * 4;
  o = (o + 15) & ~(size_t)15;
  L.off_bar = o;  o += 16 + 256 * 4;   // two mbarriers; labels of the current tile (lane = dimension update)
  L.total = o + 1024;  // slack: the tile ring is aligned to 1024 B by hand (swizzle atoms)
  return L;
}

struct KmArgs {
  const float* X; int64_t N; int d; const float* C; int k; int32_t* labels; double* part_sums;
  long long* part_counts; unsigned long long* n_changed; int update, assign, tp, ncopies;
  size_t off_acc, off_dsum, off_c, off_c2, off_cnt, off_bar;
  const int* skip;   // device flag: non-zero -> the launch is a no-op (Lloyd loop already converged, vc_kmeans_lloyd_*)
};

// 16-byte chunk j of row r of a tile.  SW: panel j/8 of [tp][128 B] rows, chunk index XORed with r%8 (what TMA wrote);
// otherwise rows are contiguous.
template <bool SW>
__device__ __forceinline__ float4 row_chunk(const unsigned char* tileb, int tp, int d, int r, int j) {
  if (SW) return *reinterpret_cast<const float4*>(tileb + (size_t)(j >> 3) * tp * 128 + tma::sw128(r, j & 7));
  return *reinterpret_cast<const float4*>(tileb + ((size_t)r * d + 4 * j) * 4);
}

// D4 = d/4 when compile-time (register-resident row), 0 = generic d.
// Centroid-sum update of one warp's 32 rows with lane = dimension (tiles in the TMA 128-byte-swizzle layout).  A row's
// label is warp-uniform (read back from shared memory), so the accumulator row it updates is too, and the only
// hazards are equal labels inside a quad of rows -- resolved in registers in row order, which keeps the summation
// order sequential while the four loads / stores of a quad stay independent (no 32-step dependent chain, no
// __syncwarp).  accw = this warp's accumulator copy + lane, rows = the warp's first row of the panel + (lane%4)*4 bytes.
__device__ __forceinline__ void update_rows_lane_dim(float* accw, const unsigned char* rows, const int* lblw, int lane,
                                                     int d) {
  const int lc = lane >> 2;
#pragma unroll 2
  for (int r4 = 0; r4 < 8; ++r4) {
    const int4 L = reinterpret_cast<const int4*>(lblw)[r4];
    const int r = r4 * 4;
    const float x0 = *reinterpret_cast<const float*>(rows + (r + 0) * 128 + ((lc ^ ((r + 0) & 7)) << 4));
    const float x1 = *reinterpret_cast<const float*>(rows + (r + 1) * 128 + ((lc ^ ((r + 1) & 7)) << 4));
    const float x2 = *reinterpret_cast<const float*>(rows + (r + 2) * 128 + ((lc ^ ((r + 2) & 7)) << 4));
    const float x3 = *reinterpret_cast<const float*>(rows + (r + 3) * 128 + ((lc ^ ((r + 3) & 7)) << 4));
    float* q0 = accw + L.x * d;
    float* q1 = accw + L.y * d;
    float* q2 = accw + L.z * d;
    float* q3 = accw + L.w * d;
    float a0 = *q0, a1 = *q1, a2 = *q2, a3 = *q3;
    const float n0 = a0 + x0;
    a1 = (L.y == L.x) ? n0 : a1;
    const float n1 = a1 + x1;
    a2 = (L.z == L.y) ? n1 : ((L.z == L.x) ? n0 : a2);
    const float n2 = a2 + x2;
    a3 = (L.w == L.z) ? n2 : ((L.w == L.y) ? n1 : ((L.w == L.x) ? n0 : a3));
    const float n3 = a3 + x3;
    *q0 = n0; *q1 = n1; *q2 = n2; *q3 = n3;
  }
}


// LD: lane = dimension centroid-sum update (update_rows_lane_dim) instead of the diagonal one.
template <int D4, bool SW, bool LD>
__global__ void __launch_bounds__(kKmThreads)
kmeans_assign_update_kernel(const KmArgs a, const __grid_constant__ CUtensorMap tmap) {
  if (a.skip && *a.skip) return;
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (tma::s32(smem_raw) & 1023u)) & 1023u);
  const int d = a.d, k = a.k, tp = a.tp, update = a.update;
  const int64_t N = a.N;
  const size_t tile_bytes = (size_t)tp * d * 4;
  float* acc = reinterpret_cast<float*>(smem + a.off_acc);
  double* dsum = reinterpret_cast<double*>(smem + a.off_dsum);
  float* cs = reinterpret_cast<float*>(smem + a.off_c);
  float* c2 = reinterpret_cast<float*>(smem + a.off_c2);
  int* cnt = reinterpret_cast<int*>(smem + a.off_cnt);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + a.off_bar);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int kd = k * d;
  const int64_t ntiles = (N + tp - 1) / tp;

  auto issue = [&](int b, int64_t t) {  // one thread: arm the barrier and start the copy of tile t into slot b
    const int64_t p0 = t * tp;
    unsigned char* dst = smem + (size_t)b * tile_bytes;
    if (SW) {
      mbar_expect_tx(&bar[b], (uint32_t)tile_bytes);  // out-of-range rows are zero-filled and still counted
      for (int p = 0; p < d / 32; ++p) tma::load_2d(dst + (size_t)p * tp * 128, &tmap, &bar[b], p * 32, (int)p0);
    } else {
      const uint32_t bytes = (uint32_t)((N - p0 < tp ? N - p0 : tp) * d * 4);
      mbar_expect_tx(&bar[b], bytes);
      bulk_g2s(dst, a.X + p0 * d, bytes, &bar[b]);
    }
  };

  if (tid == 0) {
    if (SW) tma::prefetch_map(&tmap);
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    tma::mbar_fence_init();
  }
  for (int i = tid; i < kd; i += kKmThreads) cs[i] = a.C[i];
  for (int i = tid; i < k; i += kKmThreads) cnt[i] = 0;
  if (update) {
    for (int i = tid; i < a.ncopies * kd; i += kKmThreads) acc[i] = 0.f;
    for (int i = tid; i < kd; i += kKmThreads) dsum[i] = 0.0;
  }
  __syncthreads();
  for (int c = warp; c < k; c += kKmWarps) {
    float s = 0.f;
    for (int j = lane; j < d; j += 32) s += cs[c * d + j] * cs[c * d + j];
    s = warp_sum(s);
    if (lane == 0) c2[c] = s;
  }
  if (warp == 0) {     // warp-uniform branch + elect.sync: no waterfall loop around the copy instructions
    for (int b = 0; b < 2; ++b) {
      const int64_t t = (int64_t)blockIdx.x + (int64_t)b * gridDim.x;
      if (t < ntiles && tc05::elect_one()) issue(b, t);
      __syncwarp();
    }
  }
  __syncthreads();

  unsigned long long my_changed = 0;
  int it = 0;
  for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x, ++it) {
    const int b = it & 1;
    const uint32_t parity = (it >> 1) & 1;
    const int64_t p0 = t * tp;
    const int npts = (int)(N - p0 < tp ? N - p0 : tp);
    // previous label of my point: issued before the wait so its latency hides behind the tile copy / phase A
    const int32_t old = (tid < npts) ? a.labels[p0 + tid] : -1;
    mbar_wait(&bar[b], parity);
    const unsigned char* tileb = smem + (size_t)b * tile_bytes;

    // ---------------- phase A: distances + argmin, thread = point
    int label = -1;
    if (tid < npts && !a.assign) {
      label = old;
      if ((unsigned)label >= (unsigned)k) label = -1;
      else atomicAdd(&cnt[label], 1);
    } else if (tid < npts) {
      float best = 0.f;
      // One sequential FMA chain per centroid over dimensions 0..d-1 -- the accumulation order of a BLAS sgemm
      // micro-kernel (each output element is a running rank-1 update), which is what sklearn's chunked
      // `-2 X C^T + ||c||^2` evaluates (_k_means_lloyd.pyx:196-203).  Four centroids are interleaved for ILP.
      if (D4 > 0) {
        float4 xr[D4 > 0 ? D4 : 1];
#pragma unroll
        for (int j = 0; j < D4; ++j) xr[j] = row_chunk<SW>(tileb, tp, d, tid, j);
        int c = 0;
        for (; c + 4 <= k; c += 4) {
          const float4* c0 = reinterpret_cast<const float4*>(cs + (c + 0) * d);
          const float4* c1 = reinterpret_cast<const float4*>(cs + (c + 1) * d);
          const float4* c2p = reinterpret_cast<const float4*>(cs + (c + 2) * d);
          const float4* c3 = reinterpret_cast<const float4*>(cs + (c + 3) * d);
          float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
          for (int j = 0; j < D4; ++j) {
            const float4 v0 = c0[j], v1 = c1[j], v2 = c2p[j], v3 = c3[j];
            s0 = fmaf(xr[j].x, v0.x, s0); s1 = fmaf(xr[j].x, v1.x, s1); s2 = fmaf(xr[j].x, v2.x, s2); s3 = fmaf(xr[j].x, v3.x, s3);
            s0 = fmaf(xr[j].y, v0.y, s0); s1 = fmaf(xr[j].y, v1.y, s1); s2 = fmaf(xr[j].y, v2.y, s2); s3 = fmaf(xr[j].y, v3.y, s3);
            s0 = fmaf(xr[j].z, v0.z, s0); s1 = fmaf(xr[j].z, v1.z, s1); s2 = fmaf(xr[j].z, v2.z, s2); s3 = fmaf(xr[j].z, v3.z, s3);
            s0 = fmaf(xr[j].w, v0.w, s0); s1 = fmaf(xr[j].w, v1.w, s1); s2 = fmaf(xr[j].w, v2.w, s2); s3 = fmaf(xr[j].w, v3.w, s3);
          }
          const float p0_ = fmaf(-2.f, s0, c2[c]), p1_ = fmaf(-2.f, s1, c2[c + 1]);
          const float p2_ = fmaf(-2.f, s2, c2[c + 2]), p3_ = fmaf(-2.f, s3, c2[c + 3]);
          if (c == 0 || p0_ < best) { best = p0_; label = c; }
          if (p1_ < best) { best = p1_; label = c + 1; }
          if (p2_ < best) { best = p2_; label = c + 2; }
          if (p3_ < best) { best = p3_; label = c + 3; }
        }
        for (; c < k; ++c) {
          const float4* cr = reinterpret_cast<const float4*>(cs + c * d);
          float s0 = 0.f;
#pragma unroll
          for (int j = 0; j < D4; ++j) {
            const float4 cv = cr[j];
            s0 = fmaf(xr[j].x, cv.x, s0);
            s0 = fmaf(xr[j].y, cv.y, s0);
            s0 = fmaf(xr[j].z, cv.z, s0);
            s0 = fmaf(xr[j].w, cv.w, s0);
          }
          const float pd = fmaf(-2.f, s0, c2[c]);
          if (c == 0 || pd < best) { best = pd; label = c; }
        }
      } else {
        const int d4 = d >> 2;
        for (int c = 0; c < k; ++c) {
          const float4* cr = reinterpret_cast<const float4*>(cs + c * d);
          float s0 = 0.f;
          for (int j = 0; j < d4; ++j) {
            const float4 xv = row_chunk<SW>(tileb, tp, d, tid, j), cv = cr[j];
            s0 = fmaf(xv.x, cv.x, s0);
            s0 = fmaf(xv.y, cv.y, s0);
            s0 = fmaf(xv.z, cv.z, s0);
            s0 = fmaf(xv.w, cv.w, s0);
          }
          const float pd = fmaf(-2.f, s0, c2[c]);
          if (c == 0 || pd < best) { best = pd; label = c; }
        }
      }
      a.labels[p0 + tid] = label;
      my_changed += (old != label);
      if (update) atomicAdd(&cnt[label], 1);
    }

    // ---------------- phase B: centroid-sum update
    if (update) {
      // "diagonal" update: lane = its own point; at step j it handles one dimension of that point chosen so that
      // the 32 lanes of each LDS/STS touch 32 different banks and, even when labels coincide, 32 different
      // accumulator words (a plain read-modify-write is race free; one instruction serves 32 points).
      if (SW && LD) {
        // lane = dimension of a 32-wide panel; rows past N are zero-filled by TMA, so they may add to centroid 0
        int* lbl = reinterpret_cast<int*>(smem + a.off_bar + 16);
        lbl[tid] = label < 0 ? 0 : label;
        __syncwarp();
        for (int p = 0; p < d / 32 && warp * 32 < tp; ++p)
          update_rows_lane_dim(acc + (size_t)warp * kd + p * 32 + lane,
                               tileb + (size_t)p * tp * 128 + (size_t)(warp * 32) * 128 + (lane & 3) * 4,
                               lbl + warp * 32, lane, d);
      } else if (SW) {
        // physical word w = (j + lane) % 32 of the row's 128-byte panel slice holds logical dimension
        // ((w/4) ^ (row%8))*4 + w%4 ; for a fixed j this map is injective over the lanes.
        float* my = acc + (size_t)warp * kd + (label >= 0 ? label * d : 0);
        const int r = warp * 32 + lane;
        for (int p = 0; p < d / 32; ++p) {
          const float* xrow = reinterpret_cast<const float*>(tileb + (size_t)p * tp * 128 + (size_t)r * 128);
          float* myp = my + p * 32;
          int w = lane;
#pragma unroll 4
          for (int j = 0; j < 32; ++j) {
            const int dim = ((((w >> 2) ^ (lane & 7)) << 2) | (w & 3));
            if (label >= 0) myp[dim] += xrow[w];
            w = (w + 1) & 31;
            __syncwarp();
          }
        }
      } else {
        // rows are contiguous: dimension (j + lane) % d; lanes d apart would meet, so they use separate copies
        const int cpy = d < 32 ? lane / d : 0;
        const int ncp = d < 32 ? (31 / d + 1) : 1;
        float* my = acc + (size_t)(warp * ncp + cpy) * kd + (label >= 0 ? label * d : 0);
        const float* xrow = reinterpret_cast<const float*>(tileb) + (size_t)(warp * 32 + lane) * d;
        int dim = lane % d;
        for (int j = 0; j < d; ++j) {
          if (label >= 0) my[dim] += xrow[dim];
          dim = dim + 1 == d ? 0 : dim + 1;
          __syncwarp();
        }
      }
    }
    __syncthreads();  // every warp is done with this slot
    if (update && ((it + 1) % kFlush == 0)) {
      for (int e = tid; e < kd; e += kKmThreads) {
        double s = 0.0;
        for (int cpy = 0; cpy < a.ncopies; ++cpy) {
          s += (double)acc[(size_t)cpy * kd + e];
          acc[(size_t)cpy * kd + e] = 0.f;
        }
        dsum[e] += s;
      }
      __syncthreads();
    }
    if (warp == 0) {
      const int64_t tn = t + 2 * (int64_t)gridDim.x;
      if (tn < ntiles) {
        tma::fence_proxy_async();  // order the generic-proxy reads of the slot before the async-proxy overwrite
        if (tc05::elect_one()) issue(b, tn);
        __syncwarp();
      }
    }
  }

  // ---------------- epilogue: per-CTA partials in double, fixed order
  __syncthreads();
  if (update) {
    for (int e = tid; e < kd; e += kKmThreads) {
      double s = dsum[e];
      for (int cpy = 0; cpy < a.ncopies; ++cpy) s += (double)acc[(size_t)cpy * kd + e];
      a.part_sums[(size_t)blockIdx.x * kd + e] = s;
    }
    for (int c = tid; c < k; c += kKmThreads) a.part_counts[(size_t)blockIdx.x * k + c] = cnt[c];
  }
  my_changed = (unsigned long long)__reduce_add_sync(0xffffffffu, (unsigned)my_changed);
  if (lane == 0 && my_changed) atomicAdd(a.n_changed, my_changed);
}

// ---------------------------------------------------------------------------------------------------------------
// Tensor-core E-step for d == 32 and 5 <= k <= 32 (k padded to KP = 16 or 32 with infinitely distant centroids), where the
// FFMA version is compute-bound (2kd FLOP per point).
// The 256 x 32 fp32 tile the TMA engine delivered is already a K-major, 128-byte-swizzled UMMA operand, so the
// distance GEMM  dot[256, k] = X_tile * C^T  runs on tcgen05 (kind::tf32) with the accumulator in tensor memory.
// TF32 alone (10 mantissa bits) cannot reproduce sklearn's fp32 labels, so every value is split x = hi + lo with
// hi = x with the low 13 mantissa bits cleared: dot ~= hi.chi + hi.clo + lo.chi, error <= 2^-16 |x||c|.
//  * kind::tf32 ignores the low 13 bits of each operand word, so the raw tile IS the hi operand; the workers only
//    write the lo tile (x - hi).  [chi ; clo] is one B operand of N = 2*KP, so hi.chi | hi.clo come out of a single
//    pass over A; lo.chi accumulates onto the first KP columns and the epilogue adds the two column groups.
//  * Labels stay bit-identical to kmeans_assign_update_kernel through a guard: every centroid whose tensor-core
//    distance lies within 2^-13 |x| max|c| (+ 2^-21 max|c|^2) of the best one is re-evaluated with the reference FMA
//    chain and the argmin taken over those exact values in index order; all other centroids are provably farther.
//  * Roles: warps 0-7 = workers (thread = row: split, argmin, centroid-sum update with lane = dimension), warp 8 lane 0
//    = TMA + MMA issuer.  mbarriers full/free_ (tile ring), ready/done (operands / accumulator); the MMAs of tile
//    i+1 run under the argmin and update of tile i.
struct KmTcArgs {
  KmArgs a;
  int kp;                      // k rounded up to a multiple of 16 (UMMA N)
  size_t off_lo, off_chi, off_mbar;
  uint32_t idesc, idesc2;      // UMMA instruction descriptors for N = kp and N = 2*kp
};

__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc_,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc_), "r"(accumulate)
      : "memory");
}

__device__ __forceinline__ float tf32_hi(float v) { return __uint_as_float(__float_as_uint(v) & 0xFFFFE000u); }

__host__ __device__ constexpr int tc_slots(int tp) { return tp == 128 ? 3 : 4; }   // tile ring depth

// TP rows per tile (128 or 256): TP worker threads (thread = row) + one warp whose lane 0 issues the TMA copies and
// the MMAs.  TP = 128 fits two CTAs per SM, whose pipelines drift apart and hide each other's barrier / TMEM waits.
template <int KP, int TP>
__global__ void __launch_bounds__(TP + 32, 256 / TP)
kmeans_tc_kernel(const KmTcArgs ta, const __grid_constant__ CUtensorMap tmap) {
  if (ta.a.skip && *ta.a.skip) return;
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (tma::s32(smem_raw) & 1023u)) & 1023u);
  const KmArgs& a = ta.a;
  constexpr int d = 32, tp = TP, NW = TP / 32, kTcWorkers = TP, kTcThreads = TP + 32, kTcSlots = tc_slots(TP);
  const int k = a.k, update = a.update;
  const int64_t N = a.N;
  constexpr size_t tile_bytes = (size_t)tp * d * 4;
  float* acc = reinterpret_cast<float*>(smem + a.off_acc);
  double* dsum = reinterpret_cast<double*>(smem + a.off_dsum);
  float* cs = reinterpret_cast<float*>(smem + a.off_c);
  float* c2 = reinterpret_cast<float*>(smem + a.off_c2);   // [KP], +inf beyond k
  int* cnt = reinterpret_cast<int*>(smem + a.off_cnt);
  unsigned char* lo_tile = smem + ta.off_lo;               // lo operand of the tile being multiplied
  unsigned char* chi = smem + ta.off_chi;                  // [chi ; clo]: 2*KP rows, one B operand of N = 2*KP
  unsigned char* clo = chi + (size_t)KP * 128;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + ta.off_mbar);  // [3] TMA -> workers
  uint64_t* free_ = full + kTcSlots;                                 // [3] workers -> TMA: ring slot consumed
  uint64_t* ready = free_ + kTcSlots;                                // workers -> MMA: hi/lo operands written
  uint64_t* done = ready + 1;                                        // MMA -> workers: accumulator filled
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
  float* cmax_s = reinterpret_cast<float*>(tmem_slot + 1);
  int* lbl = reinterpret_cast<int*>(smem + ta.off_mbar + 128);   // [256] labels of the current tile

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int kd = k * d;
  const int64_t ntiles = (N + tp - 1) / tp;
  const int64_t G = gridDim.x;

  if (tid == 0) {
    tma::prefetch_map(&tmap);
    for (int s = 0; s < kTcSlots; ++s) { mbar_init(&full[s], 1); mbar_init(&free_[s], NW); }
    mbar_init(ready, NW);
    mbar_init(done, 1);
    tma::mbar_fence_init();
  }
  if (warp == NW) tc05::tmem_alloc(tmem_slot, (TP / 128) * 2 * KP);
  for (int i = tid; i < kd; i += kTcThreads) cs[i] = a.C[i];
  for (int i = tid; i < k; i += kTcThreads) cnt[i] = 0;
  // centroid operands: [KP rows][32 floats], K-major with the 128-byte swizzle, split into hi / lo; padding rows zero
  for (int e = tid; e < KP * d; e += kTcThreads) {
    const int r = e >> 5, c = e & 31;
    const float v = r < k ? a.C[e] : 0.f;
    const float h = tf32_hi(v);
    const uint32_t o = tma::sw128((uint32_t)r, (uint32_t)(c >> 2)) + (uint32_t)(c & 3) * 4u;
    *reinterpret_cast<float*>(chi + o) = h;
    *reinterpret_cast<float*>(clo + o) = v - h;
  }
  if (update) {
    for (int i = tid; i < a.ncopies * kd; i += kTcThreads) acc[i] = 0.f;
    for (int i = tid; i < kd; i += kTcThreads) dsum[i] = 0.0;
  }
  tma::fence_proxy_async();
  tc05::fence_before();
  __syncthreads();
  tc05::fence_after();
  const uint32_t tmem_base = *tmem_slot;
  for (int c = warp; c < KP; c += NW + 1) {
    float s = 0.f;
    if (c < k) {
      for (int j = lane; j < d; j += 32) s += cs[c * d + j] * cs[c * d + j];
      s = warp_sum(s);
    } else {
      s = __int_as_float(0x7f800000);   // padding centroid: infinitely far
    }
    if (lane == 0) c2[c] = s;
  }
  __syncthreads();
  if (tid == 0) {
    float m = 0.f;
    for (int c = 0; c < k; ++c) m = fmaxf(m, c2[c]);
    *cmax_s = sqrtf(m);
  }
  __syncthreads();

  if (warp == NW) {
    // ================= copy + MMA issuer =================
    // The whole warp walks the loop in uniform control flow and one ELECTED lane issues: from an `if (lane == 0)` branch
    // every UTMALDG / UTCHMMA / UTCBAR is wrapped in an ELECT ... BRA.U.ANY waterfall loop (~70 clocks each).
    {
      auto issue = [&](int slot, int64_t t) {
        mbar_expect_tx(&full[slot], (uint32_t)tile_bytes);
        tma::load_2d(smem + (size_t)slot * tile_bytes, &tmap, &full[slot], 0, (int)(t * tp));
      };
      for (int s = 0; s < kTcSlots; ++s) {
        const int64_t t = (int64_t)blockIdx.x + (int64_t)s * G;
        if (t < ntiles && tc05::elect_one()) issue(s, t);
        __syncwarp();
      }
      const uint32_t sL = tma::s32(lo_tile), sCh = tma::s32(chi);
      int i = 0;
      for (int64_t t = blockIdx.x; t < ntiles; t += G, ++i) {
        tma::mbar_wait_backoff(ready, (uint32_t)(i & 1), 64);
        tc05::fence_after();
        // A = the raw fp32 tile: kind::tf32 reads the upper 19 bits of each word, i.e. exactly hi.  One N = 2*KP
        // MMA gives hi.chi | hi.clo side by side (A is fetched once); lo.chi accumulates onto the first KP columns.
        const uint32_t sX = tma::s32(smem + (size_t)(i % kTcSlots) * tile_bytes);
        if (tc05::elect_one()) {
#pragma unroll
          for (int hh = 0; hh < TP / 128; ++hh) {
            const uint32_t tacc = tmem_base + (uint32_t)(hh * 2 * KP);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              const uint32_t oa = (uint32_t)(hh * 128 * 128 + kk * 32), ob = (uint32_t)(kk * 32);
              mma_tf32(tacc, tc05::desc_k_sw128(sX + oa), tc05::desc_k_sw128(sCh + ob), ta.idesc2, kk != 0);
              mma_tf32(tacc, tc05::desc_k_sw128(sL + oa), tc05::desc_k_sw128(sCh + ob), ta.idesc, 1);
            }
          }
          tc05::commit(done);
        }
        __syncwarp();
        // refill the slot tile i-1 occupied (tile i-1+kTcSlots) once every worker warp has finished its phase B
        if (i >= 1) {
          const int64_t tn = t + (kTcSlots - 1) * G;
          if (tn < ntiles) {
            const int j = i - 1;
            tma::mbar_wait_backoff(&free_[j % kTcSlots], (uint32_t)((j / kTcSlots) & 1), 64);
            tma::fence_proxy_async();
            if (tc05::elect_one()) issue(j % kTcSlots, tn);
            __syncwarp();
          }
        }
      }
    }
  } else {
    // ================= workers: thread = row =================
    const float cmax = *cmax_s;
    const float band2 = 4.76837158203125e-7f * cmax * cmax;   // 2^-21 max|c|^2: rounding of the final fmaf
    const int q = warp & 3, hf = warp >> 2;   // row tid = hf*128 + q*32 + lane: TMEM lane quarter q of tile half hf
    // split tile i: write the lo operand tile (x - hi); returns |x|^2 of this thread's row
    auto split = [&](int i) -> float {
      const int slot = i % kTcSlots;
      mbar_wait(&full[slot], (uint32_t)((i / kTcSlots) & 1));
      const unsigned char* tileb = smem + (size_t)slot * tile_bytes;
      float nx = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint32_t o = tma::sw128((uint32_t)tid, (uint32_t)j);
        const float4 x = *reinterpret_cast<const float4*>(tileb + o);
        *reinterpret_cast<float4*>(lo_tile + o) =
            make_float4(x.x - tf32_hi(x.x), x.y - tf32_hi(x.y), x.z - tf32_hi(x.z), x.w - tf32_hi(x.w));
        nx = fmaf(x.x, x.x, nx); nx = fmaf(x.y, x.y, nx); nx = fmaf(x.z, x.z, nx); nx = fmaf(x.w, x.w, nx);
      }
      tma::fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(ready);
      return nx;
    };

    unsigned long long my_changed = 0;
    float nx_next = 0.f;
    if ((int64_t)blockIdx.x < ntiles) nx_next = split(0);
    int it = 0;
    for (int64_t t = blockIdx.x; t < ntiles; t += G, ++it) {
      const int slot = it % kTcSlots;
      const int64_t p0 = t * tp;
      const int npts = (int)(N - p0 < tp ? N - p0 : tp);
      const int32_t old = (tid < npts) ? a.labels[p0 + tid] : -1;
      const float nx = nx_next;
      const unsigned char* tileb = smem + (size_t)slot * tile_bytes;

      mbar_wait(done, (uint32_t)(it & 1));
      tc05::fence_after();
      float v[32];
      {
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(hf * 2 * KP);
        tc05::ld_32x32(taddr, v);
        if (KP == 32) {
          float v2[32];
          tc05::ld_32x32(taddr + 32, v2);
          tc05::ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] += v2[j];
        } else {
          tc05::ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] += v[16 + j];
        }
      }
      tc05::fence_before();
      // the accumulator and the operand tiles are free again: split tile it+1 so its MMAs run under the work below
      if (t + G < ntiles) nx_next = split(it + 1);

      // ---- argmin with the exactness guard
      int label = -1;
      if (tid < npts) {
        float best = __int_as_float(0x7f800000);
#pragma unroll
        for (int j4 = 0; j4 < KP / 4; ++j4) {
          const float4 cc = reinterpret_cast<const float4*>(c2)[j4];
          v[4 * j4 + 0] = fmaf(-2.f, v[4 * j4 + 0], cc.x);
          v[4 * j4 + 1] = fmaf(-2.f, v[4 * j4 + 1], cc.y);
          v[4 * j4 + 2] = fmaf(-2.f, v[4 * j4 + 2], cc.z);
          v[4 * j4 + 3] = fmaf(-2.f, v[4 * j4 + 3], cc.w);
          best = fminf(fminf(best, v[4 * j4 + 0]), fminf(v[4 * j4 + 1], fminf(v[4 * j4 + 2], v[4 * j4 + 3])));
        }
        // 2^-13 |x| max|c| covers twice the tf32x3 + fp32-chain dot error
        const float thr = best + fmaf(1.220703125e-4f * cmax, sqrtf(nx), band2);
        uint32_t cand = 0u;
#pragma unroll
        for (int j = 0; j < KP; ++j) cand |= (v[j] <= thr) ? (1u << j) : 0u;
        cand &= k >= 32 ? 0xffffffffu : (1u << k) - 1u;   // padding centroids can only pass if thr is inf (non-finite input)
        label = cand ? __ffs((int)cand) - 1 : 0;   // cand == 0 only for NaN input (FFMA kernel: label 0)
        if (cand & (cand - 1u)) {   // several centroids inside the error band: decide with the reference FMA chain
          float4 xr[8];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            xr[j] = *reinterpret_cast<const float4*>(tileb + tma::sw128((uint32_t)tid, (uint32_t)j));
          label = -1;
          float bexact = 0.f;
          while (cand) {
            const int c = __ffs((int)cand) - 1;
            cand &= cand - 1u;
            const float4* cr = reinterpret_cast<const float4*>(cs + c * d);
            float s0 = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 cv = cr[j];
              s0 = fmaf(xr[j].x, cv.x, s0);
              s0 = fmaf(xr[j].y, cv.y, s0);
              s0 = fmaf(xr[j].z, cv.z, s0);
              s0 = fmaf(xr[j].w, cv.w, s0);
            }
            const float pd = fmaf(-2.f, s0, c2[c]);
            if (label < 0 || pd < bexact) { bexact = pd; label = c; }
          }
        }
        a.labels[p0 + tid] = label;
        my_changed += (old != label);
        if (update) atomicAdd(&cnt[label], 1);
      }
      __syncwarp();

      // ---- phase B: centroid-sum update, lane = dimension (update_rows_lane_dim)
      if (update) {
        lbl[tid] = label < 0 ? 0 : label;   // rows past N are zero-filled by TMA: adding them to centroid 0 is a no-op
        __syncwarp();
        update_rows_lane_dim(acc + (size_t)warp * kd + lane, tileb + (size_t)(warp * 32) * 128 + (lane & 3) * 4,
                             lbl + warp * 32, lane, d);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&free_[slot]);
      if (update && ((it + 1) % kFlush == 0)) {
        asm volatile("bar.sync 1, %0;" ::"n"(TP) : "memory");
        for (int e = tid; e < kd; e += kTcWorkers) {
          double s = 0.0;
          for (int cpy = 0; cpy < a.ncopies; ++cpy) {
            s += (double)acc[(size_t)cpy * kd + e];
            acc[(size_t)cpy * kd + e] = 0.f;
          }
          dsum[e] += s;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(TP) : "memory");
      }
    }
    my_changed = (unsigned long long)__reduce_add_sync(0xffffffffu, (unsigned)my_changed);
    if (lane == 0 && my_changed) atomicAdd(a.n_changed, my_changed);
  }

  // ---------------- epilogue: per-CTA partials in double, fixed order
  tc05::fence_before();
  __syncthreads();
  if (update) {
    for (int e = tid; e < kd; e += kTcThreads) {
      double s = dsum[e];
      for (int cpy = 0; cpy < a.ncopies; ++cpy) s += (double)acc[(size_t)cpy * kd + e];
      a.part_sums[(size_t)blockIdx.x * kd + e] = s;
    }
    for (int c = tid; c < k; c += kTcThreads) a.part_counts[(size_t)blockIdx.x * k + c] = cnt[c];
  }
  if (warp == NW) {
    tc05::fence_after();
    tc05::tmem_dealloc(tmem_base, (TP / 128) * 2 * KP);
  }
}

__global__ void kmeans_reduce_partials_kernel(const double* __restrict__ part_sums,
                                              const long long* __restrict__ part_counts, int nblocks, int k, int d,
                                              double* __restrict__ sums, double* __restrict__ counts) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const int kd = k * d;
  if (e < kd) {
    double s = 0.0;
    for (int b = 0; b < nblocks; ++b) s += part_sums[(size_t)b * kd + e];
    sums[e] = s;
  }
  if (e < k) {
    long long c = 0;
    for (int b = 0; b < nblocks; ++b) c += part_counts[(size_t)b * k + e];
    counts[e] = (double)c;
  }
}


// Tail of one Lloyd iteration of the sync-free loop: fold the per-CTA partials into stats = [sums k*d | counts k |
// n_changed 1] (ONE buffer, so a sharded fit needs one all-reduce) and re-arm the label-change counter.
__global__ void kmeans_lloyd_tail_kernel(const double* __restrict__ part_sums, const long long* __restrict__ part_counts,
                                         int nblocks, int k, int d, double* __restrict__ stats,
                                         unsigned long long* __restrict__ nchg, const int* __restrict__ skip) {
  if (skip && *skip) return;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const int kd = k * d;
  if (e < kd) {
    double s = 0.0;
    for (int b = 0; b < nblocks; ++b) s += part_sums[(size_t)b * kd + e];
    stats[e] = s;
  }
  if (e < k) {
    long long c = 0;
    for (int b = 0; b < nblocks; ++b) c += part_counts[(size_t)b * k + e];
    stats[kd + e] = (double)c;
  }
  if (e == 0) {
    stats[kd + k] = (double)*nchg;
    *nchg = 0ull;
  }
}

// M-step + convergence test of the sync-free loop (sklearn:cluster/_kmeans.py:705-738) on the (all-reduced) stats:
// ctl = {state, n_iter}.  state 0: running, 1: strict convergence (no label changed), 2: centre shift <= tol,
// 3: an empty cluster needs the host's relocation path (centres left untouched, iteration not counted).
__global__ void kmeans_lloyd_finalize_kernel(double* __restrict__ stats, float* __restrict__ C, float* __restrict__ C_new,
                                             int k, int d, double tol, int* __restrict__ ctl) {
  __shared__ double scratch[32];
  __shared__ int big, empty;
  if (ctl[0] != 0) return;
  const double* sums = stats;
  const double* counts = stats + (size_t)k * d;
  if (threadIdx.x == 0) {
    int b = 0, em = 0;
    for (int c = 0; c < k; ++c) {
      if (counts[c] > counts[b]) b = c;
      if (counts[c] == 0.0) em = 1;
    }
    big = b;
    empty = em;
  }
  __syncthreads();
  if (empty) {
    if (threadIdx.x == 0) ctl[0] = 3;
    return;
  }
  double acc = 0.0;
  for (int e = threadIdx.x; e < k * d; e += blockDim.x) {
    const int c = e / d;
    const float nw = (float)(sums[e] / counts[c]);
    const double df = (double)nw - (double)C[e];
    acc += df * df;
    C_new[e] = nw;
  }
  acc = block_sum<double>(acc, scratch);
  __syncthreads();
  for (int e = threadIdx.x; e < k * d; e += blockDim.x) C[e] = C_new[e];   // centers, centers_new = centers_new, centers
  if (threadIdx.x == 0) {
    stats[(size_t)k * d + k + 1] = acc;
    ctl[1] += 1;
    if (tol >= 0.0) {   // tol < 0: run exactly the iterations the host enqueues (benchmarks of the steady loop)
      if (stats[(size_t)k * d + k] == 0.0) ctl[0] = 1;
      else if (acc <= tol) ctl[0] = 2;
    }
  }
}

__global__ void kmeans_finalize_kernel(const double* __restrict__ sums, const double* __restrict__ counts,
                                       const float* __restrict__ C_old, int k, int d, float* __restrict__ C_new,
                                       double* __restrict__ shift_sq) {
  __shared__ double scratch[32];
  __shared__ int big;
  if (threadIdx.x == 0) {  // first maximum, like np.argmax (_k_means_common.pyx:283)
    int b = 0;
    for (int c = 1; c < k; ++c)
      if (counts[c] > counts[b]) b = c;
    big = b;
  }
  __syncthreads();
  double acc = 0.0;
  for (int e = threadIdx.x; e < k * d; e += blockDim.x) {
    const int c = e / d, j = e % d;
    const double n = counts[c];
    // empty clusters are placed on the biggest cluster's centre (_k_means_common.pyx:286-295)
    const float nw = n > 0.0 ? (float)(sums[e] / n)
                             : (counts[big] > 0.0 ? (float)(sums[big * d + j] / counts[big]) : C_old[e]);
    C_new[e] = nw;
    const double df = (double)nw - (double)C_old[e];
    acc += df * df;
  }
  acc = block_sum<double>(acc, scratch);
  if (threadIdx.x == 0 && shift_sq) *shift_sq = acc;
}

// Per-point distance to the centre of its label.  MODE 0: squared distance in fp32 (direct differences), summed
// into acc[0]; MODE 1: Euclidean distance against double centroids, summed per cluster (Davies-Bouldin intra).
template <int MODE>
__global__ void __launch_bounds__(256)
point_dist_kernel(const float* __restrict__ X, int64_t N, int d, const float* __restrict__ Cf,
                  const double* __restrict__ Cd, int k, const int32_t* __restrict__ labels,
                  float* __restrict__ point_out, double* __restrict__ part, int nacc) {
  extern __shared__ double wacc[];  // [warps][nacc]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  for (int i = threadIdx.x; i < nwarps * nacc; i += blockDim.x) wacc[i] = 0.0;
  __syncthreads();
  double* mine = wacc + warp * nacc;
  const int64_t gw = (int64_t)blockIdx.x * nwarps + warp, nw = (int64_t)gridDim.x * nwarps;
  for (int64_t i = gw; i < N; i += nw) {
    const int lab = labels[i];
    if ((unsigned)lab >= (unsigned)k) continue;
    double v;
    if (MODE == 0) {
      float s = 0.f;
      for (int j = lane; j < d; j += 32) {
        const float df = X[i * d + j] - Cf[lab * d + j];
        s = fmaf(df, df, s);
      }
      s = warp_sum(s);
      if (point_out && lane == 0) point_out[i] = s;
      v = (double)s;
    } else {
      double s = 0.0;
      for (int j = lane; j < d; j += 32) {
        const double df = (double)X[i * d + j] - Cd[lab * d + j];
        s += df * df;
      }
      s = warp_sum(s);
      v = sqrt(s);
    }
    if (lane == 0) mine[MODE == 0 ? 0 : lab] += v;
  }
  __syncthreads();
  for (int a = threadIdx.x; a < nacc; a += blockDim.x) {
    double s = 0.0;
    for (int w = 0; w < nwarps; ++w) s += wacc[w * nacc + a];
    part[(size_t)blockIdx.x * nacc + a] = s;
  }
}

__global__ void reduce_blocks_kernel(const double* __restrict__ part, int nblocks, int nacc,
                                     double* __restrict__ out) {
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= nacc) return;
  double s = 0.0;
  for (int b = 0; b < nblocks; ++b) s += part[(size_t)b * nacc + a];
  out[a] = s;
}

// out[q, i] = ||X_i - Q_q||^2, evaluated in double from fp32 inputs and rounded to fp32 (what sklearn's
// float32 euclidean_distances upcast path stores, pairwise.py:567-638).
__global__ void sqdist_rows_kernel(const float* __restrict__ X, int64_t N, int d, const float* __restrict__ Q, int m,
                                   float* __restrict__ out) {
  extern __shared__ float qs[];
  for (int i = threadIdx.x; i < m * d; i += blockDim.x) qs[i] = Q[i];
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += stride) {
    for (int q = 0; q < m; ++q) {
      double s = 0.0;
      for (int j = 0; j < d; ++j) {
        const double df = (double)X[i * d + j] - (double)qs[q * d + j];
        s += df * df;
      }
      out[(int64_t)q * N + i] = (float)s;
    }
  }
}

}  // namespace vc

using namespace vc;

static const int kKmMaxSmem = 220 * 1024;

static int km_grid(const KmSmem& L, int64_t N) {
  int per_sm = (int)(kKmMaxSmem / (L.total + 1024));
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 2) per_sm = 2;
  int64_t g = (int64_t)sm_count() * per_sm;
  const int64_t ntiles = (N + L.tp - 1) / L.tp;
  if (g > ntiles) g = ntiles;
  if (g < 1) g = 1;
  return (int)g;
}

extern "C" size_t vc_kmeans_workspace_bytes(int k, int d) {
  // per-CTA partial sums (double) + counts (int64) for up to 2 CTAs per SM, + one 8-byte counter
  const size_t blocks = (size_t)sm_count() * 2;
  size_t pd = blocks * (size_t)(k > 0 ? k : 1) * (size_t)(d > 0 ? d : 1) * 8;
  size_t pc = blocks * (size_t)(k > 0 ? k : 1) * 8;
  return pd + pc + 64;
}

template <int D4, bool SW, bool LD = false>
static cudaError_t km_launch(int grid, const KmSmem& L, cudaStream_t s, const KmArgs& a, const CUtensorMap& tm) {
  auto kern = kmeans_assign_update_kernel<D4, SW, LD>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total);
  if (e != cudaSuccess) return e;
  kern<<<grid, kKmThreads, L.total, s>>>(a, tm);
  return cudaGetLastError();
}

// lloyd_stats != NULL selects the sync-free Lloyd form: the launch honours the device flag `skip`, the label-change
// counter is not cleared here (the tail kernel re-arms it) and the tail kernel folds everything into lloyd_stats.
static int km_run_impl(const float* X, int64_t N, int d, const float* C, int k, int32_t* labels, double* sums,
                       double* counts, int64_t* n_changed, int update, int assign, void* workspace,
                       size_t workspace_bytes, void* stream, const char* who, const int* skip, double* lloyd_stats) {
  VC_REQUIRE(X && labels && N > 0 && d > 0 && k > 0, VC_ERR_BAD_ARG, "%s: bad arguments (N=%lld d=%d k=%d)", who,
             (long long)N, d, k);
  VC_REQUIRE(d % 4 == 0 && d <= 256 && k <= 256, VC_ERR_UNSUPPORTED,
             "%s: need d %% 4 == 0, d <= 256, k <= 256 (got d=%d k=%d)", who, d, k);
  VC_REQUIRE((reinterpret_cast<uintptr_t>(X) & 15) == 0, VC_ERR_BAD_ARG, "%s: X must be 16-byte aligned", who);
  VC_REQUIRE(!update || (sums && counts) || lloyd_stats, VC_ERR_BAD_ARG, "%s: update requested without sums/counts", who);
  VC_REQUIRE(workspace && workspace_bytes >= vc_kmeans_workspace_bytes(k, d), VC_ERR_WORKSPACE,
             "%s: workspace too small", who);
  const KmSmem L = km_layout(k, d, update);
  VC_REQUIRE(L.total <= (size_t)kKmMaxSmem, VC_ERR_UNSUPPORTED,
             "%s: k*d=%d needs %zu bytes of shared memory (> %d)", who, k * d, L.total, kKmMaxSmem);
  const int grid = km_grid(L, N);
  cudaStream_t s = as_stream(stream);
  const size_t blocks = (size_t)sm_count() * 2;
  double* ps = static_cast<double*>(workspace);
  long long* pc = reinterpret_cast<long long*>(ps + blocks * (size_t)k * d);
  unsigned long long* nchg = reinterpret_cast<unsigned long long*>(pc + blocks * (size_t)k);
  if (!lloyd_stats) VC_CUDA(cudaMemsetAsync(nchg, 0, 8, s));
  KmArgs a{X, N, d, C, k, labels, ps, pc, nchg, update, assign, L.tp, L.ncopies,
           L.off_acc, L.off_dsum, L.off_c, L.off_c2, L.off_cnt, L.off_bar, skip};
  CUtensorMap tm;
  memset(&tm, 0, sizeof(tm));
  if (L.swz) {
    const uint64_t dims[2] = {(uint64_t)d, (uint64_t)N};
    const uint64_t strides[1] = {(uint64_t)d * 4};
    const uint32_t box[2] = {32u, (uint32_t)L.tp};
    int rc = tma::make_map(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, X, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc != VC_OK) return rc;
  }
  cudaError_t e;
  const char* tc_env = getenv("VAECLUST_KMEANS_TC");   // "0": keep the FFMA E-step (A/B checks)
  const int tc_min = 11;   // below, the FFMA kernel is faster (50M x 32: k=8 1.94 vs 2.21 ms, k=12 2.38 vs 2.23 ms)
  if (d == 32 && assign && k >= tc_min && k <= 32 && N < 2147483647LL - 256 && !(tc_env && tc_env[0] == '0')) {
    // tensor-core E-step (kmeans_tc_kernel): x ring 128 KB | lo operand tile 32 KB | centroid hi ; lo | accumulators
    KmTcArgs ta;
    memset(&ta, 0, sizeof(ta));
    ta.kp = k <= 16 ? 16 : 32;
    const int TP = 256;   // rows per tile; 128 (two CTAs per SM) measured 2 % slower on 50M x 32
    const size_t tile_b = (size_t)TP * 32 * 4;
    size_t o = (size_t)tc_slots(TP) * tile_b;
    ta.off_lo = o;  o += tile_b;
    ta.off_chi = o; o += (size_t)2 * ta.kp * 128;
    KmArgs a2 = a;
    a2.tp = TP;
    a2.ncopies = update ? TP / 32 : 0;
    a2.off_acc = o;  o += (size_t)a2.ncopies * k * d * 4;
    a2.off_dsum = o; o += update ? (size_t)k * d * 8 : 0;
    a2.off_c = o;    o += (size_t)k * d * 4;
    a2.off_c2 = o;   o += 32 * 4;
    a2.off_cnt = o;  o += 32 * 4;
    a2.off_bar = o;
    ta.off_mbar = o; o += 128 + 256 * 4;   // barriers, tmem slot, max|c| ; labels of the tile
    ta.a = a2;
    ta.idesc = tc05::idesc(2 /* tf32 */, 128, ta.kp, 0, 0);
    ta.idesc2 = tc05::idesc(2, 128, 2 * ta.kp, 0, 0);
    const size_t total = o + 1024;
    CUtensorMap tmt;
    {
      const uint64_t dims[2] = {(uint64_t)d, (uint64_t)N};
      const uint64_t strides[1] = {(uint64_t)d * 4};
      const uint32_t box[2] = {32u, (uint32_t)TP};
      int rc = tma::make_map(&tmt, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, X, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
      if (rc != VC_OK) return rc;
    }
    int64_t g = (int64_t)sm_count() * (256 / TP);
    const int64_t ntl = (N + TP - 1) / TP;
    if (g > ntl) g = ntl;
    auto kern = TP == 128 ? (ta.kp == 16 ? kmeans_tc_kernel<16, 128> : kmeans_tc_kernel<32, 128>)
                          : (ta.kp == 16 ? kmeans_tc_kernel<16, 256> : kmeans_tc_kernel<32, 256>);
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)total);
    if (e == cudaSuccess) {
      kern<<<(int)g, TP + 32, total, s>>>(ta, tmt);
      e = cudaGetLastError();
    }
    VC_REQUIRE(e == cudaSuccess, VC_ERR_CUDA, "%s: tensor-core launch failed: %s", who, cudaGetErrorString(e));
    if (lloyd_stats) {
      kmeans_lloyd_tail_kernel<<<(k * d + 127) / 128, 128, 0, s>>>(ps, pc, (int)g, k, d, lloyd_stats, nchg, skip);
      VC_LAUNCH_CHECK();
      return VC_OK;
    }
    if (update) {
      const int n = k * d;
      kmeans_reduce_partials_kernel<<<(n + 127) / 128, 128, 0, s>>>(ps, pc, (int)g, k, d, sums, counts);
      VC_LAUNCH_CHECK();
    }
    if (n_changed) VC_CUDA(cudaMemcpyAsync(n_changed, nchg, 8, cudaMemcpyDeviceToDevice, s));
    return VC_OK;
  }
  if (L.swz && assign) {
    // lane = dimension update; labels given by the caller (assign == 0) may be out of range, which only the diagonal
    // form's per-lane guard skips
    switch (d) {
      case 32: e = km_launch<8, true, true>(grid, L, s, a, tm); break;
      case 64: e = km_launch<16, true, true>(grid, L, s, a, tm); break;
      default: e = km_launch<0, true, true>(grid, L, s, a, tm); break;
    }
  } else if (L.swz) {
    switch (d) {
      case 32: e = km_launch<8, true>(grid, L, s, a, tm); break;
      case 64: e = km_launch<16, true>(grid, L, s, a, tm); break;
      default: e = km_launch<0, true>(grid, L, s, a, tm); break;
    }
  } else {
    switch (d) {
      case 8:  e = km_launch<2, false>(grid, L, s, a, tm); break;
      case 16: e = km_launch<4, false>(grid, L, s, a, tm); break;
      default: e = km_launch<0, false>(grid, L, s, a, tm); break;
    }
  }
  VC_REQUIRE(e == cudaSuccess, VC_ERR_CUDA, "%s: launch failed: %s", who, cudaGetErrorString(e));
  if (lloyd_stats) {
    kmeans_lloyd_tail_kernel<<<(k * d + 127) / 128, 128, 0, s>>>(ps, pc, grid, k, d, lloyd_stats, nchg, skip);
    VC_LAUNCH_CHECK();
    return VC_OK;
  }
  if (update) {
    const int n = k * d;
    kmeans_reduce_partials_kernel<<<(n + 127) / 128, 128, 0, s>>>(ps, pc, grid, k, d, sums, counts);
    VC_LAUNCH_CHECK();
  }
  if (n_changed) VC_CUDA(cudaMemcpyAsync(n_changed, nchg, 8, cudaMemcpyDeviceToDevice, s));
  return VC_OK;
}

static int km_run(const float* X, int64_t N, int d, const float* C, int k, int32_t* labels, double* sums, double* counts,
                  int64_t* n_changed, int update, int assign, void* workspace, size_t workspace_bytes, void* stream,
                  const char* who) {
  return km_run_impl(X, N, d, C, k, labels, sums, counts, n_changed, update, assign, workspace, workspace_bytes, stream,
                     who, nullptr, nullptr);
}

// ---- sync-free Lloyd loop: the host enqueues iterations without reading anything back; convergence is decided on
// the device and a converged loop turns the remaining launches into no-ops.
extern "C" int vc_kmeans_lloyd_begin(int32_t* ctl, void* workspace, size_t workspace_bytes, int k, int d, void* stream) {
  VC_REQUIRE(ctl && workspace && workspace_bytes >= vc_kmeans_workspace_bytes(k, d), VC_ERR_BAD_ARG,
             "vc_kmeans_lloyd_begin: bad arguments");
  const size_t blocks = (size_t)sm_count() * 2;
  double* ps = static_cast<double*>(workspace);
  long long* pc = reinterpret_cast<long long*>(ps + blocks * (size_t)k * d);
  VC_CUDA(cudaMemsetAsync(pc + blocks * (size_t)k, 0, 8, as_stream(stream)));
  VC_CUDA(cudaMemsetAsync(ctl, 0, 16, as_stream(stream)));
  return VC_OK;
}

extern "C" int vc_kmeans_lloyd_estep(const float* X, int64_t N, int d, const float* C, int k, int32_t* labels,
                                     double* stats, const int32_t* ctl, void* workspace, size_t workspace_bytes,
                                     void* stream) {
  VC_REQUIRE(C && stats && ctl, VC_ERR_BAD_ARG, "vc_kmeans_lloyd_estep: bad arguments");
  return km_run_impl(X, N, d, C, k, labels, nullptr, nullptr, nullptr, 1, 1, workspace, workspace_bytes, stream,
                     "vc_kmeans_lloyd_estep", ctl, stats);
}

extern "C" int vc_kmeans_lloyd_finalize(double* stats, float* C, float* C_scratch, int k, int d, double tol,
                                        int32_t* ctl, void* stream) {
  VC_REQUIRE(stats && C && C_scratch && ctl && k > 0 && d > 0, VC_ERR_BAD_ARG, "vc_kmeans_lloyd_finalize: bad arguments");
  kmeans_lloyd_finalize_kernel<<<1, 256, 0, as_stream(stream)>>>(stats, C, C_scratch, k, d, tol, ctl);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" int vc_kmeans_assign_update(const float* X, int64_t N, int d, const float* C, int k, int32_t* labels,
                                       double* sums, double* counts, int64_t* n_changed, int update, void* workspace,
                                       size_t workspace_bytes, void* stream) {
  VC_REQUIRE(C, VC_ERR_BAD_ARG, "vc_kmeans_assign_update: C is NULL");
  return km_run(X, N, d, C, k, labels, sums, counts, n_changed, update, 1, workspace, workspace_bytes, stream,
                "vc_kmeans_assign_update");
}

extern "C" int vc_cluster_sums(const float* X, int64_t N, int d, const int32_t* labels, int k, double* sums,
                               double* counts, void* workspace, size_t workspace_bytes, void* stream) {
  // centroids are not needed when labels are given; X doubles as a harmless C source for the smem preload
  VC_REQUIRE(N >= k || N > 0, VC_ERR_BAD_ARG, "vc_cluster_sums: bad N");
  VC_REQUIRE(sums && counts, VC_ERR_BAD_ARG, "vc_cluster_sums: sums/counts NULL");
  VC_REQUIRE((int64_t)k <= N, VC_ERR_BAD_ARG, "vc_cluster_sums: k=%d > N=%lld", k, (long long)N);
  return km_run(X, N, d, X, k, const_cast<int32_t*>(labels), sums, counts, nullptr, 1, 0, workspace, workspace_bytes,
                stream, "vc_cluster_sums");
}

extern "C" int vc_kmeans_finalize(const double* sums, const double* counts, const float* C_old, int k, int d,
                                  float* C_new, double* shift_sq, void* stream) {
  VC_REQUIRE(sums && counts && C_old && C_new && k > 0 && d > 0, VC_ERR_BAD_ARG, "vc_kmeans_finalize: bad arguments");
  kmeans_finalize_kernel<<<1, 256, 0, as_stream(stream)>>>(sums, counts, C_old, k, d, C_new, shift_sq);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

static int point_dist_run(int mode, const float* X, int64_t N, int d, const float* Cf, const double* Cd, int k,
                          const int32_t* labels, float* point_out, double* out, void* workspace,
                          size_t workspace_bytes, void* stream) {
  const int nacc = mode == 0 ? 1 : k;
  int grid = sm_count() * 2;
  if ((int64_t)grid * 8 > N) grid = (int)((N + 7) / 8);
  VC_REQUIRE(workspace && workspace_bytes >= (size_t)grid * nacc * 8, VC_ERR_WORKSPACE,
             "point_dist: workspace too small");
  double* part = static_cast<double*>(workspace);
  const size_t sh = (size_t)8 * nacc * 8;
  if (mode == 0)
    point_dist_kernel<0><<<grid, 256, sh, as_stream(stream)>>>(X, N, d, Cf, Cd, k, labels, point_out, part, nacc);
  else
    point_dist_kernel<1><<<grid, 256, sh, as_stream(stream)>>>(X, N, d, Cf, Cd, k, labels, point_out, part, nacc);
  VC_LAUNCH_CHECK();
  reduce_blocks_kernel<<<(nacc + 127) / 128, 128, 0, as_stream(stream)>>>(part, grid, nacc, out);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" int vc_kmeans_inertia(const float* X, int64_t N, int d, const float* C, int k, const int32_t* labels,
                                 double* inertia, float* point_sqdist, void* workspace, size_t workspace_bytes,
                                 void* stream) {
  VC_REQUIRE(X && C && labels && inertia && N > 0 && d > 0 && k > 0, VC_ERR_BAD_ARG,
             "vc_kmeans_inertia: bad arguments");
  return point_dist_run(0, X, N, d, C, nullptr, k, labels, point_sqdist, inertia, workspace, workspace_bytes, stream);
}

extern "C" int vc_dbi_intra(const float* X, int64_t N, int d, const int32_t* labels, const double* centroids, int k,
                            double* intra_sum, void* workspace, size_t workspace_bytes, void* stream) {
  VC_REQUIRE(X && labels && centroids && intra_sum && N > 0 && d > 0 && k > 0 && k <= 1024, VC_ERR_BAD_ARG,
             "vc_dbi_intra: bad arguments");
  return point_dist_run(1, X, N, d, nullptr, centroids, k, labels, nullptr, intra_sum, workspace, workspace_bytes,
                        stream);
}

extern "C" int vc_sqdist_rows(const float* X, int64_t N, int d, const float* Q, int m, float* out, void* stream) {
  VC_REQUIRE(X && Q && out && N > 0 && d > 0 && m > 0, VC_ERR_BAD_ARG, "vc_sqdist_rows: bad arguments");
  VC_REQUIRE((size_t)m * d * 4 <= 48 * 1024, VC_ERR_UNSUPPORTED, "vc_sqdist_rows: m*d too large (%d x %d)", m, d);
  int64_t blocks = (N + 255) / 256;
  if (blocks > (int64_t)sm_count() * 8) blocks = (int64_t)sm_count() * 8;
  sqdist_rows_kernel<<<(int)blocks, 256, (size_t)m * d * 4, as_stream(stream)>>>(X, N, d, Q, m, out);
  VC_LAUNCH_CHECK();
  return VC_OK;
}
