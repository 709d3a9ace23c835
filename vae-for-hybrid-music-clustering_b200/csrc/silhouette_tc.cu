// Silhouette on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), sm_100a.
//
// The N^2 d contraction x_i . x_j runs as bf16 MMAs with fp32 accumulation in tensor memory.  Plain bf16 inputs
// are not accurate enough for the 1e-5 target, so each fp32 coordinate is split x = hi + lo (two bf16) and
//   x.y = hi.hi + hi.lo + lo.hi + lo.lo     (hi + lo carries 16 mantissa bits; every product is exact in fp32)
// is accumulated by four MMAs per 16-wide k step into the same TMEM tile.  Norms stay fp32.
//
// One CTA owns a 128-row block.  Its hi/lo row operands are loaded once; 128-column tiles of the (cluster-sorted)
// matrix stream through a TMA ring.  Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane),
// warps 2..9 = epilogue.  Each epilogue thread owns one accumulator lane = one row i and half of the tile's
// columns: it pulls its 64 dot products out of TMEM (tcgen05.ld), forms d = sqrt(max(0, |x_i|^2 + |x_j|^2 - 2 x_i.x_j)) and keeps the running
// sum for the current cluster in a register -- rows are sorted by cluster, so no bucketing is needed
// (sklearn:metrics/cluster/_unsupervised.py:184-189 does np.bincount per row instead).  Two TMEM accumulator stages
// let the MMAs of tile t+1 overlap the epilogue of tile t.
#include <stdlib.h>
#include "common.cuh"
#include "tc05.cuh"

namespace vc {

constexpr int ST_ROWS = 128;      // rows per CTA (UMMA M)
// columns per tile (UMMA N) are a template parameter: 256 when the operands are 64 wide (one barrier round trip,
// TMEM load and loop overhead per 128 columns a thread instead of per 64, and the tensor pipe's widest N), 128 when
// they are 128 wide (shared memory)
// epilogue warps EW (template): EW/4 threads per accumulator lane, each owning ST_COLS/(EW/4) columns of a tile

// fp32 -> (hi, lo) bf16 split, zero-padded to dp columns (dp % 64 == 0); one warp per row.  Also emits the squared
// norm of the SPLIT row (hi + lo, 16 mantissa bits), so |x|^2 + |y|^2 - 2 x.y is formed from one consistent
// representation and the cancellation error for close pairs is fp32 rounding only.
__global__ void split_bf16_kernel(const float* __restrict__ X, int64_t N, int d, int dp, __nv_bfloat16* __restrict__ H,
                                  __nv_bfloat16* __restrict__ L, float* __restrict__ sqn) {
  const int lane = threadIdx.x & 31;
  const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = w; r < N; r += nw) {
    float s = 0.f;
    for (int c = lane; c < dp; c += 32) {
      const float v = c < d ? X[r * d + c] : 0.f;
      const __nv_bfloat16 h = __float2bfloat16_rn(v);
      const __nv_bfloat16 l = __float2bfloat16_rn(v - __bfloat162float(h));
      H[r * dp + c] = h;
      L[r * dp + c] = l;
      const float vs = __bfloat162float(h) + __bfloat162float(l);
      s = fmaf(vs, vs, s);
    }
    s = warp_sum(s);
    if (sqn && lane == 0) sqn[r] = s;
  }
}

struct SilTcArgs {
  const float* sqn; const int64_t* off; int k; int64_t N, row_begin, row_end; int nsplit; float* out;
  const float* Xs; int d;   // fp32 rows for the exact recomputation of close pairs
  int panels;   // dp / 64
  int stages;   // depth of the column-tile ring
  int mma4;     // 1: also issue the lo.lo product (exact fp32 products); 0: three MMAs per product (default)
};

// Pairs whose squared distance is below kCloseRatio * (|x|^2 + |y|^2) have lost most of their bits to cancellation in
// the norm formula (duplicates and near-duplicates): they are recomputed as sum (x_k - y_k)^2 from the fp32 rows,
// which is what sklearn's float64 upcast achieves (sklearn:metrics/pairwise.py:567-638).
constexpr float kCloseRatio = 1.f / 256.f;

__device__ __noinline__ float exact_dist(const float* __restrict__ Xs, int d, int64_t i, int64_t j) {
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

__device__ __forceinline__ float fast_sqrt(float x) {
  float y;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int ST_COLS, int EW>
__global__ void __launch_bounds__(64 + 32 * EW, 1)
silhouette_tc_kernel(const SilTcArgs a, const __grid_constant__ CUtensorMap mapH, const __grid_constant__ CUtensorMap mapL) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (tma::s32(smem_raw) & 1023u)) & 1023u);
  const int P = a.panels, S = a.stages;
  constexpr int ST_EPI = 32 * EW, PARTS = EW / 4, PCOLS = ST_COLS / PARTS;
  // columns per TMEM read: 18 warps put five on one scheduler, whose 16384 registers then allow 96 per thread
  constexpr int CW = EW == 16 ? 32 : 64;
  const uint32_t panel_bytes = ST_ROWS * 128;             // [128 rows][64 bf16]
  const uint32_t op_bytes = panel_bytes * P;              // one row operand (hi or lo)
  const uint32_t bpanel_bytes = ST_COLS * 128;            // [ST_COLS rows][64 bf16] of a column tile
  const uint32_t bop_bytes = bpanel_bytes * P;            // one column operand (hi or lo) of one tile
  unsigned char* sA = smem;                               // A_hi | A_lo
  unsigned char* sB = smem + 2 * op_bytes;                // S x (B_hi | B_lo)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + (size_t)S * 2 * bop_bytes);
  uint64_t* a_full = bars;                // 1
  uint64_t* b_full = bars + 1;            // S
  uint64_t* b_empty = b_full + S;         // S
  uint64_t* t_full = b_empty + S;         // 2
  uint64_t* t_empty = t_full + 2;         // 2
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t r0 = a.row_begin + (int64_t)blockIdx.x * ST_ROWS;
  const int split = blockIdx.y;

  if (threadIdx.x == 0) {
    tma::prefetch_map(&mapH);
    tma::prefetch_map(&mapL);
    tma::mbar_init(a_full, 1);
    for (int s = 0; s < S; ++s) { tma::mbar_init(&b_full[s], 1); tma::mbar_init(&b_empty[s], 1); }
    for (int s = 0; s < 2; ++s) { tma::mbar_init(&t_full[s], 1); tma::mbar_init(&t_empty[s], ST_EPI); }
    tma::mbar_fence_init();
  }
  if (warp == 1) tc05::tmem_alloc(tmem_slot, 2 * ST_COLS);
  tc05::fence_before();
  __syncthreads();
  tc05::fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    // whole warp in uniform control flow, one elected lane issues: from an `if (lane == 0)` branch every UTMALDG / UTCHMMA is
    // wrapped in an ELECT ... BRA.U.ANY waterfall loop (~70 clocks each, profiles/README.md round 2)
    {
      if (tc05::elect_one()) {
        tma::mbar_expect_tx(a_full, 2 * op_bytes);
        for (int p = 0; p < P; ++p) {
          tma::load_2d(sA + p * panel_bytes, &mapH, a_full, p * 64, (int)r0);
          tma::load_2d(sA + op_bytes + p * panel_bytes, &mapL, a_full, p * 64, (int)r0);
        }
      }
      __syncwarp();
      int64_t counter = 0;
      uint32_t it = 0;
      for (int c = 0; c < a.k; ++c) {
        const int64_t c0 = a.off[c], c1 = a.off[c + 1];
        for (int64_t j0 = c0 & ~(int64_t)3; j0 < c1; j0 += ST_COLS, ++counter) {   // 16-byte aligned norms
          if (counter % a.nsplit != split) continue;
          const int s = it % S;
          const uint32_t ph = (it / S) & 1;
          tma::mbar_wait(&b_empty[s], ph ^ 1);
          unsigned char* dst = sB + (size_t)s * 2 * bop_bytes;
          if (tc05::elect_one()) {
            tma::mbar_expect_tx(&b_full[s], 2 * bop_bytes);
            for (int p = 0; p < P; ++p)
              for (int h = 0; h < ST_COLS / ST_ROWS; ++h) {   // the tensor-map box is 128 rows
                const uint32_t o = p * bpanel_bytes + h * panel_bytes;
                tma::load_2d(dst + o, &mapH, &b_full[s], p * 64, (int)j0 + h * ST_ROWS);
                tma::load_2d(dst + bop_bytes + o, &mapL, &b_full[s], p * 64, (int)j0 + h * ST_ROWS);
              }
          }
          __syncwarp();
          ++it;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    {
      constexpr uint32_t kIdesc = tc05::idesc(tc05::kFmtBF16, ST_ROWS, ST_COLS, 0, 0);
      tma::mbar_wait(a_full, 0);
      tc05::fence_after();
      const uint32_t aH = tma::s32(sA), aL = aH + op_bytes;
      int64_t counter = 0;
      uint32_t it = 0;
      for (int c = 0; c < a.k; ++c) {
        const int64_t c0 = a.off[c], c1 = a.off[c + 1];
        for (int64_t j0 = c0 & ~(int64_t)3; j0 < c1; j0 += ST_COLS, ++counter) {   // 16-byte aligned norms
          if (counter % a.nsplit != split) continue;
          const int s = it % S;
          const uint32_t ph = (it / S) & 1;
          const int as = it & 1;
          const uint32_t aph = (it >> 1) & 1;
          tma::mbar_wait(&t_empty[as], aph ^ 1);   // epilogue has drained this accumulator stage
          tma::mbar_wait(&b_full[s], ph);          // column operands have landed
          tc05::fence_after();
          const uint32_t bH = tma::s32(sB + (size_t)s * 2 * bop_bytes), bL = bH + bop_bytes;
          const uint32_t tacc = tmem_base + as * ST_COLS;
          if (tc05::elect_one()) {
            uint32_t accum = 0;
            for (int p = 0; p < P; ++p) {
#pragma unroll
              for (int kk = 0; kk < 4; ++kk) {  // 64 bf16 per panel / UMMA_K 16
                const uint32_t o = p * panel_bytes + kk * 32, ob = p * bpanel_bytes + kk * 32;
                const uint64_t dAH = tc05::desc_k_sw128(aH + o), dAL = tc05::desc_k_sw128(aL + o);
                const uint64_t dBH = tc05::desc_k_sw128(bH + ob), dBL = tc05::desc_k_sw128(bL + ob);
                tc05::mma_bf16(tacc, dAH, dBH, kIdesc, accum);
                tc05::mma_bf16(tacc, dAH, dBL, kIdesc, 1);
                tc05::mma_bf16(tacc, dAL, dBH, kIdesc, 1);
                // lo.lo is 2^-16 of the product and zero-mean: it moves a squared distance that survives the close-pair
                // rule (>= (|x|^2+|y|^2)/256) by at most 2^-8 of its value and the score by < 1e-6 (validated against
                // scikit-learn at 1e-5 on every parity case); VAECLUST_SIL_MMA=4 keeps it
                if (a.mma4) tc05::mma_bf16(tacc, dAL, dBL, kIdesc, 1);
                accum = 1;
              }
            }
            tc05::commit(&b_empty[s]);   // ring slot reusable once these MMAs have read it
            tc05::commit(&t_full[as]);   // accumulator ready for the epilogue
          }
          __syncwarp();
          ++it;
        }
      }
    }
  } else {
    // ===================== epilogue: warps 2..9; thread = one row x one 64-column half of the tile ==========
    const int q = warp & 3;                      // TMEM lane quarter this warp may access
    const int half = (warp - 2) >> 2;            // which PCOLS columns of each tile
    const int rl = q * 32 + lane;                // row within the tile
    const int64_t row = r0 + rl;
    const bool row_ok = row < a.row_end;
    const float rn = row_ok ? a.sqn[row] : 0.f;
    const int64_t nrows = a.row_end - a.row_begin;
    float* o = a.out + ((int64_t)split * PARTS + half) * nrows * a.k;   // every column part is its own partial plane
    int64_t counter = 0;
    uint32_t it = 0;
    for (int c = 0; c < a.k; ++c) {
      const int64_t c0 = a.off[c], c1 = a.off[c + 1];
      float rsum = 0.f;
      for (int64_t j0 = c0 & ~(int64_t)3; j0 < c1; j0 += ST_COLS, ++counter) {   // 16-byte aligned norms
        if (counter % a.nsplit != split) continue;
        const int as = it & 1;
        const uint32_t aph = (it >> 1) & 1;
        tma::mbar_wait(&t_full[as], aph);
        tc05::fence_after();
        const bool masked = j0 < c0 || (j0 + ST_COLS > c1) || (j0 < r0 + ST_ROWS && r0 < j0 + ST_COLS);
#pragma unroll 1
        for (int ch = 0; ch < PCOLS / CW; ++ch) {
          const int cb0 = half * PCOLS + ch * CW;
          const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + as * ST_COLS + cb0;
          float v[CW];
          tc05::ld_32x32(tacc, v);          // all 32-column chunks in flight before the wait
          if (CW == 64) tc05::ld_32x32(tacc + 32, v + (CW == 64 ? 32 : 0));
          tc05::ld_wait();
          if (ch == PCOLS / CW - 1) {
            tc05::fence_before();
            tma::mbar_arrive(&t_empty[as]);   // accumulator stage is free as soon as it sits in registers
          }
          // column norms straight from global memory (every row of the tile reads the same 64 floats: L1 broadcast)
          const float* cn = a.sqn + j0 + cb0;
          if (!masked) {
            const float4* cn4 = reinterpret_cast<const float4*>(cn);
            // five instructions per pair: s = |x|^2+|y|^2, e = s - 2 x.y, running min, sqrt(|e|), running sum.  A close
            // pair (e < s/256) has | |x|-|y| | <= |x-y| < (|x|+|y|)/16, hence |y|^2 < 1.29 |x|^2 and e < |x|^2/112:
            // one comparison of the chunk minimum against |x|^2/112 finds every chunk that can hold one.
            float emin = 3.0e38f;
#pragma unroll
            for (int j = 0; j < CW / 4; ++j) {
              const float4 cc = __ldg(cn4 + j);
              const float e0 = fmaf(-2.f, v[4 * j + 0], rn + cc.x), e1 = fmaf(-2.f, v[4 * j + 1], rn + cc.y);
              const float e2 = fmaf(-2.f, v[4 * j + 2], rn + cc.z), e3 = fmaf(-2.f, v[4 * j + 3], rn + cc.w);
              emin = fminf(emin, fminf(fminf(e0, e1), fminf(e2, e3)));
              rsum += fast_sqrt(fabsf(e0));
              rsum += fast_sqrt(fabsf(e1));
              rsum += fast_sqrt(fabsf(e2));
              rsum += fast_sqrt(fabsf(e3));
            }
            if (emin < rn * (1.f / 112.f) && row_ok) {  // rare: replace the term of every close pair by the exact one
#pragma unroll
              for (int j = 0; j < CW; ++j) {
                const float s = rn + __ldg(cn + j);
                const float d2 = fmaf(-2.f, v[j], s);
                if (d2 < kCloseRatio * s)
                  rsum += exact_dist(a.Xs, a.d, row, j0 + cb0 + j) - fast_sqrt(fabsf(d2));
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < CW; ++j) {
              const int64_t col = j0 + cb0 + j;
              const bool in_cluster = col >= c0 && col < c1;
              const bool live = in_cluster && col != row && row_ok;
              const float s = rn + (in_cluster ? __ldg(cn + j) : 0.f);
              const float d2 = fmaf(-2.f, v[j], s);
              float dist = fast_sqrt(fabsf(d2));
              if (live && d2 < kCloseRatio * s) dist = exact_dist(a.Xs, a.d, row, col);
              rsum += live ? dist : 0.f;
            }
          }
        }
        ++it;
      }
      if (row_ok) o[(row - a.row_begin) * a.k + c] = rsum;
    }
  }

  tc05::fence_before();
  __syncthreads();
  if (warp == 1) {
    tc05::fence_after();
    tc05::tmem_dealloc(tmem_base, 2 * ST_COLS);
  }
}

}  // namespace vc

using namespace vc;

extern "C" int vc_silhouette_tc_planes(int dp) { return dp == 64 ? 4 : 2; }

extern "C" int vc_silhouette_split_bf16(const float* Xs, int64_t N, int d, int dp, void* H, void* L, float* sqn_split,
                                        void* stream) {
  VC_REQUIRE(Xs && H && L && N > 0 && d > 0 && dp >= d && dp % 64 == 0, VC_ERR_BAD_ARG,
             "vc_silhouette_split_bf16: bad arguments (d=%d dp=%d)", d, dp);
  int64_t blocks = (N + 7) / 8;
  if (blocks > (int64_t)sm_count() * 16) blocks = (int64_t)sm_count() * 16;
  split_bf16_kernel<<<(int)blocks, 256, 0, as_stream(stream)>>>(Xs, N, d, dp, static_cast<__nv_bfloat16*>(H),
                                                               static_cast<__nv_bfloat16*>(L), sqn_split);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" int vc_silhouette_rows_tc(const void* H, const void* L, const float* Xs, int d, const float* sqn, int64_t N, int dp,
                                     const int64_t* off_dev, int k, int64_t row_begin, int64_t row_end, int nsplit,
                                     float* cluster_sums, void* stream) {
  VC_REQUIRE(H && L && Xs && sqn && off_dev && cluster_sums && N > 0 && k > 0, VC_ERR_BAD_ARG,
             "vc_silhouette_rows_tc: bad arguments");
  VC_REQUIRE(d > 0 && d <= dp && d % 4 == 0 && (reinterpret_cast<uintptr_t>(Xs) & 15) == 0, VC_ERR_UNSUPPORTED,
             "vc_silhouette_rows_tc: fp32 rows need d %% 4 == 0 and 16-byte alignment (d=%d)", d);
  VC_REQUIRE(dp % 64 == 0 && dp >= 64 && dp <= 128, VC_ERR_UNSUPPORTED,
             "vc_silhouette_rows_tc: padded width %d not in {64, 128} (use the fp32 kernel)", dp);
  VC_REQUIRE(row_begin >= 0 && row_end <= N && row_begin < row_end, VC_ERR_BAD_ARG,
             "vc_silhouette_rows_tc: bad row range");
  VC_REQUIRE(nsplit >= 1 && nsplit <= 65535 && N < 2147483647LL, VC_ERR_BAD_ARG, "vc_silhouette_rows_tc: bad nsplit/N");
  VC_REQUIRE(((reinterpret_cast<uintptr_t>(H) | reinterpret_cast<uintptr_t>(L)) & 127) == 0, VC_ERR_BAD_ARG,
             "vc_silhouette_rows_tc: operand buffers must be 128-byte aligned");
  CUtensorMap mH, mL;
  const uint64_t dims[2] = {(uint64_t)dp, (uint64_t)N};
  const uint64_t strides[1] = {(uint64_t)dp * 2};
  const uint32_t box[2] = {64u, (uint32_t)ST_ROWS};
  int rc = tma::make_map(&mH, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, H, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc != VC_OK) return rc;
  rc = tma::make_map(&mL, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, L, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc != VC_OK) return rc;
  SilTcArgs a;
  a.sqn = sqn; a.off = off_dev; a.k = k; a.N = N; a.row_begin = row_begin; a.row_end = row_end; a.nsplit = nsplit;
  a.Xs = Xs; a.d = d;
  a.out = cluster_sums; a.panels = dp / 64; a.stages = dp == 64 ? 3 : 2;
  {
    const char* e = getenv("VAECLUST_SIL_MMA");
    a.mma4 = (e && e[0] == '4') ? 1 : 0;
  }
  const int cols = dp == 64 ? 256 : 128;
  const size_t op = (size_t)ST_ROWS * 128 * a.panels, bop = (size_t)cols * 128 * a.panels;
  const size_t smem = 2 * op + (size_t)a.stages * 2 * bop + (1 + 2 * a.stages + 4) * 8 + 16 + 1024;
  auto kern = dp == 64 ? silhouette_tc_kernel<256, 16> : silhouette_tc_kernel<128, 8>;
  const int threads = dp == 64 ? 64 + 32 * 16 : 64 + 32 * 8;
  VC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t tiles = (row_end - row_begin + ST_ROWS - 1) / ST_ROWS;
  dim3 grid((unsigned)tiles, (unsigned)nsplit);
  kern<<<grid, threads, smem, as_stream(stream)>>>(a, mH, mL);
  VC_LAUNCH_CHECK();
  return VC_OK;
}
