// Weight gradient + Adam update of the two 67M-parameter Linear layers (12:98-103,129-132) as ONE persistent kernel whose
// parameter / moment traffic is moved by TMA.  EXPERIMENT, off by default (VAECLUST_FC_ADAM_TMA=1 enables it): it reaches
// the same 5.2 TB/s as the fused epilogue of skinny_gemm_kernel (measured alone: 352-374 us vs 337 us per launch; ncu: DRAM
// active 58-64 % of cycles, warps waiting for their TMA loads ~5 us after issue, i.e. the memory system is saturated by the
// seven interleaved streams of 64-byte row pieces, not the SMs), and needs sqrt.approx / div.approx
// (VAECLUST_FC_ADAM_APPROX=1) to keep eight epilogue warps from being latency-bound on the Adam arithmetic.
//
// The op is a 24 B/parameter read-modify-write stream (p, exp_avg, exp_avg_sq in and out, + 2 B bf16 shadow out) around a
// K = batch (<= 64) GEMM that costs nothing.  Every p / m / v tile is a 32-row x 16-column fp32 TMA box (64-byte rows,
// SWIZZLE_64B):
//   * each of the eight worker warps owns a three-stage ring of {p, m, v} boxes (6 KB per stage): 144 KB of loads in
//     flight per SM with no registers behind them, issued across tile boundaries (the first ones before the first MMA);
//   * the accumulator row of a thread (TMEM lane = row of the parameter matrix, 16 consecutive columns) and its p / m / v
//     row in the swizzled box line up: four conflict-free 16-byte shared-memory reads per array, Adam in registers, four
//     writes back in place, then one TMA store per array (bulk group; the stage is reloaded once the store of the
//     PREVIOUS chunk has read its source);
//   * the GEMM is arranged so that the TMEM lane is always the row of the parameter's memory layout: for W [N][M]
//     (parameter rows = the 256-wide side) the roles of the operands are swapped (two 128 x 128 MMAs per k-step instead of
//     one 128 x 256), so no transposition through shared memory is needed in either layer;
//   * the same warps convert the fp32 activation tile of the NEXT m-tile (lanes = 8 rows x 4 k-pairs: conflict-free swizzled
//     stores; loads issued before a tile's epilogue, stored after its first chunk); the 256 x 64 bf16 small operand is
//     loaded once per CTA.
// One CTA per SM (208 KB of shared memory, all 512 TMEM columns: two accumulator buffers).
#include <string.h>
#include "common.cuh"
#include "tc05.cuh"

namespace vc {

constexpr int FA_THREADS = 288;               // warps 0-7 producers + epilogue, 8 MMA issuer
constexpr uint32_t FA_A_BYTES = 128 * 128;    // [128 m][64 k] bf16
constexpr uint32_t FA_B_BYTES = 256 * 128;    // [256 n][64 k] bf16
constexpr uint32_t FA_BOX = 32 * 64;          // one 32-row x 16-column fp32 box (64-byte rows, SWIZZLE_64B)
constexpr uint32_t FA_STAGE = 3 * FA_BOX;     // p, m, v
constexpr int FA_STAGES = 3;

struct FaArgs {
  const float* A; int64_t lda;                // fp32 activations, A(m, k) = A[k * lda + m]
  int M, K;                                   // M = the 262144-wide side, K = batch (<= 64)
  int transposed;                             // parameter element (m, n) at n * ldo + m (1) or m * ldo + n (0)
  int64_t ldo;
  __nv_bfloat16* shadow;                      // bf16 copy of the parameter (same layout) or null
  const VcAdamDeviceStep* step;
  float lr, b1, b2, eps, bc1, bc2;
  uint32_t idesc;
};

// = sk_adam (skinny_gemm.cu) = adam_multi_kernel (misc.cu): torch.optim.Adam's arithmetic.  APPROX replaces the IEEE square
// root and division (~18 of the ~35 instructions per element) by sqrt.approx / div.approx (1-2 ulp).
template <bool APPROX>
__device__ __forceinline__ void fa_adam(float g, float& p, float& m, float& v, float b1, float b2, float step_size,
                                        float inv_sqrt_bc2, float eps) {
  m = b1 * m + (1.f - b1) * g;
  v = b2 * v + (1.f - b2) * g * g;
  if (APPROX) {
    float sq;
    asm("sqrt.approx.f32 %0, %1;" : "=f"(sq) : "f"(v));
    p -= step_size * __fdividef(m, sq * inv_sqrt_bc2 + eps);
  } else {
    p -= step_size * (m / (sqrtf(v) * inv_sqrt_bc2 + eps));
  }
}
__device__ __forceinline__ uint32_t fa_pack(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ void fa_bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }

template <bool APPROX>
__global__ void __launch_bounds__(FA_THREADS, 1)
fc_adam_kernel(const FaArgs a, const __grid_constant__ CUtensorMap mapB, const __grid_constant__ CUtensorMap mapP,
               const __grid_constant__ CUtensorMap mapM, const __grid_constant__ CUtensorMap mapV) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (tma::s32(smem_raw) & 1023u)) & 1023u);
  unsigned char* sB = smem;                                   // 32 KB
  unsigned char* sA = smem + FA_B_BYTES;                      // 2 x 16 KB
  unsigned char* sP = sA + 2 * FA_A_BYTES;                    // 8 warps x 3 stages x 6 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 8 * FA_STAGES * FA_STAGE);
  uint64_t* a_full = bars;            // [2] 256 producer arrivals
  uint64_t* a_empty = bars + 2;       // [2] tcgen05.commit
  uint64_t* acc_full = bars + 4;      // [2] tcgen05.commit
  uint64_t* acc_empty = bars + 6;     // [2] 8 epilogue warps
  uint64_t* b_full = bars + 8;        // [1]
  uint64_t* pmv_full = bars + 9;      // [8][3]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9 + 8 * FA_STAGES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntiles = a.M >> 7;
  const int my_tiles = ((int)blockIdx.x < ntiles) ? (ntiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  if (threadIdx.x == 0) {
    tma::prefetch_map(&mapB); tma::prefetch_map(&mapP); tma::prefetch_map(&mapM); tma::prefetch_map(&mapV);
    for (int s = 0; s < 2; ++s) {
      tma::mbar_init(&a_full[s], 256); tma::mbar_init(&a_empty[s], 1);
      tma::mbar_init(&acc_full[s], 1); tma::mbar_init(&acc_empty[s], 8);
    }
    tma::mbar_init(b_full, 1);
    for (int s = 0; s < 8 * FA_STAGES; ++s) tma::mbar_init(&pmv_full[s], 1);
    tma::mbar_fence_init();
  }
  if (warp == 8) tc05::tmem_alloc(tmem_slot, 512);
  tc05::fence_before();
  __syncthreads();
  tc05::fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 8) {
    // ---- producer part: fp32 [k][m] -> bf16 K-major swizzled [m][k] of tile `it` (16 rows per warp).  lane = (mq = lane & 7,
    // kq = lane >> 3): a warp store covers 8 consecutive m of 4 k-pairs -> 32 distinct banks.  The loads of tile it + 1 are
    // issued before the epilogue of tile it and stored after its first chunk, so their latency hides under that chunk.
    const int mq = lane & 7, kq = lane >> 3;
    float alo[16], ahi[16];
    auto a_load = [&](int it) {
      const int m0 = ((int)blockIdx.x + it * (int)gridDim.x) << 7;
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const int k = 2 * (4 * c + kq);
        const uint32_t ok0 = k < a.K ? 1u : 0u, ok1 = k + 1 < a.K ? 1u : 0u;
        const float* p = a.A + (int64_t)k * a.lda + m0 + warp * 16 + mq;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          alo[2 * c + i] = ldg_pred(p + 8 * i, ok0);
          ahi[2 * c + i] = ldg_pred(p + a.lda + 8 * i, ok1);
        }
      }
    };
    auto a_store = [&](int it) {
      const int s = it & 1;
      tma::mbar_wait_backoff(&a_empty[s], (((uint32_t)it >> 1) & 1u) ^ 1u, 64);
      unsigned char* tile = sA + (size_t)s * FA_A_BYTES;
#pragma unroll
      for (int c = 0; c < 8; ++c)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const uint32_t r = (uint32_t)(warp * 16 + 8 * i + mq);
          *reinterpret_cast<uint32_t*>(tile + r * 128u + (((uint32_t)c ^ (r & 7u)) << 4) + ((uint32_t)kq << 2)) =
              fa_pack(alo[2 * c + i], ahi[2 * c + i]);
        }
      tma::fence_proxy_async();
      tma::mbar_arrive(&a_full[s]);
    };
    // ---- epilogue part: Adam on TMA-staged p / m / v boxes.  Warp (q = warp & 3, hh = warp >> 2) owns TMEM lanes 32 q .. and
    // the 16-column half hh of every 32-column chunk (the two warps of a quarter fetch the two halves of the same 128-byte
    // lines at about the same time).
    const int q = warp & 3, hh = warp >> 2;
    unsigned char* ring = sP + (size_t)warp * FA_STAGES * FA_STAGE;
    uint64_t* full = pmv_full + warp * FA_STAGES;
    const int total = my_tiles * 8;
    float bc1 = a.bc1, bc2 = a.bc2;
    if (a.step) { bc1 = a.step->bias_corr1; bc2 = a.step->bias_corr2; }
    const float step_size = a.lr / bc1, inv_sqrt_bc2 = 1.f / sqrtf(bc2);
    // chunk c = (tile it = c >> 3, j = c & 7) -> box coordinates (c0 = contiguous dimension, c1 = row)
    auto coords = [&](int c, int& c0, int& c1) {
      const int m0 = ((int)blockIdx.x + (c >> 3) * (int)gridDim.x) << 7, j = c & 7;
      if (a.transposed) { c0 = m0 + (j & 3) * 32 + hh * 16; c1 = (j >> 2) * 128 + q * 32; }      // rows = n, columns = m
      else { c0 = j * 32 + hh * 16; c1 = m0 + q * 32; }                                            // rows = m, columns = n
    };
    auto issue_load = [&](int c) {      // one elected lane
      const int st = c % FA_STAGES;
      int c0, c1;
      coords(c, c0, c1);
      unsigned char* dst = ring + (size_t)st * FA_STAGE;
      tma::mbar_expect_tx(&full[st], FA_STAGE);
      tma::load_2d(dst, &mapP, &full[st], c0, c1);
      tma::load_2d(dst + FA_BOX, &mapM, &full[st], c0, c1);
      tma::load_2d(dst + 2 * FA_BOX, &mapV, &full[st], c0, c1);
    };
    if (tc05::elect_one()) {
      for (int c = 0; c < FA_STAGES - 1 && c < total; ++c) issue_load(c);
    }
    __syncwarp();
    if (my_tiles > 0) { a_load(0); a_store(0); }
    int st = 0;
    uint32_t ph = 0;
    // 16-byte chunk i of this lane's 64-byte row inside a SWIZZLE_64B box: index XOR (row / 2) % 4
    const uint32_t rowoff = (uint32_t)lane * 64u, sw = ((uint32_t)lane >> 1) & 3u;
    for (int c = 0; c < total; ++c) {
      const int it = c >> 3, j = c & 7, ab = it & 1;
      if (j == 0) {
        if (it + 1 < my_tiles) a_load(it + 1);
        tma::mbar_wait_backoff(&acc_full[ab], ((uint32_t)it >> 1) & 1u, 64);
        tc05::fence_after();
      }
      float g[16];
      tc05::ld_32x16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * 256 + j * 32 + hh * 16), g);
      tma::mbar_wait(&full[st], ph);
      tc05::ld_wait();
      if (j == 7) {      // the accumulator buffer is free once its last chunk sits in registers
        tc05::fence_before();
        if (lane == 0) tma::mbar_arrive(&acc_empty[ab]);
      }
      unsigned char* box = ring + (size_t)st * FA_STAGE + rowoff;
      uint32_t pk[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t off = ((uint32_t)i ^ sw) << 4;
        float4 p4 = *reinterpret_cast<float4*>(box + off);
        float4 m4 = *reinterpret_cast<float4*>(box + FA_BOX + off);
        float4 v4 = *reinterpret_cast<float4*>(box + 2 * FA_BOX + off);
        fa_adam<APPROX>(g[4 * i], p4.x, m4.x, v4.x, a.b1, a.b2, step_size, inv_sqrt_bc2, a.eps);
        fa_adam<APPROX>(g[4 * i + 1], p4.y, m4.y, v4.y, a.b1, a.b2, step_size, inv_sqrt_bc2, a.eps);
        fa_adam<APPROX>(g[4 * i + 2], p4.z, m4.z, v4.z, a.b1, a.b2, step_size, inv_sqrt_bc2, a.eps);
        fa_adam<APPROX>(g[4 * i + 3], p4.w, m4.w, v4.w, a.b1, a.b2, step_size, inv_sqrt_bc2, a.eps);
        *reinterpret_cast<float4*>(box + off) = p4;
        *reinterpret_cast<float4*>(box + FA_BOX + off) = m4;
        *reinterpret_cast<float4*>(box + 2 * FA_BOX + off) = v4;
        pk[2 * i] = fa_pack(p4.x, p4.y);
        pk[2 * i + 1] = fa_pack(p4.z, p4.w);
      }
      int c0, c1;
      coords(c, c0, c1);
      if (a.shadow) {      // this thread's 16 consecutive elements of row c1 + lane: 32 bytes
        uint4* sh = reinterpret_cast<uint4*>(a.shadow + (int64_t)(c1 + lane) * a.ldo + c0);
        sh[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        sh[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
      }
      tma::fence_proxy_async();
      __syncwarp();
      if (tc05::elect_one()) {
        unsigned char* src = ring + (size_t)st * FA_STAGE;
        tma::store_2d(&mapP, src, c0, c1);
        tma::store_2d(&mapM, src + FA_BOX, c0, c1);
        tma::store_2d(&mapV, src + 2 * FA_BOX, c0, c1);
        tma::bulk_commit();
        // the stage of chunk c - 1 is reloaded with chunk c + 2 once its store (the previous group) has read its source
        if (c + FA_STAGES - 1 < total) {
          fa_bulk_wait_read1();
          issue_load(c + FA_STAGES - 1);
        }
      }
      __syncwarp();
      if (++st == FA_STAGES) { st = 0; ph ^= 1u; }
      if (j == 0 && it + 1 < my_tiles) a_store(it + 1);
    }
    if (tc05::elect_one()) tma::bulk_wait0();
    __syncwarp();
  } else {
    // ===================== MMA issuer (+ the one load of the small operand) =====================
    if (tc05::elect_one()) {
      tma::mbar_expect_tx(b_full, FA_B_BYTES);
      tma::load_2d(sB, &mapB, b_full, 0, 0);
    }
    __syncwarp();
    tma::mbar_wait_backoff(b_full, 0, 64);
    for (int it = 0; it < my_tiles; ++it) {
      const int s = it & 1;
      const uint32_t ph = ((uint32_t)it >> 1) & 1u;
      tma::mbar_wait_backoff(&acc_empty[s], ph ^ 1u, 64);
      tma::mbar_wait_backoff(&a_full[s], ph, 32);
      tc05::fence_after();
      if (tc05::elect_one()) {
        const uint32_t sa = tma::s32(sA + (size_t)s * FA_A_BYTES), sb = tma::s32(sB);
        const uint32_t d = tmem_base + (uint32_t)(s * 256);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          if (a.transposed) {      // D_h[n][m] = Bs_h[n][k] . Act[m][k]: rows = n (two halves), columns = m
            tc05::mma_bf16(d, tc05::desc_k_sw128(sb + kk * 32), tc05::desc_k_sw128(sa + kk * 32), a.idesc, (uint32_t)(kk != 0));
            tc05::mma_bf16(d + 128, tc05::desc_k_sw128(sb + 128 * 128 + kk * 32), tc05::desc_k_sw128(sa + kk * 32), a.idesc,
                           (uint32_t)(kk != 0));
          } else {                 // D[m][n] = Act[m][k] . Bs[n][k]
            tc05::mma_bf16(d, tc05::desc_k_sw128(sa + kk * 32), tc05::desc_k_sw128(sb + kk * 32), a.idesc, (uint32_t)(kk != 0));
          }
        }
        tc05::commit(&a_empty[s]);
        tc05::commit(&acc_full[s]);
      }
      __syncwarp();
    }
  }

  tc05::fence_before();
  __syncthreads();
  if (warp == 8) {
    tc05::fence_after();
    tc05::tmem_dealloc(tmem_base, 512);
  }
}

// ---- Variant 2 (VAECLUST_FC_ADAM_TMA=2): the accumulator tile goes to shared memory, p / m / v are swept in MEMORY ORDER.
// Both kernels above (and the fused epilogue of skinny_gemm_kernel) touch every 1 KB row of the parameter eight separate
// times per tile (once per 32-column chunk, microseconds apart): eight DRAM row activations instead of one, the suspected
// reason why this read-modify-write stream stops at 5.2 TB/s while the linear multi-tensor Adam reaches 6.1 (measured: it is
// NOT -- 382-392 us per launch against 337 us for the tile-order epilogue, profiles/fc_adam_tma_r2.txt).  Here the
// 128 x 256 (or 256 x 128) fp32 accumulator tile is drained into a 128 KB shared-memory tile G (16-byte chunks XOR-swizzled
// with the row, conflict-free for the row-per-lane drain AND the row-per-warp sweep), and the eight warps then walk the
// tile's p / m / v in address order: a warp instruction moves 512 contiguous bytes, the CTA covers 16 KB contiguous per
// batch, every DRAM page is consumed at once.  The sweep is software-pipelined across tile boundaries (the loads of the next
// batch are in flight while the drain of the next tile runs), the MMA of the next tile runs under the sweep (two TMEM buffers).
constexpr uint32_t FR_G_BYTES = 128 * 256 * 4;
constexpr int FR_NW = 16;                         // worker warps (producers + drain + sweep); warp FR_NW issues the MMAs
constexpr int FR_THREADS = (FR_NW + 1) * 32;

template <bool APPROX>
__global__ void __launch_bounds__(FR_THREADS, 1)
fc_adam_rows_kernel(const FaArgs a, const __grid_constant__ CUtensorMap mapB, float* __restrict__ P, float* __restrict__ Mo,
                    float* __restrict__ Vo) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (tma::s32(smem_raw) & 1023u)) & 1023u);
  unsigned char* sG = smem;                                   // 128 KB
  unsigned char* sB = sG + FR_G_BYTES;                        // 32 KB
  unsigned char* sA = sB + FA_B_BYTES;                        // 2 x 16 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(sA + 2 * FA_A_BYTES);
  uint64_t* a_full = bars;            // [2] FR_NW * 32 producer arrivals
  uint64_t* a_empty = bars + 2;       // [2] tcgen05.commit
  uint64_t* acc_full = bars + 4;      // [2] tcgen05.commit
  uint64_t* acc_empty = bars + 6;     // [2] FR_NW warps
  uint64_t* b_full = bars + 8;        // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntiles = a.M >> 7;
  const int my_tiles = ((int)blockIdx.x < ntiles) ? (ntiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  if (threadIdx.x == 0) {
    tma::prefetch_map(&mapB);
    for (int s = 0; s < 2; ++s) {
      tma::mbar_init(&a_full[s], FR_NW * 32); tma::mbar_init(&a_empty[s], 1);
      tma::mbar_init(&acc_full[s], 1); tma::mbar_init(&acc_empty[s], FR_NW);
    }
    tma::mbar_init(b_full, 1);
    tma::mbar_fence_init();
  }
  if (warp == FR_NW) tc05::tmem_alloc(tmem_slot, 512);
  tc05::fence_before();
  __syncthreads();
  tc05::fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < FR_NW) {
    // producer part: this warp converts rows 8 warp .. 8 warp + 7 of the fp32 [k][m] activation tile (lane = 8 rows x 4 k-pairs)
    const int mq = lane & 7, kq = lane >> 3;
    float alo[8], ahi[8];
    auto a_load = [&](int it) {
      const int m0 = ((int)blockIdx.x + it * (int)gridDim.x) << 7;
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const int k = 2 * (4 * c + kq);
        const float* p = a.A + (int64_t)k * a.lda + m0 + warp * 8 + mq;
        alo[c] = ldg_pred(p, k < a.K ? 1u : 0u);
        ahi[c] = ldg_pred(p + a.lda, k + 1 < a.K ? 1u : 0u);
      }
    };
    auto a_store = [&](int it) {
      const int s = it & 1;
      tma::mbar_wait_backoff(&a_empty[s], (((uint32_t)it >> 1) & 1u) ^ 1u, 64);
      unsigned char* tile = sA + (size_t)s * FA_A_BYTES;
      const uint32_t r = (uint32_t)(warp * 8 + mq);
#pragma unroll
      for (int c = 0; c < 8; ++c)
        *reinterpret_cast<uint32_t*>(tile + r * 128u + (((uint32_t)c ^ (r & 7u)) << 4) + ((uint32_t)kq << 2)) = fa_pack(alo[c], ahi[c]);
      tma::fence_proxy_async();
      tma::mbar_arrive(&a_full[s]);
    };
    const int q = warp & 3, hh = warp >> 2;                   // TMEM lane quarter, which two of the eight 32-column chunks
    float bc1 = a.bc1, bc2 = a.bc2;
    if (a.step) { bc1 = a.step->bias_corr1; bc2 = a.step->bias_corr2; }
    const float step_size = a.lr / bc1, inv_sqrt_bc2 = 1.f / sqrtf(bc2);
    // G: rows x cols fp32, 16-byte chunk ch of row r at r * pitch + ((ch ^ (r & 7)) << 4)
    const uint32_t pitch = a.transposed ? 512u : 1024u;       // 128 or 256 columns
    // unit u (0..255 per tile) = 128 consecutive floats of one row, in memory order: row = u / upr, column (u % upr) * 128
    const int upr_shift = a.transposed ? 0 : 1;
    auto goff = [&](int it, int u) -> int64_t {
      const int m0 = ((int)blockIdx.x + it * (int)gridDim.x) << 7;
      const int row = u >> upr_shift, c0 = (u & ((1 << upr_shift) - 1)) << 7;
      return a.transposed ? (int64_t)row * a.ldo + m0 + c0 + 4 * lane : (int64_t)(m0 + row) * a.ldo + c0 + 4 * lane;
    };
    // batch bt (0..3 per tile) of this warp: units 64 bt + 16 i + warp, i = 0..3 -> the CTA covers 64 consecutive units (32 KB
    // of each array) per batch, a warp instruction 512 contiguous bytes
    float4 pp[4], mm[4], vv[4];
    auto load_batch = [&](int it, int bt) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int64_t o = goff(it, 64 * bt + 16 * i + warp);
        pp[i] = *reinterpret_cast<const float4*>(P + o);
        mm[i] = __ldcs(reinterpret_cast<const float4*>(Mo + o));
        vv[i] = __ldcs(reinterpret_cast<const float4*>(Vo + o));
      }
    };
    auto process_batch = [&](int it, int bt) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int u = 64 * bt + 16 * i + warp;
        const uint32_t row = (uint32_t)(u >> upr_shift), ch = (uint32_t)(((u & ((1 << upr_shift) - 1)) << 5) + lane);
        const float4 g = *reinterpret_cast<const float4*>(sG + row * pitch + ((ch ^ (row & 7u)) << 4));
        fa_adam<APPROX>(g.x, pp[i].x, mm[i].x, vv[i].x, a.b1, a.b2, step_size, inv_sqrt_bc2, a.eps);
        fa_adam<APPROX>(g.y, pp[i].y, mm[i].y, vv[i].y, a.b1, a.b2, step_size, inv_sqrt_bc2, a.eps);
        fa_adam<APPROX>(g.z, pp[i].z, mm[i].z, vv[i].z, a.b1, a.b2, step_size, inv_sqrt_bc2, a.eps);
        fa_adam<APPROX>(g.w, pp[i].w, mm[i].w, vv[i].w, a.b1, a.b2, step_size, inv_sqrt_bc2, a.eps);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int64_t o = goff(it, 64 * bt + 16 * i + warp);
        *reinterpret_cast<float4*>(P + o) = pp[i];
        __stcs(reinterpret_cast<float4*>(Mo + o), mm[i]);
        __stcs(reinterpret_cast<float4*>(Vo + o), vv[i]);
        if (a.shadow)
          *reinterpret_cast<uint2*>(a.shadow + o) = make_uint2(fa_pack(pp[i].x, pp[i].y), fa_pack(pp[i].z, pp[i].w));
      }
    };
    if (my_tiles > 0) {
      a_load(0);
      a_store(0);
    }
    for (int it = 0; it < my_tiles; ++it) {
      const int ab = it & 1;
      if (it + 1 < my_tiles) a_load(it + 1);
      load_batch(it, 0);                 // in flight under the drain
      // ---- drain the accumulator of tile it into G (everybody is done reading G of tile it - 1: barrier at the loop end)
      tma::mbar_wait_backoff(&acc_full[ab], ((uint32_t)it >> 1) & 1u, 64);
      tc05::fence_after();
#pragma unroll 1
      for (int j = hh; j < 8; j += 4) {
        float g[32];
        tc05::ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * 256 + j * 32), g);
        tc05::ld_wait();
        // TMEM (lane, column j * 32 + i) -> (row, column) of the parameter tile
        const uint32_t row = a.transposed ? (uint32_t)((j >> 2) * 128 + q * 32 + lane) : (uint32_t)(q * 32 + lane);
        const uint32_t ch0 = a.transposed ? (uint32_t)((j & 3) * 8) : (uint32_t)(j * 8);
        unsigned char* gr = sG + row * pitch;
#pragma unroll
        for (int i = 0; i < 8; ++i)
          *reinterpret_cast<float4*>(gr + (((ch0 + i) ^ (row & 7u)) << 4)) = make_float4(g[4 * i], g[4 * i + 1], g[4 * i + 2], g[4 * i + 3]);
      }
      tc05::fence_before();
      if (lane == 0) tma::mbar_arrive(&acc_empty[ab]);
      asm volatile("bar.sync 1, %0;" ::"n"(FR_NW * 32) : "memory");
      if (it + 1 < my_tiles) a_store(it + 1);
      // ---- sweep in memory order
#pragma unroll 1
      for (int bt = 0; bt < 4; ++bt) {
        if (bt > 0) load_batch(it, bt);
        process_batch(it, bt);
      }
      asm volatile("bar.sync 1, %0;" ::"n"(FR_NW * 32) : "memory");      // G may be overwritten
    }
  } else {
    if (tc05::elect_one()) {
      tma::mbar_expect_tx(b_full, FA_B_BYTES);
      tma::load_2d(sB, &mapB, b_full, 0, 0);
    }
    __syncwarp();
    tma::mbar_wait_backoff(b_full, 0, 64);
    for (int it = 0; it < my_tiles; ++it) {
      const int s = it & 1;
      const uint32_t ph = ((uint32_t)it >> 1) & 1u;
      tma::mbar_wait_backoff(&acc_empty[s], ph ^ 1u, 64);
      tma::mbar_wait_backoff(&a_full[s], ph, 32);
      tc05::fence_after();
      if (tc05::elect_one()) {
        const uint32_t sa = tma::s32(sA + (size_t)s * FA_A_BYTES), sb = tma::s32(sB);
        const uint32_t d = tmem_base + (uint32_t)(s * 256);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          if (a.transposed) {
            tc05::mma_bf16(d, tc05::desc_k_sw128(sb + kk * 32), tc05::desc_k_sw128(sa + kk * 32), a.idesc, (uint32_t)(kk != 0));
            tc05::mma_bf16(d + 128, tc05::desc_k_sw128(sb + 128 * 128 + kk * 32), tc05::desc_k_sw128(sa + kk * 32), a.idesc,
                           (uint32_t)(kk != 0));
          } else {
            tc05::mma_bf16(d, tc05::desc_k_sw128(sa + kk * 32), tc05::desc_k_sw128(sb + kk * 32), a.idesc, (uint32_t)(kk != 0));
          }
        }
        tc05::commit(&a_empty[s]);
        tc05::commit(&acc_full[s]);
      }
      __syncwarp();
    }
  }

  tc05::fence_before();
  __syncthreads();
  if (warp == FR_NW) {
    tc05::fence_after();
    tc05::tmem_dealloc(tmem_base, 512);
  }
}

// ---- host -----------------------------------------------------------------------------------------------------------
bool fc_adam_tma_supported(int M, int N, int K, int64_t lda, int64_t ldo, int transposed, const void* p, const void* m,
                           const void* v, const void* shadow) {
  static const bool on = getenv("VAECLUST_FC_ADAM_TMA") && atoi(getenv("VAECLUST_FC_ADAM_TMA")) != 0;
  if (!on) return false;      // 1: TMA-staged boxes, 2: shared-memory accumulator tile + row-linear sweep
  if (N != 256 || M % 128 != 0 || K < 1 || K > 64 || lda < M || (ldo & 7)) return false;
  if (transposed ? ldo < M : ldo < N) return false;
  const uintptr_t al = reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v) |
                       reinterpret_cast<uintptr_t>(shadow);
  return (al & 15) == 0;
}

struct FcAdamParams { float* m; float* v; const VcAdamDeviceStep* step; float lr, b1, b2, eps, bc1, bc2; __nv_bfloat16* shadow; };

// act: fp32 [K][lda] (m-contiguous), Bs: bf16 [256][64] (K-major, zero padded), p: the parameter (updated in place)
int fc_adam_tma_run(const float* act, int64_t lda, const __nv_bfloat16* Bs, float* p, int M, int N, int K, int64_t ldo,
                    int transposed, const FcAdamParams& ad, cudaStream_t s) {
  CUtensorMap mb, mp, mm, mv;
  {
    const uint64_t dims[2] = {64u, 256u};
    const uint64_t str[1] = {128u};
    const uint32_t box[2] = {64u, 256u};
    const int rc = tma::make_map(&mb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<__nv_bfloat16*>(Bs), dims, str, box,
                                 CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc != VC_OK) return rc;
  }
  static const int variant = getenv("VAECLUST_FC_ADAM_TMA") ? atoi(getenv("VAECLUST_FC_ADAM_TMA")) : 0;
  static const bool approx2 = getenv("VAECLUST_FC_ADAM_APPROX") && atoi(getenv("VAECLUST_FC_ADAM_APPROX")) != 0;
  if (variant == 2) {
    FaArgs a;
    memset(&a, 0, sizeof(a));
    a.A = act; a.lda = lda; a.M = M; a.K = K; a.transposed = transposed; a.ldo = ldo; a.shadow = ad.shadow; a.step = ad.step;
    a.lr = ad.lr; a.b1 = ad.b1; a.b2 = ad.b2; a.eps = ad.eps; a.bc1 = ad.bc1; a.bc2 = ad.bc2;
    a.idesc = tc05::idesc(tc05::kFmtBF16, 128, transposed ? 128 : 256, 0, 0);
    const size_t smem = FR_G_BYTES + FA_B_BYTES + 2 * FA_A_BYTES + 9 * 8 + 16 + 1024;
    int grid = sm_count();
    if (grid > M / 128) grid = M / 128;
    if (approx2) {
      VC_CUDA(cudaFuncSetAttribute(fc_adam_rows_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      fc_adam_rows_kernel<true><<<grid, FR_THREADS, smem, s>>>(a, mb, p, ad.m, ad.v);
    } else {
      VC_CUDA(cudaFuncSetAttribute(fc_adam_rows_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      fc_adam_rows_kernel<false><<<grid, FR_THREADS, smem, s>>>(a, mb, p, ad.m, ad.v);
    }
    VC_LAUNCH_CHECK();
    return VC_OK;
  }
  const uint64_t dims[2] = {(uint64_t)(transposed ? M : N), (uint64_t)(transposed ? N : M)};
  const uint64_t str[1] = {(uint64_t)ldo * 4};
  const uint32_t box[2] = {16u, 32u};
  int rc = tma::make_map(&mp, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, p, dims, str, box, CU_TENSOR_MAP_SWIZZLE_64B);
  if (rc != VC_OK) return rc;
  rc = tma::make_map(&mm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, ad.m, dims, str, box, CU_TENSOR_MAP_SWIZZLE_64B);
  if (rc != VC_OK) return rc;
  rc = tma::make_map(&mv, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, ad.v, dims, str, box, CU_TENSOR_MAP_SWIZZLE_64B);
  if (rc != VC_OK) return rc;
  FaArgs a;
  memset(&a, 0, sizeof(a));
  a.A = act; a.lda = lda; a.M = M; a.K = K; a.transposed = transposed; a.ldo = ldo; a.shadow = ad.shadow; a.step = ad.step;
  a.lr = ad.lr; a.b1 = ad.b1; a.b2 = ad.b2; a.eps = ad.eps; a.bc1 = ad.bc1; a.bc2 = ad.bc2;
  a.idesc = tc05::idesc(tc05::kFmtBF16, 128, transposed ? 128 : 256, 0, 0);
  const size_t smem = FA_B_BYTES + 2 * FA_A_BYTES + 8 * FA_STAGES * FA_STAGE + (9 + 8 * FA_STAGES) * 8 + 16 + 1024;
  static const bool approx = getenv("VAECLUST_FC_ADAM_APPROX") && atoi(getenv("VAECLUST_FC_ADAM_APPROX")) != 0;
  int grid = sm_count();
  if (grid > M / 128) grid = M / 128;
  if (approx) {
    VC_CUDA(cudaFuncSetAttribute(fc_adam_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    fc_adam_kernel<true><<<grid, FA_THREADS, smem, s>>>(a, mb, mp, mm, mv);
  } else {
    VC_CUDA(cudaFuncSetAttribute(fc_adam_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    fc_adam_kernel<false><<<grid, FA_THREADS, smem, s>>>(a, mb, mp, mm, mv);
  }
  VC_LAUNCH_CHECK();
  return VC_OK;
}

}  // namespace vc
