// Implicit-GEMM weight gradients of Conv2d(k3,s2,p1) and ConvTranspose2d(k4,s2,p1) on tcgen05 (sm_100a).
//
// Both are   dW[cp, cq, th, tw] = sum_{b,i,j} P[b,cp,i,j] * Q[b,cq, 2i+th-1, 2j+tw-1]
//   Conv2d  (12:87-89):   P = dy [B,Cout,OH,OW], Q = x  [B,Cin,H,W],    taps 3x3 -> dW [Cout,Cin,3,3]
//   ConvT2d (12:135-137): P = x  [B,Cin,H,W],    Q = dy [B,Cout,2H,2W], taps 4x4 -> dW [Cin,Cout,4,4]
// i.e. the GEMM  C[M = cp, N = (cq,tap)] = A[M, K] * Bq[N, K]^T  whose contraction index K runs over the B*GH*GW pixels.
// Neither operand is materialised: per 64-pixel k-block the eight producer warps build both K-major, 128-byte-swizzled
// tiles in shared memory (lane = pixel pair, so loads are coalesced along W and a warp's stores cover one 128-byte row
// without bank conflicts); one elected thread issues tcgen05.mma into a 128 x BN fp32 accumulator in tensor memory.
// The pixel range is split over the grid (deterministic split-K): CTA (z, nt) reduces k-blocks [z*kbs, (z+1)*kbs) for
// column tile nt and writes its partial; a second kernel folds the partials in a fixed order into the weight layout.
#include <string.h>
#include "common.cuh"
#include "tc05.cuh"

namespace vc {

constexpr int WG_THREADS = 288;               // 8 producer/epilogue warps + 1 MMA warp
constexpr uint32_t WG_A_BYTES = 128 * 128;    // [128 rows][64 bf16]

struct WgArgs {
  const float* P; const float* Q; float* part;
  int B, Cp, GH, GW;       // P [B, Cp, GH, GW]
  int Cq, SH, SW;          // Q [B, Cq, SH, SW]
  int T, taps;             // taps = T*T, source pixel (2i + th - 1, 2j + tw - 1)
  int N, BN;               // N = Cq*taps; BN = columns per tile (multiple of 32)
  int K, nkb, kbs;         // pixels, 64-pixel k-blocks, k-blocks per split
  int stages;
  int gsize;               // producer warps that share one k-block (1, 2, 4 or 8); 8/gsize k-blocks are built at once
  uint32_t idesc;
};

__device__ __forceinline__ uint32_t wg_pack(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// predicated read-only load without a branch region: 0 when !pred (the address is then never dereferenced)
__device__ __forceinline__ float wg_ldg(const float* p, uint32_t pred) {
  float v;
  asm("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %2, 0;\n\tmov.f32 %0, 0f00000000;\n\t@q ld.global.nc.f32 %0, [%1];\n\t}"
      : "=f"(v)
      : "l"(p), "r"(pred));
  return v;
}

// 8-byte variant (the address must be 8-byte aligned when pred != 0)
__device__ __forceinline__ float2 wg_ldg2(const float* p, uint32_t pred) {
  float2 v;
  asm("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %3, 0;\n\tmov.f32 %0, 0f00000000;\n\tmov.f32 %1, 0f00000000;\n\t"
      "@q ld.global.nc.v2.f32 {%0, %1}, [%2];\n\t}"
      : "=f"(v.x), "=f"(v.y)
      : "l"(p), "r"(pred));
  return v;
}

// One real row of the B tile, (cq, tap) of Q: element offset relative to the lane's pixel, the validity bit to test in
// the lane's tap mask, and the row's byte offset in a stage with the row's swizzle phase folded in (bits 4-6), so a
// lane's store address is  soffx ^ ((chunk << 4) | 4*(lane & 3)).
struct __align__(16) WgRow { int off; uint32_t bit; uint32_t soffx; int pad; };

// VEC: both spatial widths are even, so a lane's pixel pair (j, j+1) sits in one row and its T+2 source columns
// 2j-1 .. 2j+T start one float before an 8-byte boundary: a group of T tile rows (one (cq, th), tw = 0..T-1) is fed by
// two 8-byte loads plus one or two predicated scalars instead of 2T scalar loads, and an A row by one 8-byte load.
template <int VT>   // 0: scalar producers; 3 / 4: vector producers for T = 3 / 4
__global__ void __launch_bounds__(WG_THREADS, 1) conv_wgrad_igemm_kernel(const WgArgs a) {
  constexpr bool VEC = VT != 0;
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (tma::s32(smem_raw) & 1023u)) & 1023u);
  const int S = a.stages;
  const uint32_t st_bytes = WG_A_BYTES + (uint32_t)a.BN * 128u;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)S * st_bytes);
  uint64_t* full = bars;            // S: 256 producer arrivals
  uint64_t* empty = bars + S;       // S: tcgen05.commit
  uint64_t* t_full = empty + S;     // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_full + 1);
  WgRow* rowtab = reinterpret_cast<WgRow*>(tmem_slot + 2);  // [Cp + columns of this tile]: one entry per REAL tile row

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int z = blockIdx.x, nt = blockIdx.y;
  const int kb0 = z * a.kbs, kb1 = min(a.nkb, kb0 + a.kbs);
  const int nkb = max(kb1 - kb0, 0);
  const int n_base = nt * a.BN;
  const uint32_t tmem_cols = a.BN <= 32 ? 32u : (a.BN <= 64 ? 64u : (a.BN <= 128 ? 128u : 256u));

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { tma::mbar_init(&full[s], 32 * a.gsize); tma::mbar_init(&empty[s], 1); }
    tma::mbar_init(t_full, 1);
    tma::mbar_fence_init();
  }
  if (warp == 8) tc05::tmem_alloc(tmem_slot, tmem_cols);
  const int ncols = min(a.BN, a.N - n_base);   // real columns of this tile
  if (VEC) {   // one entry per row group (cq, th): source offset of column 2j-1 and the th validity bit
    for (int gi = threadIdx.x; gi < ncols / a.T; gi += WG_THREADS) {
      const int n = n_base + gi * a.T;
      const int cq = n / a.taps, th = (n - cq * a.taps) / a.T;
      WgRow t;
      t.off = cq * a.SH * a.SW + th * a.SW; t.bit = 1u << th; t.soffx = 0; t.pad = 0;
      rowtab[gi] = t;
    }
  }
  for (int nl = threadIdx.x; !VEC && nl < ncols; nl += WG_THREADS) {
    const int n = n_base + nl;
    const int cq = n / a.taps, tap = n - cq * a.taps, th = tap / a.T, tw = tap - th * a.T;
    WgRow t;
    t.off = cq * a.SH * a.SW + th * a.SW + tw; t.bit = 1u << tap;
    t.soffx = WG_A_BYTES + (uint32_t)nl * 128u + (((uint32_t)nl & 7u) << 4); t.pad = 0;
    rowtab[nl] = t;
  }
  // padding rows of both tiles are never written by the producers: zero every stage once
  for (uint32_t o = threadIdx.x * 16u; o < (uint32_t)S * st_bytes; o += WG_THREADS * 16u)
    *reinterpret_cast<uint4*>(smem + o) = make_uint4(0u, 0u, 0u, 0u);
  tma::fence_proxy_async();
  tc05::fence_before();
  __syncthreads();
  tc05::fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 8) {
    // ===================== producers =====================
    // The eight warps form 8/gsize groups; group q builds k-blocks q, q + ngroups, ... on its own, so several ring
    // slots fill concurrently.  Inside a group the real rows of both tiles are dealt round-robin to its warps and
    // processed in "units" of 16 rows x 2 pixels = 32 independent loads per lane; the loads of the next unit are issued
    // before the current one is packed and stored (two register buffers), across k-block boundaries too.
    const int G = a.GH * a.GW;
    const int gsize = a.gsize, ngroups = 8 / gsize;
    const int group = warp / gsize, wig = warp - group * gsize;
    const int unitsA = ((a.Cp + gsize - 1) / gsize + 15) / 16;
    // VEC: a B unit is GPU row groups of T rows (4 x 4 or 5 x 3 rows = 32 / 30 values per lane)
    constexpr int T = VEC ? VT : 1, GPU = VT == 4 ? 4 : 5;
    const int ngrp = ncols / T;
    const int units = unitsA + (VEC ? ((ngrp + gsize - 1) / gsize + GPU - 1) / GPU
                                    : ((ncols + gsize - 1) / gsize + 15) / 16);
    const int nmine = nkb > group ? (nkb - group + ngroups - 1) / ngroups : 0;
    const uint32_t lane_x = (((uint32_t)lane >> 2) << 4) | ((uint32_t)(lane & 3) * 4u);

    // state of the k-block being LOADED: the lane's two pixels (b,i,j), advanced by 64 pixels per k-block
    int pb[2], pi[2], pj[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int64_t m = (int64_t)(kb0 + group) * 64 + 2 * lane + e;
      pb[e] = (int)(m / G);
      const int r = (int)(m - (int64_t)pb[e] * G);
      pi[e] = r / a.GW;
      pj[e] = r - pi[e] * a.GW;
    }
    int l_it = 0;
    const float* pp[2];
    const float* qq[2];
    uint32_t tapmask[2];
    uint32_t vmask = 0, ok_left = 0, ok_right = 0;   // VEC: valid th bits of the pair's row, column 2j-1 / 2j+4 in range
    auto pixel_state = [&]() {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int i = pi[e], j = pj[e];
        pp[e] = a.P + ((int64_t)pb[e] * a.Cp * G + i * a.GW + j);
        qq[e] = a.Q + (((int64_t)pb[e] * a.Cq * a.SH + 2 * i - 1) * a.SW + 2 * j - 1);
        const uint32_t fullT = (1u << a.T) - 1u;
        uint32_t mh = fullT, mw = fullT;
        if (i == 0) mh &= ~1u;
        if (j == 0) mw &= ~1u;
        const int lh = a.SH + 1 - 2 * i, lw = a.SW + 1 - 2 * j;   // taps t >= lh (lw) fall off the bottom (right) edge
        if (lh < a.T) mh &= (1u << max(lh, 0)) - 1u;
        if (lw < a.T) mw &= (1u << max(lw, 0)) - 1u;
        uint32_t msk = 0;
        for (int t = 0; t < a.T; ++t) msk |= ((mh >> t) & 1u) ? (mw << (t * a.T)) : 0u;
        tapmask[e] = pb[e] < a.B ? (msk | (1u << 31)) : 0u;
        if (VEC && e == 0) {
          vmask = pb[0] < a.B ? mh : 0u;
          ok_left = j > 0 ? 1u : 0u;
          ok_right = 2 * j + 4 < a.SW ? 1u : 0u;
        }
      }
    };
    pixel_state();

    auto load_unit = [&](int it, int u, float (&v)[32]) {   // `it` counts this group's own k-blocks
      if (it != l_it) {   // the load cursor entered the group's next k-block: advance both pixels
        l_it = it;
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          pj[e] += 64 * ngroups;
          while (pj[e] >= a.GW) { pj[e] -= a.GW; ++pi[e]; }
          while (pi[e] >= a.GH) { pi[e] -= a.GH; ++pb[e]; }
        }
        pixel_state();
      }
      if (u < unitsA) {   // rows of the A tile: channel cp of P at the lane's pixels
        const uint32_t ok0 = tapmask[0] >> 31, ok1 = tapmask[1] >> 31;
#pragma unroll
        for (int r = 0; r < 16; ++r) {
          const int cp = wig + gsize * (u * 16 + r);
          const uint32_t rok = cp < a.Cp ? 1u : 0u;
          if (VEC) {   // the pair is both in or both out (K is even), consecutive and 8-byte aligned
            const float2 t2 = wg_ldg2(pp[0] + cp * G, ok0 & rok);
            v[2 * r] = t2.x; v[2 * r + 1] = t2.y;
          } else {
            v[2 * r] = wg_ldg(pp[0] + cp * G, ok0 & rok);
            v[2 * r + 1] = wg_ldg(pp[1] + cp * G, ok1 & rok);
          }
        }
      } else if (VEC) {   // row groups of the B tile: (cq, th), all tw at once
#pragma unroll
        for (int r = 0; r < 5; ++r) {
          if (r < GPU) {
            const int gi = wig + gsize * ((u - unitsA) * GPU + r);
            int2 t = make_int2(0, 0);
            if (gi < ngrp) t = *reinterpret_cast<const int2*>(&rowtab[gi]);
            const uint32_t ok = vmask & (uint32_t)t.y;          // row 2i-1+th inside the image (and a real group)
            const float* src = qq[0] + t.x;                      // column 2j-1 of that row
            const float fl = wg_ldg(src, ok ? ok_left : 0u);
            const float2 m0 = wg_ldg2(src + 1, ok), m1 = wg_ldg2(src + 3, ok);
            if (T == 4) {
              const float fr = wg_ldg(src + 5, ok ? ok_right : 0u);
              float* o = v + 8 * r;   // (pixel j, pixel j+1) for tw = 0..3
              o[0] = fl;   o[1] = m0.y;
              o[2] = m0.x; o[3] = m1.x;
              o[4] = m0.y; o[5] = m1.y;
              o[6] = m1.x; o[7] = fr;
            } else {
              float* o = v + 6 * r;   // tw = 0..2
              o[0] = fl;   o[1] = m0.y;
              o[2] = m0.x; o[3] = m1.x;
              o[4] = m0.y; o[5] = m1.y;
            }
          }
        }
      } else {            // rows of the B tile: (cq, tap) of Q
#pragma unroll
        for (int r = 0; r < 16; ++r) {
          const int nl = wig + gsize * ((u - unitsA) * 16 + r);
          int2 t = make_int2(0, 0);
          if (nl < ncols) t = *reinterpret_cast<const int2*>(&rowtab[nl]);
          v[2 * r] = wg_ldg(qq[0] + t.x, tapmask[0] & (uint32_t)t.y);
          v[2 * r + 1] = wg_ldg(qq[1] + t.x, tapmask[1] & (uint32_t)t.y);
        }
      }
    };
    auto store_unit = [&](int c, int u, const float (&v)[32]) {
      const int it = group + c * ngroups;   // position in the CTA's k-block sequence
      const int s = it % S;
      unsigned char* tile = smem + (size_t)s * st_bytes;
      if (u == 0) tma::mbar_wait(&empty[s], ((uint32_t)(it / S) & 1u) ^ 1u);   // first store into this ring slot
      if (u < unitsA) {
#pragma unroll
        for (int r = 0; r < 16; ++r) {
          const uint32_t cp = (uint32_t)(wig + gsize * (u * 16 + r));
          if (cp < (uint32_t)a.Cp)
            *reinterpret_cast<uint32_t*>(tile + ((cp * 128u + ((cp & 7u) << 4)) ^ lane_x)) = wg_pack(v[2 * r], v[2 * r + 1]);
        }
      } else if (VEC) {
#pragma unroll
        for (int r = 0; r < 5; ++r) {
          if (r < GPU) {
            const int gi = wig + gsize * ((u - unitsA) * GPU + r);
            if (gi < ngrp) {
#pragma unroll
              for (int tw = 0; tw < 4; ++tw) {
                if (tw < T) {
                  const uint32_t nl = (uint32_t)(gi * T + tw);
                  const uint32_t so = WG_A_BYTES + nl * 128u + ((nl & 7u) << 4);
                  *reinterpret_cast<uint32_t*>(tile + (so ^ lane_x)) =
                      wg_pack(v[2 * (T * r + tw)], v[2 * (T * r + tw) + 1]);
                }
              }
            }
          }
        }
      } else {
#pragma unroll
        for (int r = 0; r < 16; ++r) {
          const int nl = wig + gsize * ((u - unitsA) * 16 + r);
          if (nl < ncols) *reinterpret_cast<uint32_t*>(tile + (rowtab[nl].soffx ^ lane_x)) = wg_pack(v[2 * r], v[2 * r + 1]);
        }
      }
      if (u == units - 1) {
        tma::fence_proxy_async();
        tma::mbar_arrive(&full[s]);
      }
    };

    {
      float buf0[32], buf1[32];
      const int total = nmine * units;
      int it = 0, u = 0;          // store cursor
      int it2 = 0, u2 = 0;        // load cursor (one unit ahead)
      auto advance = [&](int& ii, int& uu) { if (++uu == units) { uu = 0; ++ii; } };
      if (total > 0) { load_unit(0, 0, buf0); advance(it2, u2); }
      for (int g = 0; g < total; g += 2) {
        if (g + 1 < total) { load_unit(it2, u2, buf1); advance(it2, u2); }
        store_unit(it, u, buf0); advance(it, u);
        if (g + 1 >= total) break;
        if (g + 2 < total) { load_unit(it2, u2, buf0); advance(it2, u2); }
        store_unit(it, u, buf1); advance(it, u);
      }
    }
    // ===================== epilogue: partial tile -> workspace =====================
    const int q = warp & 3, hh = warp >> 2;
    const int cp = q * 32 + lane;
    const int nchunks = a.BN >> 5;
    if (nkb > 0) {
      tma::mbar_wait(t_full, 0);
      tc05::fence_after();
    }
    float* o = a.part + ((int64_t)z * a.Cp + cp) * a.N;
    for (int cc = hh; cc < nchunks; cc += 2) {
      float v[32];
      if (nkb > 0) {
        tc05::ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + cc * 32, v);
        tc05::ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = 0.f;
      }
      if (cp < a.Cp) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int n = n_base + cc * 32 + j;
          if (n < a.N) o[n] = v[j];
        }
      }
    }
  } else {
    // ===================== MMA issuer =====================
    if (tc05::elect_one()) {   // a single lane the compiler knows about: no waterfall loop around UTCHMMA / UTMALDG
      for (int it = 0; it < nkb; ++it) {
        const int s = it % S;
        const uint32_t ph = (uint32_t)(it / S) & 1u;
        tma::mbar_wait_backoff(&full[s], ph, 32);
        tc05::fence_after();
        const uint32_t sa = tma::s32(smem + (size_t)s * st_bytes), sb = sa + WG_A_BYTES;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          tc05::mma_bf16(tmem_base, tc05::desc_k_sw128(sa + kk * 32), tc05::desc_k_sw128(sb + kk * 32), a.idesc,
                         (uint32_t)((it | kk) != 0));
        tc05::commit(&empty[s]);
      }
      if (nkb > 0) tc05::commit(t_full);
    }
  }

  tc05::fence_before();
  __syncthreads();
  if (warp == 8) {
    tc05::fence_after();
    tc05::tmem_dealloc(tmem_base, tmem_cols);
  }
}

// ---- fast producers ------------------------------------------------------------------------------------------------
// Same GEMM, same tiles, same MMA issuer, for the shapes of the model itself: GW % 64 == 0 (a 64-pixel k-block is one
// run of a P row, so batch, row and the valid source rows are warp-uniform), SW == 2 GW, Cp == 8 NA, and a column tile
// of BN = 8 NG T columns that holds whole row groups (cq, th).  All eight producer warps build the same k-block: warp w
// owns A rows cp = w + 8r and row groups gi = w + 8g, so
//   * a row's swizzle phase (row & 7) does not depend on r / g: every shared-memory store is base register + immediate;
//   * a group's source offset and th class are per-warp constants held in registers (no table, no index arithmetic);
//   * rows outside the image are read from the nearest valid row (no predicate on the wide loads) and the packed words
//     are ANDed with a per-k-block mask; only the two scalar edge columns keep a (loop-invariant) predicate.
// A k-block is loaded in two halves into two register buffers; the loads of the next half are in flight while the
// current one is packed and stored.  About 5 instructions per tile row instead of ~35.
template <int T, int NG, int NA>
struct WgBuf { float a[NA]; float q[(NG / 2) * (T + 2)]; uint32_t mk[T == 4 ? 1 : 3]; };

template <int T, int NG, int NA>
__global__ void __launch_bounds__(WG_THREADS, 1) conv_wgrad_fast_kernel(const WgArgs a) {
  constexpr int BN = 8 * NG * T;
  constexpr int NC = T == 4 ? 1 : 3;   // th classes of a warp's groups: th(g) = (Gi0 + w + 8g) % T has period NC in g
  constexpr int V = T + 2;             // source columns 2j-1 .. 2j+T of a pixel pair
  constexpr int NGH = NG / 2, NAH = NA / 2;
  constexpr uint32_t ST_BYTES = WG_A_BYTES + (uint32_t)BN * 128u;
  constexpr uint32_t TMEM_COLS = BN <= 128 ? 128u : 256u;
  static_assert(NG % 2 == 0 && NA % 2 == 0 && BN % 16 == 0 && BN <= 256, "tile shape");
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (tma::s32(smem_raw) & 1023u)) & 1023u);
  const int S = a.stages;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)S * ST_BYTES);
  uint64_t* full = bars;            // S: 256 producer arrivals
  uint64_t* empty = bars + S;       // S: tcgen05.commit
  uint64_t* t_full = empty + S;     // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int z = blockIdx.x, nt = blockIdx.y;
  const int kb0 = z * a.kbs, kb1 = min(a.nkb, kb0 + a.kbs);
  const int nkb = max(kb1 - kb0, 0);
  const int n_base = nt * BN;

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { tma::mbar_init(&full[s], 256); tma::mbar_init(&empty[s], 1); }
    tma::mbar_init(t_full, 1);
    tma::mbar_fence_init();
  }
  if (warp == 8) tc05::tmem_alloc(tmem_slot, TMEM_COLS);
  // A rows >= Cp are never written by the producers: zero every stage once
  for (uint32_t o = threadIdx.x * 16u; o < (uint32_t)S * ST_BYTES; o += WG_THREADS * 16u)
    *reinterpret_cast<uint4*>(smem + o) = make_uint4(0u, 0u, 0u, 0u);
  tma::fence_proxy_async();
  tc05::fence_before();
  __syncthreads();
  tc05::fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 8) {
    // ===================== producers =====================
    const int w = warp;
    const int G = a.GH * a.GW;
    const int Gi0 = nt * (BN / T);   // first row group of this column tile
    int off[NG];                     // source offset of group g relative to (channel 0, row 2i-1)
#pragma unroll
    for (int g = 0; g < NG; ++g) {
      const int gi = Gi0 + w + 8 * g, cq = gi / T, th = gi - cq * T;
      off[g] = (cq * a.SH + th) * a.SW;
    }
    int thc[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) thc[c] = (Gi0 + w + 8 * c) % T;
    const uint32_t lane_x = (((uint32_t)lane >> 2) << 4) | ((uint32_t)(lane & 3) * 4u);
    const uint32_t sA = (uint32_t)w * 128u + ((((uint32_t)w & 7u) << 4) ^ lane_x);   // row w + 8r: + r * 1024
    uint32_t sB[T];                                                                 // row T(w + 8g) + tw: + g * 8T * 128
#pragma unroll
    for (int tw = 0; tw < T; ++tw) {
      const uint32_t nl = (uint32_t)(T * w + tw);
      sB[tw] = WG_A_BYTES + nl * 128u + (((nl & 7u) << 4) ^ lane_x);
    }

    // load cursor: k-block kb0 + c covers pixels (b, i, jb .. jb+63); this lane owns the pair jb + 2 lane, + 1
    int b, i, jb;
    {
      const int64_t m = (int64_t)kb0 * 64;
      b = (int)(m / G);
      const int r = (int)(m - (int64_t)b * G);
      i = r / a.GW;
      jb = r - i * a.GW;
    }
    const float* pA = nullptr;
    const float* qc[NC];
    uint32_t mk[NC];
    bool okl = false, okr = false;
    auto state = [&]() {
      const int j = jb + 2 * lane;
      pA = a.P + (((int64_t)b * a.Cp + w) * G + i * a.GW + j);
      const float* pQ = a.Q + (((int64_t)b * a.Cq * a.SH + (2 * i - 1)) * a.SW + 2 * j - 1);
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        const int row = 2 * i - 1 + thc[c];
        const int rc = min(max(row, 0), a.SH - 1);
        qc[c] = pQ + (rc - row) * a.SW;
        mk[c] = rc == row ? 0xffffffffu : 0u;
      }
      okl = j > 0;
      okr = j + 2 < a.GW;
    };
    auto advance = [&]() {
      jb += 64;
      if (jb >= a.GW) { jb = 0; if (++i >= a.GH) { i = 0; ++b; } }
      state();
    };

    using Buf = WgBuf<T, NG, NA>;
    auto load_half = [&](const int h, Buf& u) {
#pragma unroll
      for (int r = 0; r < NAH; ++r) {
        const float2 t2 = __ldg(reinterpret_cast<const float2*>(pA + (int64_t)((h * NAH + r) * 8) * G));
        u.a[2 * r] = t2.x; u.a[2 * r + 1] = t2.y;
      }
#pragma unroll
      for (int gg = 0; gg < NGH; ++gg) {
        const int g = h * NGH + gg;
        const float* src = qc[g % NC] + off[g];          // column 2j-1 of the group's source row
        float* o = u.q + gg * V;
        o[0] = okl ? __ldg(src) : 0.f;
        const float2 m0 = __ldg(reinterpret_cast<const float2*>(src + 1));
        const float2 m1 = __ldg(reinterpret_cast<const float2*>(src + 3));
        o[1] = m0.x; o[2] = m0.y; o[3] = m1.x; o[4] = m1.y;
        if (T == 4) o[5] = okr ? __ldg(src + 5) : 0.f;
      }
#pragma unroll
      for (int c = 0; c < NC; ++c) u.mk[c] = mk[c];
    };
    auto store_half = [&](const int h, const Buf& u, unsigned char* tile) {
#pragma unroll
      for (int r = 0; r < NAH; ++r)
        *reinterpret_cast<uint32_t*>(tile + sA + (uint32_t)((h * NAH + r) * 1024)) = wg_pack(u.a[2 * r], u.a[2 * r + 1]);
#pragma unroll
      for (int gg = 0; gg < NGH; ++gg) {
        const int g = h * NGH + gg;
        const float* o = u.q + gg * V;
        const uint32_t m = u.mk[g % NC];
        unsigned char* dst = tile + (uint32_t)(g * 8 * T * 128);
        // (pixel j, pixel j+1) of tap tw are source columns (2j-1+tw, 2j+1+tw) = o[tw], o[tw+2]
#pragma unroll
        for (int tw = 0; tw < T; ++tw)
          *reinterpret_cast<uint32_t*>(dst + sB[tw]) = wg_pack(o[tw], o[tw + 2]) & m;
      }
    };

    {
      Buf u0, u1;
      int s = 0;
      uint32_t ph = 1;   // parity to wait for on empty[s]: first pass through the ring never blocks
      if (nkb > 0) { state(); load_half(0, u0); }
      for (int c = 0; c < nkb; ++c) {
        load_half(1, u1);
        unsigned char* tile = smem + (size_t)s * ST_BYTES;
        tma::mbar_wait(&empty[s], ph);
        store_half(0, u0, tile);
        if (c + 1 < nkb) { advance(); load_half(0, u0); }
        store_half(1, u1, tile);
        tma::fence_proxy_async();
        tma::mbar_arrive(&full[s]);
        if (++s == S) { s = 0; ph ^= 1u; }
      }
    }
    // ===================== epilogue: partial tile -> workspace =====================
    const int q = warp & 3, hh = warp >> 2;
    const int cp = q * 32 + lane;
    constexpr int nchunks = (BN + 31) / 32;
    if (nkb > 0) {
      tma::mbar_wait(t_full, 0);
      tc05::fence_after();
    }
    float* o = a.part + ((int64_t)z * a.Cp + cp) * a.N + n_base;
    for (int cc = hh; cc < nchunks; cc += 2) {
      float v[32];
      if (nkb > 0) {
        tc05::ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + cc * 32, v);
        tc05::ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = 0.f;
      }
      if (cp < a.Cp) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (cc * 32 + j < BN) o[cc * 32 + j] = v[j];
      }
    }
  } else {
    // ===================== MMA issuer =====================
    if (tc05::elect_one()) {   // a single lane the compiler knows about: no waterfall loop around UTCHMMA / UTMALDG
      for (int it = 0; it < nkb; ++it) {
        const int s = it % S;
        const uint32_t ph = (uint32_t)(it / S) & 1u;
        tma::mbar_wait_backoff(&full[s], ph, 32);
        tc05::fence_after();
        const uint32_t sa = tma::s32(smem + (size_t)s * ST_BYTES), sb = sa + WG_A_BYTES;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          tc05::mma_bf16(tmem_base, tc05::desc_k_sw128(sa + kk * 32), tc05::desc_k_sw128(sb + kk * 32), a.idesc,
                         (uint32_t)((it | kk) != 0));
        tc05::commit(&empty[s]);
      }
      if (nkb > 0) tc05::commit(t_full);
    }
  }

  tc05::fence_before();
  __syncthreads();
  if (warp == 8) {
    tc05::fence_after();
    tc05::tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// split-K partials -> dw.  A block owns 32 consecutive outputs; its eight warps each add every eighth partial and the
// eight sub-sums are added in warp order (a fixed order: the gradient is bit-identical from run to run)
__global__ void __launch_bounds__(256) wgrad_fold_kernel(const float* __restrict__ part, int nsplit, int total,
                                                         float* __restrict__ out) {
  __shared__ float sub[8][32];
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int e = blockIdx.x * 32 + lane;
  float acc = 0.f;
  if (e < total)
    for (int zz = g; zz < nsplit; zz += 8) acc += part[(int64_t)zz * total + e];
  sub[g][lane] = acc;
  __syncthreads();
  if (g != 0 || e >= total) return;
#pragma unroll
  for (int i = 1; i < 8; ++i) acc += sub[i][lane];
  out[e] = acc;
}

// ---- host ---------------------------------------------------------------------------------------------------------
struct WgPlan { int Cp, Cq, T, GH, GW, SH, SW, N, ntiles, BN, K, nkb, nsplit, kbs; bool ok, fast; };

static WgPlan wg_plan(int op, int B, int Cin, int H, int W, int Cout, bool allow_fast = true) {
  WgPlan p;
  memset(&p, 0, sizeof(p));
  if (op == VC_OP_CONV_WGRAD) {
    p.Cp = Cout; p.Cq = Cin; p.T = 3; p.GH = (H - 1) / 2 + 1; p.GW = (W - 1) / 2 + 1; p.SH = H; p.SW = W;
  } else if (op == VC_OP_CONVT_WGRAD) {
    p.Cp = Cin; p.Cq = Cout; p.T = 4; p.GH = H; p.GW = W; p.SH = 2 * H; p.SW = 2 * W;
  } else {
    return p;
  }
  p.N = p.Cq * p.T * p.T;
  p.ntiles = (p.N + 255) / 256;
  p.BN = ((p.N + p.ntiles - 1) / p.ntiles + 31) / 32 * 32;
  // fast producers (conv_wgrad_fast_kernel): k-blocks are runs of one P row, tiles of whole row groups
  if (allow_fast && p.GW % 64 == 0 && p.SW == 2 * p.GW && (p.Cp == 64 || p.Cp == 128)) {
    const int bn = p.T == 4 ? 256 : (p.N % 192 == 0 ? 192 : 144);
    if (p.N % bn == 0) { p.fast = true; p.BN = bn; p.ntiles = p.N / bn; }
  }
  // (T = 3 tiles that hold whole row groups would need BN % 96 == 0: measured slower -- unbalanced tiles -- so the
  // vector producers below are used for T = 4 only, where BN % 4 == 0 always holds)
  const int64_t K = (int64_t)B * p.GH * p.GW;
  p.ok = p.Cp <= 128 && K < 2147483647LL - 64 && (int64_t)B * (Cin > Cout ? Cin : Cout) * 4 * H * W < 2147483647LL;
  if (!p.ok) return p;
  p.K = (int)K;
  p.nkb = (p.K + 63) / 64;
  int ns = sm_count() / p.ntiles;
  if (ns < 1) ns = 1;
  if (ns > p.nkb) ns = p.nkb;
  p.kbs = (p.nkb + ns - 1) / ns;
  p.nsplit = (p.nkb + p.kbs - 1) / p.kbs;
  return p;
}

bool conv_wgrad_igemm_supported(int op, int B, int Cin, int H, int W, int Cout) {
  return wg_plan(op, B, Cin, H, W, Cout).ok;
}

size_t conv_wgrad_igemm_workspace(int op, int B, int Cin, int H, int W, int Cout) {
  const WgPlan p = wg_plan(op, B, Cin, H, W, Cout);
  if (!p.ok) return 0;
  const WgPlan q = wg_plan(op, B, Cin, H, W, Cout, false);   // a misaligned operand falls back to the general kernel
  const int ns = p.nsplit > q.nsplit ? p.nsplit : q.nsplit;
  return (size_t)ns * p.Cp * p.N * 4 + 1024;
}

// in0 = dy, in1 = x (vc_dense_op convention for the weight-gradient ops)
int conv_wgrad_igemm_run(int op, const float* dy, const float* x, float* dw, int B, int Cin, int H, int W, int Cout,
                         void* ws, size_t wsb, cudaStream_t s) {
  const float* P0 = op == VC_OP_CONV_WGRAD ? dy : x;
  const float* Q0 = op == VC_OP_CONV_WGRAD ? x : dy;
  const bool aligned8 = ((reinterpret_cast<uintptr_t>(P0) | reinterpret_cast<uintptr_t>(Q0)) & 7) == 0;
  const WgPlan p = wg_plan(op, B, Cin, H, W, Cout, aligned8);
  VC_REQUIRE(p.ok, VC_ERR_UNSUPPORTED, "conv_wgrad_igemm: unsupported shape");
  VC_REQUIRE(ws && wsb >= conv_wgrad_igemm_workspace(op, B, Cin, H, W, Cout), VC_ERR_WORKSPACE,
             "conv_wgrad_igemm: workspace too small");
  WgArgs a;
  memset(&a, 0, sizeof(a));
  a.P = op == VC_OP_CONV_WGRAD ? dy : x;
  a.Q = op == VC_OP_CONV_WGRAD ? x : dy;
  a.part = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(ws) + 255) & ~(uintptr_t)255);
  a.B = B; a.Cp = p.Cp; a.GH = p.GH; a.GW = p.GW; a.Cq = p.Cq; a.SH = p.SH; a.SW = p.SW;
  a.T = p.T; a.taps = p.T * p.T; a.N = p.N; a.BN = p.BN; a.K = p.K; a.nkb = p.nkb; a.kbs = p.kbs;
  const size_t st_bytes = WG_A_BYTES + (size_t)p.BN * 128;
  int stages = (int)((200 * 1024) / st_bytes);
  if (stages > 10) stages = 10;
  if (stages < 2) stages = 2;
  a.stages = stages;
  {  // smallest group that keeps a warp's share of a k-block at <= 80 rows (5 full units), if the ring is deep enough
    const int ncols = p.N < p.BN ? p.N : p.BN;
    int gs = 1;
    while (gs < 8 && ((p.Cp + gs - 1) / gs + (ncols + gs - 1) / gs > 80 || 8 / gs > stages - 1)) gs *= 2;
    a.gsize = gs;
  }
  a.idesc = tc05::idesc(tc05::kFmtBF16, 128, p.BN, 0, 0);
  if (p.fast) {
    void (*fk)(const WgArgs) = nullptr;
    if (p.T == 4) fk = p.Cp == 64 ? conv_wgrad_fast_kernel<4, 8, 8> : conv_wgrad_fast_kernel<4, 8, 16>;
    else if (p.BN == 192) fk = p.Cp == 64 ? conv_wgrad_fast_kernel<3, 8, 8> : conv_wgrad_fast_kernel<3, 8, 16>;
    else fk = p.Cp == 64 ? conv_wgrad_fast_kernel<3, 6, 8> : conv_wgrad_fast_kernel<3, 6, 16>;
    const size_t fsmem = (size_t)stages * st_bytes + (2 * stages + 1) * 8 + 16 + 1024;
    VC_CUDA(cudaFuncSetAttribute(fk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsmem));
    fk<<<dim3((unsigned)p.nsplit, (unsigned)p.ntiles), WG_THREADS, fsmem, s>>>(a);
    VC_LAUNCH_CHECK();
    const int tot = p.Cp * p.N;
    wgrad_fold_kernel<<<(tot + 31) / 32, 256, 0, s>>>(a.part, p.nsplit, tot, dw);
    VC_LAUNCH_CHECK();
    return VC_OK;
  }
  const size_t smem = (size_t)stages * st_bytes + (2 * stages + 1) * 8 + 16 + (size_t)p.BN * 16 + 1024;
  // vector producers: even widths (pixel pairs never straddle a row, 8-byte aligned sources), whole row groups per tile
  const bool vec = p.T == 4 && p.GW % 2 == 0 && p.SW % 2 == 0 && p.BN % p.T == 0 && p.N % p.T == 0 &&
                   ((reinterpret_cast<uintptr_t>(a.P) | reinterpret_cast<uintptr_t>(a.Q)) & 7) == 0;
  auto kern = !vec ? conv_wgrad_igemm_kernel<0> : (p.T == 4 ? conv_wgrad_igemm_kernel<4> : conv_wgrad_igemm_kernel<3>);
  VC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((unsigned)p.nsplit, (unsigned)p.ntiles);
  kern<<<grid, WG_THREADS, smem, s>>>(a);
  VC_LAUNCH_CHECK();
  const int total = p.Cp * p.N;
  wgrad_fold_kernel<<<(total + 31) / 32, 256, 0, s>>>(a.part, p.nsplit, total, dw);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

}  // namespace vc
