// TMA-fed implicit-GEMM convolutions on bf16 NHWC activations: the four operators of the script-12 VAE's hidden conv
// layers (12:87-89,135-137) -- Conv2d k3 s2 p1 forward, ConvTranspose2d k4 s2 p1 forward (phase-merged), and their two
// data gradients -- as ONE kernel whose A operand is fetched by 4-D tensor-map loads: a k-block of the A tile is 128
// consecutive pixels of one image row x 64 channels of one tap, channel-contiguous 128-byte rows under the 128-byte
// swizzle, padding zero-filled by the copy engine, stride-2 sampling done by the map's traversal stride.  There are no
// producer warps, no bounds checks and no conversions in the kernel.  The epilogue fuses what used to be separate
// passes over fp32 NCHW tensors: bias + ReLU (forward) or the ReLU mask of the PREVIOUS layer read from its bf16 NHWC
// activation (data gradient), and writes bf16 NHWC for the next TMA-fed kernel and / or fp32 NCHW for the layers that
// still want it (the single-channel ends, the fully connected layers).
//
// 32-channel sources fill only half a 128-byte row; they are viewed as [B][H][W/2][2 x 32] (a row = two neighbouring
// pixels = two taps of a stride-2 gather) with zeros in the weight slab where a (pixel, tap) pair does not occur.
// Weight slabs that fit (<= 150 KB) are loaded once per CTA and stay resident in shared memory; larger ones stream
// through the ring with the A boxes.
//
// warps 0, 10, 11, 12: TMA producers (k-block `it` belongs to producer it % 4), warp 1: tcgen05.mma issuer (2 accumulator
// stages in TMEM), warps 2..9: epilogue.
#include <stdlib.h>
#include <string.h>
#include "common.cuh"
#include "tc05.cuh"

namespace vc {

constexpr int NH_BM = 128;
constexpr int NH_EPI = 256;                      // epilogue threads of the one-CTA-per-SM variant (8 warps); the two-CTA
                                                 // variant (EPW = 4) has 128
constexpr int NH_PROD = 4;                       // TMA-issuing warps: one thread sustains only ~1 box per 460 clocks
constexpr int NH_THREADS = 64 + NH_EPI + 32 * (NH_PROD - 1);   // (profiles/tma_rate.cu), four warps 4x that
constexpr uint32_t NH_A_BYTES = NH_BM * 128;

struct NhArgs {
  // A-tile geometry: tap t of k-block kb sits at source pixel (S*i + th + D0, S*j + tw + D0)
  int pair, nblk;            // pair view of a 32-channel source: nblk boxes per tap row (pairs j-1, j[, j+1])
  int cchunks, TW, S, D0;
  int B, GH, GW, tilesW;     // GEMM rows = grid pixels (b, i, j); a tile = 128 consecutive j
  int N, BN, OC, OH, OW, quad;
  int tiles, nkb, stages, wres;
  int G;                     // k-blocks per ring stage (one barrier round trip per G k-blocks)
  int nprod;                 // active producer warps; divides `stages` (see nh_plan)
  long long* trace;          // experiments only (vc_conv_nhwc_debug_trace): clock64 stamps of CTA 0
  int dbg;                   // VAECLUST_NH_DEBUG (experiments only): 1 = every A box from image 0 / row 0, 4 = no epilogue traffic, 8 = no MMAs, 16 = no TMA loads
  uint32_t idesc;
  // epilogue
  const float* bias;         // [OC] or null
  int relu;
  const __nv_bfloat16* mask; // [B][OH][OW][OC] or null: out = acc * (mask > 0)
  __nv_bfloat16* out_nhwc;   // [B][OH][OW][OC] or null
  float* out_nchw;           // [B][OC][OH][OW] or null
};

__device__ __forceinline__ void nh_coords(const NhArgs& a, int kb, int i, int j0, int& c, int& w, int& h) {
  if (!a.pair) {
    const int tap = kb / a.cchunks;
    const int th = tap / a.TW, tw = tap - th * a.TW;
    c = (kb - tap * a.cchunks) * 64; w = a.S * j0 + tw + a.D0; h = a.S * i + th + a.D0;
  } else {   // pair coordinates along W: pairs j0-1, j0 (, j0+1) of source row 2i + th - 1
    const int th = kb / a.nblk;
    c = 0; w = j0 + (kb - th * a.nblk) - 1; h = 2 * i + th - 1;
  }
}

__device__ __forceinline__ uint32_t nh_pack(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float nh_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float nh_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

// weight slab for the pair views, k = kb*64 + half*32 + c, kb = th*nblk + blk.
//   mode 0: Conv forward of a 32-channel input  (nblk = 2): (blk, half) = (0,1) -> kw 0, (1,0) -> kw 1, (1,1) -> kw 2
//   mode 1: ConvT data gradient of a 32-channel dy (nblk = 3): (0,1) -> 0, (1,0) -> 1, (1,1) -> 2, (2,0) -> 3
__global__ void nh_weights_pair_kernel(const float* __restrict__ w, int mode, int Cin, int Cout, int BN, int Kp,
                                       __nv_bfloat16* __restrict__ out) {
  const int total = BN * Kp;
  const int nblk = mode == 0 ? 2 : 3, TH = mode == 0 ? 3 : 4;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    const int n = e / Kp, k = e - n * Kp;
    const int kb = k >> 6, half = (k >> 5) & 1, c = k & 31;
    const int th = kb / nblk, blk = kb - th * nblk;
    const int tw = blk == 0 ? (half ? 0 : -1) : (blk == 1 ? (half ? 2 : 1) : (half ? -1 : 3));
    float v = 0.f;
    if (th < TH && tw >= 0) {
      if (mode == 0) { if (n < Cout && c < Cin) v = w[(((int64_t)n * Cin + c) * 3 + th) * 3 + tw]; }   // W[co, ci, kh, kw]
      else { if (n < Cin && c < Cout) v = w[((int64_t)n * Cout + c) * 16 + th * 4 + tw]; }            // W[ci, co, kh, kw]
    }
    out[e] = __float2bfloat16_rn(v);
  }
}

// EPW = epilogue warps (8: one CTA per SM; 4: two CTAs per SM, for operators with at most 128 GEMM columns -- two MMA issuers
// and two epilogues per SM overlap the issue latency of small-N MMAs and the per-tile accumulator hand-over)
template <int EPW, int OCC>
__global__ void __launch_bounds__(64 + 32 * EPW + 32 * (NH_PROD - 1), OCC)
conv_nhwc_kernel(const NhArgs a, const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapW) {
  constexpr int NHALF = EPW / 4;                 // warps per TMEM lane quarter = column chunks processed side by side
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (tma::s32(smem_raw) & 1023u)) & 1023u);
  const int S = a.stages;
  const uint32_t b_bytes = (uint32_t)a.BN * 128u;
  const uint32_t kb_bytes = NH_A_BYTES + (a.wres ? 0u : b_bytes);     // one k-block: A box (+ its weight rows)
  const uint32_t st_bytes = (uint32_t)a.G * kb_bytes;
  unsigned char* slab = smem;                                            // resident weight slab: nkb x [BN][64] bf16
  unsigned char* ring = smem + (a.wres ? (size_t)a.nkb * b_bytes : 0);
  uint64_t* bars = reinterpret_cast<uint64_t*>(ring + (size_t)S * st_bytes);
  uint64_t* full = bars;             // S: TMA transaction bytes
  uint64_t* empty = bars + S;        // S: tcgen05.commit
  uint64_t* t_full = empty + S;      // 2: accumulator stage complete
  uint64_t* t_empty = t_full + 2;    // 2: accumulator stage drained by the epilogue threads
  uint64_t* w_full = t_empty + 2;    // 1: resident slab has landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_full + 1);
  // per-k-block box offsets (c, dw, dh): the hot loops contain no integer division (a runtime division is ~150 clocks of
  // dependent latency for a one-warp loop: it was most of the ~500 clocks per k-block the bare pipeline used to cost)
  int* kb_c = reinterpret_cast<int*>(tmem_slot + 2);
  int* kb_w = kb_c + 32;
  int* kb_h = kb_w + 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t tmem_cols = a.BN <= 32 ? 64u : (a.BN <= 64 ? 128u : (a.BN <= 128 ? 256u : 512u));  // 2 stages

  if (threadIdx.x == 0) {
    tma::prefetch_map(&mapA);
    tma::prefetch_map(&mapW);
    for (int s = 0; s < S; ++s) { tma::mbar_init(&full[s], 1); tma::mbar_init(&empty[s], 1); }
    for (int s = 0; s < 2; ++s) { tma::mbar_init(&t_full[s], 1); tma::mbar_init(&t_empty[s], 32 * EPW); }
    tma::mbar_init(w_full, 1);
    tma::mbar_fence_init();
  }
  if (warp == 1) tc05::tmem_alloc(tmem_slot, tmem_cols);
  if (threadIdx.x >= 64 && (int)threadIdx.x - 64 < a.nkb) {
    const int kb = (int)threadIdx.x - 64;
    int c, w, h;
    nh_coords(a, kb, 0, 0, c, w, h);     // offsets relative to (S*i, S*j0)  [pair: (2i, j0)]
    kb_c[kb] = c; kb_w[kb] = w; kb_h[kb] = h;
  }
  tc05::fence_before();
  __syncthreads();
  tc05::fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int my_tiles = (a.tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 0 || warp >= 2 + EPW) {
    // ===================== TMA producers =====================
    // whole warp, uniform control flow, one elected lane issues (see the MMA issuer below for why)
    {
      const uint32_t me = warp == 0 ? 0u : (uint32_t)(warp - (1 + EPW));
      if (a.wres && me == 0) {
        if (tc05::elect_one()) {
          tma::mbar_expect_tx(w_full, (uint32_t)a.nkb * b_bytes);
          for (int kb = 0; kb < a.nkb; ++kb) tma::load_2d(slab + (size_t)kb * b_bytes, &mapW, w_full, kb * 64, 0);
        }
        __syncwarp();
      }
      // this producer owns ring stages me, me + 4, ... of the CTA's flat (tile, k-group) sequence -- a stage holds G
      // k-blocks; ring slot and phase advance by additions only
      const int wS = a.pair ? 1 : a.S, hS = a.pair ? 2 : a.S;
      const int ngr = a.nkb / a.G;                          // k-groups per tile
      const int NP = a.nprod;
      uint32_t it = me;
      int s = (int)me % S, ph = ((int)me / S) & 1;         // (once, outside the loop)
      int tl = (int)me < NP ? 0 : my_tiles, kg = (int)me;
      while (kg >= ngr) { kg -= ngr; ++tl; }
      while (tl < my_tiles) {
        const int t = (int)blockIdx.x + tl * (int)gridDim.x;
        const int tw = t % a.tilesW, r = t / a.tilesW;       // per tile only
        const int i = r % a.GH, b = r / a.GH, j0 = tw * NH_BM;
        for (; kg < ngr; kg += NP, it += NP) {
          const bool tr = a.trace && blockIdx.x == 0 && it < 64 && lane == 0;
          if (tr) a.trace[it * 8 + 0] = clock64();
          tma::mbar_wait_backoff(&empty[s], (uint32_t)ph ^ 1u, 64);
          if (tr) a.trace[it * 8 + 1] = clock64();
          unsigned char* tile = ring + (size_t)s * st_bytes;
          if (a.dbg & 16) {
            if (tc05::elect_one()) tma::mbar_arrive(&full[s]);
          } else if (tc05::elect_one()) {
            tma::mbar_expect_tx(&full[s], st_bytes);
            for (int g = 0; g < a.G; ++g) {
              const int kb = kg * a.G + g;
              const int c = kb_c[kb], w = wS * j0 + kb_w[kb], h = hS * i + kb_h[kb];
              if (a.dbg & 1) tma::load_4d(tile + g * kb_bytes, &mapA, &full[s], c, w & 255, h & 3, 0);
              else tma::load_4d(tile + g * kb_bytes, &mapA, &full[s], c, w, h, b);
              if (!a.wres) tma::load_2d(tile + g * kb_bytes + NH_A_BYTES, &mapW, &full[s], kb * 64, 0);
            }
          }
          __syncwarp();
          if (tr) a.trace[it * 8 + 2] = clock64();
          s += NP;
          while (s >= S) { s -= S; ph ^= 1; }
        }
        kg -= ngr;
        ++tl;
        while (kg >= ngr) { kg -= ngr; ++tl; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // The WHOLE warp walks the loop (uniform control flow, warp-uniform values) and one ELECTED lane issues: issued from a
    // `lane == 0` branch the compiler cannot prove a single active lane and wraps every tcgen05.mma / commit in an
    // ELECT ... BRA.U.ANY waterfall loop, which cost ~670 clocks per k-block (profiles/trace_nhwc.py).
    {
      if (a.wres) tma::mbar_wait_backoff(w_full, 0, 64);
      uint32_t it = 0;
      int s = 0;
      uint32_t ph = 0;
      for (int tl = 0; tl < my_tiles; ++tl) {
        const int as = tl & 1;
        const uint32_t aph = (uint32_t)(tl >> 1) & 1u;
        tma::mbar_wait_backoff(&t_empty[as], aph ^ 1u, 64);
        tc05::fence_after();
        const uint32_t tacc = tmem_base + (uint32_t)(as * a.BN);
        for (int kb = 0; kb < a.nkb; kb += a.G, ++it) {
          const bool tr = a.trace && blockIdx.x == 0 && it < 64 && lane == 0;
          if (tr) a.trace[it * 8 + 3] = clock64();
          tma::mbar_wait_backoff(&full[s], ph, 32);
          if (tr) a.trace[it * 8 + 4] = clock64();
          tc05::fence_after();
          const uint32_t st0 = tma::s32(ring + (size_t)s * st_bytes);
          if (tc05::elect_one()) {
            if (!(a.dbg & 8)) {
              for (int g = 0; g < a.G; ++g) {
                const uint32_t sa = st0 + (uint32_t)g * kb_bytes;
                const uint32_t sb = a.wres ? tma::s32(slab + (size_t)(kb + g) * b_bytes) : sa + NH_A_BYTES;
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
                  tc05::mma_bf16(tacc, tc05::desc_k_sw128(sa + kk * 32), tc05::desc_k_sw128(sb + kk * 32), a.idesc,
                                 (uint32_t)((kb | g | kk) != 0));
              }
            }
            tc05::commit(&empty[s]);
          }
          __syncwarp();
          if (tr) a.trace[it * 8 + 5] = clock64();
          if (++s == S) { s = 0; ph ^= 1u; }
        }
        if (tc05::elect_one()) tc05::commit(&t_full[as]);
        __syncwarp();
      }
    }
  } else {
    // ===================== epilogue: warps 2..9 =====================
    const int ew = warp - 2;
    const int q = warp & 3;          // TMEM lane quarter this warp may read
    const int hh = ew >> 2;          // even / odd 32-column chunks
    const int row = q * 32 + lane;
    const int nchunks = a.BN >> 5;
    const int64_t plane = (int64_t)a.OH * a.OW;
    for (int tl = 0; tl < my_tiles; ++tl) {
      const int as = tl & 1;
      const uint32_t aph = (uint32_t)(tl >> 1) & 1u;
      const int t = (int)blockIdx.x + tl * (int)gridDim.x;
      const int twi = t % a.tilesW, r = t / a.tilesW;
      const int pi = r % a.GH, pb = r / a.GH, pj = twi * NH_BM + row;
      // output pixel(s) of this thread: one (non-quad) or the 2x2 block (quad)
      const int64_t px0 = a.quad ? ((int64_t)pb * a.OH + 2 * pi) * a.OW + 2 * pj : ((int64_t)pb * a.OH + pi) * a.OW + pj;
      // The ReLU mask of the layer below does not depend on the accumulator: fetch it for this thread's first two chunks
      // BEFORE waiting for the main loop, so the ~1 us of memory latency hides under the MMAs of this tile (taken after the
      // wait it was serial per chunk and made the phase-merged data gradients epilogue-bound: 144 vs 45 us).
      uint4 mk[2][4];
      if (a.mask && !(a.dbg & 64)) {
#pragma unroll
        for (int ci = 0; ci < 2; ++ci) {
          const int cc = hh + NHALF * ci;
          if (cc < nchunks) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int64_t off = a.quad ? (px0 + (u >> 1) * a.OW + (u & 1)) * a.OC + cc * 8 : px0 * a.OC + cc * 32 + u * 8;
              mk[ci][u] = __ldg(reinterpret_cast<const uint4*>(a.mask + off));
            }
          }
        }
      }
      // one warp polls the accumulator barrier; the other seven park on a hardware named barrier (no issue slots)
      if (ew == 0) tma::mbar_wait_backoff(&t_full[as], aph, 128);
      asm volatile("bar.sync 1, %0;" ::"n"(32 * EPW) : "memory");
      tc05::fence_after();
      const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * a.BN);
#pragma unroll
      for (int ci = 0; ci < 4; ++ci) {
        const int cc = hh + NHALF * ci;
        if (cc < nchunks) {
          float v[32];
          tc05::ld_32x32(tacc + cc * 32, v);
          tc05::ld_wait();
          if (cc + NHALF >= nchunks) {  // last chunk of this thread: the accumulator stage is free once it sits in registers
            tc05::fence_before();
            tma::mbar_arrive(&t_empty[as]);
          }
          if (a.dbg & 4) continue;
          uint4 m4[4] = {};
          if (a.mask && !(a.dbg & 64)) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              if (ci < 2) {
                m4[u] = mk[ci][u];
              } else {
                const int64_t off = a.quad ? (px0 + (u >> 1) * a.OW + (u & 1)) * a.OC + cc * 8 : px0 * a.OC + cc * 32 + u * 8;
                m4[u] = __ldg(reinterpret_cast<const uint4*>(a.mask + off));
              }
            }
          }
          if (!a.quad) {
            const int c0 = cc * 32;
            if (a.bias) {
              const float4* b4 = reinterpret_cast<const float4*>(a.bias + c0);
#pragma unroll
              for (int j4 = 0; j4 < 8; ++j4) {
                const float4 bb = __ldg(b4 + j4);
                v[4 * j4] += bb.x; v[4 * j4 + 1] += bb.y; v[4 * j4 + 2] += bb.z; v[4 * j4 + 3] += bb.w;
              }
            }
            if (a.relu) {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
            }
            if (a.mask) {
#pragma unroll
              for (int j8 = 0; j8 < 4; ++j8) {
                const uint32_t mw[4] = {m4[j8].x, m4[j8].y, m4[j8].z, m4[j8].w};
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                  v[8 * j8 + 2 * u] = nh_lo(mw[u]) > 0.f ? v[8 * j8 + 2 * u] : 0.f;
                  v[8 * j8 + 2 * u + 1] = nh_hi(mw[u]) > 0.f ? v[8 * j8 + 2 * u + 1] : 0.f;
                }
              }
            }
            if (a.out_nhwc && !(a.dbg & 32)) {
              uint4* o4 = reinterpret_cast<uint4*>(a.out_nhwc + px0 * a.OC + c0);
#pragma unroll
              for (int j8 = 0; j8 < 4; ++j8)
                o4[j8] = make_uint4(nh_pack(v[8 * j8], v[8 * j8 + 1]), nh_pack(v[8 * j8 + 2], v[8 * j8 + 3]),
                                    nh_pack(v[8 * j8 + 4], v[8 * j8 + 5]), nh_pack(v[8 * j8 + 6], v[8 * j8 + 7]));
            }
            if (a.out_nchw) {
              float* o = a.out_nchw + ((int64_t)pb * a.OC + c0) * plane + (int64_t)pi * a.OW + pj;
#pragma unroll
              for (int j = 0; j < 32; ++j) o[j * plane] = v[j];
            }
          } else {
            const int c0 = cc * 8;
            const int oh = 2 * pi, ow = 2 * pj;
            if (a.bias) {
              const float4* b4 = reinterpret_cast<const float4*>(a.bias + c0);
              const float4 b0 = __ldg(b4), b1 = __ldg(b4 + 1);
              const float bj[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
              for (int g = 0; g < 8; ++g) {
                v[4 * g] += bj[g]; v[4 * g + 1] += bj[g]; v[4 * g + 2] += bj[g]; v[4 * g + 3] += bj[g];
              }
            }
            if (a.relu) {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
            }
#pragma unroll
            for (int p = 0; p < 4; ++p) {
              const int64_t px = px0 + (p >> 1) * a.OW + (p & 1);
              if (a.mask) {
                const uint32_t mw[4] = {m4[p].x, m4[p].y, m4[p].z, m4[p].w};
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                  v[4 * (2 * u) + p] = nh_lo(mw[u]) > 0.f ? v[4 * (2 * u) + p] : 0.f;
                  v[4 * (2 * u + 1) + p] = nh_hi(mw[u]) > 0.f ? v[4 * (2 * u + 1) + p] : 0.f;
                }
              }
              if (a.out_nhwc && !(a.dbg & 32))
                *reinterpret_cast<uint4*>(a.out_nhwc + px * a.OC + c0) =
                    make_uint4(nh_pack(v[p], v[4 + p]), nh_pack(v[8 + p], v[12 + p]), nh_pack(v[16 + p], v[20 + p]),
                               nh_pack(v[24 + p], v[28 + p]));
            }
            if (a.out_nchw) {
              float* o = a.out_nchw + ((int64_t)pb * a.OC + c0) * plane + (int64_t)oh * a.OW + ow;
#pragma unroll
              for (int g = 0; g < 8; ++g) {
                float* r0 = o + g * plane;
                *reinterpret_cast<float2*>(r0) = make_float2(v[4 * g], v[4 * g + 1]);
                *reinterpret_cast<float2*>(r0 + a.OW) = make_float2(v[4 * g + 2], v[4 * g + 3]);
              }
            }
          }
        }
      }
      if (nchunks <= hh) {  // this thread has no chunk (BN == 32, odd half): still release the stage
        tc05::fence_before();
        tma::mbar_arrive(&t_empty[as]);
      }
    }
  }

  tc05::fence_before();
  __syncthreads();
  if (warp == 1) {
    tc05::fence_after();
    tc05::tmem_dealloc(tmem_base, tmem_cols);
  }
}

// ---- host ---------------------------------------------------------------------------------------------------------
typedef CUresult (*NhEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// 4-D bf16 map over [B][SH][SWv][Cv] (innermost first: Cv, SWv, SH, B): box = 64 channels x boxW pixels, traversal
// stride along W
int nh_make_map(CUtensorMap* m, const void* base, int Cv, int SWv, int SH, int B, int boxC, int boxW, int strideW) {
  NhEncodeFn fn = reinterpret_cast<NhEncodeFn>(tma::encode_fn());
  if (!fn) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return VC_ERR_CUDA; }
  cuuint64_t gd[4] = {(cuuint64_t)Cv, (cuuint64_t)SWv, (cuuint64_t)SH, (cuuint64_t)B};
  cuuint64_t gs[3] = {(cuuint64_t)Cv * 2, (cuuint64_t)SWv * Cv * 2, (cuuint64_t)SH * SWv * Cv * 2};
  cuuint32_t bx[4] = {(cuuint32_t)boxC, (cuuint32_t)boxW, 1u, 1u};
  cuuint32_t es[4] = {1u, (cuuint32_t)strideW, 1u, 1u};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), gd, gs, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (4-D, box %d x %d, stride %d) failed with CUresult %d", boxC, boxW, strideW, (int)r);
    return VC_ERR_UNSUPPORTED;
  }
  return VC_OK;
}

struct NhPlan {
  bool ok;
  int wmode, pair, nblk, C, cchunks, TW, S, D0, taps, N, BN, nkb, Kp, GH, GW, SH, SW, OC, OH, OW, quad, wres, stages, G, nprod, occ2;
  size_t smem;
};

static NhPlan nh_plan(int op, int B, int Cin, int H, int W, int Cout) {
  NhPlan p;
  memset(&p, 0, sizeof(p));
  if (B <= 0 || H <= 0 || W <= 0) return p;
  if ((op == VC_OP_CONV_FWD || op == VC_OP_CONV_DGRAD) && (H % 2 || W % 2)) return p;
  const int OHc = H / 2, OWc = W / 2;
  switch (op) {
    case VC_OP_CONV_FWD:     // x [B,H,W,Cin] -> y [B,H/2,W/2,Cout]
      p.wmode = 0; p.C = Cin; p.TW = 3; p.taps = 9; p.S = 2; p.D0 = -1; p.N = Cout; p.GH = OHc; p.GW = OWc;
      p.SH = H; p.SW = W; p.OC = Cout; p.OH = OHc; p.OW = OWc; p.quad = 0;
      if (Cin == 32) { p.pair = 1; p.nblk = 2; }
      break;
    case VC_OP_CONVT_FWD:    // x [B,H,W,Cin] -> y [B,2H,2W,Cout], columns (co, 2x2 parity)
      p.wmode = 2; p.C = Cin; p.TW = 3; p.taps = 9; p.S = 1; p.D0 = -1; p.N = 4 * Cout; p.GH = H; p.GW = W;
      p.SH = H; p.SW = W; p.OC = Cout; p.OH = 2 * H; p.OW = 2 * W; p.quad = 1;
      break;
    case VC_OP_CONV_DGRAD:   // g [B,H/2,W/2,Cout] -> dx [B,H,W,Cin], columns (ci, 2x2 parity)
      p.wmode = 1; p.C = Cout; p.TW = 2; p.taps = 4; p.S = 1; p.D0 = 0; p.N = 4 * Cin; p.GH = OHc; p.GW = OWc;
      p.SH = OHc; p.SW = OWc; p.OC = Cin; p.OH = H; p.OW = W; p.quad = 1;
      break;
    case VC_OP_CONVT_DGRAD:  // g [B,2H,2W,Cout] -> dx [B,H,W,Cin]
      p.wmode = 3; p.C = Cout; p.TW = 4; p.taps = 16; p.S = 2; p.D0 = -1; p.N = Cin; p.GH = H; p.GW = W;
      p.SH = 2 * H; p.SW = 2 * W; p.OC = Cin; p.OH = H; p.OW = W; p.quad = 0;
      if (Cout == 32) { p.pair = 1; p.nblk = 3; }
      break;
    default: return p;
  }
  if (p.pair) { p.cchunks = 1; p.nkb = (p.taps / p.TW) * p.nblk; }
  else { if (p.C % 64) return p; p.cchunks = p.C / 64; p.nkb = p.taps * p.cchunks; }
  if (p.N % 32 || p.N > 256 || p.GW % NH_BM || p.nkb > 32) return p;
  if ((int64_t)B * p.GH * (p.GW / NH_BM) >= 2147483647LL) return p;
  p.BN = p.N;
  p.Kp = p.nkb * 64;
  // two CTAs per SM when the operator has at most 128 columns (2 x 128 TMEM columns per CTA) -- VAECLUST_NH_OCC=1 forces one
  static const char* occ_env = getenv("VAECLUST_NH_OCC");
  // measured at the script-12 shapes (profiles/README.md): it wins for the forward operators whose weight slab does not
  // stay resident anyway (ConvT 64->32: 122 -> 97 us, Conv 64->128: 41 -> 37 us), loses for the pair views and the Conv
  // data gradient (64 -> 71, 107 -> 121 us) and is neutral for the ConvT data gradients.  VAECLUST_NH_OCC=2 forces it.
  const bool by_op = p.wmode == 2 || (p.wmode == 0 && !p.pair);
  p.occ2 = (p.BN <= 128 && (by_op || (occ_env && atoi(occ_env) == 2)) && !(occ_env && atoi(occ_env) == 1)) ? 1 : 0;
  const size_t avail = (p.occ2 ? 232448 / 2 - 1024 : 232448) - 1024 - 1024;
  const size_t slab = (size_t)p.nkb * p.BN * 128;
  p.wres = slab + 4 * NH_A_BYTES <= avail ? 1 : 0;
  const size_t kb_bytes = NH_A_BYTES + (p.wres ? 0 : (size_t)p.BN * 128);
  const size_t ring = avail - (p.wres ? slab : 0);
  // k-blocks per ring stage: a barrier round trip (producer wait / arrive, issuer wait / commit) costs ~380 clocks however
  // little a stage holds, so group up to four k-blocks while at least two stages (and six k-blocks in flight) remain
  p.G = 1;
  for (int g = 4; g >= 2; --g)
    if (p.nkb % g == 0 && ring / (g * kb_bytes) >= 2 && (ring / (g * kb_bytes)) * g >= 8) { p.G = g; break; }
  {
    static const char* g_env = getenv("VAECLUST_NH_GROUP");
    if (g_env && atoi(g_env) >= 1 && p.nkb % atoi(g_env) == 0 && ring / (atoi(g_env) * kb_bytes) >= 2) p.G = atoi(g_env);
  }
  int stages = (int)(ring / (p.G * kb_bytes));
  if (stages > 8) stages = 8;
  if (stages < 2) return p;
  // Producer warp w fills stages w, w + nprod, ...  The stage count must be a multiple of nprod: then a producer meets
  // each of its ring slots in EVERY round; otherwise it can skip a round of a slot and the one-bit phase parity of the
  // slot's `empty` barrier aliases (the wait passes two phases early and the slot is overwritten while in use).
  if (stages >= 4) { stages -= stages % 4; p.nprod = 4; }
  else { stages = 2; p.nprod = 2; }
  p.stages = stages;
  p.smem = (p.wres ? slab : 0) + (size_t)stages * p.G * kb_bytes + (2 * stages + 5) * 8 + 16 + 3 * 32 * 4 + 1024;
  p.ok = true;
  return p;
}

int ig_build_weights(const float* w, int mode, int Cin, int Cout, int N, int BN, int K, int Kp, __nv_bfloat16* out,
                     cudaStream_t s);   // conv_igemm.cu: tap-major slabs

}  // namespace vc

using namespace vc;

static long long* g_nh_trace = nullptr;
extern "C" int vc_conv_nhwc_debug_trace(long long* buf) { g_nh_trace = buf; return VC_OK; }

extern "C" int vc_conv_nhwc_supported(int op, int B, int Cin, int H, int W, int Cout) {
  return nh_plan(op, B, Cin, H, W, Cout).ok ? 1 : 0;
}

extern "C" size_t vc_conv_nhwc_workspace_bytes(int op, int B, int Cin, int H, int W, int Cout) {
  const NhPlan p = nh_plan(op, B, Cin, H, W, Cout);
  return p.ok ? (size_t)p.BN * p.Kp * 2 + 2048 : 0;
}

extern "C" int vc_conv_nhwc(int op, const void* src_nhwc, const float* Wt, const float* bias, int relu,
                            const void* mask_nhwc, void* out_nhwc, float* out_nchw, int B, int Cin, int H, int W,
                            int Cout, void* workspace, size_t workspace_bytes, void* stream) {
  const NhPlan p = nh_plan(op, B, Cin, H, W, Cout);
  VC_REQUIRE(p.ok, VC_ERR_UNSUPPORTED, "vc_conv_nhwc: unsupported op / shape (op %d, B %d, Cin %d, H %d, W %d, Cout %d)", op,
             B, Cin, H, W, Cout);
  VC_REQUIRE(src_nhwc && Wt && (out_nhwc || out_nchw), VC_ERR_BAD_ARG, "vc_conv_nhwc: null pointer");
  VC_REQUIRE(workspace && workspace_bytes >= vc_conv_nhwc_workspace_bytes(op, B, Cin, H, W, Cout), VC_ERR_WORKSPACE,
             "vc_conv_nhwc: workspace too small");
  VC_REQUIRE(((reinterpret_cast<uintptr_t>(src_nhwc) | reinterpret_cast<uintptr_t>(mask_nhwc) |
               reinterpret_cast<uintptr_t>(out_nhwc) | reinterpret_cast<uintptr_t>(out_nchw) |
               reinterpret_cast<uintptr_t>(bias)) & 15) == 0, VC_ERR_BAD_ARG, "vc_conv_nhwc: pointers must be 16-byte aligned");
  cudaStream_t s = as_stream(stream);
  __nv_bfloat16* Wb = reinterpret_cast<__nv_bfloat16*>((reinterpret_cast<uintptr_t>(workspace) + 1023) & ~(uintptr_t)1023);
  CUtensorMap ma, mw;
  int rc;
  if (p.pair) rc = nh_make_map(&ma, src_nhwc, 64, p.SW / 2, p.SH, B, 64, NH_BM, 1);
  else rc = nh_make_map(&ma, src_nhwc, p.C, p.SW, p.SH, B, 64, p.S * NH_BM, p.S);
  if (rc != VC_OK) return rc;
  if (p.pair) {
    const int total = p.BN * p.Kp;
    nh_weights_pair_kernel<<<(total + 255) / 256, 256, 0, s>>>(Wt, p.wmode == 0 ? 0 : 1, Cin, Cout, p.BN, p.Kp, Wb);
    VC_LAUNCH_CHECK();
  } else {
    rc = ig_build_weights(Wt, p.wmode, Cin, Cout, p.N, p.BN, p.Kp, p.Kp, Wb, s);
    if (rc != VC_OK) return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)p.Kp, (uint64_t)p.BN};
    const uint64_t str[1] = {(uint64_t)p.Kp * 2};
    const uint32_t box[2] = {64u, (uint32_t)p.BN};
    rc = tma::make_map(&mw, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, Wb, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc != VC_OK) return rc;
  }
  NhArgs a;
  memset(&a, 0, sizeof(a));
  a.pair = p.pair; a.nblk = p.nblk; a.cchunks = p.cchunks; a.TW = p.TW; a.S = p.S; a.D0 = p.D0;
  a.B = B; a.GH = p.GH; a.GW = p.GW; a.tilesW = p.GW / NH_BM;
  a.N = p.N; a.BN = p.BN; a.OC = p.OC; a.OH = p.OH; a.OW = p.OW; a.quad = p.quad;
  a.tiles = B * p.GH * a.tilesW; a.nkb = p.nkb; a.stages = p.stages; a.wres = p.wres; a.G = p.G; a.nprod = p.nprod;
  {
    static const char* dbg_env = getenv("VAECLUST_NH_DEBUG");
    static const char* st_env = getenv("VAECLUST_NH_STAGES");
    a.dbg = dbg_env ? atoi(dbg_env) : 0;
    a.trace = g_nh_trace;
    if (st_env && atoi(st_env) >= 2 && atoi(st_env) < a.stages) a.stages = atoi(st_env);
  }
  a.idesc = tc05::idesc(tc05::kFmtBF16, NH_BM, p.BN, 0, 0);
  a.bias = bias; a.relu = relu; a.mask = static_cast<const __nv_bfloat16*>(mask_nhwc);
  a.out_nhwc = static_cast<__nv_bfloat16*>(out_nhwc); a.out_nchw = out_nchw;
  const int grid = a.tiles < sm_count() ? a.tiles : sm_count();
  if (p.occ2) {
    const int grid2 = a.tiles < 2 * sm_count() ? a.tiles : 2 * sm_count();
    VC_CUDA(cudaFuncSetAttribute(conv_nhwc_kernel<4, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
    conv_nhwc_kernel<4, 2><<<grid2, 64 + 32 * 4 + 32 * (NH_PROD - 1), p.smem, s>>>(a, ma, mw);
  } else {
    VC_CUDA(cudaFuncSetAttribute(conv_nhwc_kernel<8, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
    conv_nhwc_kernel<8, 1><<<grid, NH_THREADS, p.smem, s>>>(a, ma, mw);
  }
  VC_LAUNCH_CHECK();
  return VC_OK;
}
