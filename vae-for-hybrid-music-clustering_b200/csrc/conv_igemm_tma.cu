// TMA-fed implicit-GEMM data gradients of the script-12 VAE (12:87-89,135-137): the same GEMMs as conv_igemm.cu
// (Conv2d k3 s2 p1 data-gradient, phase-merged; ConvTranspose2d k4 s2 p1 data-gradient, a stride-2 gather), but the A
// operand is no longer gathered by eight producer warps from fp32 NCHW: the masked gradient g = dy * (y > 0) arrives as a
// bf16 NHWC copy (written by the ReLU-backward / bias-gradient pass that has to touch every element anyway,
// vc_relu_bwd_bias_nhwc), so a k-block of the A tile -- 128 consecutive pixels of one image row x 64 channels of one
// tap -- is exactly ONE 4-D TMA box: channel-contiguous 128-byte rows, 128-byte swizzle, out-of-range pixels (the
// padding) zero-filled by the copy engine, the stride-2 sampling of the transposed convolution done by the tensor
// map's traversal stride.  One elected thread issues the loads; ncu had put the old kernel's bound at producer
// instruction issue (58 % of the issue slots, tensor pipe 20 %, DRAM 22 %).
//
// 32-channel sources (the gradient of deconv.2's output) would only fill half a 128-byte row, so they are viewed as
// [B][H][W/2][2 x 32]: a row holds two neighbouring pixels, i.e. two taps of the stride-2 gather; the weight slab has
// zeros where a (pixel, tap) combination does not occur (12 k-blocks instead of 8: the tensor pipe has the room).
//
// warp 0: TMA producer (A boxes + weight-slab k-blocks), warp 1: tcgen05.mma issuer (2 accumulator stages in TMEM),
// warps 2..9: epilogue (tcgen05.ld, coalesced fp32 NCHW stores; 2x2 output blocks for the phase-merged operator).
#include <string.h>
#include "common.cuh"
#include "tc05.cuh"

namespace vc {

constexpr int IT_BM = 128;
constexpr int IT_EPI = 256;
constexpr int IT_THREADS = 64 + IT_EPI;
constexpr uint32_t IT_A_BYTES = IT_BM * 128;

enum ItMode { IT_CONV_DGRAD = 0, IT_CONVT_DGRAD = 1, IT_CONVT_DGRAD_PAIR = 2 };

struct ItArgs {
  float* out;
  int mode, C, cchunks;      // source channels, 64-channel chunks per tap
  int B, GH, GW, tilesW;     // GEMM rows = pixels (b, i, j); a tile = 128 consecutive j
  int N, BN, OC, OH, OW, quad;
  int tiles, nkb, stages;
  uint32_t idesc;
};

// k-block -> TMA coordinates (c, w, h) of its box for the tile whose first pixel is (i, j0)
__device__ __forceinline__ void it_coords(const ItArgs& a, int kb, int i, int j0, int& c, int& w, int& h) {
  if (a.mode == IT_CONV_DGRAD) {            // taps (dh, dw) in {0,1}^2 of dy at unit stride
    const int tap = kb / a.cchunks;
    c = (kb - tap * a.cchunks) * 64; w = j0 + (tap & 1); h = i + (tap >> 1);
  } else if (a.mode == IT_CONVT_DGRAD) {    // taps (th, tw) in {0..3}^2: dy[2i + th - 1][2j + tw - 1]
    const int tap = kb / a.cchunks;
    c = (kb - tap * a.cchunks) * 64; w = 2 * j0 + (tap & 3) - 1; h = 2 * i + (tap >> 2) - 1;
  } else {                                  // pixel pairs: pair j-1, j, j+1 of source row 2i + th - 1
    const int th = kb / 3;
    c = 0; w = j0 + (kb - th * 3) - 1; h = 2 * i + th - 1;
  }
}

// weight slab for the pair view: k = kb*64 + half*32 + co, kb = th*3 + blk; (blk, half) -> tw or nothing
__global__ void it_weights_pair_kernel(const float* __restrict__ w, int Cin, int Cout, int BN, int Kp,
                                       __nv_bfloat16* __restrict__ out) {
  const int total = BN * Kp;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    const int n = e / Kp, k = e - n * Kp;
    const int kb = k >> 6, half = (k >> 5) & 1, co = k & 31;
    const int th = kb / 3, blk = kb - th * 3;
    const int tw = blk == 0 ? (half ? 0 : -1) : (blk == 1 ? (half ? 2 : 1) : (half ? -1 : 3));
    float v = 0.f;
    if (n < Cin && th < 4 && tw >= 0 && co < Cout) v = w[((int64_t)n * Cout + co) * 16 + th * 4 + tw];   // W[ci, co, kh, kw]
    out[e] = __float2bfloat16_rn(v);
  }
}

__global__ void __launch_bounds__(IT_THREADS, 1)
conv_igemm_tma_kernel(const ItArgs a, const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapW) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (tma::s32(smem_raw) & 1023u)) & 1023u);
  const int S = a.stages;
  const uint32_t b_bytes = (uint32_t)a.BN * 128u, st_bytes = IT_A_BYTES + b_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)S * st_bytes);
  uint64_t* full = bars;             // S: TMA transaction bytes
  uint64_t* empty = bars + S;        // S: tcgen05.commit
  uint64_t* t_full = empty + S;      // 2: accumulator stage complete
  uint64_t* t_empty = t_full + 2;    // 2: accumulator stage drained by the epilogue threads
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t tmem_cols = a.BN <= 32 ? 64u : (a.BN <= 64 ? 128u : (a.BN <= 128 ? 256u : 512u));  // 2 stages

  if (threadIdx.x == 0) {
    tma::prefetch_map(&mapA);
    tma::prefetch_map(&mapW);
    for (int s = 0; s < S; ++s) { tma::mbar_init(&full[s], 1); tma::mbar_init(&empty[s], 1); }
    for (int s = 0; s < 2; ++s) { tma::mbar_init(&t_full[s], 1); tma::mbar_init(&t_empty[s], IT_EPI); }
    tma::mbar_fence_init();
  }
  if (warp == 1) tc05::tmem_alloc(tmem_slot, tmem_cols);
  tc05::fence_before();
  __syncthreads();
  tc05::fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int my_tiles = (a.tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (tc05::elect_one()) {   // a single lane the compiler knows about: no waterfall loop around UTCHMMA / UTMALDG
      uint32_t it = 0;
      for (int tl = 0; tl < my_tiles; ++tl) {
        const int t = (int)blockIdx.x + tl * (int)gridDim.x;
        const int tw = t % a.tilesW, r = t / a.tilesW;
        const int i = r % a.GH, b = r / a.GH, j0 = tw * IT_BM;
        for (int kb = 0; kb < a.nkb; ++kb, ++it) {
          const int s = it % S;
          tma::mbar_wait_backoff(&empty[s], ((it / S) & 1u) ^ 1u, 64);
          unsigned char* tile = smem + (size_t)s * st_bytes;
          int c, w, h;
          it_coords(a, kb, i, j0, c, w, h);
          tma::mbar_expect_tx(&full[s], IT_A_BYTES + b_bytes);
          tma::load_4d(tile, &mapA, &full[s], c, w, h, b);
          tma::load_2d(tile + IT_A_BYTES, &mapW, &full[s], kb * 64, 0);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (tc05::elect_one()) {   // a single lane the compiler knows about: no waterfall loop around UTCHMMA / UTMALDG
      uint32_t it = 0;
      for (int tl = 0; tl < my_tiles; ++tl) {
        const int as = tl & 1;
        const uint32_t aph = (uint32_t)(tl >> 1) & 1u;
        tma::mbar_wait_backoff(&t_empty[as], aph ^ 1u, 64);
        tc05::fence_after();
        const uint32_t tacc = tmem_base + (uint32_t)(as * a.BN);
        for (int kb = 0; kb < a.nkb; ++kb, ++it) {
          const int s = it % S;
          tma::mbar_wait_backoff(&full[s], (it / S) & 1u, 32);
          tc05::fence_after();
          const uint32_t sa = tma::s32(smem + (size_t)s * st_bytes), sb = sa + IT_A_BYTES;
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            tc05::mma_bf16(tacc, tc05::desc_k_sw128(sa + kk * 32), tc05::desc_k_sw128(sb + kk * 32), a.idesc,
                           (uint32_t)((kb | kk) != 0));
          tc05::commit(&empty[s]);
        }
        tc05::commit(&t_full[as]);
      }
    }
  } else {
    // ===================== epilogue: warps 2..9 =====================
    const int ew = warp - 2;
    const int q = warp & 3;          // TMEM lane quarter this warp may read
    const int hh = ew >> 2;          // even / odd 32-column chunks
    const int row = q * 32 + lane;
    const int nchunks = a.BN >> 5;
    const bool vec2 = a.quad && (a.OW % 2 == 0) && ((reinterpret_cast<uintptr_t>(a.out) & 7) == 0);
    for (int tl = 0; tl < my_tiles; ++tl) {
      const int as = tl & 1;
      const uint32_t aph = (uint32_t)(tl >> 1) & 1u;
      const int t = (int)blockIdx.x + tl * (int)gridDim.x;
      const int twi = t % a.tilesW, r = t / a.tilesW;
      const int pi = r % a.GH, pb = r / a.GH, pj = twi * IT_BM + row;
      // one warp polls the accumulator barrier; the other seven park on a hardware named barrier (no issue slots)
      if (ew == 0) tma::mbar_wait_backoff(&t_full[as], aph, 128);
      asm volatile("bar.sync 1, 256;" ::: "memory");
      tc05::fence_after();
      const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * a.BN);
      for (int cc = hh; cc < nchunks; cc += 2) {
        float v[32];
        tc05::ld_32x32(tacc + cc * 32, v);
        tc05::ld_wait();
        if (cc + 2 >= nchunks) {  // last chunk of this thread: the accumulator stage is free once it sits in registers
          tc05::fence_before();
          tma::mbar_arrive(&t_empty[as]);
        }
        if (!a.quad) {
          float* o = a.out + (((int64_t)pb * a.OC + cc * 32) * a.OH + pi) * a.OW + pj;
          const int64_t plane = (int64_t)a.OH * a.OW;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (cc * 32 + j < a.N) o[j * plane] = v[j];
        } else {
          const int oh = 2 * pi, ow = 2 * pj;
          const int64_t plane = (int64_t)a.OH * a.OW;
          float* o = a.out + ((int64_t)pb * a.OC + cc * 8) * plane + (int64_t)oh * a.OW + ow;
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            const int c = cc * 8 + g;
            if (c < a.OC) {
              float* r0 = o + g * plane;
              if (vec2) {
                *reinterpret_cast<float2*>(r0) = make_float2(v[4 * g], v[4 * g + 1]);
                if (oh + 1 < a.OH) *reinterpret_cast<float2*>(r0 + a.OW) = make_float2(v[4 * g + 2], v[4 * g + 3]);
              } else {
                r0[0] = v[4 * g];
                if (ow + 1 < a.OW) r0[1] = v[4 * g + 1];
                if (oh + 1 < a.OH) {
                  r0[a.OW] = v[4 * g + 2];
                  if (ow + 1 < a.OW) r0[a.OW + 1] = v[4 * g + 3];
                }
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

// ---- g = dy * (y > 0) as fp32 NCHW (for the weight-gradient kernels) AND as bf16 NHWC (for the kernel above), plus
// the per-channel sums (bias gradient), in one pass over dy / y: 12 B/element of mandatory traffic + 2 B for the copy.
// A block walks [64-pixel row segment] x [all channels] tiles; the transposition goes through shared memory.
constexpr int RN_PX = 64;
// PX = pixels per tile: 64 (8-byte loads, two pixels per lane) or 128 (16-byte loads, four per lane; needs W % 128 == 0)
template <int PX>
__global__ void __launch_bounds__(256)
relu_bwd_bias_nhwc_kernel(const float* __restrict__ dy, const float* __restrict__ y, float* __restrict__ g,
                          __nv_bfloat16* __restrict__ gn, int B, int C, int H, int W, int ntiles,
                          double* __restrict__ part) {
  extern __shared__ unsigned char rn_smem[];
  constexpr int PL = PX / 32;                                           // pixels per lane
  const int pitch = C + 8;                                              // bf16 elements per pixel row (16-byte multiple)
  __nv_bfloat16* tile = reinterpret_cast<__nv_bfloat16*>(rn_smem);      // [PX][pitch]
  double* csum = reinterpret_cast<double*>(rn_smem + (size_t)PX * pitch * 2);   // [C]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int c = threadIdx.x; c < C; c += blockDim.x) csum[c] = 0.0;
  __syncthreads();
  const int segs = W / PX;
  const int64_t plane = (int64_t)H * W;
  for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const int seg = t % segs, r = t / segs;
    const int h = r % H, b = r / H;
    const int64_t base = ((int64_t)b * C) * plane + (int64_t)h * W + seg * PX;
    for (int c0 = warp; c0 < C; c0 += 16) {      // two channel rows per pass: four independent loads in flight per lane
      float d[2][PL], v[2][PL];
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int c = c0 + 8 * k;
        const int64_t o = base + (int64_t)c * plane + PL * lane;
        if (c < C) {
          if (PL == 4) {
            const float4 dd = __ldcs(reinterpret_cast<const float4*>(dy + o)), vv = __ldcs(reinterpret_cast<const float4*>(y + o));
            d[k][0] = dd.x; d[k][1] = dd.y; d[k][PL - 2] = dd.z; d[k][PL - 1] = dd.w;
            v[k][0] = vv.x; v[k][1] = vv.y; v[k][PL - 2] = vv.z; v[k][PL - 1] = vv.w;
          } else {
            const float2 dd = __ldcs(reinterpret_cast<const float2*>(dy + o)), vv = __ldcs(reinterpret_cast<const float2*>(y + o));
            d[k][0] = dd.x; d[k][1] = dd.y; v[k][0] = vv.x; v[k][1] = vv.y;
          }
        }
      }
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int c = c0 + 8 * k;
        if (c >= C) continue;
        const int64_t o = base + (int64_t)c * plane + PL * lane;
        float s = 0.f;
#pragma unroll
        for (int u = 0; u < PL; ++u) {
          d[k][u] = v[k][u] > 0.f ? d[k][u] : 0.f;
          s += d[k][u];
          tile[(PL * lane + u) * pitch + c] = __float2bfloat16_rn(d[k][u]);
        }
        if (g) {
          if (PL == 4) *reinterpret_cast<float4*>(g + o) = make_float4(d[k][0], d[k][1], d[k][PL - 2], d[k][PL - 1]);
          else *reinterpret_cast<float2*>(g + o) = make_float2(d[k][0], d[k][1]);
        }
        s = warp_sum(s);
        if (lane == 0) csum[c] += (double)s;   // channel c is always handled by this warp: no race, fixed order
      }
    }
    __syncthreads();
    // the tile's PX pixels x C channels are one contiguous block of the NHWC tensor
    __nv_bfloat16* dst = gn + (((int64_t)b * H + h) * W + seg * PX) * C;
    const int vec_per_px = C / 8;
    for (int e = threadIdx.x; e < PX * vec_per_px; e += blockDim.x) {
      const int px = e / vec_per_px, q = e - px * vec_per_px;
      reinterpret_cast<uint4*>(dst)[e] = *reinterpret_cast<const uint4*>(tile + px * pitch + q * 8);
    }
    __syncthreads();
  }
  for (int c = threadIdx.x; c < C; c += blockDim.x) part[(int64_t)blockIdx.x * C + c] = csum[c];
}

// The same pass with conflict-free shared-memory traffic (C % 16 == 0, W % 128 == 0).  In the kernel above a lane owns four
// consecutive pixels of one channel, so its 2-byte stores into the [pixel][channel] tile are 1088 bytes apart: 16-way bank
// conflicts made it run at 1.9 TB/s.  Here a lane owns pixels lane, lane + 32, lane + 64, lane + 96 of a channel PAIR
// (coalesced 128-byte scalar loads), stores one bf16x2 word per pixel into rows of C/2 + 1 words (bank = pixel + pair: no
// conflicts on the way in or out), and keeps its channel sums in registers until the end (no shuffles per tile).
template <int NPASS>      // channel pairs per warp = C / 16
__global__ void __launch_bounds__(256)
relu_bwd_bias_nhwc2_kernel(const float* __restrict__ dy, const float* __restrict__ y, float* __restrict__ g,
                           __nv_bfloat16* __restrict__ gn, int B, int C, int H, int W, int ntiles, double* __restrict__ part) {
  extern __shared__ unsigned char rn_smem[];
  constexpr int PX = 128;
  const int pw = (C >> 1) + 1;                                          // words per pixel row
  uint32_t* tile = reinterpret_cast<uint32_t*>(rn_smem);                // [PX][pw]
  float* csum = reinterpret_cast<float*>(rn_smem + (size_t)PX * pw * 4);   // [8 warps][... ] reused at the end: [C][32]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int segs = W / PX;
  const int64_t plane = (int64_t)H * W;
  float acc[NPASS][2];
#pragma unroll
  for (int i = 0; i < NPASS; ++i) acc[i][0] = acc[i][1] = 0.f;
  for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const int seg = t % segs, r = t / segs;
    const int h = r % H, b = r / H;
    const int64_t base = ((int64_t)b * C) * plane + (int64_t)h * W + seg * PX + lane;
#pragma unroll
    for (int i0 = 0; i0 < NPASS; i0 += 2) {      // two channel pairs per round: 32 independent loads in flight per lane
      float d[2][2][4], v[2][2][4];
#pragma unroll
      for (int k = 0; k < 2; ++k)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int64_t o = base + (int64_t)(2 * (warp + 8 * (i0 + k)) + e) * plane;
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            d[k][e][u] = __ldcs(dy + o + 32 * u);
            v[k][e][u] = __ldcs(y + o + 32 * u);
          }
        }
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int cp = warp + 8 * (i0 + k);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const float g0 = v[k][0][u] > 0.f ? d[k][0][u] : 0.f, g1 = v[k][1][u] > 0.f ? d[k][1][u] : 0.f;
          acc[i0 + k][0] += g0;
          acc[i0 + k][1] += g1;
          __nv_bfloat162 hh = __floats2bfloat162_rn(g0, g1);
          tile[(lane + 32 * u) * pw + cp] = *reinterpret_cast<uint32_t*>(&hh);
          if (g) {
            const int64_t o = base + (int64_t)(2 * cp) * plane + 32 * u;
            g[o] = g0;
            g[o + plane] = g1;
          }
        }
      }
    }
    __syncthreads();
    // the tile's PX pixels x C channels are one contiguous block of the NHWC tensor: C / 2 words per pixel
    uint32_t* dst = reinterpret_cast<uint32_t*>(gn + (((int64_t)b * H + h) * W + seg * PX) * C);
    const int wpp = C >> 1;
    for (int e = threadIdx.x; e < PX * wpp; e += blockDim.x) {
      const int px = e / wpp, q = e - px * wpp;
      dst[e] = tile[px * pw + q];
    }
    __syncthreads();
  }
  // per-channel sums: lanes (butterfly, fixed order), then this block's partial in double
#pragma unroll
  for (int i = 0; i < NPASS; ++i)
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      float s = warp_sum(acc[i][e]);
      if (lane == 0) csum[2 * (warp + 8 * i) + e] = s;
    }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) part[(int64_t)blockIdx.x * C + c] = (double)csum[c];
}

// db[c] = sum over blocks: 256 threads per 32 channels, each warp adds every eighth partial, sub-sums added in warp order
__global__ void __launch_bounds__(256) rn_fold_kernel(const double* __restrict__ part, int nblocks, int C, float* __restrict__ db) {
  __shared__ double sub[8][32];
  const int lane = threadIdx.x & 31, gw = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  double v = 0.0;
  if (c < C) v = strided_sum_ordered(part + c, (int64_t)C, gw, nblocks, 8);
  sub[gw][lane] = v;
  __syncthreads();
  if (gw != 0 || c >= C) return;
#pragma unroll
  for (int i = 1; i < 8; ++i) v += sub[i][lane];
  db[c] = (float)v;
}

// ---- host ---------------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn4)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// 4-D bf16 map [B][SH][SWv][Cv] (innermost first: Cv, SWv, SH, B) with a traversal stride along W
static int it_make_map(CUtensorMap* m, const void* base, int Cv, int SWv, int SH, int B, int boxW, int strideW) {
  EncodeTiledFn4 fn = reinterpret_cast<EncodeTiledFn4>(tma::encode_fn());
  if (!fn) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return VC_ERR_CUDA; }
  cuuint64_t gd[4] = {(cuuint64_t)Cv, (cuuint64_t)SWv, (cuuint64_t)SH, (cuuint64_t)B};
  cuuint64_t gs[3] = {(cuuint64_t)Cv * 2, (cuuint64_t)SWv * Cv * 2, (cuuint64_t)SH * SWv * Cv * 2};
  cuuint32_t bx[4] = {64u, (cuuint32_t)boxW, 1u, 1u};
  cuuint32_t es[4] = {1u, (cuuint32_t)strideW, 1u, 1u};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), gd, gs, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (4-D, box W %d, stride %d) failed with CUresult %d", boxW, strideW, (int)r);
    return VC_ERR_UNSUPPORTED;
  }
  return VC_OK;
}

struct ItPlan { bool ok; int mode, C, cchunks, N, BN, nkb, Kp, GH, GW, SH, SW, OC, OH, OW, quad; };

static ItPlan it_plan(int op, int B, int Cin, int H, int W, int Cout) {
  ItPlan p;
  memset(&p, 0, sizeof(p));
  const int OHc = (H - 1) / 2 + 1, OWc = (W - 1) / 2 + 1;
  if (op == VC_OP_CONV_DGRAD) {          // dy [B,Cout,OHc,OWc] -> dx [B,Cin,H,W], phase-merged: N = 4 Cin
    if (H % 2 || W % 2) return p;
    p.mode = IT_CONV_DGRAD; p.C = Cout; p.N = 4 * Cin; p.GH = OHc; p.GW = OWc; p.SH = OHc; p.SW = OWc;
    p.OC = Cin; p.OH = H; p.OW = W; p.quad = 1;
    if (p.C % 64) return p;
    p.cchunks = p.C / 64; p.nkb = 4 * p.cchunks;
  } else if (op == VC_OP_CONVT_DGRAD) {  // dy [B,Cout,2H,2W] -> dx [B,Cin,H,W]
    p.C = Cout; p.N = Cin; p.GH = H; p.GW = W; p.SH = 2 * H; p.SW = 2 * W; p.OC = Cin; p.OH = H; p.OW = W; p.quad = 0;
    if (p.C == 32) { p.mode = IT_CONVT_DGRAD_PAIR; p.cchunks = 1; p.nkb = 12; }
    else if (p.C % 64 == 0) { p.mode = IT_CONVT_DGRAD; p.cchunks = p.C / 64; p.nkb = 16 * p.cchunks; }
    else return p;
  } else {
    return p;
  }
  p.BN = (p.N + 31) / 32 * 32;
  p.Kp = p.nkb * 64;
  p.ok = p.GW % IT_BM == 0 && p.BN <= 256 && (int64_t)B * p.GH * (p.GW / IT_BM) < 2147483647LL;
  return p;
}

bool conv_igemm_tma_supported(int op, int B, int Cin, int H, int W, int Cout) { return it_plan(op, B, Cin, H, W, Cout).ok; }

int ig_build_weights(const float* w, int mode, int Cin, int Cout, int N, int BN, int K, int Kp, __nv_bfloat16* out,
                     cudaStream_t s);   // conv_igemm.cu: tap-major slabs

}  // namespace vc

using namespace vc;

extern "C" int vc_conv_dgrad_nhwc_supported(int op, int B, int Cin, int H, int W, int Cout) {
  return conv_igemm_tma_supported(op, B, Cin, H, W, Cout) ? 1 : 0;
}

extern "C" size_t vc_conv_dgrad_nhwc_workspace_bytes(int op, int B, int Cin, int H, int W, int Cout) {
  const ItPlan p = it_plan(op, B, Cin, H, W, Cout);
  return p.ok ? (size_t)p.BN * p.Kp * 2 + 2048 : 0;
}

extern "C" int vc_conv_dgrad_nhwc(int op, const void* g_nhwc, const float* Wt, float* dx, int B, int Cin, int H, int W,
                                  int Cout, void* workspace, size_t workspace_bytes, void* stream) {
  const ItPlan p = it_plan(op, B, Cin, H, W, Cout);
  VC_REQUIRE(p.ok && g_nhwc && Wt && dx, VC_ERR_UNSUPPORTED, "vc_conv_dgrad_nhwc: unsupported shape");
  VC_REQUIRE(workspace && workspace_bytes >= vc_conv_dgrad_nhwc_workspace_bytes(op, B, Cin, H, W, Cout), VC_ERR_WORKSPACE,
             "vc_conv_dgrad_nhwc: workspace too small");
  VC_REQUIRE((reinterpret_cast<uintptr_t>(g_nhwc) & 15) == 0, VC_ERR_BAD_ARG, "vc_conv_dgrad_nhwc: g_nhwc must be 16-byte aligned");
  cudaStream_t s = as_stream(stream);
  __nv_bfloat16* Wb = reinterpret_cast<__nv_bfloat16*>((reinterpret_cast<uintptr_t>(workspace) + 1023) & ~(uintptr_t)1023);
  CUtensorMap ma, mw;
  int rc;
  if (p.mode == IT_CONVT_DGRAD_PAIR) rc = it_make_map(&ma, g_nhwc, 64, p.SW / 2, p.SH, B, IT_BM, 1);
  else if (p.mode == IT_CONVT_DGRAD) rc = it_make_map(&ma, g_nhwc, p.C, p.SW, p.SH, B, 2 * IT_BM, 2);
  else rc = it_make_map(&ma, g_nhwc, p.C, p.SW, p.SH, B, IT_BM, 1);
  if (rc != VC_OK) return rc;
  if (p.mode == IT_CONVT_DGRAD_PAIR) {
    const int total = p.BN * p.Kp;
    it_weights_pair_kernel<<<(total + 255) / 256, 256, 0, s>>>(Wt, Cin, Cout, p.BN, p.Kp, Wb);
    VC_LAUNCH_CHECK();
  } else {   // the tap-major slab of conv_igemm.cu: mode 1 = Conv data-grad, 3 = ConvT data-grad
    rc = ig_build_weights(Wt, p.mode == IT_CONV_DGRAD ? 1 : 3, Cin, Cout, p.N, p.BN, p.Kp, p.Kp, Wb, s);
    if (rc != VC_OK) return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)p.Kp, (uint64_t)p.BN};
    const uint64_t str[1] = {(uint64_t)p.Kp * 2};
    const uint32_t box[2] = {64u, (uint32_t)p.BN};
    rc = tma::make_map(&mw, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, Wb, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc != VC_OK) return rc;
  }
  ItArgs a;
  memset(&a, 0, sizeof(a));
  a.out = dx; a.mode = p.mode; a.C = p.C; a.cchunks = p.cchunks; a.B = B; a.GH = p.GH; a.GW = p.GW; a.tilesW = p.GW / IT_BM;
  a.N = p.N; a.BN = p.BN; a.OC = p.OC; a.OH = p.OH; a.OW = p.OW; a.quad = p.quad;
  a.tiles = B * p.GH * a.tilesW; a.nkb = p.nkb;
  const size_t st_bytes = IT_A_BYTES + (size_t)p.BN * 128;
  int stages = (int)((200 * 1024) / st_bytes);
  if (stages > 8) stages = 8;
  if (stages < 2) stages = 2;
  a.stages = stages;
  a.idesc = tc05::idesc(tc05::kFmtBF16, IT_BM, p.BN, 0, 0);
  const size_t smem = (size_t)stages * st_bytes + (2 * stages + 4) * 8 + 16 + 1024;
  const int grid = a.tiles < sm_count() ? a.tiles : sm_count();
  VC_CUDA(cudaFuncSetAttribute(conv_igemm_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  conv_igemm_tma_kernel<<<grid, IT_THREADS, smem, s>>>(a, ma, mw);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" size_t vc_relu_bwd_bias_nhwc_workspace_bytes(int C) { return (size_t)4 * 148 * (C > 0 ? C : 1) * 8 + 64; }

extern "C" int vc_relu_bwd_bias_nhwc(const float* dy, const float* y, int B, int C, int H, int W, float* g, void* g_nhwc,
                                     float* db, void* workspace, size_t workspace_bytes, void* stream) {
  VC_REQUIRE(dy && y && g_nhwc && db && B > 0 && H > 0, VC_ERR_BAD_ARG, "vc_relu_bwd_bias_nhwc: bad arguments");   // g may be null
  VC_REQUIRE(C % 8 == 0 && C <= 256 && W % RN_PX == 0, VC_ERR_UNSUPPORTED,
             "vc_relu_bwd_bias_nhwc: need C %% 8 == 0, C <= 256, W %% 64 == 0 (got C=%d W=%d)", C, W);
  VC_REQUIRE(((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(g)) & 7) == 0 &&
                 (reinterpret_cast<uintptr_t>(g_nhwc) & 15) == 0, VC_ERR_BAD_ARG, "vc_relu_bwd_bias_nhwc: misaligned pointers");
  int grid = 4 * sm_count();
  if (grid > 4 * 148) grid = 4 * 148;
  const bool wide = W % 128 == 0 && ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(g)) & 15) == 0;
  const int PX = wide ? 128 : RN_PX;
  const int64_t ntiles = (int64_t)B * H * (W / PX);
  VC_REQUIRE(ntiles < 2147483647LL, VC_ERR_UNSUPPORTED, "vc_relu_bwd_bias_nhwc: tensor too large");
  if (grid > ntiles) grid = (int)ntiles;
  VC_REQUIRE(workspace && workspace_bytes >= (size_t)grid * C * 8, VC_ERR_WORKSPACE, "vc_relu_bwd_bias_nhwc: workspace too small");
  double* part = static_cast<double*>(workspace);
  const size_t smem = (size_t)PX * (C + 8) * 2 + (size_t)C * 8;
  static const bool v2off = getenv("VAECLUST_RELU_NHWC2") && atoi(getenv("VAECLUST_RELU_NHWC2")) == 0;
  if (wide && !v2off && (C == 128 || C == 64 || C == 32) && (int64_t)B * C * H * W < (1LL << 40)) {
    const size_t smem2 = (size_t)128 * ((C >> 1) + 1) * 4 + (size_t)C * 4;
    auto* k2 = C == 128 ? relu_bwd_bias_nhwc2_kernel<8> : (C == 64 ? relu_bwd_bias_nhwc2_kernel<4> : relu_bwd_bias_nhwc2_kernel<2>);
    if (C == 128 && grid > 2 * sm_count()) grid = 2 * sm_count();      // 120 registers: two resident blocks per SM
    k2<<<grid, 256, smem2, as_stream(stream)>>>(dy, y, g, static_cast<__nv_bfloat16*>(g_nhwc), B, C, H, W, (int)ntiles, part);
  } else if (wide) {
    VC_CUDA(cudaFuncSetAttribute(relu_bwd_bias_nhwc_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    relu_bwd_bias_nhwc_kernel<128><<<grid, 256, smem, as_stream(stream)>>>(dy, y, g, static_cast<__nv_bfloat16*>(g_nhwc), B, C,
                                                                            H, W, (int)ntiles, part);
  } else {
    relu_bwd_bias_nhwc_kernel<64><<<grid, 256, smem, as_stream(stream)>>>(dy, y, g, static_cast<__nv_bfloat16*>(g_nhwc), B, C,
                                                                           H, W, (int)ntiles, part);
  }
  VC_LAUNCH_CHECK();
  rn_fold_kernel<<<(C + 31) / 32, 256, 0, as_stream(stream)>>>(part, grid, C, db);
  VC_LAUNCH_CHECK();
  return VC_OK;
}
