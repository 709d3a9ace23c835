// Implicit-GEMM convolutions on tcgen05 (sm_100a): Conv2d(k3,s2,p1) forward / data-gradient and
// ConvTranspose2d(k4,s2,p1) forward / data-gradient of the script-12 VAE (12:87-89,135-137) without ever writing an
// im2col matrix to HBM.
//
// One persistent CTA per SM walks 128-pixel row tiles of   C[M = pixels, N] = A[M, K = taps x C] * Bw[N, K]^T :
//   * warps 0-7  (producers): gather the A tile straight from the NCHW fp32 activations (lane = pixel, so every load
//     instruction is coalesced along W), convert to bf16 and store it K-major into 128-byte-swizzled shared memory --
//     the layout a TMA box would have produced, so the same UMMA descriptors apply.  K is ordered tap-major
//     (k = tap*C + c): a 32-wide k run is one tap and 32 consecutive channels, i.e. one bounds check and 32
//     independent loads that differ only by the channel stride.  The next k-block's loads are issued before the current
//     block is packed (register double buffer).
//   * thread 0 also fetches the matching [BN x 64] slab of the bf16 weight matrix by TMA.
//   * warp 8 (one elected lane): tcgen05.mma, 128 x BN fp32 accumulator in tensor memory, two accumulator stages.
//   * warps 9-16 (epilogue): tcgen05.ld, bias / ReLU, coalesced NCHW stores; overlaps the next tile's main loop.
//
// The two "scattering" operators are phase-merged so that they also become gathers with dense 2x2 output blocks:
//   ConvT forward:  rows = input pixels (i,j); taps (dh,dw) in {-1,0,1}^2 of x; columns n = co*4 + ph*2 + pw hold the
//                   four output parities y[2i+ph, 2j+pw]; the weight slab is W[ci,co,1+ph-2dh,1+pw-2dw] or 0.
//   Conv data-grad: rows = coarse pixels (i,j) of dy; taps (dh,dw) in {0,1}^2; columns n = ci*4 + p*2 + q hold
//                   dx[2i+p, 2j+q]; weights W[co,ci,p+1-2dh,q+1-2dw] or 0.
// Zero weights cost tensor-core issue slots (2.25x / 1.78x the useful FLOPs) but the gather -- the real bound -- drops
// from 16 (resp. 9) loads per source element to 9 (resp. 4), and every output row is written as full 256-byte lines.
#include <string.h>
#include "common.cuh"
#include "tc05.cuh"

namespace vc {

constexpr int IG_BM = 128;
constexpr int IG_PROD = 256;                 // producer threads (2 per row: one 32-wide k half each)
constexpr int IG_EPI = 256;                  // epilogue threads
constexpr int IG_THREADS = IG_PROD + 32 + IG_EPI;
constexpr uint32_t IG_A_BYTES = IG_BM * 128;  // [128 rows][64 bf16]

enum IgMode { IG_CONV_FWD = 0, IG_CONV_DGRAD = 1, IG_CONVT_FWD = 2, IG_CONVT_DGRAD = 3 };

struct IgArgs {
  const float* src; float* out; const float* bias;
  int B, C, SH, SW;      // source tensor [B, C, SH, SW]
  int GH, GW;            // pixel grid that indexes the GEMM rows
  int S, D0, TH, TW;     // source pixel of tap (th,tw): (S*i + th + D0, S*j + tw + D0)
  int N, BN;             // real / padded (multiple of 32) column count
  int OC, OH, OW;        // output tensor [B, OC, OH, OW]
  int quad, relu;        // quad: columns are (channel, 2x2 parity) blocks
  int M, tiles, nkb, stages;
  uint32_t idesc;
};

__device__ __forceinline__ uint32_t ig_pack(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// ---- weight slab  Bw[n][k] (bf16, [BN][Kp], zero padded), k = tap*C + c
__global__ void ig_weights_kernel(const float* __restrict__ w, int mode, int Cin, int Cout, int N, int BN, int K, int Kp,
                                  __nv_bfloat16* __restrict__ out) {
  const int total = BN * Kp;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    const int n = e / Kp, k = e - n * Kp;
    float v = 0.f;
    if (n < N && k < K) {
      if (mode == IG_CONV_FWD) {            // C = Cin, taps 3x3, W[co, ci, kh, kw]
        const int tap = k / Cin, c = k - tap * Cin;
        v = w[((int64_t)n * Cin + c) * 9 + tap];
      } else if (mode == IG_CONV_DGRAD) {   // C = Cout, taps 2x2, n = ci*4 + p*2 + q
        const int tap = k / Cout, co = k - tap * Cout;
        const int dh = tap >> 1, dw = tap & 1, ci = n >> 2, p = (n >> 1) & 1, q = n & 1;
        const int kh = p + 1 - 2 * dh, kw = q + 1 - 2 * dw;
        if (kh >= 0 && kh <= 2 && kw >= 0 && kw <= 2) v = w[(((int64_t)co * Cin + ci) * 3 + kh) * 3 + kw];
      } else if (mode == IG_CONVT_FWD) {    // C = Cin, taps 3x3 (dh = th-1), n = co*4 + ph*2 + pw, W[ci, co, kh, kw]
        const int tap = k / Cin, ci = k - tap * Cin;
        const int dh = tap / 3 - 1, dw = tap % 3 - 1, co = n >> 2, ph = (n >> 1) & 1, pw = n & 1;
        const int kh = 1 + ph - 2 * dh, kw = 1 + pw - 2 * dw;
        if (kh >= 0 && kh <= 3 && kw >= 0 && kw <= 3) v = w[(((int64_t)ci * Cout + co) * 4 + kh) * 4 + kw];
      } else {                              // IG_CONVT_DGRAD: C = Cout, taps 4x4, n = ci, W[ci, co, kh, kw]
        const int tap = k / Cout, co = k - tap * Cout;
        v = w[((int64_t)n * Cout + co) * 16 + tap];
      }
    }
    out[e] = __float2bfloat16_rn(v);
  }
}

// ---- producer helpers -------------------------------------------------------------------------------------------
struct IgPixel { int b, i, j; bool ok; };

__device__ __forceinline__ IgPixel ig_pixel(const IgArgs& a, int m) {
  IgPixel p;
  p.ok = m < a.M;
  const int mm = p.ok ? m : 0;
  p.j = mm % a.GW;
  const int t = mm / a.GW;
  p.i = t % a.GH;
  p.b = t / a.GH;
  return p;
}

// 32 k-values [k0, k0+32) of one row into v[].  kOneCh: C == 1 (k = tap); else C % 32 == 0 (one tap, 32 channels).
template <bool kOneCh>
__device__ __forceinline__ void ig_gather(const IgArgs& a, const IgPixel& px, int k0, float (&v)[32]) {
  const int ntaps = a.TH * a.TW;
  if (kOneCh) {
    const float* base = a.src + (int64_t)px.b * a.SH * a.SW;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const int tap = k0 + j;
      const int th = tap / a.TW, tw = tap - th * a.TW;
      const int sh = a.S * px.i + th + a.D0, sw = a.S * px.j + tw + a.D0;
      const bool ok = px.ok && tap < ntaps && (unsigned)sh < (unsigned)a.SH && (unsigned)sw < (unsigned)a.SW;
      v[j] = ok ? __ldg(base + (int64_t)sh * a.SW + sw) : 0.f;
    }
  } else {
    const int tap = k0 / a.C, c0 = k0 - tap * a.C;
    const int th = tap / a.TW, tw = tap - th * a.TW;
    const int sh = a.S * px.i + th + a.D0, sw = a.S * px.j + tw + a.D0;
    const bool ok = px.ok && tap < ntaps && (unsigned)sh < (unsigned)a.SH && (unsigned)sw < (unsigned)a.SW;
    const int plane = a.SH * a.SW;
    // byte address kept as one 64-bit integer and bumped by the channel stride: two integer ops per load
    uint64_t addr = reinterpret_cast<uint64_t>(a.src + ((int64_t)px.b * a.C + c0) * plane + (ok ? sh * a.SW + sw : 0));
    const uint64_t step = (uint64_t)plane * 4u;
    if (ok) {
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        asm("ld.global.nc.f32 %0, [%1];" : "=f"(v[j]) : "l"(addr));
        addr += step;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = 0.f;
    }
  }
}

// Table-driven variant of the C % 32 == 0 branch above: everything that depends only on (k-block, k half) -- tap
// offsets and the element offset c0*plane + th*SW + tw -- is read from a shared-memory table built once per CTA, and
// everything that depends only on the pixel (pixbase, hi0 = S*i + D0, wj0 = S*j + D0) is computed once per tile, so a
// k-block costs one LDS, the bounds test and one 64-bit add before its 32 loads.
struct IgPix { const float* base; int hi0, wj0; bool ok; };

__device__ __forceinline__ IgPix ig_pix(const IgArgs& a, int m) {
  const IgPixel p = ig_pixel(a, m);
  IgPix q;
  q.ok = p.ok;
  q.hi0 = a.S * p.i + a.D0;
  q.wj0 = a.S * p.j + a.D0;
  q.base = a.src + ((int64_t)p.b * a.C * a.SH + q.hi0) * a.SW + q.wj0;
  return q;
}

__device__ __forceinline__ void ig_gather_tab(const IgArgs& a, const IgPix& px, const int2 t, uint64_t step,
                                              float (&v)[32]) {
  const int th = t.y & 0xff, tw = (t.y >> 8) & 0xff;
  const bool ok = px.ok && (t.y >> 16) != 0 && (unsigned)(px.hi0 + th) < (unsigned)a.SH &&
                  (unsigned)(px.wj0 + tw) < (unsigned)a.SW;
  if (ok) {
    uint64_t addr = reinterpret_cast<uint64_t>(px.base + t.x);
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      asm("ld.global.nc.f32 %0, [%1];" : "=f"(v[j]) : "l"(addr));
      addr += step;
    }
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = 0.f;
  }
}

// pack 32 floats -> 4 x 16-byte chunks of row `row`, chunks [4*kq, 4*kq+4) of the swizzled K-major tile
__device__ __forceinline__ void ig_store(unsigned char* tile, int row, int kq, const float (&v)[32]) {
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const uint4 o = make_uint4(ig_pack(v[8 * c + 0], v[8 * c + 1]), ig_pack(v[8 * c + 2], v[8 * c + 3]),
                               ig_pack(v[8 * c + 4], v[8 * c + 5]), ig_pack(v[8 * c + 6], v[8 * c + 7]));
    *reinterpret_cast<uint4*>(tile + tma::sw128((uint32_t)row, (uint32_t)(4 * kq + c))) = o;
  }
}

// ---- the kernel ---------------------------------------------------------------------------------------------------
template <bool kOneCh>
__global__ void __launch_bounds__(IG_THREADS, 1)
conv_igemm_kernel(const IgArgs a, const __grid_constant__ CUtensorMap mapW) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (tma::s32(smem_raw) & 1023u)) & 1023u);
  const int S = a.stages;
  const uint32_t b_bytes = (uint32_t)a.BN * 128u, st_bytes = IG_A_BYTES + b_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)S * st_bytes);
  uint64_t* full = bars;             // S: 256 producer arrivals + the TMA transaction
  uint64_t* empty = bars + S;        // S: tcgen05.commit
  uint64_t* t_full = empty + S;      // 2: accumulator stage complete
  uint64_t* t_empty = t_full + 2;    // 2: accumulator stage drained by the 256 epilogue threads
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);
  int2* ktab = reinterpret_cast<int2*>(tmem_slot + 2);   // [nkb][2]: {element offset, th | tw << 8 | valid << 16}

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (!kOneCh) {
    const int ntaps = a.TH * a.TW, plane = a.SH * a.SW;
    for (int e = threadIdx.x; e < 2 * a.nkb; e += IG_THREADS) {
      const int k0 = e * 32;
      const int tap = k0 / a.C, c0 = k0 - tap * a.C;
      const int th = tap / a.TW, tw = tap - th * a.TW;
      ktab[e] = make_int2(c0 * plane + th * a.SW + tw, th | (tw << 8) | ((tap < ntaps ? 1 : 0) << 16));
    }
  }
  const uint32_t tmem_cols = a.BN <= 32 ? 64u : (a.BN <= 64 ? 128u : (a.BN <= 128 ? 256u : 512u));  // 2 stages

  if (threadIdx.x == 0) {
    tma::prefetch_map(&mapW);
    for (int s = 0; s < S; ++s) { tma::mbar_init(&full[s], IG_PROD + 1); tma::mbar_init(&empty[s], 1); }
    for (int s = 0; s < 2; ++s) { tma::mbar_init(&t_full[s], 1); tma::mbar_init(&t_empty[s], IG_EPI); }
    tma::mbar_fence_init();
  }
  if (warp == 8) tc05::tmem_alloc(tmem_slot, tmem_cols);
  tc05::fence_before();
  __syncthreads();
  tc05::fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int my_tiles = (a.tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp < 8) {
    // ===================== producers =====================
    const int row = threadIdx.x & 127, kq = threadIdx.x >> 7;
    const int total = my_tiles * a.nkb;
    const uint64_t step = (uint64_t)a.SH * (uint64_t)a.SW * 4u;
    float nxt[32];
    IgPixel px = ig_pixel(a, (int)blockIdx.x * IG_BM + row);
    IgPix pq = ig_pix(a, (int)blockIdx.x * IG_BM + row);
    if (total > 0) {
      if constexpr (kOneCh) ig_gather<true>(a, px, kq * 32, nxt);
      else ig_gather_tab(a, pq, ktab[kq], step, nxt);
    }
    int t = 0, kb = 0, s = 0;
    uint32_t ph = 1;   // parity to wait for on empty[s]
    for (int it = 0; it < total; ++it) {
      float cur[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) cur[j] = nxt[j];
      // issue the loads of the next k-block (possibly of the next tile) before touching shared memory
      int kb2 = kb + 1, t2 = t;
      if (kb2 == a.nkb) { kb2 = 0; ++t2; }
      if (it + 1 < total) {
        if (t2 != t) {
          const int m = ((int)blockIdx.x + t2 * (int)gridDim.x) * IG_BM + row;
          if constexpr (kOneCh) px = ig_pixel(a, m);
          else pq = ig_pix(a, m);
        }
        if constexpr (kOneCh) ig_gather<true>(a, px, kb2 * 64 + kq * 32, nxt);
        else ig_gather_tab(a, pq, ktab[2 * kb2 + kq], step, nxt);
      }
      tma::mbar_wait(&empty[s], ph);
      unsigned char* tile = smem + (size_t)s * st_bytes;
      if (threadIdx.x == 0) {
        tma::mbar_expect_tx(&full[s], b_bytes);
        tma::load_2d(tile + IG_A_BYTES, &mapW, &full[s], kb * 64, 0);
      }
      ig_store(tile, row, kq, cur);
      tma::fence_proxy_async();
      tma::mbar_arrive(&full[s]);
      kb = kb2; t = t2;
      if (++s == S) { s = 0; ph ^= 1u; }
    }
  } else if (warp == 8) {
    // ===================== MMA issuer =====================
    if (tc05::elect_one()) {   // a single lane the compiler knows about: no waterfall loop around UTCHMMA / UTMALDG
      int it = 0;
      for (int tl = 0; tl < my_tiles; ++tl) {
        const int as = tl & 1;
        const uint32_t aph = (uint32_t)(tl >> 1) & 1u;
        tma::mbar_wait_backoff(&t_empty[as], aph ^ 1u, 64);
        tc05::fence_after();
        const uint32_t tacc = tmem_base + (uint32_t)(as * a.BN);
        for (int kb = 0; kb < a.nkb; ++kb, ++it) {
          const int s = it % S;
          const uint32_t ph = (uint32_t)(it / S) & 1u;
          tma::mbar_wait_backoff(&full[s], ph, 32);
          tc05::fence_after();
          const uint32_t sa = tma::s32(smem + (size_t)s * st_bytes), sb = sa + IG_A_BYTES;
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
    // ===================== epilogue: warps 9..16 =====================
    const int ew = warp - 9;
    const int q = warp & 3;          // TMEM lane quarter this warp may read
    const int hh = ew >> 2;          // even / odd 32-column chunks
    const int row = q * 32 + lane;
    const int nchunks = a.BN >> 5;
    const bool vec2 = a.quad && (a.OW % 2 == 0) && ((reinterpret_cast<uintptr_t>(a.out) & 7) == 0);
    for (int tl = 0; tl < my_tiles; ++tl) {
      const int as = tl & 1;
      const uint32_t aph = (uint32_t)(tl >> 1) & 1u;
      const int m = ((int)blockIdx.x + tl * (int)gridDim.x) * IG_BM + row;
      const IgPixel px = ig_pixel(a, m);
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
        if (!px.ok) continue;
        if (!a.quad) {
          float* o = a.out + (((int64_t)px.b * a.OC + cc * 32) * a.OH + px.i) * a.OW + px.j;
          const int64_t plane = (int64_t)a.OH * a.OW;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int n = cc * 32 + j;
            if (n < a.N) {
              float x = v[j];
              if (a.bias) x += __ldg(a.bias + n);
              if (a.relu) x = fmaxf(x, 0.f);
              o[j * plane] = x;
            }
          }
        } else {
          const int oh = 2 * px.i, ow = 2 * px.j;
          const int64_t plane = (int64_t)a.OH * a.OW;
          float* o = a.out + ((int64_t)px.b * a.OC + cc * 8) * plane + (int64_t)oh * a.OW + ow;
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            const int c = cc * 8 + g;
            if (c < a.OC) {
              float x0 = v[4 * g], x1 = v[4 * g + 1], x2 = v[4 * g + 2], x3 = v[4 * g + 3];
              if (a.bias) { const float bb = __ldg(a.bias + c); x0 += bb; x1 += bb; x2 += bb; x3 += bb; }
              if (a.relu) { x0 = fmaxf(x0, 0.f); x1 = fmaxf(x1, 0.f); x2 = fmaxf(x2, 0.f); x3 = fmaxf(x3, 0.f); }
              float* r0 = o + g * plane;
              if (vec2) {
                *reinterpret_cast<float2*>(r0) = make_float2(x0, x1);
                if (oh + 1 < a.OH) *reinterpret_cast<float2*>(r0 + a.OW) = make_float2(x2, x3);
              } else {
                r0[0] = x0;
                if (ow + 1 < a.OW) r0[1] = x1;
                if (oh + 1 < a.OH) {
                  r0[a.OW] = x2;
                  if (ow + 1 < a.OW) r0[a.OW + 1] = x3;
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
  if (warp == 8) {
    tc05::fence_after();
    tc05::tmem_dealloc(tmem_base, tmem_cols);
  }
}

}  // namespace vc

// ---- host side ----------------------------------------------------------------------------------------------------
namespace vc {

// tap-major weight slab Bw[n][k] for the other implicit-GEMM translation unit (conv_igemm_tma.cu)
int ig_build_weights(const float* w, int mode, int Cin, int Cout, int N, int BN, int K, int Kp, __nv_bfloat16* out,
                     cudaStream_t s) {
  const int total = BN * Kp;
  ig_weights_kernel<<<(total + 255) / 256, 256, 0, s>>>(w, mode, Cin, Cout, N, BN, K, Kp, out);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

struct IgPlan { int mode, C, taps, N, BN, K, Kp; bool ok; };

static IgPlan ig_plan(int op, int B, int Cin, int H, int W, int Cout) {
  IgPlan p;
  memset(&p, 0, sizeof(p));
  switch (op) {
    case VC_OP_CONV_FWD: p.mode = IG_CONV_FWD; p.C = Cin; p.taps = 9; p.N = Cout; break;
    case VC_OP_CONV_DGRAD: p.mode = IG_CONV_DGRAD; p.C = Cout; p.taps = 4; p.N = 4 * Cin; break;
    case VC_OP_CONVT_FWD: p.mode = IG_CONVT_FWD; p.C = Cin; p.taps = 9; p.N = 4 * Cout; break;
    case VC_OP_CONVT_DGRAD: p.mode = IG_CONVT_DGRAD; p.C = Cout; p.taps = 16; p.N = Cin; break;
    default: return p;
  }
  p.BN = (p.N + 31) / 32 * 32;
  p.K = p.taps * p.C;
  p.Kp = (p.K + 63) / 64 * 64;
  const int64_t px = (int64_t)B * (2 * H) * (2 * W);
  p.ok = (p.C == 1 || p.C % 32 == 0) && p.BN <= 256 && px < 2147483647LL &&
         (int64_t)B * (Cin > Cout ? Cin : Cout) * (2 * H) * (2 * W) < (1LL << 40);
  return p;
}

bool conv_igemm_supported(int op, int B, int Cin, int H, int W, int Cout) { return ig_plan(op, B, Cin, H, W, Cout).ok; }

size_t conv_igemm_workspace(int op, int B, int Cin, int H, int W, int Cout) {
  const IgPlan p = ig_plan(op, B, Cin, H, W, Cout);
  return p.ok ? (size_t)p.BN * p.Kp * 2 + 2048 : 0;
}

// in0 = activations (x or dy), in1 = weights in the state_dict layout; geometry arguments as in vc_dense_op
int conv_igemm_run(int op, const float* in0, const float* in1, const float* bias, float* out, int B, int Cin, int H,
                   int W, int Cout, int relu, void* ws, size_t wsb, cudaStream_t s) {
  const IgPlan p = ig_plan(op, B, Cin, H, W, Cout);
  VC_REQUIRE(p.ok, VC_ERR_UNSUPPORTED, "conv_igemm: unsupported shape");
  VC_REQUIRE(ws && wsb >= conv_igemm_workspace(op, B, Cin, H, W, Cout), VC_ERR_WORKSPACE,
             "conv_igemm: workspace too small");
  __nv_bfloat16* Wb =
      reinterpret_cast<__nv_bfloat16*>((reinterpret_cast<uintptr_t>(ws) + 1023) & ~(uintptr_t)1023);
  {
    const int total = p.BN * p.Kp;
    ig_weights_kernel<<<(total + 255) / 256, 256, 0, s>>>(in1, p.mode, Cin, Cout, p.N, p.BN, p.K, p.Kp, Wb);
    VC_LAUNCH_CHECK();
  }
  const int OHc = (H - 1) / 2 + 1, OWc = (W - 1) / 2 + 1;  // Conv2d output dims
  IgArgs a;
  memset(&a, 0, sizeof(a));
  a.src = in0; a.out = out; a.bias = bias; a.relu = relu;
  a.B = B; a.C = p.C; a.N = p.N; a.BN = p.BN;
  switch (p.mode) {
    case IG_CONV_FWD:     // x [B,Cin,H,W] -> y [B,Cout,OHc,OWc]
      a.SH = H; a.SW = W; a.GH = OHc; a.GW = OWc; a.S = 2; a.D0 = -1; a.TH = a.TW = 3;
      a.OC = Cout; a.OH = OHc; a.OW = OWc; a.quad = 0;
      break;
    case IG_CONV_DGRAD:   // dy [B,Cout,OHc,OWc] -> dx [B,Cin,H,W]
      a.SH = OHc; a.SW = OWc; a.GH = OHc; a.GW = OWc; a.S = 1; a.D0 = 0; a.TH = a.TW = 2;
      a.OC = Cin; a.OH = H; a.OW = W; a.quad = 1;
      break;
    case IG_CONVT_FWD:    // x [B,Cin,H,W] -> y [B,Cout,2H,2W]
      a.SH = H; a.SW = W; a.GH = H; a.GW = W; a.S = 1; a.D0 = -1; a.TH = a.TW = 3;
      a.OC = Cout; a.OH = 2 * H; a.OW = 2 * W; a.quad = 1;
      break;
    default:              // IG_CONVT_DGRAD: dy [B,Cout,2H,2W] -> dx [B,Cin,H,W]
      a.SH = 2 * H; a.SW = 2 * W; a.GH = H; a.GW = W; a.S = 2; a.D0 = -1; a.TH = a.TW = 4;
      a.OC = Cin; a.OH = H; a.OW = W; a.quad = 0;
      break;
  }
  a.M = B * a.GH * a.GW;
  a.tiles = (a.M + IG_BM - 1) / IG_BM;
  a.nkb = p.Kp / 64;
  const size_t st_bytes = IG_A_BYTES + (size_t)p.BN * 128;
  int stages = (int)((200 * 1024) / st_bytes);
  if (stages > 6) stages = 6;
  if (stages < 2) stages = 2;
  a.stages = stages;
  a.idesc = tc05::idesc(tc05::kFmtBF16, IG_BM, p.BN, 0, 0);
  CUtensorMap mw;
  {
    const uint64_t dims[2] = {(uint64_t)p.Kp, (uint64_t)p.BN};
    const uint64_t str[1] = {(uint64_t)p.Kp * 2};
    const uint32_t box[2] = {64u, (uint32_t)p.BN};
    int rc = tma::make_map(&mw, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, Wb, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc != VC_OK) return rc;
  }
  const size_t smem = (size_t)stages * st_bytes + (2 * stages + 4) * 8 + 16 + (size_t)a.nkb * 16 + 1024;
  const int grid = a.tiles < sm_count() ? a.tiles : sm_count();
  if (p.C == 1) {
    VC_CUDA(cudaFuncSetAttribute(conv_igemm_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    conv_igemm_kernel<true><<<grid, IG_THREADS, smem, s>>>(a, mw);
  } else {
    VC_CUDA(cudaFuncSetAttribute(conv_igemm_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    conv_igemm_kernel<false><<<grid, IG_THREADS, smem, s>>>(a, mw);
  }
  VC_LAUNCH_CHECK();
  return VC_OK;
}

}  // namespace vc
