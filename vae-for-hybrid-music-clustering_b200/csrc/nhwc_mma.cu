// Tensor-core (mma.sync m16n8k16, bf16 x bf16 -> fp32) versions of the kernels at the single-channel ends of the script-12
// conv stacks (12:87 Conv2d(1 -> 32), 12:137 ConvTranspose2d(32 -> 1)) that work on the bf16 NHWC activations of
// convstack.py.  With one channel on one side these layers are 9- / 16-wide problems: far too small for a tcgen05 tile
// (a 128 x 16 UMMA is issue-bound at ~64 clocks per instruction), but as CUDA-core kernels they were instruction-bound
// at 0.3-0.4 of the HBM roof (up to 512 FMAs per pixel).  One warp-level 16 x 8 x 16 MMA replaces 2048 of those FMAs; the
// operands come straight from global memory in fragment layout:
//   * the K index of an MMA (and the N index, together with the matching operand) may be permuted freely, so every
//     thread reads / writes 8 or 16 contiguous bytes of an NHWC pixel row and a warp access covers whole 64-byte rows;
//   * the one-channel tensor is read as aligned 8-byte pairs plus scalars from L1 (it is a quarter of the traffic).
// Arithmetic is the one of every other kernel of the bf16 mode: operands rounded to bf16 (the packed conversion), exact
// products, fp32 accumulation.  C = 32 channels only (the reference's layer width); other widths keep the FFMA kernels.
#include "common.cuh"

namespace vc {

__device__ __forceinline__ void hm_mma(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                       uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t hm_pack(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float2 hm_ldg2_pred(const float* p, uint32_t pred) {   // 0 when pred == 0, never dereferenced then
  float2 v;
  asm("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %3, 0;\n\tmov.f32 %0, 0f00000000;\n\tmov.f32 %1, 0f00000000;\n\t"
      "@q ld.global.nc.v2.f32 {%0, %1}, [%2];\n\t}"
      : "=f"(v.x), "=f"(v.y)
      : "l"(p), "r"(pred));
  return v;
}
__device__ __forceinline__ uint4 hm_ldg16_pred(const void* p, uint32_t pred) {
  uint4 v;
  asm("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %5, 0;\n\tmov.b32 %0, 0;\n\tmov.b32 %1, 0;\n\tmov.b32 %2, 0;\n\tmov.b32 %3, 0;\n\t"
      "@q ld.global.nc.v4.b32 {%0, %1, %2, %3}, [%4];\n\t}"
      : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
      : "l"(p), "r"(pred));
  return v;
}
__device__ __forceinline__ uint2 hm_ldg8_pred(const void* p, uint32_t pred) {
  uint2 v;
  asm("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %3, 0;\n\tmov.b32 %0, 0;\n\tmov.b32 %1, 0;\n\t"
      "@q ld.global.nc.v2.b32 {%0, %1}, [%2];\n\t}"
      : "=r"(v.x), "=r"(v.y)
      : "l"(p), "r"(pred));
  return v;
}
__device__ __forceinline__ float hm_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float hm_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }

// (b, i, s) walks units of a [B][H][segs] grid in steps of `stride` units without a division in the loop
struct HmWalk {
  int b, i, s, db, di, ds;
  __device__ __forceinline__ void init(int u, int stride, int H, int segs) {
    s = u % segs; i = (u / segs) % H; b = u / segs / H;
    ds = stride % segs; di = (stride / segs) % H; db = stride / segs / H;
  }
  __device__ __forceinline__ void advance(int H, int segs) {
    s += ds;
    if (s >= segs) { s -= segs; ++i; }
    i += di;
    if (i >= H) { i -= H; ++b; }
    b += db;
  }
};

// ---- ConvTranspose2d(32 -> 1, k4 s2 p1) data gradient, ReLU mask of the layer's input applied, bf16 NHWC out (12:137):
//   g[b,i,j,c] = (act[b,i,j,c] > 0) * sum_{kh,kw} dy[b, 2i-1+kh, 2j-1+kw] * W[c, kh*4+kw]
// D[16 px][32 ch] = P[16 px][16 taps] * W[16 taps][32 ch]: one k-step, four n-tiles.  Thread (g = lane / 4, t = lane % 4) owns
// tap row kh = t: its k slots 2t, 2t+1 are the taps (t, kw 1), (t, kw 2) -- the aligned pair dy[.., 2j .. 2j+1] -- and slots
// 2t+8, 2t+9 the taps (t, kw 0), (t, kw 3) (two scalars); n-tile q, column n <-> channel 8 (n / 2) + 2 q + (n % 2), so a
// thread's accumulators of pixel row g are the eight consecutive channels 8t .. 8t+7: one 16-byte store (and mask load).
// BIAS: per-channel sums of the (masked, bf16-rounded) result as well -- the bias gradient of the layer below (12:136), which
// otherwise costs another pass over g; block partials in double, folded in block order by nm_colsum_fold_kernel.
template <bool BIAS>
__global__ void __launch_bounds__(256, 3)
hm_convT1_dgrad_kernel(const float* __restrict__ dy, const float* __restrict__ w, const __nv_bfloat16* __restrict__ act,
                       __nv_bfloat16* __restrict__ gout, int B, int H, int W, double* __restrict__ bpart) {
  constexpr int C = 32;
  __shared__ float bred[8][32];
  float bs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int SH = 2 * H, SW = 2 * W, segs = W >> 5;
  uint32_t bw[4][2];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float* wc = w + (8 * (g >> 1) + 2 * q + (g & 1)) * 16 + 4 * t;
    bw[q][0] = hm_pack(wc[1], wc[2]);
    bw[q][1] = hm_pack(wc[0], wc[3]);
  }
  const int units = B * H * segs;
  const int stride = (int)gridDim.x * 8;
  int u = (int)blockIdx.x * 8 + (threadIdx.x >> 5);
  HmWalk k;
  k.init(u, stride, H, segs);
  for (; u < units; u += stride, k.advance(H, segs)) {
    const int row = 2 * k.i - 1 + t;
    const uint32_t rok = (unsigned)row < (unsigned)SH ? 1u : 0u;
    const float* sp = dy + (int64_t)((k.b * SH + row) * SW);
    const int px0 = k.s * 32 + g;                           // this thread's pixels: px0 + 8 e, e = 0 .. 3
    const int64_t pbase = ((int64_t)(k.b * H + k.i) * W + px0) * C + 8 * t;
    float2 mid[4];
    float lo[4], hi[4];
    uint4 mk[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int c0 = 2 * (px0 + 8 * e);
      mid[e] = hm_ldg2_pred(sp + c0, rok);
      lo[e] = ldg_pred(sp + c0 - 1, rok & (c0 > 0 ? 1u : 0u));
      hi[e] = ldg_pred(sp + c0 + 2, rok & (c0 + 2 < SW ? 1u : 0u));
    }
    if (act) {
#pragma unroll
      for (int e = 0; e < 4; ++e) mk[e] = __ldg(reinterpret_cast<const uint4*>(act + pbase + (int64_t)e * 8 * C));
    }
#pragma unroll
    for (int m = 0; m < 2; ++m) {        // m-tile = pixels px0 + 16 m (row g) and px0 + 16 m + 8 (row g + 8)
      const int e0 = 2 * m, e1 = 2 * m + 1;
      const uint32_t a0 = hm_pack(mid[e0].x, mid[e0].y), a1 = hm_pack(mid[e1].x, mid[e1].y);
      const uint32_t a2 = hm_pack(lo[e0], hi[e0]), a3 = hm_pack(lo[e1], hi[e1]);
      float d[4][4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        d[q][0] = d[q][1] = d[q][2] = d[q][3] = 0.f;
        hm_mma(d[q], a0, a1, a2, a3, bw[q][0], bw[q][1]);
      }
#pragma unroll
      for (int r = 0; r < 2; ++r) {      // r = 0: pixel row g (accumulators 0, 1), r = 1: row g + 8 (accumulators 2, 3)
        const int e = 2 * m + r;
        float v[8];
#pragma unroll
        for (int q = 0; q < 4; ++q) { v[2 * q] = d[q][2 * r]; v[2 * q + 1] = d[q][2 * r + 1]; }
        if (act) {
          const uint32_t mw[4] = {mk[e].x, mk[e].y, mk[e].z, mk[e].w};
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            v[2 * q] = hm_lo(mw[q]) > 0.f ? v[2 * q] : 0.f;
            v[2 * q + 1] = hm_hi(mw[q]) > 0.f ? v[2 * q + 1] : 0.f;
          }
        }
        const uint4 o = make_uint4(hm_pack(v[0], v[1]), hm_pack(v[2], v[3]), hm_pack(v[4], v[5]), hm_pack(v[6], v[7]));
        *reinterpret_cast<uint4*>(gout + pbase + (int64_t)e * 8 * C) = o;
        if (BIAS) {
          const uint32_t ow[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
          for (int q = 0; q < 4; ++q) { bs[2 * q] += hm_lo(ow[q]); bs[2 * q + 1] += hm_hi(ow[q]); }
        }
      }
    }
  }
  if (BIAS) {      // lanes of equal t hold the same eight channels: add over g (fixed butterfly), then over the warps in order
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      float v = bs[c];
      v += __shfl_xor_sync(0xffffffffu, v, 4);
      v += __shfl_xor_sync(0xffffffffu, v, 8);
      v += __shfl_xor_sync(0xffffffffu, v, 16);
      if (g == 0) bred[threadIdx.x >> 5][8 * t + c] = v;
    }
    __syncthreads();
    if (threadIdx.x < 32) {
      double v = 0.0;
#pragma unroll
      for (int wq = 0; wq < 8; ++wq) v += (double)bred[wq][threadIdx.x];
      bpart[(int64_t)blockIdx.x * C + threadIdx.x] = v;
    }
  }
}

// ---- Conv2d(1 -> 32, k3 s2 p1) forward (+ bias, ReLU) writing bf16 NHWC (12:87): the same GEMM shape,
//   D[16 px][32 ch] = P[16 px][9 taps, padded to 16] * W[16][32]:  thread t < 3 owns tap row kh = t -- k slots 2t, 2t+1 = the
// aligned pair x[.., 2 ow .. 2 ow + 1] (kw 1, 2), slot 2t+8 = x[.., 2 ow - 1] (kw 0), slot 2t+9 and all of t = 3 are zero.
__global__ void __launch_bounds__(256, 3)
hm_conv1_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                    __nv_bfloat16* __restrict__ y, int B, int H, int W, int relu) {
  constexpr int C = 32;
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int OH = (H - 1) / 2 + 1, OW = W >> 1, segs = OW >> 5;
  uint32_t bw[4][2];
  float bv[8];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float* wc = w + (8 * (g >> 1) + 2 * q + (g & 1)) * 9 + 3 * (t < 3 ? t : 0);
    bw[q][0] = t < 3 ? hm_pack(wc[1], wc[2]) : 0u;
    bw[q][1] = t < 3 ? hm_pack(wc[0], 0.f) : 0u;
  }
#pragma unroll
  for (int c = 0; c < 8; ++c) bv[c] = bias ? __ldg(bias + 8 * t + c) : 0.f;
  const int units = B * OH * segs;
  const int stride = (int)gridDim.x * 8;
  int u = (int)blockIdx.x * 8 + (threadIdx.x >> 5);
  HmWalk k;
  k.init(u, stride, OH, segs);
  for (; u < units; u += stride, k.advance(OH, segs)) {
    const int row = 2 * k.i - 1 + t;
    const uint32_t rok = (t < 3 && (unsigned)row < (unsigned)H) ? 1u : 0u;
    const float* sp = x + (int64_t)((k.b * H + row) * W);
    const int px0 = k.s * 32 + g;                           // this thread's pixels: px0 + 8 e, e = 0 .. 3
    const int64_t pbase = ((int64_t)(k.b * OH + k.i) * OW + px0) * C + 8 * t;
    float2 mid[4];
    float lo[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int c0 = 2 * (px0 + 8 * e);
      mid[e] = hm_ldg2_pred(sp + c0, rok);
      lo[e] = ldg_pred(sp + c0 - 1, rok & (c0 > 0 ? 1u : 0u));
    }
#pragma unroll
    for (int m = 0; m < 2; ++m) {
      const int e0 = 2 * m, e1 = 2 * m + 1;
      const uint32_t a0 = hm_pack(mid[e0].x, mid[e0].y), a1 = hm_pack(mid[e1].x, mid[e1].y);
      const uint32_t a2 = hm_pack(lo[e0], 0.f), a3 = hm_pack(lo[e1], 0.f);
      float d[4][4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        d[q][0] = d[q][2] = bv[2 * q];
        d[q][1] = d[q][3] = bv[2 * q + 1];
        hm_mma(d[q], a0, a1, a2, a3, bw[q][0], bw[q][1]);
      }
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        float v[8];
#pragma unroll
        for (int q = 0; q < 4; ++q) { v[2 * q] = d[q][2 * r]; v[2 * q + 1] = d[q][2 * r + 1]; }
        if (relu) {
#pragma unroll
          for (int c = 0; c < 8; ++c) v[c] = fmaxf(v[c], 0.f);
        }
        *reinterpret_cast<uint4*>(y + pbase + (int64_t)(2 * m + r) * 8 * C) =
            make_uint4(hm_pack(v[0], v[1]), hm_pack(v[2], v[3]), hm_pack(v[4], v[5]), hm_pack(v[6], v[7]));
      }
    }
  }
}

// ---- weight gradients of both single-channel ends from a bf16 NHWC tensor A [B][PH][PW][C] and the one-channel fp32
// tensor S [B][2 PH][2 PW]:   out[c][kh][kw] = sum_{b,i,j} A[b,i,j,c] * S[b, 2i-1+kh, 2j-1+kw]
//   Conv2d(1 -> C) k3: A = g (gradient of the output), S = x;   ConvTranspose2d(C -> 1) k4: A = x (layer input), S = dy.
// D[16 taps][32 ch] += St[16 taps][16 px] * A[16 px][32 ch] per k-step, K = pixels: a unit is one 32-pixel row segment
// (two k-steps, four n-tiles).  K slots of thread t in k-step s: pixels 16 s + t, + 4, + 8, + 12 -- so the B fragments come from
// four 8-byte loads (channels 4g .. 4g+3 of four pixels; a warp access = four whole 64-byte pixel rows) and byte
// permutes; n-tile q, column n <-> channel 4 n + q.  The S patch of the segment is staged in shared memory per warp (one
// aligned 8-byte load per lane and tap row plus the two halo columns), the A fragment of tap row (th, tw) is
// srow[th][3 + 2 px + tw].  KS = 3 leaves taps 9 .. 15 free: tap 9 is fed the constant 1, which makes column sums of A
// -- the bias gradient of Conv2d(1 -> C) -- fall out of the same MMAs (BIAS).
// Loads of the next unit are issued before the MMAs of the current one.
template <int KS>
struct HmTapUnit {
  uint2 a[8];        // [k-step][slot]: channels 4g .. 4g+3 of pixel 16 s + 4 q + t
  float2 sv[KS];
  float sh[KS];
};
template <int KS>
__device__ __forceinline__ void hm_tap_unit_load(HmTapUnit<KS>& q, const __nv_bfloat16* __restrict__ A,
                                                 const float* __restrict__ S, int b, int i, int seg, int C, int PH, int PW,
                                                 int SH, int SW, int coff, int t, int lane) {
  const int j0 = seg * 32;
  const __nv_bfloat16* ap = A + ((int64_t)((b * PH + i) * PW + j0 + t) * C + coff);      // coff = first channel + 4 g
#pragma unroll
  for (int e = 0; e < 8; ++e) q.a[e] = __ldg(reinterpret_cast<const uint2*>(ap + (int64_t)(4 * e) * C));
  const int hcol = lane == 0 ? -1 : 64;                     // halo column of lanes 0 / 1 (relative to 2 j0)
  const uint32_t hcok = lane < 2 && (unsigned)(2 * j0 + hcol) < (unsigned)SW ? 1u : 0u;
  const int h0 = 2 * i - 1;
  const float* sb = S + (int64_t)((b * SH + h0) * SW + 2 * j0);
#pragma unroll
  for (int th = 0; th < KS; ++th) {
    const uint32_t hok = (unsigned)(h0 + th) < (unsigned)SH ? 1u : 0u;
    const float* rowp = sb + th * SW;
    q.sv[th] = hm_ldg2_pred(rowp + 2 * lane, hok);
    q.sh[th] = ldg_pred(rowp + hcol, hok & hcok);
  }
}

template <int KS, bool BIAS>
__global__ void __launch_bounds__(256, 3)
hm_tap_wgrad_kernel(const __nv_bfloat16* __restrict__ A, const float* __restrict__ S, double* __restrict__ part, int B, int C,
                    int PH, int PW, int SH, int SW, double* __restrict__ bpart) {
  constexpr int T = KS * KS, RW = 72;
  // srow[w][th][4 + c] = S[row th of the patch][2 j0 + c], c = -1 .. 64
  __shared__ __align__(16) float srow[8][KS][RW];
  __shared__ float red[8][16][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int segs = PW >> 5;
  const int units = B * PH * segs;
  const int stride = (int)gridDim.x * 8;
  const int coff = (int)blockIdx.y * 32 + 4 * g;
  float d[4][4];
#pragma unroll
  for (int q = 0; q < 4; ++q) d[q][0] = d[q][1] = d[q][2] = d[q][3] = 0.f;
  const int tap0 = g, tap1 = g + 8;
  const int th0 = tap0 / KS, tw0 = tap0 % KS, th1 = tap1 < T ? tap1 / KS : 0, tw1 = tap1 < T ? tap1 % KS : 0;
  const float* p0 = &srow[warp][th0][3 + 2 * t + tw0];
  const float* p1 = &srow[warp][th1][3 + 2 * t + tw1];
  const uint32_t ones = (BIAS && tap1 == T) ? 0x3f803f80u : 0u;      // tap 9 of KS = 3: bf16 (1, 1)
  HmTapUnit<KS> un;
  int u = (int)blockIdx.x * 8 + warp;
  HmWalk k;
  k.init(u, stride, PH, segs);
  if (u < units) hm_tap_unit_load<KS>(un, A, S, k.b, k.i, k.s, C, PH, PW, SH, SW, coff, t, lane);
  k.advance(PH, segs);
  for (; u < units; u += stride) {
    __syncwarp();                    // the previous unit's fragment reads are done
#pragma unroll
    for (int th = 0; th < KS; ++th) {
      *reinterpret_cast<float2*>(&srow[warp][th][4 + 2 * lane]) = un.sv[th];
      if (lane < 2) srow[warp][th][lane == 0 ? 3 : 68] = un.sh[th];
    }
    uint32_t bf[2][4][2];            // [k-step][n-tile][b0 / b1]
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      const uint2 l0 = un.a[4 * s], l1 = un.a[4 * s + 1], l2 = un.a[4 * s + 2], l3 = un.a[4 * s + 3];
      bf[s][0][0] = __byte_perm(l0.x, l1.x, 0x5410); bf[s][0][1] = __byte_perm(l2.x, l3.x, 0x5410);
      bf[s][1][0] = __byte_perm(l0.x, l1.x, 0x7632); bf[s][1][1] = __byte_perm(l2.x, l3.x, 0x7632);
      bf[s][2][0] = __byte_perm(l0.y, l1.y, 0x5410); bf[s][2][1] = __byte_perm(l2.y, l3.y, 0x5410);
      bf[s][3][0] = __byte_perm(l0.y, l1.y, 0x7632); bf[s][3][1] = __byte_perm(l2.y, l3.y, 0x7632);
    }
    if (u + stride < units) hm_tap_unit_load<KS>(un, A, S, k.b, k.i, k.s, C, PH, PW, SH, SW, coff, t, lane);
    k.advance(PH, segs);
    __syncwarp();
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      // A fragments: rows = taps g / g + 8, k = this thread's pixels 16 s + t + {0, 4 | 8, 12}
      const uint32_t a0 = hm_pack(p0[32 * s], p0[32 * s + 8]), a2 = hm_pack(p0[32 * s + 16], p0[32 * s + 24]);
      uint32_t a1 = hm_pack(p1[32 * s], p1[32 * s + 8]), a3 = hm_pack(p1[32 * s + 16], p1[32 * s + 24]);
      if (T < 16 && tap1 >= T) a1 = a3 = ones;
#pragma unroll
      for (int q = 0; q < 4; ++q) hm_mma(d[q], a0, a1, a2, a3, bf[s][q][0], bf[s][q][1]);
    }
  }
  // accumulators: rows g / g + 8 = taps, columns 2t, 2t + 1 of n-tile q = channels 4 (2t) + q, 4 (2t + 1) + q
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    red[warp][g][8 * t + q] = d[q][0];
    red[warp][g][8 * t + 4 + q] = d[q][1];
    red[warp][g + 8][8 * t + q] = d[q][2];
    red[warp][g + 8][8 * t + 4 + q] = d[q][3];
  }
  __syncthreads();
  for (int e = threadIdx.x; e < 32 * T; e += blockDim.x) {
    const int c = e / T, tp = e - c * T;
    double v = 0.0;
#pragma unroll
    for (int wq = 0; wq < 8; ++wq) v += (double)red[wq][tp][c];
    part[((int64_t)blockIdx.x * C + blockIdx.y * 32 + c) * T + tp] = v;
  }
  if (BIAS) {
    for (int c = threadIdx.x; c < 32; c += blockDim.x) {
      double v = 0.0;
#pragma unroll
      for (int wq = 0; wq < 8; ++wq) v += (double)red[wq][T][c];
      bpart[(int64_t)blockIdx.x * C + blockIdx.y * 32 + c] = v;
    }
  }
}

// ---- ConvTranspose2d(32 -> 1, k4 s2 p1) forward from the bf16 NHWC input (12:137), fp32 output [B][2H][2W].
// Output row 2i + ph takes input rows i + dh with kernel row kh = ph + 1 - 2 dh (dh = -1: ph 0 / kh 3; dh = 0: kh 1 / 2;
// dh = +1: ph 1 / kh 0).  Per input row source dh:  D[16 px][8 = (ph, kw)] += X_{i+dh}[16 px][32 ch] * Wdh[32 ch][8]
// (two k-steps; columns whose kh falls outside 0..3 hold zeros): six MMAs per tile into ONE accumulator tile.  The column
// part of the transposed convolution is then a neighbour sum,
//   y[2i+ph][2j] = D[j][ph, kw 1] + D[j-1][ph, kw 3],      y[2i+ph][2j+1] = D[j][ph, kw 2] + D[j+1][ph, kw 0],
// done with three warp shuffles.  A tile's 16 MMA rows are the pixels 14 jt - 1 .. 14 jt + 14: rows 0 and 15 only feed
// their neighbours, rows 1 .. 14 are written (tiles overlap by two pixels instead of exchanging edges).  The eight warps
// of a block take eight consecutive rows i of the same two column tiles, so rows i - 1 .. i + 8 are shared through L1.
__global__ void __launch_bounds__(256, 3)
hm_convT1_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                     float* __restrict__ y, int B, int H, int W) {
  constexpr int C = 32, TP = 14;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  uint32_t bw[3][2][2];     // [dh + 1][k-step][b0 / b1]; k slots 2t, 2t+1 <-> channels 8t + 4s + {0, 1}, slots + 8 <-> + {2, 3}
  {
    const int ph = g >> 2, kw = g & 3;
#pragma unroll
    for (int dh = -1; dh <= 1; ++dh) {
      const int kh = ph + 1 - 2 * dh;
      const bool ok = kh >= 0 && kh <= 3;
#pragma unroll
      for (int s = 0; s < 2; ++s)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int c = 8 * t + 4 * s + 2 * h;
          bw[dh + 1][s][h] = ok ? hm_pack(w[c * 16 + kh * 4 + kw], w[(c + 1) * 16 + kh * 4 + kw]) : 0u;
        }
    }
  }
  const float b0 = bias ? __ldg(bias) : 0.f;
  const int ntj = (W + TP - 1) / TP, npair = (ntj + 1) >> 1, hg = (H + 7) >> 3;
  const int total = B * hg * npair;
  const int OW = 2 * W;
  // shuffle sources of the neighbour sums (see above): even t takes (row - 1, kw 3) from lane (g - 1, t + 1), odd t takes
  // (row + 1, kw 0) from lane (g + 1, t - 1); rows 7 / 8 wrap between the two accumulator halves
  const bool even = (t & 1) == 0;
  const int src = even ? (((g - 1) & 7) << 2) + t + 1 : (((g + 1) & 7) << 2) + t - 1;
  // block units (b, ig, jp) walked without divisions; element offsets are 32-bit (host: B H W C < 2^31)
  HmWalk k3;
  k3.init((int)blockIdx.x, (int)gridDim.x, hg, npair);
  const int rowstride = W * C;
  for (int bu = blockIdx.x; bu < total; bu += gridDim.x, k3.advance(hg, npair)) {
    const int jp = k3.s, b = k3.b;
    const int i = k3.i * 8 + warp;
    if (i >= H) continue;                                   // warp-uniform
    uint4 xa[2][3][2];
    const int pxb = 2 * TP * jp - 1 + g;                    // pixel of (tile 0, row g)
    const __nv_bfloat16* xb = x + (uint32_t)(((b * H + i) * W) * C + 8 * t);
    const uint32_t up = i > 0 ? 1u : 0u, dn = i + 1 < H ? 1u : 0u;
    // interior tiles (no image border within the 30 pixels x 3 rows: warp-uniform) take plain loads; the predicated form
    // costs six instructions per load
    if (up && dn && jp > 0 && 2 * TP * jp + TP + 14 < W) {
#pragma unroll
      for (int k = 0; k < 2; ++k) {
#pragma unroll
        for (int r2 = 0; r2 < 2; ++r2) {
          const __nv_bfloat16* pp = xb + (pxb + TP * k + 8 * r2) * C;
          xa[k][0][r2] = __ldg(reinterpret_cast<const uint4*>(pp - rowstride));
          xa[k][1][r2] = __ldg(reinterpret_cast<const uint4*>(pp));
          xa[k][2][r2] = __ldg(reinterpret_cast<const uint4*>(pp + rowstride));
        }
      }
    } else {
#pragma unroll
      for (int k = 0; k < 2; ++k) {
#pragma unroll
        for (int r2 = 0; r2 < 2; ++r2) {
          const int px = pxb + TP * k + 8 * r2;
          const uint32_t pok = (unsigned)px < (unsigned)W && 2 * jp + k < ntj ? 1u : 0u;
          const __nv_bfloat16* pp = xb + px * C;
          xa[k][0][r2] = hm_ldg16_pred(pp - rowstride, pok & up);
          xa[k][1][r2] = hm_ldg16_pred(pp, pok);
          xa[k][2][r2] = hm_ldg16_pred(pp + rowstride, pok & dn);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int jt = 2 * jp + k;
      if (jt >= ntj) break;                                 // warp-uniform
      float d[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int q = 0; q < 3; ++q) {
        hm_mma(d, xa[k][q][0].x, xa[k][q][1].x, xa[k][q][0].y, xa[k][q][1].y, bw[q][0][0], bw[q][0][1]);
        hm_mma(d, xa[k][q][0].z, xa[k][q][1].z, xa[k][q][0].w, xa[k][q][1].w, bw[q][1][0], bw[q][1][1]);
      }
      // d[0], d[1] = (row g; kw 2(t&1), 2(t&1)+1 of phase t >> 1), d[2], d[3] the same of row g + 8
      const float send1 = even ? (g == 0 ? d[2] : d[0]) : d[1];
      const float send2 = even ? d[2] : (g == 7 ? d[1] : d[3]);
      const float recv1 = __shfl_sync(0xffffffffu, send1, src);
      const float recv2 = __shfl_sync(0xffffffffu, send2, src);
      const float yA = (even ? d[1] : d[0]) + recv1 + b0;    // row g:     even -> column 2j, odd -> column 2j + 1
      const float yB = (even ? d[3] : d[2]) + recv2 + b0;    // row g + 8
      const float other = __shfl_xor_sync(0xffffffffu, even ? yB : yA, 1);
      // even lanes write the pair of row g, odd lanes the pair of row g + 8
      const int r = even ? g : g + 8;
      const int j = TP * jt - 1 + r;
      if (r >= 1 && r <= TP && j < W) {
        float* yo = y + ((int64_t)(b * 2 * H + 2 * i + (t >> 1)) * OW + 2 * j);
        *reinterpret_cast<float2*>(yo) = even ? make_float2(yA, other) : make_float2(other, yB);
      }
    }
  }
}

// ---- host -----------------------------------------------------------------------------------------------------------
static bool hm_enabled() {
  static const bool off = getenv("VAECLUST_NHWC_MMA") && atoi(getenv("VAECLUST_NHWC_MMA")) == 0;
  return !off;
}
bool nhwc_mma_convT1_dgrad_supported(int B, int Cin, int H, int W) {
  return hm_enabled() && Cin == 32 && W % 32 == 0 && (int64_t)B * H * W * 4 < 2147483647LL;
}
// bpart (optional): [grid][32] block partials of the per-channel sums; *grid_out = blocks launched (<= 3 * 148)
int nhwc_mma_convT1_dgrad_run(const float* dy, const float* w, const __nv_bfloat16* act, __nv_bfloat16* g, int B, int H, int W,
                              double* bpart, int* grid_out, cudaStream_t s) {
  const int64_t units = (int64_t)B * H * (W / 32);
  int grid = 3 * sm_count();
  if (grid > 3 * 148) grid = 3 * 148;
  if ((int64_t)grid * 8 > units) grid = (int)((units + 7) / 8);
  if (bpart)
    hm_convT1_dgrad_kernel<true><<<grid, 256, 0, s>>>(dy, w, act, g, B, H, W, bpart);
  else
    hm_convT1_dgrad_kernel<false><<<grid, 256, 0, s>>>(dy, w, act, g, B, H, W, nullptr);
  VC_LAUNCH_CHECK();
  if (grid_out) *grid_out = grid;
  return VC_OK;
}
bool nhwc_mma_conv1_fwd_supported(int B, int H, int W, int Cout) {
  return hm_enabled() && Cout == 32 && W % 64 == 0 && (int64_t)B * H * W < 2147483647LL;
}
int nhwc_mma_conv1_fwd_run(const float* x, const float* w, const float* bias, __nv_bfloat16* y, int B, int H, int W, int relu,
                           cudaStream_t s) {
  const int64_t units = (int64_t)B * ((H - 1) / 2 + 1) * (W / 64);
  int grid = 3 * sm_count();
  if ((int64_t)grid * 8 > units) grid = (int)((units + 7) / 8);
  hm_conv1_fwd_kernel<<<grid, 256, 0, s>>>(x, w, bias, y, B, H, W, relu);
  VC_LAUNCH_CHECK();
  return VC_OK;
}
bool nhwc_mma_convT1_fwd_supported(int B, int Cin, int H, int W) {
  return hm_enabled() && Cin == 32 && (int64_t)B * H * W < (2147483647LL >> 5);
}
int nhwc_mma_convT1_fwd_run(const __nv_bfloat16* x, const float* w, const float* bias, float* y, int B, int H, int W,
                            cudaStream_t s) {
  const int ntj = (W + 13) / 14;
  const int64_t total = (int64_t)B * ((H + 7) / 8) * ((ntj + 1) / 2);
  int grid = 3 * sm_count();
  if (grid > total) grid = (int)total;
  hm_convT1_fwd_kernel<<<grid, 256, 0, s>>>(x, w, bias, y, B, H, W);
  VC_LAUNCH_CHECK();
  return VC_OK;
}
bool nhwc_mma_tap_wgrad_supported(int B, int C, int PH, int PW) {
  return hm_enabled() && C % 32 == 0 && PW % 32 == 0 && (int64_t)B * PH * PW * 4 < 2147483647LL &&
         (int64_t)B * PH * PW * C < (1LL << 40);
}
// part [nblk][C][KS*KS] (+ bpart [nblk][C] when the bias gradient is wanted, KS = 3 only); returns the block count used
int nhwc_mma_tap_wgrad_run(int KS, const __nv_bfloat16* a, const float* S, double* part, double* bpart, int B, int C, int PH,
                           int PW, int max_blocks, int* nblk_out, cudaStream_t s) {
  const int64_t units = (int64_t)B * PH * (PW / 32);
  int nblk = 3 * sm_count() / (C / 32);
  if (nblk > max_blocks) nblk = max_blocks;
  if (nblk < 1) nblk = 1;
  if ((int64_t)nblk * 8 > units) nblk = (int)((units + 7) / 8);
  const dim3 grid(nblk, C / 32);
  if (KS == 4)
    hm_tap_wgrad_kernel<4, false><<<grid, 256, 0, s>>>(a, S, part, B, C, PH, PW, 2 * PH, 2 * PW, nullptr);
  else if (bpart)
    hm_tap_wgrad_kernel<3, true><<<grid, 256, 0, s>>>(a, S, part, B, C, PH, PW, 2 * PH, 2 * PW, bpart);
  else
    hm_tap_wgrad_kernel<3, false><<<grid, 256, 0, s>>>(a, S, part, B, C, PH, PW, 2 * PH, 2 * PW, nullptr);
  VC_LAUNCH_CHECK();
  *nblk_out = nblk;
  return VC_OK;
}

}  // namespace vc
