// Operand-gather and epilogue functors shared by the fp32 SIMT GEMM (gemm_simt.cu) and the bf16 tcgen05 GEMM
// (tc_gemm.cu).  A(m,k) / B(k,n) express Linear, Conv2d(k3,s2,p1) and ConvTranspose2d(k4,s2,p1) forward,
// data-gradient and weight-gradient as GEMMs over torch's NCHW activations and state_dict weight layouts.
#pragma once
#include "common.cuh"

namespace vc {

// ------------------------------------------------------------------------------------------- functors
struct RowMajorK {  // element (r, k) at p[r*ld + k]  (k contiguous)
  const float* p; int64_t ld;
  static constexpr bool kFastK = true;
};
struct ARowK : RowMajorK { __device__ float operator()(int m, int k) const { return __ldg(p + (int64_t)m * ld + k); } };
struct BRowK : RowMajorK { __device__ float operator()(int k, int n) const { return __ldg(p + (int64_t)n * ld + k); } };
struct AColK {  // A(m,k) = p[k*ld + m]  (m contiguous)
  const float* p; int64_t ld;
  static constexpr bool kFastK = false;
  __device__ float operator()(int m, int k) const { return __ldg(p + (int64_t)k * ld + m); }
};
struct BColK {  // B(k,n) = p[k*ld + n]  (n contiguous)
  const float* p; int64_t ld;
  static constexpr bool kFastK = false;
  __device__ float operator()(int k, int n) const { return __ldg(p + (int64_t)k * ld + n); }
};

struct EpRowMajor {  // C[m*ld + n] = act(acc + bias[n]) (+= if accumulate)
  float* c; int64_t ld; const float* bias; int relu; int accumulate;
  __device__ void operator()(int m, int n, float v, int) const {
    if (bias) v += bias[n];
    if (relu) v = fmaxf(v, 0.f);
    float* q = c + (int64_t)m * ld + n;
    *q = accumulate ? *q + v : v;
  }
};
struct EpSplitK {  // partial[z][m*N + n]
  float* part; int64_t MN; int N;
  __device__ void operator()(int m, int n, float v, int z) const { part[(int64_t)z * MN + (int64_t)m * N + n] = v; }
};

struct ConvGeom { int B, Cin, H, W, Cout, OH, OW; };

// ---- Conv2d k3 s2 p1
struct ConvFwdA {  // m -> (b,oh,ow), k -> (ci,kh,kw)
  const float* x; ConvGeom g;
  static constexpr bool kFastK = false;
  __device__ float operator()(int m, int k) const {
    const int ow = m % g.OW, t = m / g.OW, oh = t % g.OH, b = t / g.OH;
    const int kw = k % 3, t2 = k / 3, kh = t2 % 3, ci = t2 / 3;
    const int ih = 2 * oh + kh - 1, iw = 2 * ow + kw - 1;
    if ((unsigned)ih >= (unsigned)g.H || (unsigned)iw >= (unsigned)g.W) return 0.f;
    return __ldg(x + (((int64_t)b * g.Cin + ci) * g.H + ih) * g.W + iw);
  }
};
struct EpNCHW {  // m -> (b,oh,ow) of an output with OHxOW spatial dims and C channels
  float* y; int C, OH, OW; const float* bias; int relu;
  __device__ void operator()(int m, int n, float v, int) const {
    const int ow = m % OW, t = m / OW, oh = t % OH, b = t / OH;
    if (bias) v += bias[n];
    if (relu) v = fmaxf(v, 0.f);
    y[(((int64_t)b * C + n) * OH + oh) * OW + ow] = v;
  }
};
struct ConvDgradA {  // m -> (b,h,w) input pixel, k -> (co,kh,kw); gathers dy
  const float* dy; ConvGeom g;
  static constexpr bool kFastK = false;
  __device__ float operator()(int m, int k) const {
    const int w = m % g.W, t = m / g.W, h = t % g.H, b = t / g.H;
    const int kw = k % 3, t2 = k / 3, kh = t2 % 3, co = t2 / 3;
    const int th = h + 1 - kh, tw = w + 1 - kw;
    if ((th | tw) & 1) return 0.f;
    const int oh = th >> 1, ow = tw >> 1;
    if (th < 0 || tw < 0 || oh >= g.OH || ow >= g.OW) return 0.f;
    return __ldg(dy + (((int64_t)b * g.Cout + co) * g.OH + oh) * g.OW + ow);
  }
};
struct ConvDgradB {  // B(k,n) = W[co, n, kh, kw], k = co*9 + r
  const float* w; int Cin;
  static constexpr bool kFastK = false;
  __device__ float operator()(int k, int n) const {
    const int r = k % 9, co = k / 9;
    return __ldg(w + ((int64_t)co * Cin + n) * 9 + r);
  }
};
struct ConvWgradA {  // A(m=co, k=(b,oh,ow)) = dy
  const float* dy; ConvGeom g;
  static constexpr bool kFastK = true;
  __device__ float operator()(int m, int k) const {
    const int hw = g.OH * g.OW;
    const int b = k / hw, r = k % hw;
    return __ldg(dy + ((int64_t)b * g.Cout + m) * hw + r);
  }
};
struct ConvWgradB {  // B(k=(b,oh,ow), n=(ci,kh,kw)) = x gather
  const float* x; ConvGeom g;
  static constexpr bool kFastK = true;
  __device__ float operator()(int k, int n) const {
    const int ow = k % g.OW, t = k / g.OW, oh = t % g.OH, b = t / g.OH;
    const int kw = n % 3, t2 = n / 3, kh = t2 % 3, ci = t2 / 3;
    const int ih = 2 * oh + kh - 1, iw = 2 * ow + kw - 1;
    if ((unsigned)ih >= (unsigned)g.H || (unsigned)iw >= (unsigned)g.W) return 0.f;
    return __ldg(x + (((int64_t)b * g.Cin + ci) * g.H + ih) * g.W + iw);
  }
};

// ---- ConvTranspose2d k4 s2 p1 (geometry: x [B,Cin,H,W] -> y [B,Cout,2H,2W]); forward by output parity phase:
// oh = 2*oh2+ph uses kh = 1-ph+2a, ih = oh2+ph-a (a in {0,1}); same along w.
struct ConvTFwdA {  // m -> (b,oh2,ow2), k -> (ci,a,c)
  const float* x; ConvGeom g; int ph, pw;
  static constexpr bool kFastK = false;
  __device__ float operator()(int m, int k) const {
    const int ow2 = m % g.W, t = m / g.W, oh2 = t % g.H, b = t / g.H;
    const int c = k & 1, a = (k >> 1) & 1, ci = k >> 2;
    const int ih = oh2 + ph - a, iw = ow2 + pw - c;
    if ((unsigned)ih >= (unsigned)g.H || (unsigned)iw >= (unsigned)g.W) return 0.f;
    return __ldg(x + (((int64_t)b * g.Cin + ci) * g.H + ih) * g.W + iw);
  }
};
struct ConvTFwdB {  // B(k=(ci,a,c), n=co) = W[ci,co,kh,kw]
  const float* w; int Cout, ph, pw;
  static constexpr bool kFastK = false;
  __device__ float operator()(int k, int n) const {
    const int c = k & 1, a = (k >> 1) & 1, ci = k >> 2;
    const int kh = 1 - ph + 2 * a, kw = 1 - pw + 2 * c;
    return __ldg(w + (((int64_t)ci * Cout + n) * 4 + kh) * 4 + kw);
  }
};
struct EpNCHWPhase {  // m -> (b,oh2,ow2); writes y[b,n,2oh2+ph,2ow2+pw]
  float* y; int C, H, W, ph, pw; const float* bias; int relu;
  __device__ void operator()(int m, int n, float v, int) const {
    const int ow2 = m % W, t = m / W, oh2 = t % H, b = t / H;
    if (bias) v += bias[n];
    if (relu) v = fmaxf(v, 0.f);
    y[(((int64_t)b * C + n) * (2 * H) + 2 * oh2 + ph) * (2 * W) + 2 * ow2 + pw] = v;
  }
};
struct ConvTDgradA {  // m -> (b,ih,iw), k -> (co,kh,kw); gathers dy (a plain k4 s2 p1 convolution of dy)
  const float* dy; ConvGeom g;
  static constexpr bool kFastK = false;
  __device__ float operator()(int m, int k) const {
    const int iw = m % g.W, t = m / g.W, ih = t % g.H, b = t / g.H;
    const int kw = k & 3, kh = (k >> 2) & 3, co = k >> 4;
    const int oh = 2 * ih - 1 + kh, ow = 2 * iw - 1 + kw;
    if ((unsigned)oh >= (unsigned)g.OH || (unsigned)ow >= (unsigned)g.OW) return 0.f;
    return __ldg(dy + (((int64_t)b * g.Cout + co) * g.OH + oh) * g.OW + ow);
  }
};
struct ConvTWgradA {  // A(m=ci, k=(b,ih,iw)) = x
  const float* x; ConvGeom g;
  static constexpr bool kFastK = true;
  __device__ float operator()(int m, int k) const {
    const int hw = g.H * g.W;
    const int b = k / hw, r = k % hw;
    return __ldg(x + ((int64_t)b * g.Cin + m) * hw + r);
  }
};
struct ConvTWgradB {  // B(k=(b,ih,iw), n=(co,kh,kw)) = dy gather
  const float* dy; ConvGeom g;
  static constexpr bool kFastK = true;
  __device__ float operator()(int k, int n) const {
    const int iw = k % g.W, t = k / g.W, ih = t % g.H, b = t / g.H;
    const int kw = n & 3, kh = (n >> 2) & 3, co = n >> 4;
    const int oh = 2 * ih - 1 + kh, ow = 2 * iw - 1 + kw;
    if ((unsigned)oh >= (unsigned)g.OH || (unsigned)ow >= (unsigned)g.OW) return 0.f;
    return __ldg(dy + (((int64_t)b * g.Cout + co) * g.OH + oh) * g.OW + ow);
  }
};

}  // namespace vc
