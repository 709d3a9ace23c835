// Direct (CUDA-core) kernels for the single-channel ends of the script-12 VAE (12:87, 12:137):
//   Conv2d(1 -> Cout, k3 s2 p1) forward, ConvTranspose2d(Cin -> 1, k4 s2 p1) forward and its data-gradient.
// With one channel on one side the GEMM has K = 9 (or N = 1, or K = 16): a tensor-core tile would be >85 % padding and
// the layer is bound by the 268 MB activation on the many-channel side, so these run as streaming FFMA kernels:
// thread = pixel (coalesced along W), the few input values of the pixel live in registers, the small weight block is
// staged in shared memory and read as broadcast float4s, every output plane is written as full 128-byte lines.
// In bf16 mode operands are rounded to bf16 first (fp32 accumulate), i.e. the arithmetic of the tensor-core path.
#include "common.cuh"

namespace vc {

__device__ __forceinline__ float rb(float v, int round_bf16) {
  return round_bf16 ? __bfloat162float(__float2bfloat16_rn(v)) : v;
}
// two values through one packed conversion (F2FP.BF16.F32.PACK_AB on the ALU pipe + two bit operations) instead of
// two scalar F2F conversions, which issue at a quarter of the rate and were 0.045 ms of the Cin -> 1 deconvolution
__device__ __forceinline__ void rb2(float& a, float& b, int round_bf16) {
  if (round_bf16) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);   // .x = a (low half), .y = b (high half)
    const uint32_t u = *reinterpret_cast<const uint32_t*>(&h);
    a = __uint_as_float(u << 16);
    b = __uint_as_float(u & 0xffff0000u);
  }
}

// ---- Conv2d(1 -> Cout): y[b,co,oh,ow] = act(bias[co] + sum_{kh,kw} x[b,0,2oh+kh-1,2ow+kw-1] * W[co,0,kh,kw])
// shared weights transposed to wt[tap][co] (co padded to a multiple of 4) so four channels come from one LDS.128
template <int CT>   // channels per register tile (multiple of 4)
__global__ void __launch_bounds__(256)
conv1ch_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                   float* __restrict__ y, int B, int H, int W, int Cout, int OH, int OW, int relu, int round_bf16) {
  extern __shared__ float sm[];
  const int Cp = (Cout + 3) & ~3;
  float* wt = sm;              // [9][Cp]
  float* bs = sm + 9 * Cp;     // [Cp]
  for (int e = threadIdx.x; e < 9 * Cp; e += blockDim.x) {
    const int t = e / Cp, co = e - t * Cp;
    wt[e] = co < Cout ? rb(w[co * 9 + t], round_bf16) : 0.f;
  }
  for (int e = threadIdx.x; e < Cp; e += blockDim.x) bs[e] = (bias && e < Cout) ? bias[e] : 0.f;
  __syncthreads();
  const int64_t total = (int64_t)B * OH * OW;
  const int64_t plane = (int64_t)OH * OW;
  for (int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; m < total; m += (int64_t)gridDim.x * blockDim.x) {
    const int ow = (int)(m % OW);
    const int64_t t2 = m / OW;
    const int oh = (int)(t2 % OH), b = (int)(t2 / OH);
    float xin[9];
    const float* xb = x + (int64_t)b * H * W;
#pragma unroll
    for (int kh = 0; kh < 3; ++kh)
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int ih = 2 * oh + kh - 1, iw = 2 * ow + kw - 1;
        const uint32_t ok = ((unsigned)ih < (unsigned)H && (unsigned)iw < (unsigned)W) ? 1u : 0u;
        xin[kh * 3 + kw] = rb(ldg_pred(xb + (int64_t)ih * W + iw, ok), round_bf16);
      }
    float* yo = y + (int64_t)b * Cout * plane + (int64_t)oh * OW + ow;
    for (int c0 = 0; c0 < Cout; c0 += CT) {
      float acc[CT];
#pragma unroll
      for (int c = 0; c < CT; c += 4) {
        const float4 bb = *reinterpret_cast<const float4*>(bs + min(c0 + c, Cp - 4));
        acc[c] = bb.x; acc[c + 1] = bb.y; acc[c + 2] = bb.z; acc[c + 3] = bb.w;
      }
#pragma unroll
      for (int t = 0; t < 9; ++t)
#pragma unroll
        for (int c = 0; c < CT; c += 4) {
          const float4 ww = *reinterpret_cast<const float4*>(wt + t * Cp + min(c0 + c, Cp - 4));
          acc[c] = fmaf(xin[t], ww.x, acc[c]);
          acc[c + 1] = fmaf(xin[t], ww.y, acc[c + 1]);
          acc[c + 2] = fmaf(xin[t], ww.z, acc[c + 2]);
          acc[c + 3] = fmaf(xin[t], ww.w, acc[c + 3]);
        }
#pragma unroll
      for (int c = 0; c < CT; ++c)
        if (c0 + c < Cout) yo[(int64_t)(c0 + c) * plane] = relu ? fmaxf(acc[c], 0.f) : acc[c];
    }
  }
}

// ---- ConvTranspose2d(Cin -> 1) forward.  Thread = two vertically adjacent input pixels (2*i2, 2*i2+1; j): a 4x3 patch
// of x per channel feeds the 4x2 output block y[4*i2 .. 4*i2+3, 2j .. 2j+1].  Output row 2i+ph uses input rows
// i+ph-1 (kernel row 3-... see below) and i+ph:  y[2i+ph, 2j+pw] += x[i+dh, j+dw] * W[1+ph-2dh, 1+pw-2dw].
__global__ void __launch_bounds__(256, 4)
convT_to1_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                     float* __restrict__ y, int B, int Cin, int H, int W, int relu, int round_bf16) {
  extern __shared__ float sm[];   // [Cin][16]
  for (int e = threadIdx.x; e < Cin * 16; e += blockDim.x) sm[e] = rb(w[e], round_bf16);
  __syncthreads();
  const int H2 = (H + 1) / 2;
  const int64_t total = (int64_t)B * H2 * W;
  const int64_t plane = (int64_t)H * W;
  const float b0 = bias ? bias[0] : 0.f;
  for (int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; m < total; m += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(m % W);
    const int64_t t2 = m / W;
    const int i2 = (int)(t2 % H2), b = (int)(t2 / H2);
    const int i = 2 * i2;
    // patch rows i-1 .. i+2, columns j-1 .. j+1
    uint32_t rok[4], cok[3];
#pragma unroll
    for (int r = 0; r < 4; ++r) rok[r] = (unsigned)(i - 1 + r) < (unsigned)H ? 1u : 0u;
#pragma unroll
    for (int c = 0; c < 3; ++c) cok[c] = (unsigned)(j - 1 + c) < (unsigned)W ? 1u : 0u;
    float acc[4][2];
#pragma unroll
    for (int r = 0; r < 4; ++r) { acc[r][0] = b0; acc[r][1] = b0; }
    const float* xb = x + (int64_t)b * Cin * plane + (int64_t)(i - 1) * W + (j - 1);
#pragma unroll 2
    for (int ci = 0; ci < Cin; ++ci) {
      float p[4][3];
      const float* xc = xb + (int64_t)ci * plane;
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) p[r][c] = ldg_pred(xc + r * W + c, rok[r] & cok[c]);
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) p[r][c] = rb(p[r][c], round_bf16);
      const float4* wq = reinterpret_cast<const float4*>(sm + ci * 16);
      const float4 w0 = wq[0], w1 = wq[1], w2 = wq[2], w3 = wq[3];   // kernel rows kh = 0..3, .x-.w = kw 0..3
      // output row 2*(i+a)+ph, a in {0,1}; patch row index of input row i+a+dh is 1+a+dh
#pragma unroll
      for (int a = 0; a < 2; ++a) {
        const float* up = p[a];       // input row i+a-1
        const float* md = p[a + 1];   // input row i+a
        const float* dn = p[a + 2];   // input row i+a+1
        // ph = 0: rows (dh=-1 -> kh=3), (dh=0 -> kh=1);  ph = 1: (dh=0 -> kh=2), (dh=1 -> kh=0)
        // pw = 0: cols (dw=-1 -> kw=3), (dw=0 -> kw=1);  pw = 1: (dw=0 -> kw=2), (dw=1 -> kw=0)
        float s00 = acc[2 * a][0], s01 = acc[2 * a][1], s10 = acc[2 * a + 1][0], s11 = acc[2 * a + 1][1];
        s00 = fmaf(up[0], w3.w, s00); s00 = fmaf(up[1], w3.y, s00); s00 = fmaf(md[0], w1.w, s00); s00 = fmaf(md[1], w1.y, s00);
        s01 = fmaf(up[1], w3.z, s01); s01 = fmaf(up[2], w3.x, s01); s01 = fmaf(md[1], w1.z, s01); s01 = fmaf(md[2], w1.x, s01);
        s10 = fmaf(md[0], w2.w, s10); s10 = fmaf(md[1], w2.y, s10); s10 = fmaf(dn[0], w0.w, s10); s10 = fmaf(dn[1], w0.y, s10);
        s11 = fmaf(md[1], w2.z, s11); s11 = fmaf(md[2], w2.x, s11); s11 = fmaf(dn[1], w0.z, s11); s11 = fmaf(dn[2], w0.x, s11);
        acc[2 * a][0] = s00; acc[2 * a][1] = s01; acc[2 * a + 1][0] = s10; acc[2 * a + 1][1] = s11;
      }
    }
    const int OW = 2 * W;
    float* yo = y + (int64_t)b * 4 * plane + (int64_t)(2 * i) * OW + 2 * j;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      if (i + (r >> 1) < H) {
        float v0 = acc[r][0], v1 = acc[r][1];
        if (relu) { v0 = fmaxf(v0, 0.f); v1 = fmaxf(v1, 0.f); }
        *reinterpret_cast<float2*>(yo + (int64_t)r * OW) = make_float2(v0, v1);
      }
    }
  }
}

// ---- the same operator, thread = 2 input rows x 4 input columns (W % 4 == 0, 16-byte aligned x): the kernel above
// issues 12 scalar loads per channel for two input pixels and is bound by the load/store unit's one warp-load per clock
// (50 M warp loads = 0.18 ms of its 0.21 ms at 64 x 32 x 64 x 512); here a 4 x 6 patch is one 16-byte load plus two
// halo scalars per row, i.e. the same 12 load instructions feed four times the pixels.  The FMA order of every output
// is the one of the kernel above, so both produce identical bits.
__global__ void __launch_bounds__(256, 2)
convT_to1_fwd_quad_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                          float* __restrict__ y, int B, int Cin, int H, int W, int relu, int round_bf16) {
  extern __shared__ float sm[];   // [Cin][16]
  for (int e = threadIdx.x; e < Cin * 16; e += blockDim.x) sm[e] = rb(w[e], round_bf16);
  __syncthreads();
  const int H2 = (H + 1) / 2, W4 = W >> 2;
  const int64_t total = (int64_t)B * H2 * W4;
  const int64_t plane = (int64_t)H * W;
  const float b0 = bias ? bias[0] : 0.f;
  for (int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; m < total; m += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(m % W4) * 4;
    const int64_t t2 = m / W4;
    const int i2 = (int)(t2 % H2), b = (int)(t2 / H2);
    const int i = 2 * i2;
    bool rok[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) rok[r] = (unsigned)(i - 1 + r) < (unsigned)H;
    const uint32_t lok = j > 0 ? 1u : 0u, hok = j + 4 < W ? 1u : 0u;
    float acc[4][8];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 8; ++c) acc[r][c] = b0;
    const float* xb = x + (int64_t)b * Cin * plane + (int64_t)(i - 1) * W + j;
    // (an explicit register pipeline over channels -- loads of ci+1 under the FMAs of ci -- measured slower than letting
    // the compiler overlap two unrolled channels: 0.127 vs 0.115 ms)
#pragma unroll 2
    for (int ci = 0; ci < Cin; ++ci) {
      float p[4][6];   // input rows i-1 .. i+2, columns j-1 .. j+4
      const float* xc = xb + (int64_t)ci * plane;
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const float* xr = xc + r * W;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (rok[r]) v = __ldg(reinterpret_cast<const float4*>(xr));
        p[r][0] = ldg_pred(xr - 1, (rok[r] ? 1u : 0u) & lok);
        p[r][1] = v.x; p[r][2] = v.y; p[r][3] = v.z; p[r][4] = v.w;
        p[r][5] = ldg_pred(xr + 4, (rok[r] ? 1u : 0u) & hok);
      }
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 6; c += 2) rb2(p[r][c], p[r][c + 1], round_bf16);
      const float4* wq = reinterpret_cast<const float4*>(sm + ci * 16);
      const float4 w0 = wq[0], w1 = wq[1], w2 = wq[2], w3 = wq[3];   // kernel rows kh = 0..3, .x-.w = kw 0..3
#pragma unroll
      for (int a = 0; a < 2; ++a) {
        const float* up = p[a];       // input row i+a-1
        const float* md = p[a + 1];   // input row i+a
        const float* dn = p[a + 2];   // input row i+a+1
#pragma unroll
        for (int c = 0; c < 4; ++c) {   // input column j+c: patch columns c (left), c+1 (centre), c+2 (right)
          float s00 = acc[2 * a][2 * c], s01 = acc[2 * a][2 * c + 1];
          float s10 = acc[2 * a + 1][2 * c], s11 = acc[2 * a + 1][2 * c + 1];
          s00 = fmaf(up[c], w3.w, s00); s00 = fmaf(up[c + 1], w3.y, s00); s00 = fmaf(md[c], w1.w, s00); s00 = fmaf(md[c + 1], w1.y, s00);
          s01 = fmaf(up[c + 1], w3.z, s01); s01 = fmaf(up[c + 2], w3.x, s01); s01 = fmaf(md[c + 1], w1.z, s01); s01 = fmaf(md[c + 2], w1.x, s01);
          s10 = fmaf(md[c], w2.w, s10); s10 = fmaf(md[c + 1], w2.y, s10); s10 = fmaf(dn[c], w0.w, s10); s10 = fmaf(dn[c + 1], w0.y, s10);
          s11 = fmaf(md[c + 1], w2.z, s11); s11 = fmaf(md[c + 2], w2.x, s11); s11 = fmaf(dn[c + 1], w0.z, s11); s11 = fmaf(dn[c + 2], w0.x, s11);
          acc[2 * a][2 * c] = s00; acc[2 * a][2 * c + 1] = s01; acc[2 * a + 1][2 * c] = s10; acc[2 * a + 1][2 * c + 1] = s11;
        }
      }
    }
    const int OW = 2 * W;
    float* yo = y + (int64_t)b * 4 * plane + (int64_t)(2 * i) * OW + 2 * j;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      if (i + (r >> 1) < H) {
        if (relu) {
#pragma unroll
          for (int c = 0; c < 8; ++c) acc[r][c] = fmaxf(acc[r][c], 0.f);
        }
        float4* o = reinterpret_cast<float4*>(yo + (int64_t)r * OW);
        o[0] = make_float4(acc[r][0], acc[r][1], acc[r][2], acc[r][3]);
        o[1] = make_float4(acc[r][4], acc[r][5], acc[r][6], acc[r][7]);
      }
    }
  }
}

// ---- ConvTranspose2d(Cin -> 1) data-gradient: dx[b,ci,i,j] = sum_{kh,kw} dy[b,0,2i-1+kh,2j-1+kw] * W[ci,0,kh,kw]
__global__ void __launch_bounds__(256)
convT_to1_dgrad_kernel(const float* __restrict__ dy, const float* __restrict__ w, float* __restrict__ dx, int B, int Cin,
                       int H, int W, int round_bf16) {
  extern __shared__ float sm[];   // [Cin][16]
  for (int e = threadIdx.x; e < Cin * 16; e += blockDim.x) sm[e] = rb(w[e], round_bf16);
  __syncthreads();
  const int64_t plane = (int64_t)H * W;
  const int64_t total = (int64_t)B * plane;
  const int OH = 2 * H, OW = 2 * W;
  for (int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; m < total; m += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(m % W);
    const int64_t t2 = m / W;
    const int i = (int)(t2 % H), b = (int)(t2 / H);
    float g[16];
    const float* db = dy + (int64_t)b * OH * OW;
#pragma unroll
    for (int kh = 0; kh < 4; ++kh)
#pragma unroll
      for (int kw = 0; kw < 4; ++kw) {
        const int oh = 2 * i - 1 + kh, ow = 2 * j - 1 + kw;
        const uint32_t ok = ((unsigned)oh < (unsigned)OH && (unsigned)ow < (unsigned)OW) ? 1u : 0u;
        g[kh * 4 + kw] = rb(ldg_pred(db + (int64_t)oh * OW + ow, ok), round_bf16);
      }
    float* o = dx + (int64_t)b * Cin * plane + (int64_t)i * W + j;
#pragma unroll 4
    for (int ci = 0; ci < Cin; ++ci) {
      const float4* wq = reinterpret_cast<const float4*>(sm + ci * 16);
      float acc = 0.f;
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const float4 ww = wq[r];
        acc = fmaf(g[4 * r], ww.x, acc); acc = fmaf(g[4 * r + 1], ww.y, acc);
        acc = fmaf(g[4 * r + 2], ww.z, acc); acc = fmaf(g[4 * r + 3], ww.w, acc);
      }
      o[(int64_t)ci * plane] = acc;
    }
  }
}

// ---- weight gradients of the two single-channel ends.  Both are a correlation of a C-channel tensor A [B,C,PH,PW]
// with KS x KS strided taps of a ONE-channel tensor S [B,1,SH,SW], summed over every pixel:
//   out[c][kh][kw] = sum_{b,ph,pw} A[b,c,ph,pw] * S[b,0,2ph-1+kh,2pw-1+kw]
// Conv2d(1 -> C) k3:          A = dy, S = x  (KS = 3);   ConvTranspose2d(C -> 1) k4: A = x, S = dy (KS = 4).
// As an implicit GEMM this is a 9- or 16-wide problem on a 128-wide tensor-core tile (0.13 ms each); here it is a
// streaming reduction bound by reading A once (268 MB): thread = four horizontally adjacent pixels of CG channels,
// A as one 16-byte load per channel, a row of S as two 16-byte loads plus halo scalars, CG*KS*KS register accumulators
// kept across the thread's whole pixel range, then a fixed-order warp / grid reduction in double.
template <int KS, int CG, bool BIAS>
__global__ void __launch_bounds__(256, 2)
tap_wgrad_kernel(const float* __restrict__ A, const float* __restrict__ S, double* __restrict__ part, int B, int C,
                 int PH, int PW, int SH, int SW, int round_bf16, double* __restrict__ bpart) {
  constexpr int T = KS * KS, NS = KS + 6;
  // warp = channel group, lane = pixel quad: the eight warps of a block walk the SAME 32 quads, so the S patch (160 of
  // the 224 bytes a thread loads per step) comes from L2 once per block and from L1 for the other seven warps.  With
  // the group on blockIdx.y instead, S was re-read from L2 by every group and the kernel sat at the L2 bandwidth
  // (7.3 TB/s of L2->SM traffic for 2.3 TB/s of useful HBM traffic).
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c0 = (blockIdx.y * 8 + warp) * CG;
  if (c0 >= C) return;
  const int PW4 = PW >> 2;
  const int64_t total = (int64_t)B * PH * PW4;
  const int64_t aplane = (int64_t)PH * PW, splane = (int64_t)SH * SW;
  float acc[CG][T];
#pragma unroll
  for (int c = 0; c < CG; ++c)
#pragma unroll
    for (int t = 0; t < T; ++t) acc[c][t] = 0.f;
  float bsum[CG];   // per-channel sums of A (unrounded): the bias gradient when A is the output gradient (bpart != null)
#pragma unroll
  for (int c = 0; c < CG; ++c) bsum[c] = 0.f;
  for (int64_t m = (int64_t)blockIdx.x * 32 + lane; m < total; m += (int64_t)gridDim.x * 32) {
    const int pw0 = (int)(m % PW4) * 4;
    const int64_t t2 = m / PW4;
    const int ph = (int)(t2 % PH), b = (int)(t2 / PH);
    // every load of the iteration is issued before the first FMA (one memory round trip per iteration, not one per
    // tap row): A as CG 16-byte loads, each tap row of S as two 16-byte loads plus the two halo scalars
    const float* ap = A + ((int64_t)b * C + c0) * aplane + (int64_t)ph * PW + pw0;
    const float* sp = S + (int64_t)b * splane + 2 * pw0;
    const uint32_t lok = pw0 > 0 ? 1u : 0u, hok = 2 * pw0 + 8 < SW ? 1u : 0u;
    float4 av[CG], s0[KS], s1[KS];
    float sl[KS], sh[KS];
#pragma unroll
    for (int c = 0; c < CG; ++c) av[c] = __ldcs(reinterpret_cast<const float4*>(ap + c * aplane));
#pragma unroll
    for (int r = 0; r < KS; ++r) {
      const int sr = 2 * ph - 1 + r;
      const bool rok = (unsigned)sr < (unsigned)SH;
      const float* srow = sp + (int64_t)sr * SW;
      s0[r] = make_float4(0.f, 0.f, 0.f, 0.f);
      s1[r] = s0[r];
      if (rok) {
        s0[r] = __ldg(reinterpret_cast<const float4*>(srow));
        s1[r] = __ldg(reinterpret_cast<const float4*>(srow + 4));
      }
      sl[r] = ldg_pred(srow - 1, (rok ? 1u : 0u) & lok);
      sh[r] = KS == 4 ? ldg_pred(srow + 8, (rok ? 1u : 0u) & hok) : 0.f;
    }
    float a[CG][4];
#pragma unroll
    for (int c = 0; c < CG; ++c) {
      a[c][0] = av[c].x; a[c][1] = av[c].y; a[c][2] = av[c].z; a[c][3] = av[c].w;
      if (BIAS) bsum[c] += (av[c].x + av[c].y) + (av[c].z + av[c].w);
      rb2(a[c][0], a[c][1], round_bf16);
      rb2(a[c][2], a[c][3], round_bf16);
    }
#pragma unroll
    for (int r = 0; r < KS; ++r) {
      float s[NS + 1];   // S columns 2*pw0-1 .. 2*pw0+KS+4 (padded to an even count for the paired rounding)
      s[0] = sl[r];
      s[1] = s0[r].x; s[2] = s0[r].y; s[3] = s0[r].z; s[4] = s0[r].w;
      s[5] = s1[r].x; s[6] = s1[r].y; s[7] = s1[r].z; s[8] = s1[r].w;
      s[9] = sh[r];
#pragma unroll
      for (int q = 0; q < 10; q += 2) rb2(s[q], s[q + 1], round_bf16);
#pragma unroll
      for (int c = 0; c < CG; ++c)
#pragma unroll
        for (int kw = 0; kw < KS; ++kw) {
          float v = acc[c][r * KS + kw];
#pragma unroll
          for (int px = 0; px < 4; ++px) v = fmaf(a[c][px], s[2 * px + kw], v);
          acc[c][r * KS + kw] = v;
        }
    }
  }
  // fixed-order reduction: lanes (xor butterfly) here, blocks in order in the fold kernel, in double
  double* pp = part + ((int64_t)blockIdx.x * C + c0) * T;
#pragma unroll
  for (int c = 0; c < CG; ++c)
#pragma unroll
    for (int t = 0; t < T; ++t) {
      float v = acc[c][t];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) pp[c * T + t] = (double)v;
    }
  if (BIAS) {
#pragma unroll
    for (int c = 0; c < CG; ++c) {
      float v = bsum[c];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) bpart[(int64_t)blockIdx.x * C + c0 + c] = (double)v;
    }
  }
}

// ---- the same correlation on the tensor cores (bf16 mode): per 32-pixel row segment and 16 channels,
//   D[16 ch][16 taps] += A[16 ch x 32 px] * Sim2col[32 px x 16 taps]      (2 k-steps x 2 tap tiles of mma.sync m16n8k16)
// instead of 16 x 32 x T scalar FMAs -- the FFMA kernel above is instruction-issue bound (25 instructions per element of A).
// The K index (pixel) may be permuted freely as long as both fragments agree: thread (g, t) owns the EIGHT consecutive
// pixels 8t .. 8t+7 of the segment, so A comes as two 16-byte loads per channel row (rows g and g+8); pixel 8t+e sits in
// k-step e >> 2 at fragment column 2t + (e & 1) + 8 ((e >> 1) & 1).  The S patch of the segment (KS rows x 66..67 values,
// bf16-rounded) is staged in shared memory per warp; B fragment element (k, n = tap) = S[2 p - 1 + tw] of row th.
// BIAS: per-channel sums of the unrounded A (Conv2d(1 -> C): A is the output gradient).
__device__ __forceinline__ void tap_mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t tap_pack(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

__device__ __forceinline__ float2 ldg_pred2(const float* p, uint32_t pred) {   // 8-byte variant of ldg_pred
  float2 v;
  asm("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %3, 0;\n\tmov.f32 %0, 0f00000000;\n\tmov.f32 %1, 0f00000000;\n\t"
      "@q ld.global.nc.v2.f32 {%0, %1}, [%2];\n\t}"
      : "=f"(v.x), "=f"(v.y)
      : "l"(p), "r"(pred));
  return v;
}

// One unit = one 32-pixel row segment x 16 channels.  Its loads (A: four 16-byte loads per thread; S: per tap row one
// 8-byte load per lane for columns 2 j0 .. 2 j0 + 63 and one scalar for the halo columns 2 j0 - 1 (lane 0) and
// 2 j0 + 64 (lane 1)) are all issued before anything waits on them, and the loads of the NEXT unit go out before the
// MMAs of this one (one memory round trip per unit, overlapped with the previous unit's arithmetic).
template <int KS>
struct TapUnit {
  float4 x0, x1, y0, y1;
  float2 sv[KS];
  float sh[KS];
};
// (b, i, seg) = image, row and 32-pixel segment of the unit; offsets are 32-bit (host: both tensors < 2^31 elements)
template <int KS>
__device__ __forceinline__ void tap_unit_load(TapUnit<KS>& q, const float* __restrict__ A, const float* __restrict__ S, int b,
                                              int i, int seg, int C, int PH, int PW, int SH, int SW, int cg, int t, int lane) {
  const int j0 = seg * 32;
  const int aplane = PH * PW;
  const float* ap = A + (uint32_t)(((b * C + cg) * PH + i) * PW + j0 + 8 * t);      // cg = first channel of the tile + g
  q.x0 = __ldcs(reinterpret_cast<const float4*>(ap));
  q.x1 = __ldcs(reinterpret_cast<const float4*>(ap) + 1);
  q.y0 = __ldcs(reinterpret_cast<const float4*>(ap + (uint32_t)(8 * aplane)));
  q.y1 = __ldcs(reinterpret_cast<const float4*>(ap + (uint32_t)(8 * aplane)) + 1);
  // row 2i - 1 of the patch may lie above the image (i = 0), row 2i - 1 + (KS - 1) below it (KS = 4, i = PH - 1): the
  // offset of a row that does not exist is never dereferenced (predicated loads), so it may wrap
  const int hcol = lane == 0 ? -1 : 64;                     // halo column of lanes 0 / 1 (relative to 2 j0)
  const uint32_t hcok = lane < 2 && (unsigned)(2 * j0 + hcol) < (unsigned)SW ? 1u : 0u;
  const int h0 = 2 * i - 1;
  const float* sb = S + (int64_t)((b * SH + h0) * SW + 2 * j0);
#pragma unroll
  for (int th = 0; th < KS; ++th) {
    const uint32_t hok = (unsigned)(h0 + th) < (unsigned)SH ? 1u : 0u;
    const float* rowp = sb + th * SW;
    q.sv[th] = ldg_pred2(rowp + 2 * lane, hok);
    q.sh[th] = ldg_pred(rowp + hcol, hok & hcok);
  }
}

template <int KS, bool BIAS>
__global__ void __launch_bounds__(256, 3)
tap_wgrad_mma_kernel(const float* __restrict__ A, const float* __restrict__ S, double* __restrict__ part, int B, int C,
                     int PH, int PW, int SH, int SW, double* __restrict__ bpart) {
  constexpr int T = KS * KS, RW = 72;
  // srow[w][th][4 + c] = S[row th of the patch][2 j0 + c], c = -1 .. 64 (the +4 keeps the 8-byte stores aligned)
  __shared__ __align__(16) float srow[8][KS][RW];
  __shared__ float red[8][16][17];
  __shared__ float bred[8][16];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int CT = C >> 4;                                   // 16-channel tiles
  const int gw = blockIdx.x * 8 + warp;                    // consecutive warps share a segment (L1 hits on S), differ in tile
  const int ct = gw % CT;
  const int segs = PW >> 5;
  const int units = B * PH * segs;                         // (host: < 2^31)
  const int ustride = (int)gridDim.x * 8 / CT;
  float d[2][4];
#pragma unroll
  for (int nt = 0; nt < 2; ++nt)
#pragma unroll
    for (int e = 0; e < 4; ++e) d[nt][e] = 0.f;
  float bs0 = 0.f, bs1 = 0.f;
  // tap (n = g for tile 0, g + 8 for tile 1) -> patch row / column offset; taps >= T contribute zeros
  const int tap0 = g, tap1 = g + 8;
  const int th0 = tap0 / KS, tw0 = tap0 % KS, th1 = tap1 < T ? tap1 / KS : 0, tw1 = tap1 < T ? tap1 % KS : 0;
  // S index of pixel 8t + e of the segment for tap column tw: 2 (8t + e) - 1 + tw  ->  srow[..][3 + 16 t + 2 e + tw]
  const float* p0 = &srow[warp][th0][3 + 16 * t + tw0];
  const float* p1 = &srow[warp][th1][3 + 16 * t + tw1];
  TapUnit<KS> q;
  // (nb, ni, ns): the unit whose loads go out next; advanced by ustride units as a mixed-radix addition (no division in the loop)
  int u = gw / CT;
  int ns = u % segs, ni = (u / segs) % PH, nb = u / segs / PH;
  const int ds = ustride % segs, di = (ustride / segs) % PH, db = ustride / segs / PH;
  const int cg = ct * 16 + g;
  auto advance = [&]() {
    ns += ds;
    if (ns >= segs) { ns -= segs; ++ni; }
    ni += di;
    if (ni >= PH) { ni -= PH; ++nb; }
    nb += db;
  };
  if (u < units) tap_unit_load<KS>(q, A, S, nb, ni, ns, C, PH, PW, SH, SW, cg, t, lane);
  advance();
  for (; u < units; u += ustride) {
    __syncwarp();                    // the previous unit's fragment reads are done
#pragma unroll
    for (int th = 0; th < KS; ++th) {
      *reinterpret_cast<float2*>(&srow[warp][th][4 + 2 * lane]) = q.sv[th];
      if (lane < 2) srow[warp][th][lane == 0 ? 3 : 68] = q.sh[th];
    }
    if (BIAS) {
      bs0 += ((q.x0.x + q.x0.y) + (q.x0.z + q.x0.w)) + ((q.x1.x + q.x1.y) + (q.x1.z + q.x1.w));
      bs1 += ((q.y0.x + q.y0.y) + (q.y0.z + q.y0.w)) + ((q.y1.x + q.y1.y) + (q.y1.z + q.y1.w));
    }
    // A fragments: rows g / g + 8 of the channel tile, k = this thread's pixels 8t + {0,1 | 2,3} and {4,5 | 6,7}
    const uint32_t a0[4] = {tap_pack(q.x0.x, q.x0.y), tap_pack(q.y0.x, q.y0.y), tap_pack(q.x0.z, q.x0.w), tap_pack(q.y0.z, q.y0.w)};
    const uint32_t a1[4] = {tap_pack(q.x1.x, q.x1.y), tap_pack(q.y1.x, q.y1.y), tap_pack(q.x1.z, q.x1.w), tap_pack(q.y1.z, q.y1.w)};
    if (u + ustride < units) tap_unit_load<KS>(q, A, S, nb, ni, ns, C, PH, PW, SH, SW, cg, t, lane);
    advance();
    __syncwarp();
    // B fragments (the packed conversion rounds to bf16): k rows = the same pixels, n = tap
    tap_mma16816(d[0], a0, tap_pack(p0[0], p0[2]), tap_pack(p0[4], p0[6]));
    tap_mma16816(d[0], a1, tap_pack(p0[8], p0[10]), tap_pack(p0[12], p0[14]));
    // every lane executes every mma (.aligned); lanes whose second-tile tap does not exist (KS = 3: taps 9..15) feed zeros
    uint32_t q0 = tap_pack(p1[0], p1[2]), q1 = tap_pack(p1[4], p1[6]), q2 = tap_pack(p1[8], p1[10]), q3 = tap_pack(p1[12], p1[14]);
    if (T < 16 && tap1 >= T) q0 = q1 = q2 = q3 = 0u;
    tap_mma16816(d[1], a0, q0, q1);
    tap_mma16816(d[1], a1, q2, q3);
  }
  // accumulator (row g / g + 8 = channel, column 2t / 2t + 1 = tap within the tile) -> block reduction in a fixed order
#pragma unroll
  for (int nt = 0; nt < 2; ++nt) {
    red[warp][g][nt * 8 + 2 * t] = d[nt][0];
    red[warp][g][nt * 8 + 2 * t + 1] = d[nt][1];
    red[warp][g + 8][nt * 8 + 2 * t] = d[nt][2];
    red[warp][g + 8][nt * 8 + 2 * t + 1] = d[nt][3];
  }
  if (BIAS) {
    bs0 += __shfl_xor_sync(0xffffffffu, bs0, 1); bs0 += __shfl_xor_sync(0xffffffffu, bs0, 2);
    bs1 += __shfl_xor_sync(0xffffffffu, bs1, 1); bs1 += __shfl_xor_sync(0xffffffffu, bs1, 2);
    if (t == 0) { bred[warp][g] = bs0; bred[warp][g + 8] = bs1; }
  }
  __syncthreads();
  // warps w with (blockIdx.x * 8 + w) % CT == tile hold that tile's partials
  for (int e = threadIdx.x; e < C * T; e += blockDim.x) {
    const int c = e / T, tp = e - c * T;
    const int tile = c >> 4;
    double v = 0.0;
    for (int w = 0; w < 8; ++w)
      if ((blockIdx.x * 8 + w) % CT == tile) v += (double)red[w][c & 15][tp];
    part[((int64_t)blockIdx.x * C + c) * T + tp] = v;
  }
  if (BIAS) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      const int tile = c >> 4;
      double v = 0.0;
      for (int w = 0; w < 8; ++w)
        if ((blockIdx.x * 8 + w) % CT == tile) v += (double)bred[w][c & 15];
      bpart[(int64_t)blockIdx.x * C + c] = v;
    }
  }
}

// out[e] = sum over the nblk block partials part[blk][C*T].  A block owns 32 consecutive outputs; its eight warps each
// add every eighth partial and the eight sub-sums are added in warp order (fixed order: bit-identical from run to run)
__global__ void __launch_bounds__(256) tap_wgrad_fold_kernel(const double* __restrict__ part, int nblk, int n,
                                                             float* __restrict__ out) {
  __shared__ double sub[8][32];
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int e = blockIdx.x * 32 + lane;
  double v = 0.0;
  if (e < n) v = strided_sum_ordered(part + e, (int64_t)n, g, nblk, 8);
  sub[g][lane] = v;
  __syncthreads();
  if (g != 0 || e >= n) return;
#pragma unroll
  for (int i = 1; i < 8; ++i) v += sub[i][lane];
  out[e] = (float)v;
}

// ---- host -----------------------------------------------------------------------------------------------------------
bool conv_direct_supported(int op, int B, int Cin, int H, int W, int Cout) {
  (void)B; (void)H;
  if (op == VC_OP_CONV_FWD) return Cin == 1 && Cout <= 256;
  if (op == VC_OP_CONVT_FWD) return Cout == 1 && Cin <= 512;
  if (op == VC_OP_CONVT_DGRAD) return Cout == 1 && Cin <= 512;
  return false;
}

// weight gradients of the single-channel ends (tap_wgrad_kernel): in0 = dy, in1 = x as in vc_dense_op
constexpr int kTapBlocks = 2 * 148 * 2;   // upper bound of blockIdx.x per launch (workspace sizing: [blk][C][T])
bool conv_direct_wgrad_supported(int op, int B, int Cin, int H, int W, int Cout) {
  (void)B; (void)H;
  if (op == VC_OP_CONV_WGRAD) return Cin == 1 && Cout % 4 == 0 && Cout <= 512 && W % 8 == 0;
  if (op == VC_OP_CONVT_WGRAD) return Cout == 1 && Cin % 4 == 0 && Cin <= 512 && W % 4 == 0;
  return false;
}
size_t conv_direct_wgrad_workspace(int op, int Cin, int Cout) {
  const int C = op == VC_OP_CONV_WGRAD ? Cout : Cin, T = op == VC_OP_CONV_WGRAD ? 9 : 16;
  return sizeof(double) * (size_t)kTapBlocks * C * (T + 1);   // + one column of bias-gradient partials
}
int conv_direct_wgrad_bias_run(int op, const float* dy, const float* x, float* dw, float* db, int B, int Cin, int H, int W,
                               int Cout, int round_bf16, void* ws, size_t wsb, cudaStream_t s);
int conv_direct_wgrad_run(int op, const float* dy, const float* x, float* dw, int B, int Cin, int H, int W, int Cout,
                          int round_bf16, void* ws, size_t wsb, cudaStream_t s) {
  return conv_direct_wgrad_bias_run(op, dy, x, dw, nullptr, B, Cin, H, W, Cout, round_bf16, ws, wsb, s);
}
// db (Conv2d(1 -> C) only): the bias gradient sum(dy) from the same pass over dy (12:87)
int conv_direct_wgrad_bias_run(int op, const float* dy, const float* x, float* dw, float* db, int B, int Cin, int H, int W,
                               int Cout, int round_bf16, void* ws, size_t wsb, cudaStream_t s) {
  VC_REQUIRE(ws && wsb >= conv_direct_wgrad_workspace(op, Cin, Cout), VC_ERR_WORKSPACE,
             "conv_direct_wgrad_run: workspace too small");
  double* part = static_cast<double*>(ws);
  int nblk = 2 * sm_count();   // two 256-thread blocks per SM, every block walks all channels of its blockIdx.y
  if (nblk > kTapBlocks) nblk = kTapBlocks;
  {  // bf16 mode, 16 / 32 / 64 / 128 channels, rows of 32-pixel segments: the tensor-core kernel (three blocks per SM)
    const int C = op == VC_OP_CONV_WGRAD ? Cout : Cin;
    const int PH = op == VC_OP_CONV_WGRAD ? (H - 1) / 2 + 1 : H, PW = op == VC_OP_CONV_WGRAD ? W / 2 : W;
    const int SH = op == VC_OP_CONV_WGRAD ? H : 2 * H, SW = op == VC_OP_CONV_WGRAD ? W : 2 * W;
    const float* A = op == VC_OP_CONV_WGRAD ? dy : x;
    const float* S = op == VC_OP_CONV_WGRAD ? x : dy;
    static const bool off = getenv("VAECLUST_TAP_MMA") && atoi(getenv("VAECLUST_TAP_MMA")) == 0;
    const bool ct_ok = C == 16 || C == 32 || C == 64 || C == 128;
    if (!off && round_bf16 && ct_ok && PW % 32 == 0 && (reinterpret_cast<uintptr_t>(A) & 15) == 0 &&
        (op == VC_OP_CONVT_WGRAD ? !db : true)) {
      const int T = op == VC_OP_CONV_WGRAD ? 9 : 16, CT = C / 16;
      int nb = 3 * sm_count();
      if (nb > kTapBlocks) nb = kTapBlocks;
      const int64_t units = (int64_t)B * PH * (PW / 32);
      VC_REQUIRE((int64_t)B * C * PH * PW < 2147483647LL, VC_ERR_UNSUPPORTED,
                 "conv_direct_wgrad_bias_run: tensor too large for 32-bit element offsets");
      if ((int64_t)nb * 8 / CT > units) nb = (int)((units * CT + 7) / 8);
      double* bpart = db ? part + (size_t)kTapBlocks * C * T : nullptr;
      if (op == VC_OP_CONVT_WGRAD)
        tap_wgrad_mma_kernel<4, false><<<nb, 256, 0, s>>>(A, S, part, B, C, PH, PW, SH, SW, nullptr);
      else if (db)
        tap_wgrad_mma_kernel<3, true><<<nb, 256, 0, s>>>(A, S, part, B, C, PH, PW, SH, SW, bpart);
      else
        tap_wgrad_mma_kernel<3, false><<<nb, 256, 0, s>>>(A, S, part, B, C, PH, PW, SH, SW, nullptr);
      VC_LAUNCH_CHECK();
      tap_wgrad_fold_kernel<<<(C * T + 31) / 32, 256, 0, s>>>(part, nb, C * T, dw);
      if (db) {
        VC_LAUNCH_CHECK();
        tap_wgrad_fold_kernel<<<(C + 31) / 32, 256, 0, s>>>(bpart, nb, C, db);
      }
      VC_LAUNCH_CHECK();
      return VC_OK;
    }
  }
  if (op == VC_OP_CONV_WGRAD) {   // A = dy [B,Cout,OH,OW], S = x [B,1,H,W]
    constexpr int CG = 4, T = 9;
    const int OH = (H - 1) / 2 + 1, OW = W / 2, groups = Cout / CG;
    const int64_t quads = (int64_t)B * OH * (OW / 4);
    if ((int64_t)nblk * 32 > quads) nblk = (int)((quads + 31) / 32);
    double* bpart = db ? part + (size_t)kTapBlocks * Cout * T : nullptr;
    if (db)
      tap_wgrad_kernel<3, CG, true><<<dim3(nblk, (groups + 7) / 8), 256, 0, s>>>(dy, x, part, B, Cout, OH, OW, H, W, round_bf16, bpart);
    else
      tap_wgrad_kernel<3, CG, false><<<dim3(nblk, (groups + 7) / 8), 256, 0, s>>>(dy, x, part, B, Cout, OH, OW, H, W, round_bf16, nullptr);
    VC_LAUNCH_CHECK();
    tap_wgrad_fold_kernel<<<(Cout * T + 31) / 32, 256, 0, s>>>(part, nblk, Cout * T, dw);
    if (db) {
      VC_LAUNCH_CHECK();
      tap_wgrad_fold_kernel<<<(Cout + 31) / 32, 256, 0, s>>>(bpart, nblk, Cout, db);
    }
  } else {                        // A = x [B,Cin,H,W], S = dy [B,1,2H,2W]
    constexpr int CG = 4, T = 16;
    const int groups = Cin / CG;
    const int64_t quads = (int64_t)B * H * (W / 4);
    if ((int64_t)nblk * 32 > quads) nblk = (int)((quads + 31) / 32);
    VC_REQUIRE(!db, VC_ERR_UNSUPPORTED, "conv_direct_wgrad_bias_run: the fused bias gradient exists for Conv2d(1 -> C) only");
    tap_wgrad_kernel<4, CG, false><<<dim3(nblk, (groups + 7) / 8), 256, 0, s>>>(x, dy, part, B, Cin, H, W, 2 * H, 2 * W,
                                                                          round_bf16, nullptr);
    VC_LAUNCH_CHECK();
    tap_wgrad_fold_kernel<<<(Cin * T + 31) / 32, 256, 0, s>>>(part, nblk, Cin * T, dw);
  }
  VC_LAUNCH_CHECK();
  return VC_OK;
}

static int direct_grid(int64_t total) {
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  return (int)(blocks > cap ? cap : blocks);
}

int conv_direct_run(int op, const float* in0, const float* in1, const float* bias, float* out, int B, int Cin, int H,
                    int W, int Cout, int relu, int round_bf16, cudaStream_t s) {
  if (op == VC_OP_CONV_FWD) {
    const int OH = (H - 1) / 2 + 1, OW = (W - 1) / 2 + 1;
    const int Cp = (Cout + 3) & ~3;
    const size_t smem = (size_t)10 * Cp * sizeof(float);
    const int grid = direct_grid((int64_t)B * OH * OW);
    if (Cout > 16)
      conv1ch_fwd_kernel<32><<<grid, 256, smem, s>>>(in0, in1, bias, out, B, H, W, Cout, OH, OW, relu, round_bf16);
    else
      conv1ch_fwd_kernel<16><<<grid, 256, smem, s>>>(in0, in1, bias, out, B, H, W, Cout, OH, OW, relu, round_bf16);
  } else if (op == VC_OP_CONVT_FWD) {
    const bool quad = (W % 4) == 0 && ((reinterpret_cast<uintptr_t>(in0) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
    if (quad) {
      const int grid = direct_grid((int64_t)B * ((H + 1) / 2) * (W / 4));
      convT_to1_fwd_quad_kernel<<<grid, 256, (size_t)Cin * 16 * sizeof(float), s>>>(in0, in1, bias, out, B, Cin, H, W,
                                                                                      relu, round_bf16);
    } else {
      const int grid = direct_grid((int64_t)B * ((H + 1) / 2) * W);
      convT_to1_fwd_kernel<<<grid, 256, (size_t)Cin * 16 * sizeof(float), s>>>(in0, in1, bias, out, B, Cin, H, W, relu,
                                                                                 round_bf16);
    }
  } else if (op == VC_OP_CONVT_DGRAD) {
    const int grid = direct_grid((int64_t)B * H * W);
    convT_to1_dgrad_kernel<<<grid, 256, (size_t)Cin * 16 * sizeof(float), s>>>(in0, in1, out, B, Cin, H, W, round_bf16);
  } else {
    set_error("conv_direct_run: unsupported op %d", op);
    return VC_ERR_UNSUPPORTED;
  }
  VC_LAUNCH_CHECK();
  return VC_OK;
}

}  // namespace vc

using namespace vc;

extern "C" size_t vc_convT1_wgrad_workspace_bytes(int Cin) { return conv_direct_wgrad_workspace(VC_OP_CONVT_WGRAD, Cin, 1); }

extern "C" int vc_convT1_wgrad(const float* dy, const float* x, float* dW, int B, int Cin, int H, int W, int round_bf16,
                               void* workspace, size_t workspace_bytes, void* stream) {
  VC_REQUIRE(dy && x && dW && B > 0, VC_ERR_BAD_ARG, "vc_convT1_wgrad: bad arguments");
  VC_REQUIRE(conv_direct_wgrad_supported(VC_OP_CONVT_WGRAD, B, Cin, H, W, 1) &&
                 ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(x)) & 15) == 0,
             VC_ERR_UNSUPPORTED, "vc_convT1_wgrad: need Cin %% 4 == 0, Cin <= 512, W %% 4 == 0, 16-byte aligned tensors");
  return conv_direct_wgrad_bias_run(VC_OP_CONVT_WGRAD, dy, x, dW, nullptr, B, Cin, H, W, 1, round_bf16, workspace,
                                    workspace_bytes, as_stream(stream));
}

extern "C" size_t vc_conv1_wgrad_bias_workspace_bytes(int Cout) { return conv_direct_wgrad_workspace(VC_OP_CONV_WGRAD, 1, Cout); }

extern "C" int vc_conv1_wgrad_bias(const float* dy, const float* x, float* dW, float* db, int B, int H, int W, int Cout,
                                   int round_bf16, void* workspace, size_t workspace_bytes, void* stream) {
  VC_REQUIRE(dy && x && dW && db && B > 0, VC_ERR_BAD_ARG, "vc_conv1_wgrad_bias: bad arguments");
  VC_REQUIRE(conv_direct_wgrad_supported(VC_OP_CONV_WGRAD, B, 1, H, W, Cout) &&
                 ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(x)) & 15) == 0,
             VC_ERR_UNSUPPORTED, "vc_conv1_wgrad_bias: need Cout %% 4 == 0, Cout <= 512, W %% 8 == 0, 16-byte aligned tensors");
  return conv_direct_wgrad_bias_run(VC_OP_CONV_WGRAD, dy, x, dW, db, B, 1, H, W, Cout, round_bf16, workspace, workspace_bytes,
                                    as_stream(stream));
}
