// The small kernels around the TMA-fed bf16-NHWC convolution path of script 12 (conv_nhwc.cu, conv_nhwc_wgrad.cu):
//   * Conv2d(1 -> C, k3 s2 p1) forward (12:87) writing bf16 NHWC directly (thread = output pixel: its C channels are one
//     contiguous run of the NHWC tensor),
//   * ConvTranspose2d(C -> 1, k4 s2 p1) data gradient (12:137) with the ReLU mask of the layer below applied and the result
//     written as bf16 NHWC (what used to be a gather kernel + a 12 B/element ReLU-backward pass),
//   * fp32 NCHW -> bf16 NHWC conversion (the decoder's fully connected output, 12:129-133),
//   * per-channel sums of a bf16 NHWC tensor (bias gradients, fixed summation order).
// All HBM-bound streaming kernels; operands are rounded to bf16 exactly as the tensor-core kernels do.
#include "common.cuh"

namespace vc {

__device__ __forceinline__ float nm_rb(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
__device__ __forceinline__ uint32_t nm_pack(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float nm_ldg_pred(const float* p, uint32_t pred) {
  float v;
  asm("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %2, 0;\n\tmov.f32 %0, 0f00000000;\n\t@q ld.global.nc.f32 %0, [%1];\n\t}"
      : "=f"(v)
      : "l"(p), "r"(pred));
  return v;
}

// ---- Conv2d(1 -> Cout) forward, y as bf16 NHWC [B][OH][OW][Cout]; Cout % 8 == 0
__global__ void __launch_bounds__(256)
nm_conv1_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                    __nv_bfloat16* __restrict__ y, int B, int H, int W, int Cout, int OH, int OW, int relu) {
  extern __shared__ float sm[];
  float* wt = sm;                // [9][Cout]
  float* bs = sm + 9 * Cout;     // [Cout]
  for (int e = threadIdx.x; e < 9 * Cout; e += blockDim.x) {
    const int t = e / Cout, co = e - t * Cout;
    wt[e] = nm_rb(w[co * 9 + t]);
  }
  for (int e = threadIdx.x; e < Cout; e += blockDim.x) bs[e] = bias ? bias[e] : 0.f;
  __syncthreads();
  const int64_t total = (int64_t)B * OH * OW;
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
        xin[kh * 3 + kw] = nm_rb(nm_ldg_pred(xb + (int64_t)ih * W + iw, ok));
      }
    uint4* yo = reinterpret_cast<uint4*>(y + m * Cout);
    for (int c0 = 0; c0 < Cout; c0 += 8) {
      float acc[8];
      const float4 b0 = *reinterpret_cast<const float4*>(bs + c0), b1 = *reinterpret_cast<const float4*>(bs + c0 + 4);
      acc[0] = b0.x; acc[1] = b0.y; acc[2] = b0.z; acc[3] = b0.w; acc[4] = b1.x; acc[5] = b1.y; acc[6] = b1.z; acc[7] = b1.w;
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const float4 w0 = *reinterpret_cast<const float4*>(wt + t * Cout + c0);
        const float4 w1 = *reinterpret_cast<const float4*>(wt + t * Cout + c0 + 4);
        acc[0] = fmaf(xin[t], w0.x, acc[0]); acc[1] = fmaf(xin[t], w0.y, acc[1]);
        acc[2] = fmaf(xin[t], w0.z, acc[2]); acc[3] = fmaf(xin[t], w0.w, acc[3]);
        acc[4] = fmaf(xin[t], w1.x, acc[4]); acc[5] = fmaf(xin[t], w1.y, acc[5]);
        acc[6] = fmaf(xin[t], w1.z, acc[6]); acc[7] = fmaf(xin[t], w1.w, acc[7]);
      }
      if (relu) {
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[c] = fmaxf(acc[c], 0.f);
      }
      yo[c0 >> 3] = make_uint4(nm_pack(acc[0], acc[1]), nm_pack(acc[2], acc[3]), nm_pack(acc[4], acc[5]),
                               nm_pack(acc[6], acc[7]));
    }
  }
}

// ---- ConvTranspose2d(Cin -> 1) data gradient with the ReLU mask of the layer's INPUT activation (bf16 NHWC) applied:
//   g[b,i,j,ci] = (act[b,i,j,ci] > 0) * sum_{kh,kw} dy[b,0,2i-1+kh,2j-1+kw] * W[ci,0,kh,kw]      (bf16 NHWC out)
__global__ void __launch_bounds__(256)
nm_convT1_dgrad_kernel(const float* __restrict__ dy, const float* __restrict__ w, const __nv_bfloat16* __restrict__ act,
                       __nv_bfloat16* __restrict__ g, int B, int Cin, int H, int W) {
  extern __shared__ float sm[];   // [16][Cin]: tap-major so eight channels of one tap are two LDS.128
  for (int e = threadIdx.x; e < Cin * 16; e += blockDim.x) {
    const int t = e / Cin, ci = e - t * Cin;
    sm[e] = nm_rb(w[ci * 16 + t]);
  }
  __syncthreads();
  const int64_t total = (int64_t)B * H * W;
  const int OH = 2 * H, OW = 2 * W;
  for (int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; m < total; m += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(m % W);
    const int64_t t2 = m / W;
    const int i = (int)(t2 % H), b = (int)(t2 / H);
    float d[16];
    const float* db = dy + (int64_t)b * OH * OW;
#pragma unroll
    for (int kh = 0; kh < 4; ++kh)
#pragma unroll
      for (int kw = 0; kw < 4; ++kw) {
        const int oh = 2 * i - 1 + kh, ow = 2 * j - 1 + kw;
        const uint32_t ok = ((unsigned)oh < (unsigned)OH && (unsigned)ow < (unsigned)OW) ? 1u : 0u;
        d[kh * 4 + kw] = nm_rb(nm_ldg_pred(db + (int64_t)oh * OW + ow, ok));
      }
    const uint4* a4 = reinterpret_cast<const uint4*>(act + m * Cin);
    uint4* g4 = reinterpret_cast<uint4*>(g + m * Cin);
    for (int c0 = 0; c0 < Cin; c0 += 8) {
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int t = 0; t < 16; ++t) {
        const float4 w0 = *reinterpret_cast<const float4*>(sm + t * Cin + c0);
        const float4 w1 = *reinterpret_cast<const float4*>(sm + t * Cin + c0 + 4);
        acc[0] = fmaf(d[t], w0.x, acc[0]); acc[1] = fmaf(d[t], w0.y, acc[1]);
        acc[2] = fmaf(d[t], w0.z, acc[2]); acc[3] = fmaf(d[t], w0.w, acc[3]);
        acc[4] = fmaf(d[t], w1.x, acc[4]); acc[5] = fmaf(d[t], w1.y, acc[5]);
        acc[6] = fmaf(d[t], w1.z, acc[6]); acc[7] = fmaf(d[t], w1.w, acc[7]);
      }
      if (act) {
        const uint4 mm = __ldg(a4 + (c0 >> 3));
        const uint32_t mw[4] = {mm.x, mm.y, mm.z, mm.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          acc[2 * u] = __uint_as_float(mw[u] << 16) > 0.f ? acc[2 * u] : 0.f;
          acc[2 * u + 1] = __uint_as_float(mw[u] & 0xffff0000u) > 0.f ? acc[2 * u + 1] : 0.f;
        }
      }
      g4[c0 >> 3] = make_uint4(nm_pack(acc[0], acc[1]), nm_pack(acc[2], acc[3]), nm_pack(acc[4], acc[5]),
                               nm_pack(acc[6], acc[7]));
    }
  }
}

// ---- fp32 NCHW -> bf16 NHWC.  A block takes (b, h, 64-pixel segment) x all channels; the transposition goes through
// shared memory (reads coalesced along W, writes are one contiguous 64 x C block of the NHWC tensor).
constexpr int NM_PX = 64;
// A lane owns pixels lane and lane + 32 of a channel PAIR (coalesced 128-byte loads) and stores one bf16x2 word per pixel
// into rows of C/2 + 1 words: bank = pixel + pair, no conflicts on the way in or out (a lane owning two adjacent pixels of one
// channel made its 2-byte stores 544 bytes apart: 8-way conflicts, 2.7 TB/s).  C % 16 == 0.
__global__ void __launch_bounds__(256)
nm_from_nchw_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int B, int C, int H, int W, int ntiles) {
  extern __shared__ unsigned char nm_smem[];
  const int pw = (C >> 1) + 1;
  uint32_t* tile = reinterpret_cast<uint32_t*>(nm_smem);      // [NM_PX][pw]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int segs = W / NM_PX;
  const int64_t plane = (int64_t)H * W;
  const int npairs = C >> 1;
  for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const int seg = t % segs, r = t / segs;
    const int h = r % H, b = r / H;
    const int64_t base = ((int64_t)b * C) * plane + (int64_t)h * W + seg * NM_PX + lane;
    for (int cp0 = warp; cp0 < npairs; cp0 += 32) {      // four pairs per round: 16 independent loads per lane
      float v[4][2][2];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int cp = cp0 + 8 * k;
        if (cp < npairs) {
          const float* p = src + base + (int64_t)(2 * cp) * plane;
          v[k][0][0] = __ldcs(p); v[k][0][1] = __ldcs(p + 32);
          v[k][1][0] = __ldcs(p + plane); v[k][1][1] = __ldcs(p + plane + 32);
        }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int cp = cp0 + 8 * k;
        if (cp < npairs) {
          tile[lane * pw + cp] = nm_pack(v[k][0][0], v[k][1][0]);
          tile[(lane + 32) * pw + cp] = nm_pack(v[k][0][1], v[k][1][1]);
        }
      }
    }
    __syncthreads();
    uint32_t* o = reinterpret_cast<uint32_t*>(dst + (((int64_t)b * H + h) * W + seg * NM_PX) * C);
    for (int e = threadIdx.x; e < NM_PX * npairs; e += blockDim.x) {
      const int px = e / npairs, q = e - px * npairs;
      o[e] = tile[px * pw + q];
    }
    __syncthreads();
  }
}

// ---- per-channel sums of a bf16 [rows][C] matrix (C % 8 == 0, C <= 2048): thread = eight channels of a row; the
// block's row groups meet in shared memory in a fixed order, block partials (double) are folded by nm_colsum_fold_kernel
__global__ void __launch_bounds__(256)
nm_colsum_kernel(const __nv_bfloat16* __restrict__ g, int64_t rows, int C, double* __restrict__ part) {
  extern __shared__ float nm_red[];          // [groups][C]
  const int vecs = C >> 3;                   // threads per row
  const int groups = 256 / vecs;             // rows per pass of the block
  const int grp = threadIdx.x / vecs, q = threadIdx.x - grp * vecs;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (grp < groups) {
    // four independent 16-byte loads per round (one dependent load per iteration left the kernel latency-bound at 2.5 TB/s)
    const int64_t step = (int64_t)gridDim.x * groups;
    int64_t r = (int64_t)blockIdx.x * groups + grp;
    auto add = [&](const uint4& v) {
      const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        acc[2 * k] += __uint_as_float(u[k] << 16);
        acc[2 * k + 1] += __uint_as_float(u[k] & 0xffff0000u);
      }
    };
    for (; r + 3 * step < rows; r += 4 * step) {
      const uint4 v0 = __ldcs(reinterpret_cast<const uint4*>(g + r * C) + q);
      const uint4 v1 = __ldcs(reinterpret_cast<const uint4*>(g + (r + step) * C) + q);
      const uint4 v2 = __ldcs(reinterpret_cast<const uint4*>(g + (r + 2 * step) * C) + q);
      const uint4 v3 = __ldcs(reinterpret_cast<const uint4*>(g + (r + 3 * step) * C) + q);
      add(v0); add(v1); add(v2); add(v3);
    }
    for (; r < rows; r += step) add(__ldcs(reinterpret_cast<const uint4*>(g + r * C) + q));
#pragma unroll
    for (int k = 0; k < 8; ++k) nm_red[grp * C + q * 8 + k] = acc[k];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    double s = 0.0;
    for (int gi = 0; gi < groups; ++gi) s += (double)nm_red[gi * C + c];
    part[(int64_t)blockIdx.x * C + c] = s;
  }
}

// db[c] = sum over blocks, eight warps per 32 channels, sub-sums added in warp order
__global__ void __launch_bounds__(256)
nm_colsum_fold_kernel(const double* __restrict__ part, int nblk, int C, float* __restrict__ db) {
  __shared__ double sub[8][32];
  const int lane = threadIdx.x & 31, gw = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  double v = 0.0;
  if (c < C) v = strided_sum_ordered(part + c, (int64_t)C, gw, nblk, 8);
  sub[gw][lane] = v;
  __syncthreads();
  if (gw != 0 || c >= C) return;
#pragma unroll
  for (int i = 1; i < 8; ++i) v += sub[i][lane];
  db[c] = (float)v;
}

// ---- ConvTranspose2d(C -> 1, k4 s2 p1) forward from a bf16 NHWC input (12:137): thread = coarse pixel (i, j) -> the 2x2
// output block y[2i..2i+1][2j..2j+1].  Output row 2i+ph takes input rows i+dh with kernel row kh = 1 + ph - 2 dh
// (dh = -1: ph 0 only, kh 3; dh = 0: kh 1 / 2; dh = +1: ph 1 only, kh 0); columns likewise: 16 products per channel.
__global__ void __launch_bounds__(256)
nm_convT1_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                     float* __restrict__ y, int B, int C, int H, int W) {
  extern __shared__ float sm[];   // [C][16] (kh, kw), bf16-rounded
  for (int e = threadIdx.x; e < C * 16; e += blockDim.x) sm[e] = nm_rb(w[e]);
  __syncthreads();
  const float b0 = bias ? __ldg(bias) : 0.f;
  const int64_t total = (int64_t)B * H * W;
  const int OW = 2 * W;
  for (int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; m < total; m += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(m % W);
    const int64_t t2 = m / W;
    const int i = (int)(t2 % H), b = (int)(t2 / H);
    float o00 = b0, o01 = b0, o10 = b0, o11 = b0;
#pragma unroll
    for (int dh = -1; dh <= 1; ++dh) {
      const int ih = i + dh;
      if ((unsigned)ih >= (unsigned)H) continue;
#pragma unroll
      for (int dw = -1; dw <= 1; ++dw) {
        const int iw = j + dw;
        if ((unsigned)iw >= (unsigned)W) continue;
        const uint4* px = reinterpret_cast<const uint4*>(x + (((int64_t)b * H + ih) * W + iw) * C);
        for (int c0 = 0; c0 < C; c0 += 8) {
          const uint4 v = __ldg(px + (c0 >> 3));
          const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const float xv = (k & 1) ? __uint_as_float(u[k >> 1] & 0xffff0000u) : __uint_as_float(u[k >> 1] << 16);
            const float* wc = sm + (c0 + k) * 16;
            // (ph, kh) pairs for this dh, (pw, kw) pairs for this dw -- resolved at compile time by the unrolling
            if (dh <= 0 && dw <= 0) o00 = fmaf(xv, wc[(1 - 2 * dh) * 4 + (1 - 2 * dw)], o00);
            if (dh <= 0 && dw >= 0) o01 = fmaf(xv, wc[(1 - 2 * dh) * 4 + (2 - 2 * dw)], o01);
            if (dh >= 0 && dw <= 0) o10 = fmaf(xv, wc[(2 - 2 * dh) * 4 + (1 - 2 * dw)], o10);
            if (dh >= 0 && dw >= 0) o11 = fmaf(xv, wc[(2 - 2 * dh) * 4 + (2 - 2 * dw)], o11);
          }
        }
      }
    }
    float* yo = y + ((int64_t)b * 2 * H + 2 * i) * OW + 2 * j;
    *reinterpret_cast<float2*>(yo) = make_float2(o00, o01);
    *reinterpret_cast<float2*>(yo + OW) = make_float2(o10, o11);
  }
}

// ---- weight gradients of the two single-channel ends from a bf16 NHWC tensor A [B][PH][PW][C] and the ONE-channel fp32
// tensor S [B][SH][SW] it is correlated with (conv_direct.cu has the fp32-NCHW version):
//   out[c][kh][kw] = sum_{b,i,j} A[b,i,j,c] * S[b, 2i-1+kh, 2j-1+kw]
//   Conv2d(1 -> C) k3: A = g (output gradient), S = x;   ConvTranspose2d(C -> 1) k4: A = x (layer input), S = dy.
// lane = channel (a warp owns 32 channels: one coalesced 64-byte read per pixel), a warp walks 32-pixel row segments: the
// KS rows of S it needs (66..67 values each) are staged in shared memory once per segment, the taps of a pixel are two
// 8-byte broadcast reads per row.  KS*KS register accumulators per lane over the warp's whole range, then a fixed-order
// block / grid reduction in double.
template <int KS>
__global__ void __launch_bounds__(256)
nm_tap_wgrad_kernel(const __nv_bfloat16* __restrict__ A, const float* __restrict__ S, double* __restrict__ part, int B, int C,
                    int PH, int PW, int SH, int SW) {
  constexpr int T = KS * KS, RW = 72;
  __shared__ __align__(16) float srow[8][KS][RW];
  __shared__ float red[8][32][T + 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cg = blockIdx.y;
  const int segs = PW >> 5;
  const int64_t units = (int64_t)B * PH * segs;
  float acc[T];
#pragma unroll
  for (int t = 0; t < T; ++t) acc[t] = 0.f;
  for (int64_t u = (int64_t)blockIdx.x * 8 + warp; u < units; u += (int64_t)gridDim.x * 8) {
    const int seg = (int)(u % segs);
    const int64_t r = u / segs;
    const int i = (int)(r % PH), b = (int)(r / PH), j0 = seg * 32;
    const float* sb = S + (int64_t)b * SH * SW;
    __syncwarp();
#pragma unroll
    for (int th = 0; th < KS; ++th) {
      const int h = 2 * i - 1 + th;
      const bool hok = (unsigned)h < (unsigned)SH;
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const int idx = lane + 32 * k;            // srow[..][idx] = S[h][2*j0 - 1 + idx]
        const int wcol = 2 * j0 - 1 + idx;
        if (idx < 64 + KS - 1)
          srow[warp][th][idx] = nm_rb(nm_ldg_pred(sb + (int64_t)h * SW + wcol, (hok && (unsigned)wcol < (unsigned)SW) ? 1u : 0u));
      }
    }
    __syncwarp();
    const __nv_bfloat16* ap = A + ((((int64_t)b * PH + i) * PW + j0) * C + cg * 32 + lane);
#pragma unroll 4
    for (int p = 0; p < 32; ++p) {
      const float a = __bfloat162float(ap[(int64_t)p * C]);
#pragma unroll
      for (int th = 0; th < KS; ++th) {
        const float2 s01 = *reinterpret_cast<const float2*>(&srow[warp][th][2 * p]);
        acc[th * KS] = fmaf(a, s01.x, acc[th * KS]);
        acc[th * KS + 1] = fmaf(a, s01.y, acc[th * KS + 1]);
        if (KS == 3) {
          acc[th * KS + 2] = fmaf(a, srow[warp][th][2 * p + 2], acc[th * KS + 2]);
        } else {
          const float2 s23 = *reinterpret_cast<const float2*>(&srow[warp][th][2 * p + 2]);
          acc[th * KS + 2] = fmaf(a, s23.x, acc[th * KS + 2]);
          acc[th * KS + 3] = fmaf(a, s23.y, acc[th * KS + 3]);
        }
      }
    }
  }
#pragma unroll
  for (int t = 0; t < T; ++t) red[warp][lane][t] = acc[t];
  __syncthreads();
  for (int e = threadIdx.x; e < 32 * T; e += blockDim.x) {
    const int c = e / T, t = e - c * T;
    double v = 0.0;
#pragma unroll
    for (int wq = 0; wq < 8; ++wq) v += (double)red[wq][c][t];
    part[((int64_t)blockIdx.x * C + cg * 32 + c) * T + t] = v;
  }
}

// out[e] = sum over blocks of part[blk][C*T]: eight warps per 32 outputs, sub-sums added in warp order
__global__ void __launch_bounds__(256)
nm_tap_fold_kernel(const double* __restrict__ part, int nblk, int n, float* __restrict__ out) {
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
constexpr int kTapBlocksNhwc = 148 * 4;

static int nm_grid(int64_t total) {
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 8;
  return (int)(blocks > cap ? cap : blocks);
}
constexpr int kColsumBlocks = 148 * 4;

// nhwc_mma.cu: the tensor-core (mma.sync) versions for 32-channel layers
bool nhwc_mma_convT1_dgrad_supported(int B, int Cin, int H, int W);
int nhwc_mma_convT1_dgrad_run(const float* dy, const float* w, const __nv_bfloat16* act, __nv_bfloat16* g, int B, int H, int W,
                              double* bpart, int* grid_out, cudaStream_t s);
bool nhwc_mma_conv1_fwd_supported(int B, int H, int W, int Cout);
int nhwc_mma_conv1_fwd_run(const float* x, const float* w, const float* bias, __nv_bfloat16* y, int B, int H, int W, int relu,
                           cudaStream_t s);
bool nhwc_mma_convT1_fwd_supported(int B, int Cin, int H, int W);
int nhwc_mma_convT1_fwd_run(const __nv_bfloat16* x, const float* w, const float* bias, float* y, int B, int H, int W,
                            cudaStream_t s);
bool nhwc_mma_tap_wgrad_supported(int B, int C, int PH, int PW);
int nhwc_mma_tap_wgrad_run(int KS, const __nv_bfloat16* a, const float* S, double* part, double* bpart, int B, int C, int PH,
                           int PW, int max_blocks, int* nblk_out, cudaStream_t s);

}  // namespace vc

using namespace vc;

extern "C" int vc_nhwc_conv1_fwd(const float* x, const float* Wt, const float* bias, int relu, void* y_nhwc, int B, int H,
                                 int W, int Cout, void* stream) {
  VC_REQUIRE(x && Wt && y_nhwc && B > 0 && H > 0 && W > 0, VC_ERR_BAD_ARG, "vc_nhwc_conv1_fwd: bad arguments");
  VC_REQUIRE(Cout % 8 == 0 && Cout > 0 && Cout <= 256, VC_ERR_UNSUPPORTED, "vc_nhwc_conv1_fwd: Cout must be a multiple of 8, <= 256");
  VC_REQUIRE((reinterpret_cast<uintptr_t>(y_nhwc) & 15) == 0, VC_ERR_BAD_ARG, "vc_nhwc_conv1_fwd: y must be 16-byte aligned");
  if (nhwc_mma_conv1_fwd_supported(B, H, W, Cout))
    return nhwc_mma_conv1_fwd_run(x, Wt, bias, static_cast<__nv_bfloat16*>(y_nhwc), B, H, W, relu, as_stream(stream));
  const int OH = (H - 1) / 2 + 1, OW = (W - 1) / 2 + 1;
  nm_conv1_fwd_kernel<<<nm_grid((int64_t)B * OH * OW), 256, (size_t)10 * Cout * sizeof(float), as_stream(stream)>>>(
      x, Wt, bias, static_cast<__nv_bfloat16*>(y_nhwc), B, H, W, Cout, OH, OW, relu);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" int vc_nhwc_convT1_dgrad(const float* dy, const float* Wt, const void* act_nhwc, void* g_nhwc, int B, int Cin,
                                    int H, int W, void* stream) {
  VC_REQUIRE(dy && Wt && g_nhwc && B > 0 && H > 0 && W > 0, VC_ERR_BAD_ARG, "vc_nhwc_convT1_dgrad: bad arguments");
  VC_REQUIRE(Cin % 8 == 0 && Cin > 0 && Cin <= 512, VC_ERR_UNSUPPORTED, "vc_nhwc_convT1_dgrad: Cin must be a multiple of 8, <= 512");
  VC_REQUIRE(((reinterpret_cast<uintptr_t>(g_nhwc) | reinterpret_cast<uintptr_t>(act_nhwc)) & 15) == 0, VC_ERR_BAD_ARG,
             "vc_nhwc_convT1_dgrad: NHWC tensors must be 16-byte aligned");
  if (nhwc_mma_convT1_dgrad_supported(B, Cin, H, W))
    return nhwc_mma_convT1_dgrad_run(dy, Wt, static_cast<const __nv_bfloat16*>(act_nhwc), static_cast<__nv_bfloat16*>(g_nhwc), B,
                                     H, W, nullptr, nullptr, as_stream(stream));
  nm_convT1_dgrad_kernel<<<nm_grid((int64_t)B * H * W), 256, (size_t)16 * Cin * sizeof(float), as_stream(stream)>>>(
      dy, Wt, static_cast<const __nv_bfloat16*>(act_nhwc), static_cast<__nv_bfloat16*>(g_nhwc), B, Cin, H, W);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

/* ConvTranspose2d(Cin -> 1) data gradient AND the bias gradient of the layer below (column sums of the masked result) */
extern "C" size_t vc_nhwc_convT1_dgrad_bias_workspace_bytes(int Cin) { return vc_nhwc_bias_grad_workspace_bytes(Cin); }

extern "C" int vc_nhwc_convT1_dgrad_bias(const float* dy, const float* Wt, const void* act_nhwc, void* g_nhwc, float* db, int B,
                                         int Cin, int H, int W, void* workspace, size_t workspace_bytes, void* stream) {
  VC_REQUIRE(db, VC_ERR_BAD_ARG, "vc_nhwc_convT1_dgrad_bias: db is required");
  VC_REQUIRE(workspace && workspace_bytes >= vc_nhwc_convT1_dgrad_bias_workspace_bytes(Cin), VC_ERR_WORKSPACE,
             "vc_nhwc_convT1_dgrad_bias: workspace too small");
  if (dy && Wt && g_nhwc && B > 0 && H > 0 && W > 0 && nhwc_mma_convT1_dgrad_supported(B, Cin, H, W) &&
      ((reinterpret_cast<uintptr_t>(g_nhwc) | reinterpret_cast<uintptr_t>(act_nhwc)) & 15) == 0) {
    int grid = 0;
    double* bpart = static_cast<double*>(workspace);
    const int rc = nhwc_mma_convT1_dgrad_run(dy, Wt, static_cast<const __nv_bfloat16*>(act_nhwc),
                                             static_cast<__nv_bfloat16*>(g_nhwc), B, H, W, bpart, &grid, as_stream(stream));
    if (rc != VC_OK) return rc;
    nm_colsum_fold_kernel<<<(Cin + 31) / 32, 256, 0, as_stream(stream)>>>(bpart, grid, Cin, db);
    VC_LAUNCH_CHECK();
    return VC_OK;
  }
  const int rc = vc_nhwc_convT1_dgrad(dy, Wt, act_nhwc, g_nhwc, B, Cin, H, W, stream);
  if (rc != VC_OK) return rc;
  return vc_nhwc_bias_grad(g_nhwc, (int64_t)B * H * W, Cin, db, workspace, workspace_bytes, stream);
}

extern "C" int vc_nhwc_from_nchw(const float* src, void* dst_nhwc, int B, int C, int H, int W, void* stream) {
  VC_REQUIRE(src && dst_nhwc && B > 0 && H > 0, VC_ERR_BAD_ARG, "vc_nhwc_from_nchw: bad arguments");
  VC_REQUIRE(C % 16 == 0 && C > 0 && C <= 512 && W % NM_PX == 0, VC_ERR_UNSUPPORTED,
             "vc_nhwc_from_nchw: need C %% 16 == 0, C <= 512, W %% 64 == 0 (got C=%d W=%d)", C, W);
  VC_REQUIRE((reinterpret_cast<uintptr_t>(src) & 7) == 0 && (reinterpret_cast<uintptr_t>(dst_nhwc) & 15) == 0, VC_ERR_BAD_ARG,
             "vc_nhwc_from_nchw: misaligned pointers");
  const int64_t ntiles = (int64_t)B * H * (W / NM_PX);
  VC_REQUIRE(ntiles < 2147483647LL, VC_ERR_UNSUPPORTED, "vc_nhwc_from_nchw: tensor too large");
  int grid = 4 * sm_count();
  if (grid > ntiles) grid = (int)ntiles;
  nm_from_nchw_kernel<<<grid, 256, (size_t)NM_PX * ((C >> 1) + 1) * 4, as_stream(stream)>>>(
      src, static_cast<__nv_bfloat16*>(dst_nhwc), B, C, H, W, (int)ntiles);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" size_t vc_nhwc_bias_grad_workspace_bytes(int C) { return (size_t)kColsumBlocks * (C > 0 ? C : 1) * 8; }

extern "C" int vc_nhwc_bias_grad(const void* g_nhwc, int64_t rows, int C, float* db, void* workspace, size_t workspace_bytes,
                                 void* stream) {
  VC_REQUIRE(g_nhwc && db && rows > 0, VC_ERR_BAD_ARG, "vc_nhwc_bias_grad: bad arguments");
  VC_REQUIRE(C % 8 == 0 && C > 0 && C <= 2048, VC_ERR_UNSUPPORTED, "vc_nhwc_bias_grad: C must be a multiple of 8, <= 2048");
  VC_REQUIRE(workspace && workspace_bytes >= vc_nhwc_bias_grad_workspace_bytes(C), VC_ERR_WORKSPACE,
             "vc_nhwc_bias_grad: workspace too small");
  VC_REQUIRE((reinterpret_cast<uintptr_t>(g_nhwc) & 15) == 0, VC_ERR_BAD_ARG, "vc_nhwc_bias_grad: g must be 16-byte aligned");
  const int groups = 256 / (C >> 3);
  VC_REQUIRE(groups >= 1, VC_ERR_UNSUPPORTED, "vc_nhwc_bias_grad: C too large");
  int grid = kColsumBlocks;
  if ((int64_t)grid * groups > rows) grid = (int)((rows + groups - 1) / groups);
  double* part = static_cast<double*>(workspace);
  nm_colsum_kernel<<<grid, 256, (size_t)groups * C * sizeof(float), as_stream(stream)>>>(
      static_cast<const __nv_bfloat16*>(g_nhwc), rows, C, part);
  VC_LAUNCH_CHECK();
  nm_colsum_fold_kernel<<<(C + 31) / 32, 256, 0, as_stream(stream)>>>(part, grid, C, db);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" int vc_nhwc_convT1_fwd(const void* x_nhwc, const float* Wt, const float* bias, float* y, int B, int Cin, int H,
                                  int W, void* stream) {
  VC_REQUIRE(x_nhwc && Wt && y && B > 0 && H > 0 && W > 0, VC_ERR_BAD_ARG, "vc_nhwc_convT1_fwd: bad arguments");
  VC_REQUIRE(Cin % 8 == 0 && Cin > 0 && Cin <= 512, VC_ERR_UNSUPPORTED, "vc_nhwc_convT1_fwd: Cin must be a multiple of 8, <= 512");
  VC_REQUIRE((reinterpret_cast<uintptr_t>(x_nhwc) & 15) == 0 && (reinterpret_cast<uintptr_t>(y) & 7) == 0, VC_ERR_BAD_ARG,
             "vc_nhwc_convT1_fwd: misaligned pointers");
  if (nhwc_mma_convT1_fwd_supported(B, Cin, H, W))
    return nhwc_mma_convT1_fwd_run(static_cast<const __nv_bfloat16*>(x_nhwc), Wt, bias, y, B, H, W, as_stream(stream));
  nm_convT1_fwd_kernel<<<nm_grid((int64_t)B * H * W), 256, (size_t)16 * Cin * sizeof(float), as_stream(stream)>>>(
      static_cast<const __nv_bfloat16*>(x_nhwc), Wt, bias, y, B, Cin, H, W);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" size_t vc_nhwc_tap_wgrad_workspace_bytes(int C, int KS) {
  return (size_t)kTapBlocksNhwc * (C > 0 ? C : 1) * (KS * KS + 1) * sizeof(double);   // + one column of bias partials
}

// dW (and db = column sums of A when given: Conv2d(1 -> C), KS = 3 only) from the bf16 NHWC tensor A and the fp32 S
static int nm_tap_wgrad_run(const char* who, int KS, const void* a_nhwc, const float* S, float* dW, float* db, int B, int C,
                            int PH, int PW, void* workspace, size_t workspace_bytes, void* stream) {
  VC_REQUIRE(a_nhwc && S && dW && B > 0 && PH > 0 && PW > 0, VC_ERR_BAD_ARG, "%s: bad arguments", who);
  VC_REQUIRE((KS == 3 || KS == 4) && C % 32 == 0 && C > 0 && C <= 512 && PW % 32 == 0, VC_ERR_UNSUPPORTED,
             "%s: need KS 3 or 4, C %% 32 == 0, row width %% 32 == 0 (got KS=%d C=%d PW=%d)", who, KS, C, PW);
  VC_REQUIRE(!db || KS == 3, VC_ERR_UNSUPPORTED, "%s: the fused bias gradient exists for KS = 3 only", who);
  VC_REQUIRE(workspace && workspace_bytes >= vc_nhwc_tap_wgrad_workspace_bytes(C, KS), VC_ERR_WORKSPACE,
             "%s: workspace too small", who);
  VC_REQUIRE((reinterpret_cast<uintptr_t>(a_nhwc) & 15) == 0 && (reinterpret_cast<uintptr_t>(S) & 7) == 0, VC_ERR_BAD_ARG,
             "%s: misaligned pointers", who);
  const int T = KS * KS;
  double* part = static_cast<double*>(workspace);
  double* bpart = part + (size_t)kTapBlocksNhwc * C * T;
  const __nv_bfloat16* a = static_cast<const __nv_bfloat16*>(a_nhwc);
  cudaStream_t s = as_stream(stream);
  int nblk = 0;
  if (nhwc_mma_tap_wgrad_supported(B, C, PH, PW)) {
    const int rc = nhwc_mma_tap_wgrad_run(KS, a, S, part, db ? bpart : nullptr, B, C, PH, PW, kTapBlocksNhwc, &nblk, s);
    if (rc != VC_OK) return rc;
    if (db) {
      nm_tap_fold_kernel<<<(C + 31) / 32, 256, 0, s>>>(bpart, nblk, C, db);
      VC_LAUNCH_CHECK();
    }
  } else {
    const int64_t units = (int64_t)B * PH * (PW / 32);
    nblk = kTapBlocksNhwc / (C / 32);
    if (nblk < 1) nblk = 1;
    if ((int64_t)nblk * 8 > units) nblk = (int)((units + 7) / 8);
    const dim3 grid(nblk, C / 32);
    if (KS == 3)
      nm_tap_wgrad_kernel<3><<<grid, 256, 0, s>>>(a, S, part, B, C, PH, PW, 2 * PH, 2 * PW);
    else
      nm_tap_wgrad_kernel<4><<<grid, 256, 0, s>>>(a, S, part, B, C, PH, PW, 2 * PH, 2 * PW);
    VC_LAUNCH_CHECK();
    if (db) {      // column sums of A by the generic NHWC bias-gradient kernels (their partials fit the same slot)
      const int groups = 256 / (C >> 3);
      int grid2 = kColsumBlocks;
      const int64_t rows = (int64_t)B * PH * PW;
      if ((int64_t)grid2 * groups > rows) grid2 = (int)((rows + groups - 1) / groups);
      nm_colsum_kernel<<<grid2, 256, (size_t)groups * C * sizeof(float), s>>>(a, rows, C, bpart);
      VC_LAUNCH_CHECK();
      nm_colsum_fold_kernel<<<(C + 31) / 32, 256, 0, s>>>(bpart, grid2, C, db);
      VC_LAUNCH_CHECK();
    }
  }
  nm_tap_fold_kernel<<<(C * T + 31) / 32, 256, 0, s>>>(part, nblk, C * T, dW);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" int vc_nhwc_tap_wgrad(int KS, const void* a_nhwc, const float* S, float* dW, int B, int C, int PH, int PW,
                                 void* workspace, size_t workspace_bytes, void* stream) {
  return nm_tap_wgrad_run("vc_nhwc_tap_wgrad", KS, a_nhwc, S, dW, nullptr, B, C, PH, PW, workspace, workspace_bytes, stream);
}

extern "C" size_t vc_nhwc_conv1_wgrad_bias_workspace_bytes(int Cout) { return vc_nhwc_tap_wgrad_workspace_bytes(Cout, 3); }

extern "C" int vc_nhwc_conv1_wgrad_bias(const void* g_nhwc, const float* x, float* dW, float* db, int B, int H, int W, int Cout,
                                        void* workspace, size_t workspace_bytes, void* stream) {
  VC_REQUIRE(db && H % 2 == 0 && W % 2 == 0, VC_ERR_BAD_ARG, "vc_nhwc_conv1_wgrad_bias: db is required, H and W must be even");
  return nm_tap_wgrad_run("vc_nhwc_conv1_wgrad_bias", 3, g_nhwc, x, dW, db, B, Cout, H / 2, W / 2, workspace, workspace_bytes,
                          stream);
}
