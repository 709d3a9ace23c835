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

// ---- host -----------------------------------------------------------------------------------------------------------
bool conv_direct_supported(int op, int B, int Cin, int H, int W, int Cout) {
  (void)B; (void)H;
  if (op == VC_OP_CONV_FWD) return Cin == 1 && Cout <= 256;
  if (op == VC_OP_CONVT_FWD) return Cout == 1 && Cin <= 512;
  if (op == VC_OP_CONVT_DGRAD) return Cout == 1 && Cin <= 512;
  return false;
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
    const int grid = direct_grid((int64_t)B * ((H + 1) / 2) * W);
    convT_to1_fwd_kernel<<<grid, 256, (size_t)Cin * 16 * sizeof(float), s>>>(in0, in1, bias, out, B, Cin, H, W, relu,
                                                                               round_bf16);
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
