// Small Linear layers of the VAE latent head and lyrics branch (12:88-96,104-111: 384->256, 161->256, 256->128,
// 256->32 mu / logvar, 32->256 ...), batch <= 256.
//
// Each op is a few MFLOP -- its cost is launch + memory latency, not arithmetic -- so the tcgen05 path (operand
// conversion kernels + TMA descriptors + a 128-row MMA tile for a 64-row problem: 20-30 us) is the wrong tool.  This is
// ONE launch per op: every CTA owns a 16 x 8 output tile, its four warps split the reduction dimension, read both
// operands straight from global memory (fp32, any stride, no staging), round them to bf16 in registers and run warp-level
// mma.sync m16n8k16 (bf16 operands, fp32 accumulate: the same arithmetic as the tensor-memory path), then fold their
// four partial tiles through shared memory in a fixed order and apply bias / ReLU.
//
//   C[i][j] = sum_r A(i,r) * B(j,r)      A(i,r) = A[i*sai + r*sar],  B(j,r) = B[j*sbj + r*sbr]
//   forward : i = sample, j = out feature, r = in feature     (x [B,Cin] , W [Cout,Cin])
//   dgrad   : i = sample, j = in feature,  r = out feature    (dy [B,Cout], W read column-wise)
//   wgrad   : i = out feature, j = in feature, r = sample     (dy and x read column-wise)
#include <cuda_bf16.h>
#include "common.cuh"

namespace vc {

struct SlArgs {
  const float* A; const float* B; const float* bias; float* C;
  int I, J, R;
  int64_t sai, sar, sbj, sbr;
  int relu;
};

__device__ __forceinline__ uint32_t sl_pack(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);   // .x (low half) = element with the smaller reduction index
  return *reinterpret_cast<uint32_t*>(&v);
}

__device__ __forceinline__ void sl_mma(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

__global__ void __launch_bounds__(128) small_linear_kernel(const SlArgs a) {
  __shared__ float red[3][128];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int tiles_j = (a.J + 7) >> 3;
  const int ti = blockIdx.x / tiles_j, tj = blockIdx.x - ti * tiles_j;
  const int i0 = ti * 16 + g, i1 = i0 + 8, j = tj * 8 + g;
  // rows / columns past the edge are clamped (their results are never stored), so only r needs predicates
  const float* A0 = a.A + (int64_t)(i0 < a.I ? i0 : a.I - 1) * a.sai;
  const float* A1 = a.A + (int64_t)(i1 < a.I ? i1 : a.I - 1) * a.sai;
  const float* Bj = a.B + (int64_t)(j < a.J ? j : a.J - 1) * a.sbj;
  const int nsteps = (a.R + 15) >> 4;
  float c[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 2
  for (int s = warp; s < nsteps; s += 4) {
    const int r = s * 16 + t * 2;
    float av[8], bv[4];
    if (s * 16 + 16 <= a.R) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {   // u: (r, r+1, r+8, r+9)
        const int64_t rr = r + (u & 1) + (u >> 1) * 8;
        av[u] = __ldg(A0 + rr * a.sar);
        av[4 + u] = __ldg(A1 + rr * a.sar);
        bv[u] = __ldg(Bj + rr * a.sbr);
      }
    } else {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int64_t rr = r + (u & 1) + (u >> 1) * 8;
        const uint32_t ok = rr < a.R;
        av[u] = ldg_pred(A0 + rr * a.sar, ok);
        av[4 + u] = ldg_pred(A1 + rr * a.sar, ok);
        bv[u] = ldg_pred(Bj + rr * a.sbr, ok);
      }
    }
    // fragments: a0 (row g, k 2t..2t+1), a1 (row g+8, same k), a2 (row g, k+8), a3 (row g+8, k+8); b0 (k 2t..), b1 (k+8)
    const uint32_t af[4] = {sl_pack(av[0], av[1]), sl_pack(av[4], av[5]), sl_pack(av[2], av[3]), sl_pack(av[6], av[7])};
    const uint32_t bf[2] = {sl_pack(bv[0], bv[1]), sl_pack(bv[2], bv[3])};
    sl_mma(c, af, bf);
  }
  if (warp > 0) {
#pragma unroll
    for (int u = 0; u < 4; ++u) red[warp - 1][u * 32 + lane] = c[u];
  }
  __syncthreads();
  if (warp == 0) {
#pragma unroll
    for (int u = 0; u < 4; ++u) c[u] += red[0][u * 32 + lane] + red[1][u * 32 + lane] + red[2][u * 32 + lane];
    // c0 (row g, col 2t), c1 (row g, col 2t+1), c2 (row g+8, col 2t), c3 (row g+8, col 2t+1)
    const int jc = tj * 8 + t * 2;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int ii = (u & 2) ? i1 : i0, jj = jc + (u & 1);
      if (ii < a.I && jj < a.J) {
        float v = c[u] + (a.bias ? __ldg(a.bias + jj) : 0.f);
        if (a.relu) v = fmaxf(v, 0.f);
        a.C[(int64_t)ii * a.J + jj] = v;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------ fused backward
// The whole backward of one small Linear(+ReLU) layer in ONE launch: g = dy * (y > 0) is formed on the fly while the
// fragments are loaded (y may be null: no activation), CTAs [0, tiles_w) compute dW = g^T x and -- those of the first
// column block -- db = column sums of g; CTAs [tiles_w, tiles_w + tiles_d) compute dx = g W.  It replaces four to five
// launches (ReLU-mask / bias partials / bias fold / wgrad / dgrad) on layers whose cost is launch latency.
struct SlBwdArgs {
  const float* dy; const float* y; const float* x; const float* W;
  float* dW; float* db; float* dx;
  int B, Cin, Cout, tiles_w, tiles_d;
};

__device__ __forceinline__ float sl_g(const float* __restrict__ dy, const float* __restrict__ y, int64_t idx, uint32_t ok) {
  const float d = ldg_pred(dy + idx, ok);
  if (y == nullptr) return d;
  return ldg_pred(y + idx, ok) > 0.f ? d : 0.f;
}

__global__ void __launch_bounds__(128) small_linear_bwd_kernel(const SlBwdArgs a) {
  __shared__ float red[3][128];
  __shared__ float bsum[4][16];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const bool wg = (int)blockIdx.x < a.tiles_w;
  // wgrad: C[i = out feature][j = in feature], r = sample;  dgrad: C[i = sample][j = in feature], r = out feature
  const int I = wg ? a.Cout : a.B, J = a.Cin, R = wg ? a.B : a.Cout;
  const int bid = wg ? blockIdx.x : blockIdx.x - a.tiles_w;
  const int tiles_j = (J + 7) >> 3;
  const int ti = bid / tiles_j, tj = bid - ti * tiles_j;
  const int i0 = ti * 16 + g, i1 = i0 + 8, j = tj * 8 + g;
  const int ic0 = i0 < I ? i0 : I - 1, ic1 = i1 < I ? i1 : I - 1, jc = j < J ? j : J - 1;
  // g(i, r): wgrad reads dy[r*Cout + i], dgrad dy[i*Cout + r];  B(j, r): wgrad x[r*Cin + j], dgrad W[r*Cin + j]
  const int64_t sai = wg ? 1 : a.Cout, sar = wg ? a.Cout : 1;
  const float* Bm = (wg ? a.x : a.W) + jc;
  const int nsteps = (R + 15) >> 4;
  float c[4] = {0.f, 0.f, 0.f, 0.f};
  float s0 = 0.f, s1 = 0.f;   // row sums of g (bias gradient), rows i0 / i1
#pragma unroll 2
  for (int s = warp; s < nsteps; s += 4) {
    const int r = s * 16 + t * 2;
    float av[8], bv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {   // u: (r, r+1, r+8, r+9)
      const int64_t rr = r + (u & 1) + (u >> 1) * 8;
      const uint32_t ok = rr < R;
      av[u] = sl_g(a.dy, a.y, ic0 * sai + rr * sar, ok);
      av[4 + u] = sl_g(a.dy, a.y, ic1 * sai + rr * sar, ok);
      bv[u] = ldg_pred(Bm + rr * a.Cin, ok);
    }
    s0 += (av[0] + av[1]) + (av[2] + av[3]);
    s1 += (av[4] + av[5]) + (av[6] + av[7]);
    const uint32_t af[4] = {sl_pack(av[0], av[1]), sl_pack(av[4], av[5]), sl_pack(av[2], av[3]), sl_pack(av[6], av[7])};
    const uint32_t bf[2] = {sl_pack(bv[0], bv[1]), sl_pack(bv[2], bv[3])};
    sl_mma(c, af, bf);
  }
  // bias gradient: lanes t = 0..3 of a row hold disjoint r; fold them, then the four warps, in a fixed order
  s0 += __shfl_xor_sync(0xffffffffu, s0, 1); s0 += __shfl_xor_sync(0xffffffffu, s0, 2);
  s1 += __shfl_xor_sync(0xffffffffu, s1, 1); s1 += __shfl_xor_sync(0xffffffffu, s1, 2);
  if (t == 0) { bsum[warp][g] = s0; bsum[warp][g + 8] = s1; }
  if (warp > 0) {
#pragma unroll
    for (int u = 0; u < 4; ++u) red[warp - 1][u * 32 + lane] = c[u];
  }
  __syncthreads();
  if (warp == 0) {
#pragma unroll
    for (int u = 0; u < 4; ++u) c[u] += red[0][u * 32 + lane] + red[1][u * 32 + lane] + red[2][u * 32 + lane];
    float* C = wg ? a.dW : a.dx;
    const int jcol = tj * 8 + t * 2;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int ii = (u & 2) ? i1 : i0, jj = jcol + (u & 1);
      if (ii < I && jj < J) C[(int64_t)ii * J + jj] = c[u];
    }
    if (wg && tj == 0 && a.db && lane < 16) {
      const int ii = ti * 16 + lane;
      if (ii < I) a.db[ii] = (bsum[0][lane] + bsum[1][lane]) + (bsum[2][lane] + bsum[3][lane]);
    }
  }
}

bool small_linear_supported(int op, int B, int Cin, int Cout) {
  if (op != VC_OP_LINEAR_FWD && op != VC_OP_LINEAR_DGRAD && op != VC_OP_LINEAR_WGRAD) return false;
  return B <= 256 && (int64_t)Cin * Cout <= (1 << 20);
}

int small_linear_run(int op, const float* in0, const float* in1, const float* bias, float* out, int B, int Cin, int Cout,
                     int relu, cudaStream_t s) {
  SlArgs a;
  a.bias = nullptr; a.relu = 0; a.C = out;
  switch (op) {
    case VC_OP_LINEAR_FWD:    // in0 = x [B,Cin], in1 = W [Cout,Cin] -> y [B,Cout]
      a.A = in0; a.sai = Cin; a.sar = 1; a.B = in1; a.sbj = Cin; a.sbr = 1;
      a.I = B; a.J = Cout; a.R = Cin; a.bias = bias; a.relu = relu;
      break;
    case VC_OP_LINEAR_DGRAD:  // in0 = dy [B,Cout], in1 = W [Cout,Cin] -> dx [B,Cin]
      a.A = in0; a.sai = Cout; a.sar = 1; a.B = in1; a.sbj = 1; a.sbr = Cin;
      a.I = B; a.J = Cin; a.R = Cout;
      break;
    default:                  // in0 = dy [B,Cout], in1 = x [B,Cin] -> dW [Cout,Cin]
      a.A = in0; a.sai = 1; a.sar = Cout; a.B = in1; a.sbj = 1; a.sbr = Cin;
      a.I = Cout; a.J = Cin; a.R = B;
      break;
  }
  const int tiles = ((a.I + 15) / 16) * ((a.J + 7) / 8);
  small_linear_kernel<<<tiles, 128, 0, s>>>(a);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

}  // namespace vc

using namespace vc;

extern "C" int vc_linear_bwd_small_supported(int B, int Cin, int Cout) {
  return small_linear_supported(VC_OP_LINEAR_WGRAD, B, Cin, Cout) ? 1 : 0;
}

extern "C" int vc_linear_bwd_small(const float* dy, const float* y, const float* x, const float* W, float* dW, float* db,
                                   float* dx, int B, int Cin, int Cout, void* stream) {
  VC_REQUIRE(dy && x && W && dW && B > 0 && Cin > 0 && Cout > 0, VC_ERR_BAD_ARG, "vc_linear_bwd_small: bad arguments");
  VC_REQUIRE(small_linear_supported(VC_OP_LINEAR_WGRAD, B, Cin, Cout), VC_ERR_UNSUPPORTED,
             "vc_linear_bwd_small: B=%d Cin=%d Cout=%d is not a small layer (use vc_dense_op)", B, Cin, Cout);
  SlBwdArgs a;
  a.dy = dy; a.y = y; a.x = x; a.W = W; a.dW = dW; a.db = db; a.dx = dx;
  a.B = B; a.Cin = Cin; a.Cout = Cout;
  a.tiles_w = ((Cout + 15) / 16) * ((Cin + 7) / 8);
  a.tiles_d = dx ? ((B + 15) / 16) * ((Cin + 7) / 8) : 0;
  small_linear_bwd_kernel<<<a.tiles_w + a.tiles_d, 128, 0, as_stream(stream)>>>(a);
  VC_LAUNCH_CHECK();
  return VC_OK;
}
