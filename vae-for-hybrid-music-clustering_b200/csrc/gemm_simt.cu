// fp32 "parity mode" dense layers: one register-tiled SIMT GEMM body, specialised by operand-gather functors into
// Linear / Conv2d(k3,s2,p1) / ConvTranspose2d(k4,s2,p1) forward, data-gradient and weight-gradient.
// Accumulation is fp32 FFMA in a fixed k order (deterministic; split-K partials are reduced in a fixed order).
// The bf16 tcgen05 path lives in tc_gemm.cu; this file is the fp32 reference-precision path the north star asks
// for ("loss curves within 1e-3 relative in fp32").  Reference layers: scripts/12:87-89,98-103,114-117,129-137,
// 159-164; 19:79-95; 06:149-157.
#include "common.cuh"
#include "gemm_functors.cuh"

namespace vc {

constexpr int BK = 16;

// TX*TY threads; block tile (8*TX) x (4*TY); thread (tx,ty) owns rows tx+TX*i (i<8) and cols 4*ty+j (j<4).
template <int TX, int TY, class AL, class BL, class EP>
__global__ void __launch_bounds__(TX * TY)
gemm_simt_kernel(const AL al, const BL bl, const EP ep, int M, int N, int K, int kchunk) {
  constexpr int BM = 8 * TX, BN = 4 * TY, NT = TX * TY;
  constexpr int A_PER = BM * BK / NT, B_PER = (BN * BK + NT - 1) / NT;
  __shared__ float As[BK][BM + 4];
  __shared__ __align__(16) float Bs[BK][BN + 4];
  const int tid = threadIdx.x, tx = tid % TX, ty = tid / TX;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int kbeg = blockIdx.z * kchunk;
  const int kend = min(K, kbeg + kchunk);
  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  float ar[A_PER], br[B_PER];

  auto gload = [&](int k0) {
#pragma unroll
    for (int i = 0; i < A_PER; ++i) {
      const int e = tid + NT * i;
      int mm, kk;
      if (AL::kFastK) { kk = e % BK; mm = e / BK; } else { mm = e % BM; kk = e / BM; }
      const int m = m0 + mm, k = k0 + kk;
      ar[i] = (m < M && k < kend) ? al(m, k) : 0.f;
    }
#pragma unroll
    for (int i = 0; i < B_PER; ++i) {
      const int e = tid + NT * i;
      int nn, kk;
      if (BL::kFastK) { kk = e % BK; nn = e / BK; } else { nn = e % BN; kk = e / BN; }
      const int n = n0 + nn, k = k0 + kk;
      br[i] = (e < BN * BK && n < N && k < kend) ? bl(k, n) : 0.f;
    }
  };
  auto sstore = [&]() {
#pragma unroll
    for (int i = 0; i < A_PER; ++i) {
      const int e = tid + NT * i;
      int mm, kk;
      if (AL::kFastK) { kk = e % BK; mm = e / BK; } else { mm = e % BM; kk = e / BM; }
      As[kk][mm] = ar[i];
    }
#pragma unroll
    for (int i = 0; i < B_PER; ++i) {
      const int e = tid + NT * i;
      if (e < BN * BK) {
        int nn, kk;
        if (BL::kFastK) { kk = e % BK; nn = e / BK; } else { nn = e % BN; kk = e / BN; }
        Bs[kk][nn] = br[i];
      }
    }
  };

  if (kbeg < kend) gload(kbeg);
  for (int k0 = kbeg; k0 < kend; k0 += BK) {
    __syncthreads();
    sstore();
    __syncthreads();
    if (k0 + BK < kend) gload(k0 + BK);
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = As[kk][tx + TX * i];
      const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][4 * ty]);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        acc[i][0] = fmaf(a[i], b.x, acc[i][0]);
        acc[i][1] = fmaf(a[i], b.y, acc[i][1]);
        acc[i][2] = fmaf(a[i], b.z, acc[i][2]);
        acc[i][3] = fmaf(a[i], b.w, acc[i][3]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int n = n0 + 4 * ty + j;
    if (n >= N) continue;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int m = m0 + tx + TX * i;
      if (m < M) ep(m, n, acc[i][j], (int)blockIdx.z);
    }
  }
}

__global__ void splitk_reduce_kernel(const float* __restrict__ part, int nsplit, int64_t MN, float* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= MN) return;
  float acc = 0.f;
  for (int z = 0; z < nsplit; ++z) acc += part[(int64_t)z * MN + i];
  out[i] = acc;
}

template <class EP>
__global__ void splitk_reduce_ep_kernel(const float* __restrict__ part, int nsplit, int M, int N, const EP ep) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t MN = (int64_t)M * N;
  if (i >= MN) return;
  float acc = 0.f;
  for (int z = 0; z < nsplit; ++z) acc += part[(int64_t)z * MN + i];
  ep((int)(i / N), (int)(i % N), acc, 0);
}

template <class AL, class BL, class EP>
static int launch_gemm(const AL& al, const BL& bl, const EP& ep, int M, int N, int K, int nsplit, cudaStream_t s) {
  if (M <= 0 || N <= 0 || K <= 0) return VC_OK;
  const int kchunk = ((K + nsplit - 1) / nsplit + BK - 1) / BK * BK;
  if (N <= 16) {
    dim3 grid((M + 511) / 512, (N + 15) / 16, nsplit);
    gemm_simt_kernel<64, 4, AL, BL, EP><<<grid, 256, 0, s>>>(al, bl, ep, M, N, K, kchunk);
  } else {
    dim3 grid((M + 127) / 128, (N + 63) / 64, nsplit);
    gemm_simt_kernel<16, 16, AL, BL, EP><<<grid, 256, 0, s>>>(al, bl, ep, M, N, K, kchunk);
  }
  VC_LAUNCH_CHECK();
  return VC_OK;
}

// Skinny GEMMs with a huge reduction (the script-12 FC layers: M=64, K or N = 262144): split the reduction over
// enough CTAs to fill the GPU, partials to the workspace, fixed-order reduce fused with the epilogue.
static int linear_splits(int M, int N, int K) {
  const int tiles = ((M + 127) / 128) * ((N + 63) / 64);
  if (tiles >= 148 || K < 4096) return 1;
  int s = (2 * 148 + tiles - 1) / tiles;
  while (s > 1 && K / s < 512) --s;
  return s;
}

template <class AL, class BL, class EP>
static int launch_gemm_auto(const AL& al, const BL& bl, const EP& ep, int M, int N, int K, void* ws, size_t ws_bytes,
                            cudaStream_t s) {
  const int ns = linear_splits(M, N, K);
  if (ns <= 1) return launch_gemm(al, bl, ep, M, N, K, 1, s);
  const int64_t MN = (int64_t)M * N;
  if (!ws || ws_bytes < sizeof(float) * (size_t)ns * MN) {
    set_error("linear: workspace too small for split-K (%d splits of %lld floats)", ns, (long long)MN);
    return VC_ERR_WORKSPACE;
  }
  EpSplitK eps{static_cast<float*>(ws), MN, N};
  int rc = launch_gemm(al, bl, eps, M, N, K, ns, s);
  if (rc != VC_OK) return rc;
  splitk_reduce_ep_kernel<EP><<<(int)((MN + 255) / 256), 256, 0, s>>>(static_cast<float*>(ws), ns, M, N, ep);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

constexpr int kWgradSplits = 296;  // 2 x 148 SMs: weight-gradient GEMMs reduce over B*OH*OW with tiny M x N

static int wgrad_splits(int64_t K) {
  int s = kWgradSplits;
  while (s > 1 && K / s < 4 * BK) s >>= 1;
  return s;
}

}  // namespace vc

using namespace vc;

static inline ConvGeom conv_geom(int B, int Cin, int H, int W, int Cout) {
  return ConvGeom{B, Cin, H, W, Cout, (H - 1) / 2 + 1, (W - 1) / 2 + 1};
}
static inline ConvGeom convT_geom(int B, int Cin, int H, int W, int Cout) {
  return ConvGeom{B, Cin, H, W, Cout, 2 * H, 2 * W};
}
static inline bool fits_int(int64_t v) { return v > 0 && v < (int64_t)2147483647; }

extern "C" size_t vc_linear_workspace_bytes(int M, int N, int K) {
  // worst case over fwd (reduce over K) and dgrad (reduce over N)
  size_t a = (size_t)linear_splits(M, N, K) * M * N, b = (size_t)linear_splits(M, K, N) * M * K;
  if (linear_splits(M, N, K) <= 1) a = 0;
  if (linear_splits(M, K, N) <= 1) b = 0;
  return sizeof(float) * (a > b ? a : b);
}

extern "C" int vc_linear_fwd(const float* x, const float* W, const float* b, float* y, int M, int N, int K, int relu,
                             void* workspace, size_t workspace_bytes, void* stream) {
  VC_REQUIRE(x && W && y && M > 0 && N > 0 && K > 0, VC_ERR_BAD_ARG, "vc_linear_fwd: bad arguments M=%d N=%d K=%d", M,
             N, K);
  ARowK a; a.p = x; a.ld = K;
  BRowK bb; bb.p = W; bb.ld = K;
  EpRowMajor ep{y, N, b, relu, 0};
  return launch_gemm_auto(a, bb, ep, M, N, K, workspace, workspace_bytes, as_stream(stream));
}

extern "C" int vc_linear_dgrad(const float* dy, const float* W, float* dx, int M, int N, int K, void* workspace,
                               size_t workspace_bytes, void* stream) {
  VC_REQUIRE(dy && W && dx && M > 0 && N > 0 && K > 0, VC_ERR_BAD_ARG, "vc_linear_dgrad: bad arguments");
  ARowK a; a.p = dy; a.ld = N;           // A(m, n') over reduction index n' = out feature
  BColK bb{W, K};                         // B(n', k) = W[n'*K + k]
  EpRowMajor ep{dx, K, nullptr, 0, 0};
  return launch_gemm_auto(a, bb, ep, M, K, N, workspace, workspace_bytes, as_stream(stream));
}

extern "C" int vc_linear_wgrad(const float* dy, const float* x, float* dW, int M, int N, int K, int accumulate,
                               void* stream) {
  VC_REQUIRE(dy && x && dW && M > 0 && N > 0 && K > 0, VC_ERR_BAD_ARG, "vc_linear_wgrad: bad arguments");
  AColK a{dy, N};                         // A(n, m) = dy[m*N + n]
  BColK bb{x, K};                         // B(m, k) = x[m*K + k]
  EpRowMajor ep{dW, K, nullptr, 0, accumulate};
  return launch_gemm(a, bb, ep, N, K, M, 1, as_stream(stream));
}

extern "C" size_t vc_conv_wgrad_workspace_bytes(int Cout, int Cin, int ksize) {
  return sizeof(float) * (size_t)kWgradSplits * Cout * Cin * ksize * ksize;
}

extern "C" int vc_conv2d_s2k3_fwd(const float* x, const float* W, const float* b, float* y, int B, int Cin, int H,
                                  int Wd, int Cout, int relu, void* stream) {
  VC_REQUIRE(x && W && y && B > 0 && Cin > 0 && H > 0 && Wd > 0 && Cout > 0, VC_ERR_BAD_ARG,
             "vc_conv2d_s2k3_fwd: bad arguments");
  const ConvGeom g = conv_geom(B, Cin, H, Wd, Cout);
  VC_REQUIRE(fits_int((int64_t)B * g.OH * g.OW) && fits_int((int64_t)B * Cin * H * Wd), VC_ERR_UNSUPPORTED,
             "vc_conv2d_s2k3_fwd: tensor too large for 32-bit pixel index");
  ConvFwdA a{x, g};
  BRowK bb; bb.p = W; bb.ld = Cin * 9;
  EpNCHW ep{y, Cout, g.OH, g.OW, b, relu};
  return launch_gemm(a, bb, ep, B * g.OH * g.OW, Cout, Cin * 9, 1, as_stream(stream));
}

extern "C" int vc_conv2d_s2k3_dgrad(const float* dy, const float* W, float* dx, int B, int Cin, int H, int Wd,
                                    int Cout, void* stream) {
  VC_REQUIRE(dy && W && dx && B > 0 && Cin > 0 && H > 0 && Wd > 0 && Cout > 0, VC_ERR_BAD_ARG,
             "vc_conv2d_s2k3_dgrad: bad arguments");
  const ConvGeom g = conv_geom(B, Cin, H, Wd, Cout);
  VC_REQUIRE(fits_int((int64_t)B * H * Wd), VC_ERR_UNSUPPORTED, "vc_conv2d_s2k3_dgrad: tensor too large");
  ConvDgradA a{dy, g};
  ConvDgradB bb{W, Cin};
  EpNCHW ep{dx, Cin, H, Wd, nullptr, 0};
  return launch_gemm(a, bb, ep, B * H * Wd, Cin, Cout * 9, 1, as_stream(stream));
}

extern "C" int vc_conv2d_s2k3_wgrad(const float* dy, const float* x, float* dW, int B, int Cin, int H, int Wd,
                                    int Cout, void* workspace, size_t workspace_bytes, void* stream) {
  VC_REQUIRE(dy && x && dW && B > 0 && Cin > 0 && H > 0 && Wd > 0 && Cout > 0, VC_ERR_BAD_ARG,
             "vc_conv2d_s2k3_wgrad: bad arguments");
  VC_REQUIRE(workspace && workspace_bytes >= vc_conv_wgrad_workspace_bytes(Cout, Cin, 3), VC_ERR_WORKSPACE,
             "vc_conv2d_s2k3_wgrad: workspace too small");
  const ConvGeom g = conv_geom(B, Cin, H, Wd, Cout);
  const int64_t K = (int64_t)B * g.OH * g.OW;
  VC_REQUIRE(fits_int(K), VC_ERR_UNSUPPORTED, "vc_conv2d_s2k3_wgrad: tensor too large");
  const int ns = wgrad_splits(K);
  const int64_t MN = (int64_t)Cout * Cin * 9;
  ConvWgradA a{dy, g};
  ConvWgradB bb{x, g};
  EpSplitK ep{static_cast<float*>(workspace), MN, Cin * 9};
  int rc = launch_gemm(a, bb, ep, Cout, Cin * 9, (int)K, ns, as_stream(stream));
  if (rc != VC_OK) return rc;
  splitk_reduce_kernel<<<(int)((MN + 255) / 256), 256, 0, as_stream(stream)>>>(static_cast<float*>(workspace), ns, MN,
                                                                              dW);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" int vc_convT2d_s2k4_fwd(const float* x, const float* W, const float* b, float* y, int B, int Cin, int H,
                                   int Wd, int Cout, int relu, void* stream) {
  VC_REQUIRE(x && W && y && B > 0 && Cin > 0 && H > 0 && Wd > 0 && Cout > 0, VC_ERR_BAD_ARG,
             "vc_convT2d_s2k4_fwd: bad arguments");
  const ConvGeom g = convT_geom(B, Cin, H, Wd, Cout);
  VC_REQUIRE(fits_int((int64_t)B * H * Wd) && fits_int((int64_t)B * Cout * g.OH * g.OW / 4 + 1), VC_ERR_UNSUPPORTED,
             "vc_convT2d_s2k4_fwd: tensor too large");
  for (int ph = 0; ph < 2; ++ph)
    for (int pw = 0; pw < 2; ++pw) {
      ConvTFwdA a{x, g, ph, pw};
      ConvTFwdB bb{W, Cout, ph, pw};
      EpNCHWPhase ep{y, Cout, H, Wd, ph, pw, b, relu};
      int rc = launch_gemm(a, bb, ep, B * H * Wd, Cout, Cin * 4, 1, as_stream(stream));
      if (rc != VC_OK) return rc;
    }
  return VC_OK;
}

extern "C" int vc_convT2d_s2k4_dgrad(const float* dy, const float* W, float* dx, int B, int Cin, int H, int Wd,
                                     int Cout, void* stream) {
  VC_REQUIRE(dy && W && dx && B > 0 && Cin > 0 && H > 0 && Wd > 0 && Cout > 0, VC_ERR_BAD_ARG,
             "vc_convT2d_s2k4_dgrad: bad arguments");
  const ConvGeom g = convT_geom(B, Cin, H, Wd, Cout);
  VC_REQUIRE(fits_int((int64_t)B * H * Wd), VC_ERR_UNSUPPORTED, "vc_convT2d_s2k4_dgrad: tensor too large");
  ConvTDgradA a{dy, g};
  BRowK bb; bb.p = W; bb.ld = Cout * 16;  // B(k=(co,kh,kw), n=ci) = W[ci*Cout*16 + k]
  EpNCHW ep{dx, Cin, H, Wd, nullptr, 0};
  return launch_gemm(a, bb, ep, B * H * Wd, Cin, Cout * 16, 1, as_stream(stream));
}

extern "C" int vc_convT2d_s2k4_wgrad(const float* dy, const float* x, float* dW, int B, int Cin, int H, int Wd,
                                     int Cout, void* workspace, size_t workspace_bytes, void* stream) {
  VC_REQUIRE(dy && x && dW && B > 0 && Cin > 0 && H > 0 && Wd > 0 && Cout > 0, VC_ERR_BAD_ARG,
             "vc_convT2d_s2k4_wgrad: bad arguments");
  VC_REQUIRE(workspace && workspace_bytes >= vc_conv_wgrad_workspace_bytes(Cout, Cin, 4), VC_ERR_WORKSPACE,
             "vc_convT2d_s2k4_wgrad: workspace too small");
  const ConvGeom g = convT_geom(B, Cin, H, Wd, Cout);
  const int64_t K = (int64_t)B * H * Wd;
  VC_REQUIRE(fits_int(K), VC_ERR_UNSUPPORTED, "vc_convT2d_s2k4_wgrad: tensor too large");
  const int ns = wgrad_splits(K);
  const int64_t MN = (int64_t)Cin * Cout * 16;
  ConvTWgradA a{x, g};
  ConvTWgradB bb{dy, g};
  EpSplitK ep{static_cast<float*>(workspace), MN, Cout * 16};
  int rc = launch_gemm(a, bb, ep, Cin, Cout * 16, (int)K, ns, as_stream(stream));
  if (rc != VC_OK) return rc;
  splitk_reduce_kernel<<<(int)((MN + 255) / 256), 256, 0, as_stream(stream)>>>(static_cast<float*>(workspace), ns, MN,
                                                                              dW);
  VC_LAUNCH_CHECK();
  return VC_OK;
}
