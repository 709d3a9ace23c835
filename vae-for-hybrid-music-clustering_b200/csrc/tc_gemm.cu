// bf16 tensor-core path of the dense layers (tcgen05 + TMEM + TMA, sm_100a).
//
// Every layer op (Linear / Conv2d k3 s2 p1 / ConvTranspose2d k4 s2 p1; forward, data-gradient, weight-gradient) is the
// GEMM  C[M,N] = A[M,K] * B[N,K]^T  described by the gather functors of gemm_functors.cuh.  This file
//   1. materialises A and B as K-major bf16 matrices (K padded to 64) with the same functors the fp32 path uses,
//   2. runs ONE warp-specialised tcgen05 kernel: warp 0 = TMA producer (2-D maps, 128-byte swizzle) into a
//      multi-stage mbarrier ring, warp 1 = single-thread MMA issuer accumulating 128 x BN fp32 in tensor memory,
//      warps 2..5 = epilogue (tcgen05.ld, thread = output row) applying the op's epilogue functor (bias, ReLU,
//      NCHW scatter, or split-K partial),
//   3. for reductions over B*OH*OW with a tiny M x N (weight gradients, the 262144-wide FC layers) splits K over
//      blockIdx.z and folds the partials in a fixed order.
// Round-1 note: operands are materialised (explicit im2col); the implicit-GEMM variant that feeds the same
// MMA loop straight from NHWC activations through TMA boxes is the next step.
#include <string.h>
#include "common.cuh"
#include "gemm_functors.cuh"
#include "tc05.cuh"

namespace vc {

constexpr int TG_BM = 128;
constexpr int TG_BK = 64;          // bf16 elements per k-block = one 128-byte swizzle atom
constexpr int TG_THREADS = 192;

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// out[m*Kp + k] = bf16(A(m,k)), zero for k >= K.  One thread = 8 consecutive k (16-byte store).
template <class AL>
__global__ void materialize_a_kernel(const AL al, int M, int K, int Kp, __nv_bfloat16* __restrict__ out) {
  const int g8 = Kp >> 3;
  const int64_t total = (int64_t)M * g8;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += stride) {
    const int m = (int)(g / g8), k0 = (int)(g % g8) * 8;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = (k0 + j < K) ? al(m, k0 + j) : 0.f;
    uint4 o = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
    *reinterpret_cast<uint4*>(out + (int64_t)m * Kp + k0) = o;
  }
}
template <class BL>
__global__ void materialize_b_kernel(const BL bl, int N, int K, int Kp, __nv_bfloat16* __restrict__ out) {
  const int g8 = Kp >> 3;
  const int64_t total = (int64_t)N * g8;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += stride) {
    const int n = (int)(g / g8), k0 = (int)(g % g8) * 8;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = (k0 + j < K) ? bl(k0 + j, n) : 0.f;
    uint4 o = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
    *reinterpret_cast<uint4*>(out + (int64_t)n * Kp + k0) = o;
  }
}

// C tile 128 x BN; k-blocks [z*kbs, min(kb_total, (z+1)*kbs)).
template <int BN, int STAGES, class EP>
__global__ void __launch_bounds__(TG_THREADS)
gemm_tc_kernel(const EP ep, int M, int N, int kb_total, int kbs, const __grid_constant__ CUtensorMap mapA,
               const __grid_constant__ CUtensorMap mapB) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (tma::s32(smem_raw) & 1023u)) & 1023u);
  constexpr uint32_t A_BYTES = TG_BM * 128, B_BYTES = BN * 128, ST_BYTES = A_BYTES + B_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)STAGES * ST_BYTES);
  uint64_t* full = bars;
  uint64_t* empty = bars + STAGES;
  uint64_t* t_full = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * TG_BM, n0 = blockIdx.y * BN, z = blockIdx.z;
  const int kb0 = z * kbs, kb1 = min(kb_total, kb0 + kbs);
  const int nkb = max(kb1 - kb0, 0);

  if (threadIdx.x == 0) {
    tma::prefetch_map(&mapA);
    tma::prefetch_map(&mapB);
    for (int s = 0; s < STAGES; ++s) { tma::mbar_init(&full[s], 1); tma::mbar_init(&empty[s], 1); }
    tma::mbar_init(t_full, 1);
    tma::mbar_fence_init();
  }
  if (warp == 1) tc05::tmem_alloc(tmem_slot, BN < 32 ? 32 : BN);
  tc05::fence_before();
  __syncthreads();
  tc05::fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (tc05::elect_one()) {   // a single lane the compiler knows about: no waterfall loop around UTCHMMA / UTMALDG
      for (int i = 0; i < nkb; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        tma::mbar_wait(&empty[s], ph ^ 1);
        unsigned char* dst = smem + (size_t)s * ST_BYTES;
        tma::mbar_expect_tx(&full[s], ST_BYTES);
        tma::load_2d(dst, &mapA, &full[s], (kb0 + i) * TG_BK, m0);
        tma::load_2d(dst + A_BYTES, &mapB, &full[s], (kb0 + i) * TG_BK, n0);
      }
    }
  } else if (warp == 1) {
    if (tc05::elect_one()) {   // a single lane the compiler knows about: no waterfall loop around UTCHMMA / UTMALDG
      constexpr uint32_t kIdesc = tc05::idesc(tc05::kFmtBF16, TG_BM, BN, 0, 0);
      for (int i = 0; i < nkb; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        tma::mbar_wait(&full[s], ph);
        tc05::fence_after();
        const uint32_t sa = tma::s32(smem + (size_t)s * ST_BYTES), sb = sa + A_BYTES;
#pragma unroll
        for (int kk = 0; kk < TG_BK / 16; ++kk)
          tc05::mma_bf16(tmem_base, tc05::desc_k_sw128(sa + kk * 32), tc05::desc_k_sw128(sb + kk * 32), kIdesc,
                         (i | kk) != 0);
        tc05::commit(&empty[s]);
      }
      tc05::commit(t_full);
    }
  } else {
    const int q = warp & 3;
    const int m = m0 + q * 32 + lane;
    if (nkb > 0) {
      tma::mbar_wait(t_full, 0);
      tc05::fence_after();
    }
    const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
    for (int cb = 0; cb < BN; cb += 32) {
      float v[32];
      if (nkb > 0) {
        tc05::ld_32x32(tacc + cb, v);
        tc05::ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = 0.f;
      }
      if (m < M) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int n = n0 + cb + j;
          if (n < N) ep(m, n, v[j], z);
        }
      }
    }
  }
  tc05::fence_before();
  __syncthreads();
  if (warp == 1) {
    tc05::fence_after();
    tc05::tmem_dealloc(tmem_base, BN < 32 ? 32 : BN);
  }
}

template <class EP>
__global__ void splitk_reduce_tc_kernel(const float* __restrict__ part, int nsplit, int M, int N, const EP ep) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t MN = (int64_t)M * N;
  if (i >= MN) return;
  float acc = 0.f;
  for (int zz = 0; zz < nsplit; ++zz) acc += part[(int64_t)zz * MN + i];
  ep((int)(i / N), (int)(i % N), acc, 0);
}

}  // namespace vc

namespace vc {

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
static inline int tc_kp(int K) { return (K + TG_BK - 1) / TG_BK * TG_BK; }
static inline int tc_bn(int N) { return N <= 32 ? 32 : (N <= 64 ? 64 : 128); }

static int tc_splits(int M, int N, int K) {
  const int bn = tc_bn(N);
  const int tiles = ((M + TG_BM - 1) / TG_BM) * ((N + bn - 1) / bn);
  const int kb = tc_kp(K) / TG_BK;
  if (tiles >= 148 || kb < 16) return 1;
  int s = (2 * 148 + tiles - 1) / tiles;
  if (s > kb / 4) s = kb / 4;
  return s < 1 ? 1 : s;
}

size_t tc_gemm_workspace(int M, int N, int K) {
  const int Kp = tc_kp(K);
  const int ns = tc_splits(M, N, K);
  return align_up((size_t)M * Kp * 2, 1024) + align_up((size_t)N * Kp * 2, 1024) +
         (ns > 1 ? align_up((size_t)ns * M * N * 4, 1024) : 0) + 1024;
}

template <int BN, int STAGES, class EP>
static int tc_launch(const EP& ep, int M, int N, int kb_total, int kbs, int ns, const CUtensorMap& ma,
                     const CUtensorMap& mb, cudaStream_t s) {
  auto kern = gemm_tc_kernel<BN, STAGES, EP>;
  const size_t smem = (size_t)STAGES * (TG_BM * 128 + BN * 128) + (2 * STAGES + 1) * 8 + 16 + 1024;
  VC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((M + TG_BM - 1) / TG_BM, (N + BN - 1) / BN, ns);
  kern<<<grid, TG_THREADS, smem, s>>>(ep, M, N, kb_total, kbs, ma, mb);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

template <class EP>
static int tc_launch_bn(const EP& ep, int M, int N, int kb_total, int kbs, int ns, const CUtensorMap& ma,
                        const CUtensorMap& mb, cudaStream_t s) {
  switch (tc_bn(N)) {
    case 32: return tc_launch<32, 4, EP>(ep, M, N, kb_total, kbs, ns, ma, mb, s);
    case 64: return tc_launch<64, 4, EP>(ep, M, N, kb_total, kbs, ns, ma, mb, s);
    default: return tc_launch<128, 3, EP>(ep, M, N, kb_total, kbs, ns, ma, mb, s);
  }
}

// materialise + GEMM (+ split-K fold) for one (AL, BL, EP) triple
template <class AL, class BL, class EP>
static int tc_gemm(const AL& al, const BL& bl, const EP& ep, int M, int N, int K, void* ws, size_t ws_bytes,
                   cudaStream_t s) {
  if (M <= 0 || N <= 0 || K <= 0) return VC_OK;
  VC_REQUIRE(ws && ws_bytes >= tc_gemm_workspace(M, N, K), VC_ERR_WORKSPACE,
             "tc_gemm: workspace too small (%zu < %zu)", ws_bytes, tc_gemm_workspace(M, N, K));
  const int Kp = tc_kp(K);
  const int ns = tc_splits(M, N, K);
  unsigned char* p = reinterpret_cast<unsigned char*>(align_up(reinterpret_cast<uintptr_t>(ws), 1024));
  __nv_bfloat16* Ab = reinterpret_cast<__nv_bfloat16*>(p);
  p += align_up((size_t)M * Kp * 2, 1024);
  __nv_bfloat16* Bb = reinterpret_cast<__nv_bfloat16*>(p);
  p += align_up((size_t)N * Kp * 2, 1024);
  float* part = reinterpret_cast<float*>(p);
  const int cap = sm_count() * 16;
  int64_t ga = ((int64_t)M * (Kp / 8) + 255) / 256, gb = ((int64_t)N * (Kp / 8) + 255) / 256;
  materialize_a_kernel<AL><<<(int)(ga > cap ? cap : ga), 256, 0, s>>>(al, M, K, Kp, Ab);
  VC_LAUNCH_CHECK();
  materialize_b_kernel<BL><<<(int)(gb > cap ? cap : gb), 256, 0, s>>>(bl, N, K, Kp, Bb);
  VC_LAUNCH_CHECK();
  CUtensorMap ma, mb;
  const int bn = tc_bn(N);
  {
    const uint64_t dims[2] = {(uint64_t)Kp, (uint64_t)M};
    const uint64_t str[1] = {(uint64_t)Kp * 2};
    const uint32_t box[2] = {(uint32_t)TG_BK, (uint32_t)TG_BM};
    int rc = tma::make_map(&ma, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, Ab, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc != VC_OK) return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)Kp, (uint64_t)N};
    const uint64_t str[1] = {(uint64_t)Kp * 2};
    const uint32_t box[2] = {(uint32_t)TG_BK, (uint32_t)bn};
    int rc = tma::make_map(&mb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, Bb, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc != VC_OK) return rc;
  }
  const int kb_total = Kp / TG_BK;
  if (ns <= 1) return tc_launch_bn(ep, M, N, kb_total, kb_total, 1, ma, mb, s);
  const int kbs = (kb_total + ns - 1) / ns;
  const int ns_eff = (kb_total + kbs - 1) / kbs;
  EpSplitK eps{part, (int64_t)M * N, N};
  int rc = tc_launch_bn(eps, M, N, kb_total, kbs, ns_eff, ma, mb, s);
  if (rc != VC_OK) return rc;
  const int64_t MN = (int64_t)M * N;
  splitk_reduce_tc_kernel<EP><<<(int)((MN + 255) / 256), 256, 0, s>>>(part, ns_eff, M, N, ep);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

}  // namespace vc

namespace vc {  // conv_igemm.cu
bool conv_igemm_supported(int op, int B, int Cin, int H, int W, int Cout);
size_t conv_igemm_workspace(int op, int B, int Cin, int H, int W, int Cout);
int conv_igemm_run(int op, const float* in0, const float* in1, const float* bias, float* out, int B, int Cin, int H,
                   int W, int Cout, int relu, void* ws, size_t wsb, cudaStream_t s);
bool conv_direct_supported(int op, int B, int Cin, int H, int W, int Cout);   // conv_direct.cu
int conv_direct_run(int op, const float* in0, const float* in1, const float* bias, float* out, int B, int Cin, int H,
                    int W, int Cout, int relu, int round_bf16, cudaStream_t s);
bool conv_direct_wgrad_supported(int op, int B, int Cin, int H, int W, int Cout);
size_t conv_direct_wgrad_workspace(int op, int Cin, int Cout);
int conv_direct_wgrad_run(int op, const float* dy, const float* x, float* dw, int B, int Cin, int H, int W, int Cout,
                          int round_bf16, void* ws, size_t wsb, cudaStream_t s);
bool skinny_gemm_supported(int op, int B, int Cin, int Cout, const void* in0, const void* in1);   // skinny_gemm.cu
size_t skinny_gemm_workspace(int op, int B, int Cin, int Cout);
int skinny_gemm_run(int op, const float* in0, const float* in1, const float* bias, float* out, int B, int Cin, int Cout,
                    int relu, void* ws, size_t wsb, cudaStream_t s);
bool small_linear_supported(int op, int B, int Cin, int Cout);   // small_linear.cu
int small_linear_run(int op, const float* in0, const float* in1, const float* bias, float* out, int B, int Cin, int Cout,
                     int relu, cudaStream_t s);
bool conv_wgrad_igemm_supported(int op, int B, int Cin, int H, int W, int Cout);   // conv_wgrad_igemm.cu
size_t conv_wgrad_igemm_workspace(int op, int B, int Cin, int H, int W, int Cout);
int conv_wgrad_igemm_run(int op, const float* dy, const float* x, float* dw, int B, int Cin, int H, int W, int Cout,
                         void* ws, size_t wsb, cudaStream_t s);
}  // namespace vc

using namespace vc;

namespace {
struct OpDims { int M, N, K; };

// GEMM shape of each op (convT forward: per output-parity phase)
OpDims op_dims(int op, int B, int Cin, int H, int W, int Cout) {
  const int OH = (H - 1) / 2 + 1, OW = (W - 1) / 2 + 1;
  switch (op) {
    case VC_OP_LINEAR_FWD: return {B, Cout, Cin};
    case VC_OP_LINEAR_DGRAD: return {B, Cin, Cout};
    case VC_OP_LINEAR_WGRAD: return {Cout, Cin, B};
    case VC_OP_CONV_FWD: return {B * OH * OW, Cout, Cin * 9};
    case VC_OP_CONV_DGRAD: return {B * H * W, Cin, Cout * 9};
    case VC_OP_CONV_WGRAD: return {Cout, Cin * 9, B * OH * OW};
    case VC_OP_CONVT_FWD: return {B * H * W, Cout, Cin * 4};
    case VC_OP_CONVT_DGRAD: return {B * H * W, Cin, Cout * 16};
    case VC_OP_CONVT_WGRAD: return {Cin, Cout * 16, B * H * W};
    default: return {0, 0, 0};
  }
}
}  // namespace

extern "C" size_t vc_dense_op_workspace_bytes(int op, int precision, int B, int Cin, int H, int W, int Cout) {
  if (precision == VC_PREC_BF16_TC) {
    if (conv_igemm_supported(op, B, Cin, H, W, Cout)) return conv_igemm_workspace(op, B, Cin, H, W, Cout);
    if (conv_direct_wgrad_supported(op, B, Cin, H, W, Cout)) {   // a misaligned pointer falls back to the GEMM below
      const size_t a = conv_direct_wgrad_workspace(op, Cin, Cout);
      const size_t b = conv_wgrad_igemm_supported(op, B, Cin, H, W, Cout)
                           ? conv_wgrad_igemm_workspace(op, B, Cin, H, W, Cout)
                           : tc_gemm_workspace(op_dims(op, B, Cin, H, W, Cout).M, op_dims(op, B, Cin, H, W, Cout).N,
                                               op_dims(op, B, Cin, H, W, Cout).K);
      return a > b ? a : b;
    }
    if (conv_wgrad_igemm_supported(op, B, Cin, H, W, Cout)) return conv_wgrad_igemm_workspace(op, B, Cin, H, W, Cout);
    const OpDims g = op_dims(op, B, Cin, H, W, Cout);
    const size_t a = tc_gemm_workspace(g.M, g.N, g.K), b = skinny_gemm_workspace(op, B, Cin, Cout);
    return a > b ? a : b;   // the skinny path may be refused at run time for a misaligned pointer
  }
  switch (op) {
    case VC_OP_LINEAR_FWD: case VC_OP_LINEAR_DGRAD: return vc_linear_workspace_bytes(B, Cout, Cin);
    case VC_OP_CONV_WGRAD: return vc_conv_wgrad_workspace_bytes(Cout, Cin, 3);
    case VC_OP_CONVT_WGRAD: return vc_conv_wgrad_workspace_bytes(Cout, Cin, 4);
    default: return 0;
  }
}

static int dense_op_tc(int op, const float* in0, const float* in1, const float* bias, float* out, int B, int Cin, int H,
                       int W, int Cout, int relu, void* ws, size_t wsb, cudaStream_t s) {
  const OpDims g = op_dims(op, B, Cin, H, W, Cout);
  switch (op) {
    case VC_OP_LINEAR_FWD: {  // in0 = x [B,Cin], in1 = W [Cout,Cin]
      ARowK a; a.p = in0; a.ld = Cin;
      BRowK b; b.p = in1; b.ld = Cin;
      return tc_gemm(a, b, EpRowMajor{out, Cout, bias, relu, 0}, g.M, g.N, g.K, ws, wsb, s);
    }
    case VC_OP_LINEAR_DGRAD: {  // in0 = dy [B,Cout], in1 = W [Cout,Cin] -> dx [B,Cin]
      ARowK a; a.p = in0; a.ld = Cout;
      return tc_gemm(a, BColK{in1, Cin}, EpRowMajor{out, Cin, nullptr, 0, 0}, g.M, g.N, g.K, ws, wsb, s);
    }
    case VC_OP_LINEAR_WGRAD:  // in0 = dy [B,Cout], in1 = x [B,Cin] -> dW [Cout,Cin]
      return tc_gemm(AColK{in0, Cout}, BColK{in1, Cin}, EpRowMajor{out, Cin, nullptr, 0, 0}, g.M, g.N, g.K, ws, wsb, s);
    case VC_OP_CONV_FWD: {  // in0 = x, in1 = W [Cout,Cin,3,3]
      const ConvGeom cg{B, Cin, H, W, Cout, (H - 1) / 2 + 1, (W - 1) / 2 + 1};
      BRowK b; b.p = in1; b.ld = Cin * 9;
      return tc_gemm(ConvFwdA{in0, cg}, b, EpNCHW{out, Cout, cg.OH, cg.OW, bias, relu}, g.M, g.N, g.K, ws, wsb, s);
    }
    case VC_OP_CONV_DGRAD: {  // in0 = dy, in1 = W -> dx
      const ConvGeom cg{B, Cin, H, W, Cout, (H - 1) / 2 + 1, (W - 1) / 2 + 1};
      return tc_gemm(ConvDgradA{in0, cg}, ConvDgradB{in1, Cin}, EpNCHW{out, Cin, H, W, nullptr, 0}, g.M, g.N, g.K, ws,
                     wsb, s);
    }
    case VC_OP_CONV_WGRAD: {  // in0 = dy, in1 = x -> dW [Cout,Cin,3,3]
      const ConvGeom cg{B, Cin, H, W, Cout, (H - 1) / 2 + 1, (W - 1) / 2 + 1};
      return tc_gemm(ConvWgradA{in0, cg}, ConvWgradB{in1, cg}, EpRowMajor{out, Cin * 9, nullptr, 0, 0}, g.M, g.N, g.K,
                     ws, wsb, s);
    }
    case VC_OP_CONVT_FWD: {  // in0 = x, in1 = W [Cin,Cout,4,4] -> y [B,Cout,2H,2W], one GEMM per output parity
      const ConvGeom cg{B, Cin, H, W, Cout, 2 * H, 2 * W};
      for (int ph = 0; ph < 2; ++ph)
        for (int pw = 0; pw < 2; ++pw) {
          int rc = tc_gemm(ConvTFwdA{in0, cg, ph, pw}, ConvTFwdB{in1, Cout, ph, pw},
                           EpNCHWPhase{out, Cout, H, W, ph, pw, bias, relu}, g.M, g.N, g.K, ws, wsb, s);
          if (rc != VC_OK) return rc;
        }
      return VC_OK;
    }
    case VC_OP_CONVT_DGRAD: {  // in0 = dy, in1 = W -> dx [B,Cin,H,W]
      const ConvGeom cg{B, Cin, H, W, Cout, 2 * H, 2 * W};
      BRowK b; b.p = in1; b.ld = Cout * 16;
      return tc_gemm(ConvTDgradA{in0, cg}, b, EpNCHW{out, Cin, H, W, nullptr, 0}, g.M, g.N, g.K, ws, wsb, s);
    }
    case VC_OP_CONVT_WGRAD: {  // in0 = dy, in1 = x -> dW [Cin,Cout,4,4]
      const ConvGeom cg{B, Cin, H, W, Cout, 2 * H, 2 * W};
      return tc_gemm(ConvTWgradA{in1, cg}, ConvTWgradB{in0, cg}, EpRowMajor{out, Cout * 16, nullptr, 0, 0}, g.M, g.N,
                     g.K, ws, wsb, s);
    }
  }
  set_error("vc_dense_op: unknown op %d", op);
  return VC_ERR_BAD_ARG;
}

extern "C" int vc_dense_op(int op, int precision, const float* in0, const float* in1, const float* bias, float* out,
                           int B, int Cin, int H, int W, int Cout, int relu, void* workspace, size_t workspace_bytes,
                           void* stream) {
  VC_REQUIRE(in0 && in1 && out && B > 0 && Cin > 0 && H > 0 && W > 0 && Cout > 0, VC_ERR_BAD_ARG,
             "vc_dense_op(op=%d): bad arguments", op);
  VC_REQUIRE(precision == VC_PREC_FP32 || precision == VC_PREC_BF16_TC, VC_ERR_BAD_ARG,
             "vc_dense_op: unknown precision %d", precision);
  if (precision == VC_PREC_BF16_TC) {
    const OpDims g = op_dims(op, B, Cin, H, W, Cout);
    VC_REQUIRE((int64_t)g.M > 0 && (int64_t)B * H * W * 4 < 2147483647LL, VC_ERR_UNSUPPORTED,
               "vc_dense_op: tensor too large for 32-bit pixel index");
    if (conv_direct_supported(op, B, Cin, H, W, Cout))  // single-channel ends: streaming FFMA kernels
      return conv_direct_run(op, in0, in1, bias, out, B, Cin, H, W, Cout, relu, 1, as_stream(stream));
    if (conv_direct_wgrad_supported(op, B, Cin, H, W, Cout) &&
        ((reinterpret_cast<uintptr_t>(in0) | reinterpret_cast<uintptr_t>(in1)) & 15) == 0)  // streaming tap reduction
      // FFMA on the fp32 values: a weight gradient gains nothing from rounding dy and x to bf16 first
      return conv_direct_wgrad_run(op, in0, in1, out, B, Cin, H, W, Cout, 0, workspace, workspace_bytes,
                                   as_stream(stream));
    if (conv_igemm_supported(op, B, Cin, H, W, Cout))  // implicit GEMM: no im2col matrix in HBM
      return conv_igemm_run(op, in0, in1, bias, out, B, Cin, H, W, Cout, relu, workspace, workspace_bytes,
                            as_stream(stream));
    if (H == 1 && W == 1 && skinny_gemm_supported(op, B, Cin, Cout, in0, in1))  // huge fp32 matrix read in place
      return skinny_gemm_run(op, in0, in1, bias, out, B, Cin, Cout, relu, workspace, workspace_bytes, as_stream(stream));
    if (H == 1 && W == 1 && small_linear_supported(op, B, Cin, Cout))  // latency-bound: one mma.sync launch, no staging
      return small_linear_run(op, in0, in1, bias, out, B, Cin, Cout, relu, as_stream(stream));
    if (conv_wgrad_igemm_supported(op, B, Cin, H, W, Cout))
      return conv_wgrad_igemm_run(op, in0, in1, out, B, Cin, H, W, Cout, workspace, workspace_bytes, as_stream(stream));
    return dense_op_tc(op, in0, in1, bias, out, B, Cin, H, W, Cout, relu, workspace, workspace_bytes, as_stream(stream));
  }
  switch (op) {
    case VC_OP_LINEAR_FWD: return vc_linear_fwd(in0, in1, bias, out, B, Cout, Cin, relu, workspace, workspace_bytes, stream);
    case VC_OP_LINEAR_DGRAD: return vc_linear_dgrad(in0, in1, out, B, Cout, Cin, workspace, workspace_bytes, stream);
    case VC_OP_LINEAR_WGRAD: return vc_linear_wgrad(in0, in1, out, B, Cout, Cin, 0, stream);
    case VC_OP_CONV_FWD: return vc_conv2d_s2k3_fwd(in0, in1, bias, out, B, Cin, H, W, Cout, relu, stream);
    case VC_OP_CONV_DGRAD: return vc_conv2d_s2k3_dgrad(in0, in1, out, B, Cin, H, W, Cout, stream);
    case VC_OP_CONV_WGRAD: return vc_conv2d_s2k3_wgrad(in0, in1, out, B, Cin, H, W, Cout, workspace, workspace_bytes, stream);
    case VC_OP_CONVT_FWD: return vc_convT2d_s2k4_fwd(in0, in1, bias, out, B, Cin, H, W, Cout, relu, stream);
    case VC_OP_CONVT_DGRAD: return vc_convT2d_s2k4_dgrad(in0, in1, out, B, Cin, H, W, Cout, stream);
    case VC_OP_CONVT_WGRAD: return vc_convT2d_s2k4_wgrad(in0, in1, out, B, Cin, H, W, Cout, workspace, workspace_bytes, stream);
  }
  set_error("vc_dense_op: unknown op %d", op);
  return VC_ERR_BAD_ARG;
}
