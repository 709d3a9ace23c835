// TMA-fed weight gradients of the hidden conv layers of script 12 (12:87-89,135-137) on bf16 NHWC tensors.
//
// Both layer types are the same contraction over pixels.  With P the tensor indexed at the coarse grid pixel (i, j)
// and Q the one sampled at (2i + th - 1, 2j + tw - 1):
//     D[cp][th][tw][cq] = sum_{b,i,j} P[b,i,j,cp] * Q[b, 2i+th-1, 2j+tw-1, cq]
//   Conv2d k3 s2 p1:          P = g (output gradient, Cout), Q = x (input, Cin),  dW[co][ci][kh][kw] = D[co][kh][kw][ci]
//   ConvTranspose2d k4 s2 p1: P = x (input, Cin),  Q = g (output gradient, Cout), dW[ci][co][kh][kw] = D[ci][kh][kw][co]
// The pixel index is the K dimension of the MMA, so BOTH operands are MN-major: a [64 pixels][64 channels] TMA box of an
// NHWC tensor (channel-contiguous 128-byte rows, 128-byte swizzle) is exactly the canonical MN-major SW128 tile, no
// transposition anywhere.  The stride-2 sampling of Q is the tensor map's traversal stride, the padding its zero fill.
// 32-channel Q tensors are viewed as pixel pairs [B][H][W/2][2 x 32]: a box row holds two neighbouring pixels = two taps;
// the fold kernel picks the (box, half) combinations that are real taps.
//
// One accumulator tile [Cp rows][taps x 64 columns] lives in tensor memory for the whole kernel (up to all 512 columns);
// when the taps of a layer do not fit, tap rows are split over blockIdx.y.  blockIdx.x splits the pixels (split-K), the
// partials are folded in a fixed order.  warps 0, 6, 7, 8: TMA producers (box x of a chunk belongs to producer x % 4, each
// announces its own bytes on the stage barrier), warp 1: MMA issuer, warps 2..5: TMEM -> partials.
#include <string.h>
#include "common.cuh"
#include "tc05.cuh"

namespace vc {

constexpr int NW_PROD = 4;               // TMA-issuing warps (one thread sustains ~1 box per 460 clocks, profiles/tma_rate.cu)
constexpr int NW_THREADS = 192 + 32 * (NW_PROD - 1);
constexpr uint32_t NW_BOX = 64 * 128;    // [64 pixels][64 bf16]

struct NwArgs {
  int Cp, pboxes;            // dense operand: channels, 64-channel boxes (1 or 2)
  int pair, KS, g, bpt;      // Q pair view; taps per side; tap rows per CTA group; Q boxes per tap row
  int nblk;                  // Q boxes per stage = g * bpt
  int B, GH, GW, segs;       // grid pixels; 64-pixel segments per row
  int nch, cps;              // pixel chunks in total / per split
  int stages;
  float* part;               // [nsplit][ngroups][Cp][nblk*64]
};

__global__ void __launch_bounds__(NW_THREADS, 1)
conv_nhwc_wgrad_kernel(const NwArgs a, const __grid_constant__ CUtensorMap mapP, const __grid_constant__ CUtensorMap mapQ) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (tma::s32(smem_raw) & 1023u)) & 1023u);
  const int S = a.stages;
  const uint32_t st_bytes = (uint32_t)(a.pboxes + a.nblk) * NW_BOX;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)S * st_bytes);
  uint64_t* empty = full + S;
  uint64_t* accb = empty + S;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accb + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int z = blockIdx.x, gy = blockIdx.y;
  const int c0 = z * a.cps, c1 = min(a.nch, c0 + a.cps);
  const int n = max(c1 - c0, 0);
  const int cols = a.nblk * 64;
  const uint32_t tmem_cols = cols <= 64 ? 64u : (cols <= 128 ? 128u : (cols <= 256 ? 256u : 512u));

  if (threadIdx.x == 0) {
    tma::prefetch_map(&mapP);
    tma::prefetch_map(&mapQ);
    const int nprod = min(NW_PROD, a.pboxes + a.nblk);
    for (int s = 0; s < S; ++s) { tma::mbar_init(&full[s], nprod); tma::mbar_init(&empty[s], 1); }
    tma::mbar_init(accb, 1);
    tma::mbar_fence_init();
  }
  if (warp == 1) tc05::tmem_alloc(tmem_slot, tmem_cols);
  tc05::fence_before();
  __syncthreads();
  tc05::fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 || warp >= 6) {
    const int me = warp == 0 ? 0 : warp - 5;
    const int nbox = a.pboxes + a.nblk;
    // whole warp, uniform control flow, one elected lane issues: from a `lane == 0` branch every UTMALDG / UTCHMMA sits in
    // an ELECT ... BRA.U.ANY waterfall loop (~460 clocks per box, profiles/tma_rate.cu)
    if (me < nbox) {
      const int mine = (nbox - me + NW_PROD - 1) / NW_PROD;      // boxes me, me + 4, ...
      for (int it = 0; it < n; ++it) {
        const int st = it % S;
        tma::mbar_wait_backoff(&empty[st], (((uint32_t)it / S) & 1u) ^ 1u, 64);
        unsigned char* sp = smem + (size_t)st * st_bytes;
        const int ch = c0 + it;
        const int seg = ch % a.segs, r = ch / a.segs;
        const int i = r % a.GH, b = r / a.GH, j0 = seg * 64;
        if (tc05::elect_one()) {
          tma::mbar_expect_tx(&full[st], (uint32_t)mine * NW_BOX);
          for (int x = me; x < nbox; x += NW_PROD) {
            if (x < a.pboxes) {
              tma::load_4d(sp + x * NW_BOX, &mapP, &full[st], x * 64, j0, i, b);
            } else {
              const int lb = x - a.pboxes;
              const int th = gy * a.g + lb / a.bpt, blk = lb % a.bpt;
              const int w = a.pair ? j0 + blk - 1 : 2 * j0 + blk - 1;
              tma::load_4d(sp + x * NW_BOX, &mapQ, &full[st], 0, w, 2 * i + th - 1, b);
            }
          }
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    {
      // M = 128 always: with a 64-channel P the second 64-row atom re-reads the first box (LBO = 0), rows 64..127 of the
      // accumulator are never read back
      const uint32_t lbo_a = a.pboxes == 2 ? NW_BOX : 0u;
      for (int it = 0; it < n; ++it) {
        const int st = it % S;
        tma::mbar_wait_backoff(&full[st], ((uint32_t)it / S) & 1u, 32);
        tc05::fence_after();
        const uint32_t sa = tma::s32(smem + (size_t)st * st_bytes), sb = sa + a.pboxes * NW_BOX;
        if (tc05::elect_one()) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            const uint64_t da = tc05::desc_mn_sw128(sa + kk * 2048, lbo_a);
            for (int b0 = 0; b0 < a.nblk; b0 += 4) {      // up to four 64-column boxes (N = 256) per instruction
              const int nb = min(4, a.nblk - b0);
              const uint64_t db = tc05::desc_mn_sw128(sb + b0 * NW_BOX + kk * 2048, NW_BOX);
              tc05::mma_bf16(tmem_base + b0 * 64, da, db, tc05::idesc(tc05::kFmtBF16, 128, nb * 64, 1, 1),
                             (uint32_t)((it | kk) != 0));
            }
          }
          tc05::commit(&empty[st]);
        }
        __syncwarp();
      }
      if (n > 0 && tc05::elect_one()) tc05::commit(accb);
      __syncwarp();
    }
  } else {
    const int q = warp & 3;
    const int m = q * 32 + lane;
    if (n > 0) {
      tma::mbar_wait_backoff(accb, 0, 128);
      tc05::fence_after();
    }
    float* prow = a.part + (((int64_t)z * gridDim.y + gy) * a.Cp + m) * cols;
    if (q * 32 < a.Cp) {      // warp-uniform: lane quarters beyond Cp hold the duplicated rows
      for (int cc = 0; cc < (cols >> 5); ++cc) {
        float v[32];
        if (n > 0) {
          tc05::ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + cc * 32, v);
          tc05::ld_wait();
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0.f;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j)
          reinterpret_cast<float4*>(prow + cc * 32)[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
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

// partials [nsplit][ngroups][Cp][cols] -> dW[cp][cq][th][tw].  A block owns 32 consecutive accumulator elements (coalesced
// reads); its eight warps each add every eighth split and the sub-sums are added in warp order (fixed order).
__global__ void __launch_bounds__(256)
nw_fold_kernel(const float* __restrict__ part, int nsplit, int ngroups, int Cp, int Cq, int KS, int g, int bpt, int pair,
               float* __restrict__ dw) {
  __shared__ float sub[8][32];
  const int lane = threadIdx.x & 31, wq = threadIdx.x >> 5;
  const int cols = g * bpt * 64;
  const int64_t per_split = (int64_t)ngroups * Cp * cols;
  const int64_t e = (int64_t)blockIdx.x * 32 + lane;
  float acc = 0.f;
  if (e < per_split) acc = strided_sum_ordered(part + e, per_split, wq, nsplit, 8);
  sub[wq][lane] = acc;
  __syncthreads();
  if (wq != 0 || e >= per_split) return;
#pragma unroll
  for (int i = 1; i < 8; ++i) acc += sub[i][lane];
  const int col = (int)(e % cols);
  const int64_t r = e / cols;
  const int cp = (int)(r % Cp), gy = (int)(r / Cp);
  const int lb = col >> 6, nn = col & 63;
  const int th = gy * g + lb / bpt, blk = lb % bpt;
  int tw, cq;
  if (!pair) { tw = blk; cq = nn; }
  else {
    const int half = nn >> 5;
    cq = nn & 31;
    tw = blk == 0 ? (half ? 0 : -1) : (blk == 1 ? (half ? 2 : 1) : (half ? -1 : 3));
  }
  if (tw < 0 || tw >= KS || th >= KS || cq >= Cq) return;
  dw[(((int64_t)cp * Cq + cq) * KS + th) * KS + tw] = acc;
}

// ---- host ---------------------------------------------------------------------------------------------------------
int nh_make_map(CUtensorMap* m, const void* base, int Cv, int SWv, int SH, int B, int boxC, int boxW, int strideW);  // conv_nhwc.cu

struct NwPlan { bool ok; int Cp, Cq, KS, pair, g, bpt, nblk, ngroups, GH, GW, SH, SW, nch, nsplit, cps, stages; size_t smem, ws; };

static NwPlan nw_plan(int op, int B, int Cin, int H, int W, int Cout) {
  NwPlan p;
  memset(&p, 0, sizeof(p));
  if (B <= 0 || H <= 0 || W <= 0) return p;
  if (op == VC_OP_CONV_WGRAD) {
    if (H % 2 || W % 2) return p;
    p.Cp = Cout; p.Cq = Cin; p.KS = 3; p.GH = H / 2; p.GW = W / 2; p.SH = H; p.SW = W;
  } else if (op == VC_OP_CONVT_WGRAD) {
    p.Cp = Cin; p.Cq = Cout; p.KS = 4; p.GH = H; p.GW = W; p.SH = 2 * H; p.SW = 2 * W;
  } else {
    return p;
  }
  if (p.Cp != 64 && p.Cp != 128) return p;
  if (p.Cq == 32) { p.pair = 1; p.bpt = p.KS == 3 ? 2 : 3; }
  else if (p.Cq == 64) { p.pair = 0; p.bpt = p.KS; }
  else return p;
  if (p.GW % 64) return p;
  p.g = 1;
  for (int g = p.KS; g >= 1; --g)
    if (p.KS % g == 0 && g * p.bpt * 64 <= 512) { p.g = g; break; }
  p.nblk = p.g * p.bpt;
  p.ngroups = p.KS / p.g;
  const int64_t nch = (int64_t)B * p.GH * (p.GW / 64);
  if (nch >= 2147483647LL) return p;
  p.nch = (int)nch;
  int ns = sm_count() / p.ngroups;
  if (ns < 1) ns = 1;
  if (ns > p.nch) ns = p.nch;
  p.cps = (p.nch + ns - 1) / ns;
  p.nsplit = (p.nch + p.cps - 1) / p.cps;
  const size_t st_bytes = (size_t)(p.Cp / 64 + p.nblk) * NW_BOX;
  int stages = (int)((232448 - 1024 - 512) / st_bytes);
  if (stages > 6) stages = 6;
  if (stages < 2) return p;
  p.stages = stages;
  p.smem = (size_t)stages * st_bytes + (2 * stages + 1) * 8 + 16 + 1024;
  p.ws = (size_t)p.nsplit * p.ngroups * p.Cp * p.nblk * 64 * sizeof(float) + 256;
  p.ok = true;
  return p;
}

}  // namespace vc

using namespace vc;

extern "C" int vc_conv_nhwc_wgrad_supported(int op, int B, int Cin, int H, int W, int Cout) {
  return nw_plan(op, B, Cin, H, W, Cout).ok ? 1 : 0;
}

extern "C" size_t vc_conv_nhwc_wgrad_workspace_bytes(int op, int B, int Cin, int H, int W, int Cout) {
  const NwPlan p = nw_plan(op, B, Cin, H, W, Cout);
  return p.ok ? p.ws : 0;
}

extern "C" int vc_conv_nhwc_wgrad(int op, const void* g_nhwc, const void* x_nhwc, float* dW, int B, int Cin, int H, int W,
                                  int Cout, void* workspace, size_t workspace_bytes, void* stream) {
  const NwPlan p = nw_plan(op, B, Cin, H, W, Cout);
  VC_REQUIRE(p.ok, VC_ERR_UNSUPPORTED, "vc_conv_nhwc_wgrad: unsupported op / shape (op %d, B %d, Cin %d, H %d, W %d, Cout %d)",
             op, B, Cin, H, W, Cout);
  VC_REQUIRE(g_nhwc && x_nhwc && dW, VC_ERR_BAD_ARG, "vc_conv_nhwc_wgrad: null pointer");
  VC_REQUIRE(workspace && workspace_bytes >= p.ws, VC_ERR_WORKSPACE, "vc_conv_nhwc_wgrad: workspace too small");
  VC_REQUIRE(((reinterpret_cast<uintptr_t>(g_nhwc) | reinterpret_cast<uintptr_t>(x_nhwc)) & 15) == 0, VC_ERR_BAD_ARG,
             "vc_conv_nhwc_wgrad: NHWC tensors must be 16-byte aligned");
  cudaStream_t s = as_stream(stream);
  const void* P = op == VC_OP_CONV_WGRAD ? g_nhwc : x_nhwc;
  const void* Q = op == VC_OP_CONV_WGRAD ? x_nhwc : g_nhwc;
  CUtensorMap mp, mq;
  int rc = nh_make_map(&mp, P, p.Cp, p.GW, p.GH, B, 64, 64, 1);
  if (rc != VC_OK) return rc;
  if (p.pair) rc = nh_make_map(&mq, Q, 64, p.SW / 2, p.SH, B, 64, 64, 1);
  else rc = nh_make_map(&mq, Q, p.Cq, p.SW, p.SH, B, 64, 128, 2);
  if (rc != VC_OK) return rc;
  NwArgs a;
  memset(&a, 0, sizeof(a));
  a.Cp = p.Cp; a.pboxes = p.Cp / 64; a.pair = p.pair; a.KS = p.KS; a.g = p.g; a.bpt = p.bpt; a.nblk = p.nblk;
  a.B = B; a.GH = p.GH; a.GW = p.GW; a.segs = p.GW / 64; a.nch = p.nch; a.cps = p.cps; a.stages = p.stages;
  a.part = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
  VC_CUDA(cudaFuncSetAttribute(conv_nhwc_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
  conv_nhwc_wgrad_kernel<<<dim3(p.nsplit, p.ngroups), NW_THREADS, p.smem, s>>>(a, mp, mq);
  VC_LAUNCH_CHECK();
  const int64_t per_split = (int64_t)p.ngroups * p.Cp * p.nblk * 64;
  nw_fold_kernel<<<(unsigned)((per_split + 31) / 32), 256, 0, s>>>(a.part, p.nsplit, p.ngroups, p.Cp, p.Cq, p.KS, p.g, p.bpt,
                                                                    p.pair, dW);
  VC_LAUNCH_CHECK();
  return VC_OK;
}
