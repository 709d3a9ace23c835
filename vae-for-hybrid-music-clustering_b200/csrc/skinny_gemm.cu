// Skinny GEMMs of the two 67M-parameter fully connected layers of the script-12 VAE (12:98-103,129-132) on tcgen05.
//
// With a batch of 64 every Linear op (forward, data-gradient, weight-gradient) streams one huge fp32 matrix -- the
// weight, or its gradient -- exactly once and is HBM-bound; the other operand is tiny.  Instead of materialising a bf16
// copy of the big matrix, this kernel reads it in place:
//   D[M, N] = sum_k A(m,k) * Bs[n][k]        N <= 256 (batch, or the 256-wide hidden layer)
//   * A is the big fp32 matrix, either k-contiguous (A(m,k) = A[m*lda + k]) or m-contiguous (A(m,k) = A[k*lda + m]).
//     Eight producer warps load it coalesced, convert to bf16 and store K-major into 128-byte-swizzled shared memory
//     (the m-contiguous case transposes on the way: lanes run along m, two consecutive k are packed per 4-byte store).
//   * Bs is the small operand, pre-converted to bf16 [BN][Kp] (K-major) and fetched by TMA.
//   * one elected thread issues tcgen05.mma (128 x BN fp32 accumulator in tensor memory); the producer warps then read
//     it back and store either D[m][n] or its transpose (lanes = consecutive m, so transposed stores are coalesced),
//     with bias (per m) and ReLU; or write a split-K partial that sk_fold_kernel reduces in a fixed order.
#include <string.h>
#include "common.cuh"
#include "tc05.cuh"

namespace vc {

constexpr int SK_THREADS = 288;              // 8 producer/epilogue warps + 1 MMA warp
constexpr uint32_t SK_A_BYTES = 128 * 128;   // [128 rows][64 bf16]

struct SkArgs {
  const float* A; int64_t lda; int a_mcontig;
  int M, N, K, BN;
  float* out; int64_t ldo; int transposed;   // out[m*ldo + n] or out[n*ldo + m]
  const float* bias; int relu;               // bias indexed by m
  float* part;                               // split-K partials [nsplit][N][M] (transposed: coalesced), or null
  int nkb, kbs, stages;
  uint32_t idesc;
  // fused Adam (weight-gradient ops only): `out` is the PARAMETER, updated in place from the accumulator; the gradient
  // is never written (saves its 4 B/element write and the optimizer's 4 B/element read)
  // bf16 shadow of the big matrix (a_bf16): the A tile is one TMA load (K-major box, or two MN-major boxes when the
  // matrix is m-contiguous) instead of eight producer warps reading fp32 and converting -- half the bytes, no instructions
  int a_bf16;
  // fused Adam on a parameter stored [N][M] (transposed): the MMA operands swap roles (D_h[n][m] = Bs_h . A^T, two 128-row
  // halves of n), so the TMEM lane is the parameter ROW in both layouts and the epilogue below needs no second code path
  // (the m-lane variant read its float4s as four scalar shared-memory loads with 8-way bank conflicts: 422 vs 348 us)
  int swap;
  int adam;
  float* ad_m; float* ad_v;
  __nv_bfloat16* ad_shadow;                  // fused Adam also refreshes the shadow (same layout as the parameter) or null
  const VcAdamDeviceStep* ad_step;          // device-side bias corrections (CUDA-graph replay), or null: use ad_bc1/2
  float ad_lr, ad_b1, ad_b2, ad_eps, ad_bc1, ad_bc2;
};

// torch.optim.Adam's arithmetic in the order adam_multi_kernel (misc.cu) uses, so fused and unfused steps agree
// APPROX (the default of the fused update, which exists in bf16 mode only; VAECLUST_FC_ADAM_IEEE=1 restores IEEE): the square
// root and the division as sqrt.approx / div.approx (1 + 2 ulp) -- 4 instructions instead of ~17, and no special-case paths.
// The fused epilogue is instruction-latency bound (six more instructions per element cost 12 %), and the IEEE sequences
// leave their fast path for zero / tiny inputs as soon as ONE lane of a warp holds one, which training data is full of
// (weights that feed from dead ReLU units have gradient, exp_avg and exp_avg_sq exactly 0): 337 us per launch with dense random
// gradients, 470 us with 30 % zero rows / columns (profiles/prof_fc_adam.py sparse).  The relative error of the update
// (<= 4e-7) is four orders below the bf16 rounding of the operands that produced the gradient.
template <bool APPROX>
__device__ __forceinline__ void sk_adam(float g, float& p, float& m, float& v, float b1, float b2, float step_size,
                                        float inv_sqrt_bc2, float eps) {
  m = b1 * m + (1.f - b1) * g;
  v = b2 * v + (1.f - b2) * g * g;
  if (APPROX) {
    float sq;
    asm("sqrt.approx.f32 %0, %1;" : "=f"(sq) : "f"(v));
    p -= step_size * __fdividef(m, sq * inv_sqrt_bc2 + eps);
  } else {
    p -= step_size * (m / (sqrtf(v) * inv_sqrt_bc2 + eps));
  }
}

__device__ __forceinline__ uint32_t sk_pack(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float sk_ldg(const float* p, uint32_t pred) {
  float v;
  asm("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %2, 0;\n\tmov.f32 %0, 0f00000000;\n\t@q ld.global.nc.f32 %0, [%1];\n\t}"
      : "=f"(v)
      : "l"(p), "r"(pred));
  return v;
}
__device__ __forceinline__ float4 sk_ldg4(const float* p, uint32_t pred) {
  float4 v;
  asm("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %5, 0;\n\tmov.f32 %0, 0f00000000;\n\tmov.f32 %1, 0f00000000;\n\t"
      "mov.f32 %2, 0f00000000;\n\tmov.f32 %3, 0f00000000;\n\t@q ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];\n\t}"
      : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
      : "l"(p), "r"(pred));
  return v;
}

// small operand -> bf16 [BN][Kp], zero padded.  colmajor == 0: B(n,k) = src[n*ld + k]; else B(n,k) = src[k*ld + n]
__global__ void sk_small_operand_kernel(const float* __restrict__ src, int64_t ld, int colmajor, int N, int K, int BN,
                                        int Kp, __nv_bfloat16* __restrict__ out) {
  const int64_t total = (int64_t)BN * Kp;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
    int n, k;
    if (colmajor) { k = (int)(e / BN); n = (int)(e - (int64_t)k * BN); }   // consecutive threads -> consecutive n (coalesced reads)
    else { n = (int)(e / Kp); k = (int)(e - (int64_t)n * Kp); }
    float v = 0.f;
    if (n < N && k < K) v = colmajor ? src[(int64_t)k * ld + n] : src[(int64_t)n * ld + k];
    out[(int64_t)n * Kp + k] = __float2bfloat16_rn(v);
  }
}

// the same for a row-major source with 16-byte aligned rows (the 64 x 262144 activations of the two wide layers): eight
// consecutive k per thread, two 16-byte loads and one 16-byte store, so the pass runs at copy speed instead of one
// division and one 2-byte store per element
__global__ void sk_small_operand_rows_kernel(const float* __restrict__ src, int64_t ld, int N, int K, int BN, int Kp,
                                             __nv_bfloat16* __restrict__ out) {
  const int g8 = Kp >> 3;
  const int64_t total = (int64_t)BN * g8;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
    const int n = (int)(e / g8), k = (int)(e - (int64_t)n * g8) << 3;
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    if (n < N && k + 8 <= K) {
      const float4* p = reinterpret_cast<const float4*>(src + (int64_t)n * ld + k);
      const float4 a = __ldg(p), b = __ldg(p + 1);
      o = make_uint4(sk_pack(a.x, a.y), sk_pack(a.z, a.w), sk_pack(b.x, b.y), sk_pack(b.z, b.w));
    } else if (n < N && k < K) {
      float v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = k + j < K ? src[(int64_t)n * ld + k + j] : 0.f;
      o = make_uint4(sk_pack(v[0], v[1]), sk_pack(v[2], v[3]), sk_pack(v[4], v[5]), sk_pack(v[6], v[7]));
    }
    *reinterpret_cast<uint4*>(out + (int64_t)n * Kp + k) = o;
  }
}

__global__ void __launch_bounds__(SK_THREADS, 2)
skinny_gemm_kernel(const SkArgs a, const __grid_constant__ CUtensorMap mapB, const __grid_constant__ CUtensorMap mapA) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (tma::s32(smem_raw) & 1023u)) & 1023u);
  const int S = a.stages;
  const uint32_t b_bytes = (uint32_t)a.BN * 128u, st_bytes = SK_A_BYTES + b_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)S * st_bytes);
  uint64_t* full = bars;            // S: 256 producer arrivals + the TMA transaction
  uint64_t* empty = bars + S;       // S: tcgen05.commit
  uint64_t* t_full = empty + S;     // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * 128, z = blockIdx.y;
  const int kb0 = z * a.kbs, kb1 = min(a.nkb, kb0 + a.kbs);
  const int nkb = max(kb1 - kb0, 0);
  const uint32_t tmem_cols = a.BN <= 32 ? 32u : (a.BN <= 64 ? 64u : (a.BN <= 128 ? 128u : 256u));

  if (threadIdx.x == 0) {
    tma::prefetch_map(&mapB);
    if (a.a_bf16) tma::prefetch_map(&mapA);
    for (int s = 0; s < S; ++s) { tma::mbar_init(&full[s], a.a_bf16 ? 1 : 256 + 1); tma::mbar_init(&empty[s], 1); }
    tma::mbar_init(t_full, 1);
    tma::mbar_fence_init();
  }
  if (warp == 8) tc05::tmem_alloc(tmem_slot, tmem_cols);
  tc05::fence_before();
  __syncthreads();
  tc05::fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 8) {
    // ===================== producers =====================
    auto load_kb = [&](int it, float (&v)[32]) {
      const int k0 = (kb0 + it) * 64;
      if (!a.a_mcontig) {
        // rows are k-contiguous: 16 lanes x float4 cover one 64-wide row slice; 8 passes of 16 rows
        const int half = lane >> 4, k4 = lane & 15;
        const int k = k0 + 4 * k4;
        const uint32_t kok = k < a.K ? 1u : 0u;   // K % 4 == 0 is required for this mode
        const float* p = a.A + (int64_t)(m0 + warp * 2 + half) * a.lda + k;
#pragma unroll
        for (int ps = 0; ps < 8; ++ps) {
          const int r = ps * 16 + warp * 2 + half;
          const float4 f = sk_ldg4(p + (int64_t)ps * 16 * a.lda, kok & (m0 + r < a.M ? 1u : 0u));
          v[4 * ps] = f.x; v[4 * ps + 1] = f.y; v[4 * ps + 2] = f.z; v[4 * ps + 3] = f.w;
        }
      } else {
        // rows are m-contiguous: lanes run along m (coalesced), this warp owns k-pairs 4*warp .. 4*warp+3
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int k = k0 + 2 * (warp * 4 + i);
          const uint32_t ok0 = k < a.K ? 1u : 0u, ok1 = k + 1 < a.K ? 1u : 0u;
          const float* p = a.A + (int64_t)k * a.lda + m0 + lane;
#pragma unroll
          for (int mi = 0; mi < 4; ++mi) {
            const uint32_t mok = m0 + lane + 32 * mi < a.M ? 1u : 0u;
            v[8 * i + 2 * mi] = sk_ldg(p + 32 * mi, ok0 & mok);
            v[8 * i + 2 * mi + 1] = sk_ldg(p + a.lda + 32 * mi, ok1 & mok);
          }
        }
      }
    };
    auto store_kb = [&](int it, const float (&v)[32]) {
      const int s = it % S;
      tma::mbar_wait(&empty[s], ((uint32_t)(it / S) & 1u) ^ 1u);
      unsigned char* tile = smem + (size_t)s * st_bytes;
      if (threadIdx.x == 0) {
        tma::mbar_expect_tx(&full[s], b_bytes);
        tma::load_2d(tile + SK_A_BYTES, &mapB, &full[s], (kb0 + it) * 64, 0);
      }
      if (!a.a_mcontig) {
        const uint32_t half = (uint32_t)lane >> 4, k4 = (uint32_t)lane & 15u;
#pragma unroll
        for (int ps = 0; ps < 8; ++ps) {
          const uint32_t r = (uint32_t)(ps * 16 + warp * 2) + half;
          const uint2 o = make_uint2(sk_pack(v[4 * ps], v[4 * ps + 1]), sk_pack(v[4 * ps + 2], v[4 * ps + 3]));
          *reinterpret_cast<uint2*>(tile + r * 128u + ((((k4 >> 1) ^ (r & 7u)) << 4) | ((k4 & 1u) << 3))) = o;
        }
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint32_t kp = (uint32_t)(warp * 4 + i);
#pragma unroll
          for (int mi = 0; mi < 4; ++mi) {
            const uint32_t r = (uint32_t)(lane + 32 * mi);
            *reinterpret_cast<uint32_t*>(tile + r * 128u + ((((kp >> 2) ^ (r & 7u)) << 4) | ((kp & 3u) << 2))) =
                sk_pack(v[8 * i + 2 * mi], v[8 * i + 2 * mi + 1]);
          }
        }
      }
      tma::fence_proxy_async();
      tma::mbar_arrive(&full[s]);
    };
    if (a.a_bf16) {
      // shadow mode: warp 0 alone feeds the ring by TMA (uniform control flow, elected lane), the other warps go straight
      // to the epilogue wait
      if (warp == 0) {
        for (int it = 0; it < nkb; ++it) {
          const int s = it % S;
          tma::mbar_wait_backoff(&empty[s], ((uint32_t)(it / S) & 1u) ^ 1u, 64);
          unsigned char* tile = smem + (size_t)s * st_bytes;
          const int k0 = (kb0 + it) * 64;
          if (tc05::elect_one()) {
            tma::mbar_expect_tx(&full[s], st_bytes);
            if (!a.a_mcontig) {
              tma::load_2d(tile, &mapA, &full[s], k0, m0);                    // [128 m rows][64 k]: K-major
            } else {
              tma::load_2d(tile, &mapA, &full[s], m0, k0);                    // [64 k rows][64 m] x 2: MN-major
              tma::load_2d(tile + SK_A_BYTES / 2, &mapA, &full[s], m0 + 64, k0);
            }
            tma::load_2d(tile + SK_A_BYTES, &mapB, &full[s], k0, 0);
          }
          __syncwarp();
        }
      }
    } else {
      float buf0[32], buf1[32];
      if (nkb > 0) load_kb(0, buf0);
      for (int it = 0; it < nkb; it += 2) {
        if (it + 1 < nkb) load_kb(it + 1, buf1);
        store_kb(it, buf0);
        if (it + 1 >= nkb) break;
        if (it + 2 < nkb) load_kb(it + 2, buf0);
        store_kb(it + 1, buf1);
      }
    }
    // ===================== epilogue =====================
    const int q = warp & 3, hh = warp >> 2;
    const int m = m0 + q * 32 + lane;
    const int nchunks = a.BN >> 5;
    if (nkb > 0) {
      tma::mbar_wait(t_full, 0);
      tc05::fence_after();
    }
    const float bm = (a.bias && !a.part && m < a.M) ? __ldg(a.bias + m) : 0.f;
    for (int cc = hh; cc < nchunks; cc += 2) {
      float v[32];
      if (nkb > 0) {
        tc05::ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + cc * 32, v);
        tc05::ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = 0.f;
      }
      if (m >= a.M && !a.adam) continue;
      if (a.part) {
        float* o = a.part + ((int64_t)z * a.N + cc * 32) * a.M + m;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (cc * 32 + j < a.N) o[(int64_t)j * a.M] = v[j];
        continue;
      }
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        v[j] += bm;
        if (a.relu) v[j] = fmaxf(v[j], 0.f);
      }
      if (a.adam) {
        float bc1 = a.ad_bc1, bc2 = a.ad_bc2;
        if (a.ad_step) { bc1 = a.ad_step->bias_corr1; bc2 = a.ad_step->bias_corr2; }
        const float step_size = a.ad_lr / bc1, inv_sqrt_bc2 = 1.f / sqrtf(bc2);
        // The warp's 32 (m) x 32 (n) block goes through shared memory (the operand ring is idle once the accumulator is
        // complete) so that a lane owns FOUR consecutive elements of the contiguous direction: every p / m / v access is a
        // 16-byte load or store, 512 bytes per warp instruction, and a batch of 12 loads is issued before its first store
        // (the three arrays are distinct allocations, but the compiler cannot know and would serialise load -> store).
        float* tb = reinterpret_cast<float*>(smem) + warp * (32 * 36);
        __syncwarp();
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4)
          *reinterpret_cast<float4*>(tb + lane * 36 + 4 * j4) = make_float4(v[4 * j4], v[4 * j4 + 1], v[4 * j4 + 2], v[4 * j4 + 3]);
        __syncwarp();
        const int sub = lane >> 3, l4 = (lane & 7) * 4;
        const int64_t mrow0 = (int64_t)(m0 + q * 32);
#pragma unroll
        for (int hb = 0; hb < 32; hb += 16) {
          float4 gg[4], pp[4], mm[4], vv[4];
          int64_t off[4];
          bool ok[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int r = hb + 4 * u + sub;           // the strided index of this lane in group u
            if (a.swap) {                             // lanes of the accumulator = n: row (cc / 4) * 128 + q * 32 + r, columns = m
              gg[u] = *reinterpret_cast<const float4*>(tb + r * 36 + l4);
              const int nrow = (cc >> 2) * 128 + q * 32 + r, mcol = m0 + (cc & 3) * 32 + l4;
              off[u] = (int64_t)nrow * a.ldo + mcol;
              ok[u] = nrow < a.N && mcol + 3 < a.M;
            } else if (a.transposed) {                // element (m, n) at n*ldo + m: contiguous in m; r is n, l4.. are m
              gg[u] = make_float4(tb[(l4 + 0) * 36 + r], tb[(l4 + 1) * 36 + r], tb[(l4 + 2) * 36 + r], tb[(l4 + 3) * 36 + r]);
              off[u] = (int64_t)(cc * 32 + r) * a.ldo + mrow0 + l4;
              ok[u] = cc * 32 + r < a.N && mrow0 + l4 + 3 < a.M;
            } else {                                  // element (m, n) at m*ldo + n: contiguous in n; r is m, l4.. are n
              gg[u] = *reinterpret_cast<const float4*>(tb + r * 36 + l4);
              off[u] = (mrow0 + r) * a.ldo + cc * 32 + l4;
              ok[u] = mrow0 + r < a.M && cc * 32 + l4 + 3 < a.N;
            }
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
            pp[u] = ok[u] ? *reinterpret_cast<const float4*>(a.out + off[u]) : z;
            mm[u] = ok[u] ? __ldcs(reinterpret_cast<const float4*>(a.ad_m + off[u])) : z;
            vv[u] = ok[u] ? __ldcs(reinterpret_cast<const float4*>(a.ad_v + off[u])) : z;
          }
          if (a.adam == 2) {      // (uniform)
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              sk_adam<true>(gg[u].x, pp[u].x, mm[u].x, vv[u].x, a.ad_b1, a.ad_b2, step_size, inv_sqrt_bc2, a.ad_eps);
              sk_adam<true>(gg[u].y, pp[u].y, mm[u].y, vv[u].y, a.ad_b1, a.ad_b2, step_size, inv_sqrt_bc2, a.ad_eps);
              sk_adam<true>(gg[u].z, pp[u].z, mm[u].z, vv[u].z, a.ad_b1, a.ad_b2, step_size, inv_sqrt_bc2, a.ad_eps);
              sk_adam<true>(gg[u].w, pp[u].w, mm[u].w, vv[u].w, a.ad_b1, a.ad_b2, step_size, inv_sqrt_bc2, a.ad_eps);
            }
          } else {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              sk_adam<false>(gg[u].x, pp[u].x, mm[u].x, vv[u].x, a.ad_b1, a.ad_b2, step_size, inv_sqrt_bc2, a.ad_eps);
              sk_adam<false>(gg[u].y, pp[u].y, mm[u].y, vv[u].y, a.ad_b1, a.ad_b2, step_size, inv_sqrt_bc2, a.ad_eps);
              sk_adam<false>(gg[u].z, pp[u].z, mm[u].z, vv[u].z, a.ad_b1, a.ad_b2, step_size, inv_sqrt_bc2, a.ad_eps);
              sk_adam<false>(gg[u].w, pp[u].w, mm[u].w, vv[u].w, a.ad_b1, a.ad_b2, step_size, inv_sqrt_bc2, a.ad_eps);
            }
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            if (ok[u]) {
              *reinterpret_cast<float4*>(a.out + off[u]) = pp[u];
              __stcs(reinterpret_cast<float4*>(a.ad_m + off[u]), mm[u]);
              __stcs(reinterpret_cast<float4*>(a.ad_v + off[u]), vv[u]);
              if (a.ad_shadow)
                *reinterpret_cast<uint2*>(a.ad_shadow + off[u]) = make_uint2(sk_pack(pp[u].x, pp[u].y), sk_pack(pp[u].z, pp[u].w));
            }
          }
        }
      } else if (a.transposed) {
        float* o = a.out + (int64_t)(cc * 32) * a.ldo + m;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (cc * 32 + j < a.N) o[(int64_t)j * a.ldo] = v[j];
      } else {
        float* o = a.out + (int64_t)m * a.ldo + cc * 32;
        if (cc * 32 + 32 <= a.N && (a.ldo & 3) == 0 && (reinterpret_cast<uintptr_t>(a.out) & 15) == 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            reinterpret_cast<float4*>(o)[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (cc * 32 + j < a.N) o[j] = v[j];
        }
      }
    }
  } else {
    // ===================== MMA issuer =====================
    if (tc05::elect_one()) {   // a single lane the compiler knows about: no waterfall loop around UTCHMMA / UTMALDG
      for (int it = 0; it < nkb; ++it) {
        const int s = it % S;
        tma::mbar_wait_backoff(&full[s], (uint32_t)(it / S) & 1u, 32);
        tc05::fence_after();
        const uint32_t sa = tma::s32(smem + (size_t)s * st_bytes), sb = sa + SK_A_BYTES;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const uint64_t da = (a.a_bf16 && a.a_mcontig) ? tc05::desc_mn_sw128(sa + kk * 2048, SK_A_BYTES / 2)
                                                        : tc05::desc_k_sw128(sa + kk * 32);
          if (a.swap) {      // rows of D = n (128-row halves of the small operand), columns = the 128 m of this tile
            for (int h = 0; h < (a.BN >> 7); ++h)
              tc05::mma_bf16(tmem_base + h * 128, tc05::desc_k_sw128(sb + h * (128 * 128) + kk * 32), da, a.idesc,
                             (uint32_t)((it | kk) != 0));
          } else {
            tc05::mma_bf16(tmem_base, da, tc05::desc_k_sw128(sb + kk * 32), a.idesc, (uint32_t)((it | kk) != 0));
          }
        }
        tc05::commit(&empty[s]);
      }
      if (nkb > 0) tc05::commit(t_full);
    }
  }

  tc05::fence_before();
  __syncthreads();
  if (warp == 8) {
    tc05::fence_after();
    tc05::tmem_dealloc(tmem_base, tmem_cols);
  }
}

// fold split-K partials [nsplit][N][M] in a fixed order, then bias (per m) / ReLU / layout.  A block owns 32 consecutive
// outputs; its eight warps each add every eighth partial (coalesced 128-byte rows), and the eight sub-sums are added in
// warp order -- the same order on every run.
__global__ void __launch_bounds__(256) sk_fold_kernel(const float* __restrict__ part, int nsplit, int M, int N,
                                                      const float* __restrict__ bias, int relu, int transposed,
                                                      int64_t ldo, float* __restrict__ out) {
  __shared__ float sub[8][32];
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int64_t e = (int64_t)blockIdx.x * 32 + lane;
  const int64_t MN = (int64_t)M * N;
  float acc = 0.f;
  if (e < MN) acc = strided_sum_ordered(part + e, MN, g, nsplit, 8);
  sub[g][lane] = acc;
  __syncthreads();
  if (g != 0 || e >= MN) return;
#pragma unroll
  for (int i = 1; i < 8; ++i) acc += sub[i][lane];
  const int n = (int)(e / M), m = (int)(e - (int64_t)n * M);
  if (bias) acc += bias[m];
  if (relu) acc = fmaxf(acc, 0.f);
  out[transposed ? (int64_t)n * ldo + m : (int64_t)m * ldo + n] = acc;
}

// ---- host ---------------------------------------------------------------------------------------------------------
struct SkPlan {
  bool ok;
  int M, N, K, BN, Kp, nkb, nsplit, kbs, stages;
  int a_mcontig, a_from;        // a_from: which input is the big operand (0 = in0, 1 = in1)
  int64_t lda, ldo, ld_small;
  int small_colmajor, transposed, has_bias;
};

static SkPlan sk_plan(int op, int B, int Cin, int Cout, const void* in0, const void* in1) {
  SkPlan p;
  memset(&p, 0, sizeof(p));
  if (B > 256) return p;
  switch (op) {
    case VC_OP_LINEAR_FWD:    // y[b][n] = sum_k x[b][k] W[n][k]:  A = W (k-contiguous), Bs = x, D[n][b] -> y transposed
      p.M = Cout; p.N = B; p.K = Cin; p.a_from = 1; p.lda = Cin; p.a_mcontig = 0;
      p.ld_small = Cin; p.small_colmajor = 0; p.ldo = Cout; p.transposed = 1; p.has_bias = 1;
      break;
    case VC_OP_LINEAR_DGRAD:  // dx[b][k] = sum_n dy[b][n] W[n][k]: A(m=k, kk=n) = W[kk*Cin + m] (m-contiguous)
      p.M = Cin; p.N = B; p.K = Cout; p.a_from = 1; p.lda = Cin; p.a_mcontig = 1;
      p.ld_small = Cout; p.small_colmajor = 0; p.ldo = Cin; p.transposed = 1;
      break;
    case VC_OP_LINEAR_WGRAD:  // dW[n][k] = sum_b dy[b][n] x[b][k]   (in0 = dy, in1 = x)
      if (Cout <= 256 && Cin >= Cout) {   // rows of D = k: A = x (m-contiguous), Bs[n][b] = dy[b][n], dW = D^T
        p.M = Cin; p.N = Cout; p.K = B; p.a_from = 1; p.lda = Cin; p.a_mcontig = 1;
        p.ld_small = Cout; p.small_colmajor = 1; p.ldo = Cin; p.transposed = 1;
      } else if (Cin <= 256) {            // rows of D = n: A = dy (m-contiguous), Bs[k][b] = x[b][k], dW = D
        p.M = Cout; p.N = Cin; p.K = B; p.a_from = 0; p.lda = Cout; p.a_mcontig = 1;
        p.ld_small = Cin; p.small_colmajor = 1; p.ldo = Cin; p.transposed = 0;
      } else {
        return p;
      }
      break;
    default: return p;
  }
  if (p.N > 256 || (int64_t)p.M * p.K < (1LL << 22)) return p;   // only worth it when the big matrix dominates
  const void* big = p.a_from ? in1 : in0;
  if (!p.a_mcontig && ((p.lda & 3) || (p.K & 3) || (reinterpret_cast<uintptr_t>(big) & 15))) return p;
  p.BN = (p.N + 31) / 32 * 32;
  p.Kp = (p.K + 63) / 64 * 64;
  p.nkb = p.Kp / 64;
  const int mt = (p.M + 127) / 128;
  int ns = 1;
  if (mt < sm_count()) {
    ns = (2 * sm_count() + mt - 1) / mt;
    if (ns > p.nkb / 4) ns = p.nkb / 4;
    if (ns < 1) ns = 1;
  }
  p.kbs = (p.nkb + ns - 1) / ns;
  p.nsplit = (p.nkb + p.kbs - 1) / p.kbs;
  const int st_bytes = (int)SK_A_BYTES + p.BN * 128;
  int stages = 98304 / st_bytes;
  if (stages > p.kbs) stages = p.kbs;
  if (stages < 1) stages = 1;
  p.stages = stages;
  p.ok = true;
  return p;
}

static size_t sk_align(size_t v) { return (v + 1023) & ~(size_t)1023; }

bool skinny_gemm_supported(int op, int B, int Cin, int Cout, const void* in0, const void* in1) {
  return sk_plan(op, B, Cin, Cout, in0, in1).ok;
}

// the alignment test needs the pointers; sized for the worst case so the query does not
size_t skinny_gemm_workspace(int op, int B, int Cin, int Cout) {
  const SkPlan p = sk_plan(op, B, Cin, Cout, nullptr, nullptr);
  if (!p.ok) return 0;
  return sk_align((size_t)p.BN * p.Kp * 2) + (p.nsplit > 1 ? sk_align((size_t)p.nsplit * p.M * p.N * 4) : 0) + 2048;
}

struct SkAdam { float* m; float* v; const VcAdamDeviceStep* step; float lr, b1, b2, eps, bc1, bc2; __nv_bfloat16* shadow; };
// fc_adam.cu: the persistent TMA-staged version of the fused weight gradient + Adam (N = 256, batch <= 64, M % 128 == 0)
struct FcAdamParams { float* m; float* v; const VcAdamDeviceStep* step; float lr, b1, b2, eps, bc1, bc2; __nv_bfloat16* shadow; };
bool fc_adam_tma_supported(int M, int N, int K, int64_t lda, int64_t ldo, int transposed, const void* p, const void* m,
                           const void* v, const void* shadow);
int fc_adam_tma_run(const float* act, int64_t lda, const __nv_bfloat16* Bs, float* p, int M, int N, int K, int64_t ldo,
                    int transposed, const FcAdamParams& ad, cudaStream_t s);
static int skinny_gemm_run_impl(int op, const float* in0, const float* in1, const float* bias, float* out, int B, int Cin,
                                int Cout, int relu, void* ws, size_t wsb, cudaStream_t s, const SkAdam* adam,
                                const __nv_bfloat16* w_shadow);

int skinny_gemm_run(int op, const float* in0, const float* in1, const float* bias, float* out, int B, int Cin, int Cout,
                    int relu, void* ws, size_t wsb, cudaStream_t s) {
  return skinny_gemm_run_impl(op, in0, in1, bias, out, B, Cin, Cout, relu, ws, wsb, s, nullptr, nullptr);
}

// w_shadow != null (forward / data gradient only): the weight operand is read from its bf16 copy [Cout][Cin] by TMA
static int skinny_gemm_run_impl(int op, const float* in0, const float* in1, const float* bias, float* out, int B, int Cin,
                                int Cout, int relu, void* ws, size_t wsb, cudaStream_t s, const SkAdam* adam,
                                const __nv_bfloat16* w_shadow) {
  const SkPlan p = sk_plan(op, B, Cin, Cout, w_shadow ? nullptr : in0, w_shadow ? nullptr : in1);
  if (w_shadow) VC_REQUIRE(p.ok && op != VC_OP_LINEAR_WGRAD && p.a_from == 1 && Cin % 8 == 0 &&
                               (reinterpret_cast<uintptr_t>(w_shadow) & 15) == 0,
                           VC_ERR_UNSUPPORTED, "skinny_gemm: the bf16-shadow path serves forward / data gradient of aligned weights");
  if (adam) VC_REQUIRE(p.ok && p.nsplit == 1 && op == VC_OP_LINEAR_WGRAD && p.M % 4 == 0 && p.N % 4 == 0 && p.ldo % 4 == 0 &&
                           ((reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(adam->m) |
                             reinterpret_cast<uintptr_t>(adam->v)) & 15) == 0,
                       VC_ERR_UNSUPPORTED, "skinny_gemm: the fused Adam epilogue needs an unsplit weight-gradient shape, "
                       "dimensions that are multiples of 4 and 16-byte aligned tensors");
  VC_REQUIRE(p.ok, VC_ERR_UNSUPPORTED, "skinny_gemm: unsupported shape");
  VC_REQUIRE(ws && wsb >= skinny_gemm_workspace(op, B, Cin, Cout), VC_ERR_WORKSPACE, "skinny_gemm: workspace too small");
  unsigned char* w = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(ws) + 1023) & ~(uintptr_t)1023);
  __nv_bfloat16* Bs = reinterpret_cast<__nv_bfloat16*>(w);
  float* part = p.nsplit > 1 ? reinterpret_cast<float*>(w + sk_align((size_t)p.BN * p.Kp * 2)) : nullptr;
  const float* big = p.a_from ? in1 : in0;
  const float* small = p.a_from ? in0 : in1;
  {
    const int64_t total = (int64_t)p.BN * p.Kp;
    int64_t blocks = (total + 255) / 256;
    if (blocks > (int64_t)sm_count() * 16) blocks = (int64_t)sm_count() * 16;
    if (!p.small_colmajor && (p.ld_small & 3) == 0 && (reinterpret_cast<uintptr_t>(small) & 15) == 0) {
      blocks = (total / 8 + 255) / 256;
      if (blocks > (int64_t)sm_count() * 16) blocks = (int64_t)sm_count() * 16;
      sk_small_operand_rows_kernel<<<(int)blocks, 256, 0, s>>>(small, p.ld_small, p.N, p.K, p.BN, p.Kp, Bs);
    } else {
      sk_small_operand_kernel<<<(int)blocks, 256, 0, s>>>(small, p.ld_small, p.small_colmajor, p.N, p.K, p.BN, p.Kp, Bs);
    }
    VC_LAUNCH_CHECK();
  }
  if (adam && p.a_mcontig && p.nkb == 1 && p.BN == 256 && p.Kp == 64 &&
      fc_adam_tma_supported(p.M, p.N, p.K, p.lda, p.ldo, p.transposed, out, adam->m, adam->v, adam->shadow)) {
    const FcAdamParams fp = {adam->m, adam->v, adam->step, adam->lr, adam->b1, adam->b2, adam->eps, adam->bc1, adam->bc2,
                             adam->shadow};
    return fc_adam_tma_run(big, p.lda, Bs, out, p.M, p.N, p.K, p.ldo, p.transposed, fp, s);
  }
  CUtensorMap mb, ma;
  memset(&ma, 0, sizeof(ma));
  if (w_shadow) {   // bf16 [Cout][Cin]: K-major boxes [128 m][64 k] (forward) or MN-major boxes [64 k][64 m] (data gradient)
    const uint64_t dims[2] = {(uint64_t)Cin, (uint64_t)Cout};
    const uint64_t str[1] = {(uint64_t)Cin * 2};
    const uint32_t box[2] = {64u, p.a_mcontig ? 64u : 128u};
    int rc = tma::make_map(&ma, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<__nv_bfloat16*>(w_shadow), dims, str, box,
                           CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc != VC_OK) return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)p.Kp, (uint64_t)p.BN};
    const uint64_t str[1] = {(uint64_t)p.Kp * 2};
    const uint32_t box[2] = {64u, (uint32_t)p.BN};
    int rc = tma::make_map(&mb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, Bs, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc != VC_OK) return rc;
  }
  SkArgs a;
  memset(&a, 0, sizeof(a));
  a.A = big; a.lda = p.lda; a.a_mcontig = p.a_mcontig;
  a.M = p.M; a.N = p.N; a.K = p.K; a.BN = p.BN;
  a.out = out; a.ldo = p.ldo; a.transposed = p.transposed;
  a.bias = p.has_bias ? bias : nullptr; a.relu = relu;
  a.part = part; a.nkb = p.nkb; a.kbs = p.kbs; a.stages = p.stages;
  a.a_bf16 = w_shadow ? 1 : 0;
  a.idesc = tc05::idesc(tc05::kFmtBF16, 128, p.BN, (w_shadow && p.a_mcontig) ? 1 : 0, 0);
  size_t ring = (size_t)p.stages * (SK_A_BYTES + (size_t)p.BN * 128);
  if (adam && p.transposed && p.BN % 128 == 0 && !getenv("VAECLUST_SK_NOSWAP")) {
    a.swap = 1;
    a.idesc = tc05::idesc(tc05::kFmtBF16, 128, 128, 0, 0);
  }
  if (adam) {
    static const bool ieee = getenv("VAECLUST_FC_ADAM_IEEE") && atoi(getenv("VAECLUST_FC_ADAM_IEEE")) != 0;
    a.adam = ieee ? 1 : 2; a.ad_m = adam->m; a.ad_v = adam->v; a.ad_step = adam->step; a.ad_shadow = adam->shadow;
    a.ad_lr = adam->lr; a.ad_b1 = adam->b1; a.ad_b2 = adam->b2; a.ad_eps = adam->eps; a.ad_bc1 = adam->bc1; a.ad_bc2 = adam->bc2;
    // the non-transposed epilogue transposes 8 x (32 x 33) floats through the operand ring; the barriers sit behind it
    VC_REQUIRE(ring >= 8 * 32 * 36 * 4, VC_ERR_UNSUPPORTED, "skinny_gemm: operand ring too small for the Adam epilogue");
  }
  const size_t smem = ring + (2 * p.stages + 1) * 8 + 16 + 1024;
  VC_CUDA(cudaFuncSetAttribute(skinny_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((unsigned)((p.M + 127) / 128), (unsigned)p.nsplit);
  skinny_gemm_kernel<<<grid, SK_THREADS, smem, s>>>(a, mb, ma);
  VC_LAUNCH_CHECK();
  if (part) {
    const int64_t MN = (int64_t)p.M * p.N;
    sk_fold_kernel<<<(int)((MN + 31) / 32), 256, 0, s>>>(part, p.nsplit, p.M, p.N, a.bias, relu, p.transposed, p.ldo, out);
    VC_LAUNCH_CHECK();
  }
  return VC_OK;
}

}  // namespace vc

using namespace vc;

extern "C" int vc_linear_wgrad_adam_supported(int B, int Cin, int Cout) {
  const SkPlan p = sk_plan(VC_OP_LINEAR_WGRAD, B, Cin, Cout, nullptr, nullptr);
  return (p.ok && p.nsplit == 1 && p.M % 4 == 0 && p.N % 4 == 0 && (size_t)p.stages * (SK_A_BYTES + (size_t)p.BN * 128) >= 8 * 32 * 36 * 4) ? 1 : 0;
}

extern "C" size_t vc_linear_wgrad_adam_workspace_bytes(int B, int Cin, int Cout) {
  return skinny_gemm_workspace(VC_OP_LINEAR_WGRAD, B, Cin, Cout);
}

extern "C" int vc_linear_wgrad_adam(const float* dy, const float* x, float* W, float* exp_avg, float* exp_avg_sq,
                                    void* w_shadow, int B, int Cin, int Cout, float lr, float beta1, float beta2, float eps,
                                    float bias_corr1, float bias_corr2, const VcAdamDeviceStep* step_state,
                                    void* workspace, size_t workspace_bytes, void* stream) {
  VC_REQUIRE(dy && x && W && exp_avg && exp_avg_sq, VC_ERR_BAD_ARG, "vc_linear_wgrad_adam: null pointer");
  VC_REQUIRE(vc_linear_wgrad_adam_supported(B, Cin, Cout), VC_ERR_UNSUPPORTED,
             "vc_linear_wgrad_adam: unsupported shape (B %d, Cin %d, Cout %d)", B, Cin, Cout);
  VC_REQUIRE((reinterpret_cast<uintptr_t>(w_shadow) & 7) == 0, VC_ERR_BAD_ARG, "vc_linear_wgrad_adam: misaligned shadow");
  SkAdam ad = {exp_avg, exp_avg_sq, step_state, lr, beta1, beta2, eps, bias_corr1, bias_corr2,
               static_cast<__nv_bfloat16*>(w_shadow)};
  return skinny_gemm_run_impl(VC_OP_LINEAR_WGRAD, dy, x, nullptr, W, B, Cin, Cout, 0, workspace, workspace_bytes,
                              as_stream(stream), &ad, nullptr);
}

extern "C" int vc_linear_bf16w_supported(int op, int B, int Cin, int Cout) {
  if (op != VC_OP_LINEAR_FWD && op != VC_OP_LINEAR_DGRAD) return 0;
  const SkPlan p = sk_plan(op, B, Cin, Cout, nullptr, nullptr);
  return (p.ok && p.a_from == 1 && Cin % 8 == 0) ? 1 : 0;
}

extern "C" size_t vc_linear_bf16w_workspace_bytes(int op, int B, int Cin, int Cout) {
  return skinny_gemm_workspace(op, B, Cin, Cout);
}

extern "C" int vc_linear_bf16w(int op, const float* in, const void* w_shadow, const float* bias, float* out, int B, int Cin,
                               int Cout, int relu, void* workspace, size_t workspace_bytes, void* stream) {
  VC_REQUIRE(in && w_shadow && out, VC_ERR_BAD_ARG, "vc_linear_bf16w: null pointer");
  VC_REQUIRE(vc_linear_bf16w_supported(op, B, Cin, Cout), VC_ERR_UNSUPPORTED,
             "vc_linear_bf16w: unsupported op / shape (op %d, B %d, Cin %d, Cout %d)", op, B, Cin, Cout);
  // in1 of the fp32 path is the weight; only its bf16 shadow is read here
  return skinny_gemm_run_impl(op, in, nullptr, bias, out, B, Cin, Cout, relu, workspace, workspace_bytes, as_stream(stream),
                              nullptr, static_cast<const __nv_bfloat16*>(w_shadow));
}
