// VAE loss kernels: reparameterisation + KL rows, crop-aware reconstruction term with its gradient in the same
// pass, and the three-scalar finalize.  HBM-bound: one read of xhat and x, one write of dxhat.
// Reference semantics: scripts/12:168-172,189-190,260-264; 19:105-108,124-126,226-228; 06:164-168,182-188.
#include "common.cuh"

namespace vc {

__global__ void reparam_kl_fwd_kernel(const float* __restrict__ mu, const float* __restrict__ lv,
                                      const float* __restrict__ eps, int B, int z, float* __restrict__ z_out,
                                      float* __restrict__ kl_rows) {
  // one warp per row
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= B) return;
  float acc = 0.f;
  for (int j = lane; j < z; j += 32) {
    const float m = mu[warp * z + j], l = lv[warp * z + j];
    const float e = expf(l);
    if (z_out) z_out[warp * z + j] = m + eps[warp * z + j] * expf(0.5f * l);
    acc += 1.f + l - m * m - e;
  }
  acc = warp_sum(acc);
  if (lane == 0 && kl_rows) kl_rows[warp] = -0.5f * acc;
}

__global__ void reparam_kl_bwd_kernel(const float* __restrict__ dz, const float* __restrict__ dkl,
                                      const float* __restrict__ mu, const float* __restrict__ lv,
                                      const float* __restrict__ eps, int B, int z, float* __restrict__ dmu,
                                      float* __restrict__ dlv) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * z) return;
  const int r = i / z;
  const float g = dz ? dz[i] : 0.f;
  const float gk = dkl ? dkl[r] : 0.f;
  const float l = lv[i];
  const float e = eps ? eps[i] : 0.f;
  dmu[i] = g + gk * mu[i];
  dlv[i] = g * e * 0.5f * expf(0.5f * l) + gk * 0.5f * (expf(l) - 1.f);
}

constexpr int kReconThreads = 256;
constexpr int kReconMaxBlocks = 148 * 8;

// Fast path: no crop, n % 4 == 0, 16-byte aligned.
__global__ void __launch_bounds__(kReconThreads)
recon_vec_kernel(const float4* __restrict__ xhat, const float4* __restrict__ x, int64_t n4, float gs,
                 float4* __restrict__ dxhat, double* __restrict__ partials) {
  __shared__ double scratch[32];
  float acc0 = 0.f, acc1 = 0.f;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  // two independent 16-byte loads per operand in flight
  for (; i + stride < n4; i += 2 * stride) {
    const float4 a0 = __ldcs(xhat + i), b0 = __ldcs(x + i);
    const float4 a1 = __ldcs(xhat + i + stride), b1 = __ldcs(x + i + stride);
    float4 d0 = make_float4(a0.x - b0.x, a0.y - b0.y, a0.z - b0.z, a0.w - b0.w);
    float4 d1 = make_float4(a1.x - b1.x, a1.y - b1.y, a1.z - b1.z, a1.w - b1.w);
    acc0 += d0.x * d0.x + d0.y * d0.y + d0.z * d0.z + d0.w * d0.w;
    acc1 += d1.x * d1.x + d1.y * d1.y + d1.z * d1.z + d1.w * d1.w;
    if (dxhat) {
      __stcs(dxhat + i, make_float4(gs * d0.x, gs * d0.y, gs * d0.z, gs * d0.w));
      __stcs(dxhat + i + stride, make_float4(gs * d1.x, gs * d1.y, gs * d1.z, gs * d1.w));
    }
  }
  for (; i < n4; i += stride) {
    const float4 a0 = __ldcs(xhat + i), b0 = __ldcs(x + i);
    float4 d0 = make_float4(a0.x - b0.x, a0.y - b0.y, a0.z - b0.z, a0.w - b0.w);
    acc0 += d0.x * d0.x + d0.y * d0.y + d0.z * d0.z + d0.w * d0.w;
    if (dxhat) __stcs(dxhat + i, make_float4(gs * d0.x, gs * d0.y, gs * d0.z, gs * d0.w));
  }
  double t = block_sum<double>((double)acc0 + (double)acc1, scratch);
  if (threadIdx.x == 0) partials[blockIdx.x] = t;
}

// General path: xhat [outer,Hh,Wh], x [outer,H,W] cropped top-left.
__global__ void __launch_bounds__(kReconThreads)
recon_crop_kernel(const float* __restrict__ xhat, const float* __restrict__ x, int64_t outer, int Hh, int Wh,
                  int H, int W, float gs, float* __restrict__ dxhat, double* __restrict__ partials) {
  __shared__ double scratch[32];
  const int64_t n = outer * Hh * Wh;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  float acc = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int w = (int)(i % Wh);
    const int64_t r = i / Wh;
    const int h = (int)(r % Hh);
    const int64_t o = r / Hh;
    float g = 0.f;
    if (w < W && h < H) {
      const float dlt = xhat[i] - x[(o * H + h) * W + w];
      acc += dlt * dlt;
      g = gs * dlt;
    }
    if (dxhat) dxhat[i] = g;
  }
  double t = block_sum<double>((double)acc, scratch);
  if (threadIdx.x == 0) partials[blockIdx.x] = t;
}

__global__ void sum_partials_kernel(const double* __restrict__ partials, int n, double* __restrict__ out) {
  __shared__ double scratch[32];
  double acc = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) acc += partials[i];
  acc = block_sum<double>(acc, scratch);
  if (threadIdx.x == 0) *out = acc;
}

__global__ void loss_finalize_kernel(const double* __restrict__ sq_sum, const float* __restrict__ kl_rows, int B,
                                     int z, int64_t n_recon, float beta, int mode, float* __restrict__ out3) {
  __shared__ double scratch[32];
  double acc = 0.0;
  for (int i = threadIdx.x; i < B; i += blockDim.x) acc += (double)kl_rows[i];
  acc = block_sum<double>(acc, scratch);
  if (threadIdx.x == 0) {
    double recon, kl;
    if (mode == VC_LOSS_MEAN_MEAN) {
      recon = *sq_sum / (double)n_recon;
      kl = acc / ((double)B * (double)z);
    } else {
      recon = *sq_sum / (double)B;
      kl = acc / (double)B;
    }
    out3[0] = (float)(recon + (double)beta * kl);
    out3[1] = (float)recon;
    out3[2] = (float)kl;
  }
}

}  // namespace vc

using namespace vc;

extern "C" int vc_reparam_kl_fwd(const float* mu, const float* lv, const float* eps, int B, int z, float* z_out,
                                 float* kl_rows, void* stream) {
  VC_REQUIRE(mu && lv && B > 0 && z > 0, VC_ERR_BAD_ARG, "vc_reparam_kl_fwd: bad arguments (B=%d z=%d)", B, z);
  VC_REQUIRE(!z_out || eps, VC_ERR_BAD_ARG, "vc_reparam_kl_fwd: z_out requested without eps");
  const int threads = 256, rows_per_block = threads / 32;
  reparam_kl_fwd_kernel<<<(B + rows_per_block - 1) / rows_per_block, threads, 0, as_stream(stream)>>>(
      mu, lv, eps, B, z, z_out, kl_rows);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" int vc_reparam_kl_bwd(const float* dz, const float* dkl_rows, const float* mu, const float* lv,
                                 const float* eps, int B, int z, float* dmu, float* dlv, void* stream) {
  VC_REQUIRE(mu && lv && dmu && dlv && B > 0 && z > 0, VC_ERR_BAD_ARG, "vc_reparam_kl_bwd: bad arguments");
  VC_REQUIRE(!dz || eps, VC_ERR_BAD_ARG, "vc_reparam_kl_bwd: dz given without eps");
  const int n = B * z;
  reparam_kl_bwd_kernel<<<(n + 255) / 256, 256, 0, as_stream(stream)>>>(dz, dkl_rows, mu, lv, eps, B, z, dmu, dlv);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" size_t vc_recon_loss_workspace_bytes(void) { return sizeof(double) * kReconMaxBlocks; }

extern "C" int vc_recon_loss_fwd_bwd(const float* xhat, const float* x, int64_t outer, int Hh, int Wh, int H,
                                     int W, float grad_scale, float* dxhat, double* sq_sum, void* workspace,
                                     size_t workspace_bytes, void* stream) {
  VC_REQUIRE(xhat && x && sq_sum && outer > 0 && H > 0 && W > 0 && H <= Hh && W <= Wh, VC_ERR_BAD_ARG,
             "vc_recon_loss_fwd_bwd: bad shape outer=%lld xhat=%dx%d x=%dx%d", (long long)outer, Hh, Wh, H, W);
  VC_REQUIRE(workspace && workspace_bytes >= vc_recon_loss_workspace_bytes(), VC_ERR_WORKSPACE,
             "vc_recon_loss_fwd_bwd: workspace too small");
  double* partials = static_cast<double*>(workspace);
  const float gs = 2.f * grad_scale;
  const int64_t n = outer * Hh * Wh;
  const bool vec = (H == Hh && W == Wh && n % 4 == 0 && (reinterpret_cast<uintptr_t>(xhat) & 15) == 0 &&
                    (reinterpret_cast<uintptr_t>(x) & 15) == 0 &&
                    (!dxhat || (reinterpret_cast<uintptr_t>(dxhat) & 15) == 0));
  int blocks;
  if (vec) {
    const int64_t n4 = n / 4;
    blocks = (int)((n4 + kReconThreads * 2 - 1) / (kReconThreads * 2));
    if (blocks > kReconMaxBlocks) blocks = kReconMaxBlocks;
    if (blocks < 1) blocks = 1;
    recon_vec_kernel<<<blocks, kReconThreads, 0, as_stream(stream)>>>(
        reinterpret_cast<const float4*>(xhat), reinterpret_cast<const float4*>(x), n4, gs,
        reinterpret_cast<float4*>(dxhat), partials);
  } else {
    blocks = (int)((n + kReconThreads * 4 - 1) / (kReconThreads * 4));
    if (blocks > kReconMaxBlocks) blocks = kReconMaxBlocks;
    if (blocks < 1) blocks = 1;
    recon_crop_kernel<<<blocks, kReconThreads, 0, as_stream(stream)>>>(xhat, x, outer, Hh, Wh, H, W, gs, dxhat,
                                                                      partials);
  }
  VC_LAUNCH_CHECK();
  sum_partials_kernel<<<1, 256, 0, as_stream(stream)>>>(partials, blocks, sq_sum);
  VC_LAUNCH_CHECK();
  return VC_OK;
}

extern "C" int vc_loss_finalize(const double* sq_sum, const float* kl_rows, int B, int z, int64_t n_recon_elems,
                                float beta, int mode, float* out3, void* stream) {
  VC_REQUIRE(sq_sum && kl_rows && out3 && B > 0 && z > 0 && n_recon_elems > 0, VC_ERR_BAD_ARG,
             "vc_loss_finalize: bad arguments");
  VC_REQUIRE(mode == VC_LOSS_MEAN_MEAN || mode == VC_LOSS_SUM_MEAN, VC_ERR_BAD_ARG, "vc_loss_finalize: bad mode %d",
             mode);
  loss_finalize_kernel<<<1, 256, 0, as_stream(stream)>>>(sq_sum, kl_rows, B, z, n_recon_elems, beta, mode, out3);
  VC_LAUNCH_CHECK();
  return VC_OK;
}
