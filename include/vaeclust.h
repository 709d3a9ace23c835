/* libvaeclust.so -- C-ABI of the B200 (sm_100a) kernels behind the VAE-training and latent-clustering hot path
 * of NawrozHaseen/VAE-for-Hybrid-Music-Clustering.
 *
 * The reference ships no native code and no FFI; every entry point below therefore cites the
 * reference *call site* (file:line under /root/reference/scripts, or sklearn:<path> for the un-vendored
 * scikit-learn arithmetic those scripts reach) that it replaces.  INTEGRATION.md shows the ctypes stubs a
 * maintainer of the reference would add.
 *
 * Conventions
 *   - every function returns VC_OK (0) or a negative VC_ERR_*; vc_last_error_string() describes the last failure
 *     of the calling thread;
 *   - all data pointers are DEVICE pointers owned by the caller unless the comment says "host"; nothing is
 *     allocated, freed or synchronised inside the library; `stream` is a cudaStream_t passed as void*;
 *   - dense tensors are contiguous, row-major, fp32 unless stated; activations at the boundary are NCHW and
 *     weights use torch's state_dict layouts (Linear [out,in], Conv2d [Cout,Cin,kh,kw], ConvTranspose2d
 *     [Cin,Cout,kh,kw]) so checkpoints interchange with the reference;
 *   - workspace sizes come from the matching vc_*_workspace_bytes().
 */
#ifndef VAECLUST_H_
#define VAECLUST_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VC_OK 0
#define VC_ERR_BAD_ARG (-1)
#define VC_ERR_UNSUPPORTED (-2)
#define VC_ERR_CUDA (-3)
#define VC_ERR_WORKSPACE (-4)

/* loss reduction conventions */
#define VC_LOSS_MEAN_MEAN 0 /* scripts 06/12: mean over all elements for both terms (12:262-264, 06:182-188) */
#define VC_LOSS_SUM_MEAN 1  /* script 19: feature-sum then batch-mean (19:226-228, 19:124-126)            */

const char* vc_last_error_string(void);
int vc_version(void);
int vc_sm_count(void);

/* ------------------------------------------------------------------------------------------------ VAE loss
 * reparameterisation + KL (12:168-172,189-190; 19:105-108,124-126; 06:164-168,186):
 *   z = mu + eps*exp(0.5*lv);  kl_rows[i] = -0.5 * sum_j (1 + lv - mu^2 - exp(lv))
 * eps is supplied by the caller (torch.randn on the host side keeps the reference's RNG stream). */
int vc_reparam_kl_fwd(const float* mu, const float* lv, const float* eps, int B, int z, float* z_out,
                      float* kl_rows, void* stream);
/* dmu = dz + dkl_rows[i]*mu ; dlv = dz*eps*0.5*exp(0.5 lv) + dkl_rows[i]*0.5*(exp(lv)-1).
 * dz or dkl_rows may be NULL (treated as zero). */
int vc_reparam_kl_bwd(const float* dz, const float* dkl_rows, const float* mu, const float* lv, const float* eps,
                      int B, int z, float* dmu, float* dlv, void* stream);

/* Reconstruction term with the crop of 12:260 folded in: xhat is [outer, Hh, Wh], x is [outer, H, W] with
 * H<=Hh, W<=Wh; one pass computes sq_sum = sum (xhat-x)^2 over the cropped window (double, deterministic
 * two-stage reduction) AND dxhat = grad_scale * 2 (xhat-x) (zero outside the window).  For 2-d inputs
 * (scripts 06/19) pass outer=B, Hh=H=1, Wh=W=features.  dxhat may be NULL (forward only). */
size_t vc_recon_loss_workspace_bytes(void);
int vc_recon_loss_fwd_bwd(const float* xhat, const float* x, int64_t outer, int Hh, int Wh, int H, int W,
                          float grad_scale, float* dxhat, double* sq_sum, void* workspace, size_t workspace_bytes,
                          void* stream);
/* out3 = {loss, recon, kl} (device floats) from sq_sum and kl_rows under `mode` (VC_LOSS_*). */
int vc_loss_finalize(const double* sq_sum, const float* kl_rows, int B, int z, int64_t n_recon_elems, float beta,
                     int mode, float* out3, void* stream);

/* ------------------------------------------------------------------------------------------ dense layers
 * nn.Linear (+ReLU) of 12:98-103,114-117,129-132,159-164; 19:79-95; 06:149-157.
 * y[M,N] = act(x[M,K] W[N,K]^T + b[N]).  Skinny shapes with a huge reduction (the 262144-wide FC layers of script 12
 * at batch 64) are split over the reduction into the workspace and reduced in a fixed order. */
size_t vc_linear_workspace_bytes(int M, int N, int K);
int vc_linear_fwd(const float* x, const float* W, const float* b, float* y, int M, int N, int K, int relu,
                  void* workspace, size_t workspace_bytes, void* stream);
int vc_linear_dgrad(const float* dy, const float* W, float* dx, int M, int N, int K, void* workspace,
                    size_t workspace_bytes, void* stream);
int vc_linear_wgrad(const float* dy, const float* x, float* dW, int M, int N, int K, int accumulate,
                    void* stream);
/* db[c] (+)= sum over outer,inner of dy[outer, C, inner]  (bias gradient of Linear: inner=1; Conv: inner=H*W) */
size_t vc_bias_grad_workspace_bytes(int C);
int vc_bias_grad(const float* dy, int64_t outer, int C, int64_t inner, float* db, int accumulate, void* workspace,
                 size_t workspace_bytes, void* stream);
/* vc_bias_grad with NO workspace argument and no allocation of any kind (no counterpart in the reference: added
 * for callers that cannot call cudaMalloc -- CUDA graph capture, serving loops).  The
 * partial sums live in a 2 MiB file-scope device array that every GPU loading the library holds until exit; C is limited
 * to vc_bias_grad_noalloc_max_channels() (4096), larger C returns VC_ERR_UNSUPPORTED.  Results are bit-identical to
 * vc_bias_grad.  All calls on a device share the array: calls that overlap on the GPU (different streams without a
 * dependency) race; calls ordered on one stream are safe and nothing needs clearing between them. */
int vc_bias_grad_noalloc_max_channels(void);
int vc_bias_grad_noalloc(const float* dy, int64_t outer, int C, int64_t inner, float* db, int accumulate, void* stream);
/* dx = dy * (y > 0) */
int vc_relu_bwd(const float* dy, const float* y, float* dx, int64_t n, void* stream);
/* Both of the above in one pass over dy/y ([outer, C, inner] layout): g = dy * (y > 0) and db[c] = sum g, the two
 * things autograd computes for `ReLU(conv(x) + b)` (12:87-89,135-137).  Workspace as vc_bias_grad. */
int vc_relu_bwd_bias(const float* dy, const float* y, int64_t outer, int C, int64_t inner, float* g, float* db,
                     void* workspace, size_t workspace_bytes, void* stream);
/* out[b,:] = [a[b,0:na], l[b,0:nl]*gate[b], gate[b] if append_gate]  -- `l*m` and torch.cat of 12:178-180 and
 * the cat([x,c]) / cat([z,c]) of 19:97-115 (gate=NULL, append_gate=0). */
int vc_cat_gate_fwd(const float* a, int na, const float* l, int nl, const float* gate, int append_gate, int B,
                    float* out, void* stream);
int vc_cat_gate_bwd(const float* dout, int na, int nl, const float* gate, int append_gate, int B, float* da,
                    float* dl, void* stream);

/* nn.Conv2d(Cin,Cout,3,stride=2,padding=1) (+ReLU) of 12:87-89.  x [B,Cin,H,W] -> y [B,Cout,OH,OW],
 * OH=(H-1)/2+1, OW=(W-1)/2+1. */
size_t vc_conv_wgrad_workspace_bytes(int Cout, int Cin, int ksize);
int vc_conv2d_s2k3_fwd(const float* x, const float* W, const float* b, float* y, int B, int Cin, int H, int Wd,
                       int Cout, int relu, void* stream);
int vc_conv2d_s2k3_dgrad(const float* dy, const float* W, float* dx, int B, int Cin, int H, int Wd, int Cout,
                         void* stream);
int vc_conv2d_s2k3_wgrad(const float* dy, const float* x, float* dW, int B, int Cin, int H, int Wd, int Cout,
                         void* workspace, size_t workspace_bytes, void* stream);
/* nn.ConvTranspose2d(Cin,Cout,4,stride=2,padding=1) (+ReLU) of 12:135-137.  x [B,Cin,H,W] -> y [B,Cout,2H,2W]. */
int vc_convT2d_s2k4_fwd(const float* x, const float* W, const float* b, float* y, int B, int Cin, int H, int Wd,
                        int Cout, int relu, void* stream);
int vc_convT2d_s2k4_dgrad(const float* dy, const float* W, float* dx, int B, int Cin, int H, int Wd, int Cout,
                          void* stream);
int vc_convT2d_s2k4_wgrad(const float* dy, const float* x, float* dW, int B, int Cin, int H, int Wd, int Cout,
                          void* workspace, size_t workspace_bytes, void* stream);

/* Weight AND bias gradient of nn.Conv2d(1, Cout, 3, 2, 1) (12:87) in one pass over dy [B,Cout,H/2,W/2] (fp32 NCHW) and
 * x [B,1,H,W]: dW[co][kh][kw] = sum dy * x(tap), db[co] = sum dy.  round_bf16 = 1 rounds the multiplied operands as the
 * tensor-core path does. */
size_t vc_conv1_wgrad_bias_workspace_bytes(int Cout);
int vc_conv1_wgrad_bias(const float* dy, const float* x, float* dW, float* db, int B, int H, int Wd, int Cout, int round_bf16,
                        void* workspace, size_t workspace_bytes, void* stream);
/* Weight gradient of nn.ConvTranspose2d(Cin, 1, 4, 2, 1) (12:137) in one pass over x [B,Cin,H,W] and dy [B,1,2H,2W]
 * (fp32 NCHW): dW[ci][0][kh][kw] = sum x * dy(tap).  round_bf16 = 1 multiplies bf16-rounded operands (fp32 accumulation)
 * and, for Cin in {16, 32, 64, 128} with W a multiple of 32, runs on the tensor cores (mma.sync); the same holds for
 * vc_conv1_wgrad_bias above with Cout / W/2 in those roles. */
size_t vc_convT1_wgrad_workspace_bytes(int Cin);
int vc_convT1_wgrad(const float* dy, const float* x, float* dW, int B, int Cin, int H, int Wd, int round_bf16,
                    void* workspace, size_t workspace_bytes, void* stream);
/* TMA-fed weight gradients of the same layers (csrc/conv_nhwc_wgrad.cu; autograd of nn.Conv2d / nn.ConvTranspose2d at
 * 12:87-89,135-137).  op = VC_OP_CONV_WGRAD or VC_OP_CONVT_WGRAD, (B, Cin, H, W, Cout) the layer's forward geometry;
 * g_nhwc = bf16 NHWC gradient of the layer's output (ReLU mask already applied), x_nhwc = bf16 NHWC layer input; dW in the
 * state_dict layout.  Supported: the channel count on the coarse side (Cout for Conv, Cin for ConvT) 64 or 128, on the
 * sampled side 32 or 64, coarse row width a multiple of 64.  Deterministic (fixed-order split-K fold). */
int vc_conv_nhwc_wgrad_supported(int op, int B, int Cin, int H, int W, int Cout);
size_t vc_conv_nhwc_wgrad_workspace_bytes(int op, int B, int Cin, int H, int W, int Cout);
int vc_conv_nhwc_wgrad(int op, const void* g_nhwc, const void* x_nhwc, float* dW, int B, int Cin, int H, int Wd, int Cout,
                       void* workspace, size_t workspace_bytes, void* stream);
/* Small kernels around the bf16-NHWC convolution path (csrc/nhwc_misc.cu).
 *  vc_nhwc_conv1_fwd    : nn.Conv2d(1, Cout, 3, 2, 1) (+ReLU) of 12:87 with y written as bf16 NHWC [B,H/2,W/2,Cout].
 *  vc_nhwc_convT1_dgrad : data gradient of nn.ConvTranspose2d(Cin, 1, 4, 2, 1) (12:137) with the ReLU mask of its input
 *                         activation act_nhwc (bf16 [B,H,W,Cin], may be NULL) applied; g written as bf16 NHWC.
 *  vc_nhwc_from_nchw    : fp32 [B,C,H,W] -> bf16 [B,H,W,C]  (the decoder's Linear output reshaped, 12:133); C % 16 == 0, W % 64 == 0.
 *  vc_nhwc_bias_grad    : db[c] = sum over rows of a bf16 [rows, C] matrix (bias gradient of a conv layer whose output
 *                         gradient lives in NHWC), fixed summation order. */
int vc_nhwc_conv1_fwd(const float* x, const float* W, const float* bias, int relu, void* y_nhwc, int B, int H, int Wd,
                      int Cout, void* stream);
int vc_nhwc_convT1_dgrad(const float* dy, const float* W, const void* act_nhwc, void* g_nhwc, int B, int Cin, int H,
                         int Wd, void* stream);
/* vc_nhwc_convT1_dgrad plus db[c] = sum over pixels of the (masked, bf16) result: the bias gradient of the layer below
 * (12:136) from the same pass. */
size_t vc_nhwc_convT1_dgrad_bias_workspace_bytes(int Cin);
int vc_nhwc_convT1_dgrad_bias(const float* dy, const float* W, const void* act_nhwc, void* g_nhwc, float* db, int B, int Cin,
                              int H, int Wd, void* workspace, size_t workspace_bytes, void* stream);
int vc_nhwc_from_nchw(const float* src, void* dst_nhwc, int B, int C, int H, int Wd, void* stream);
/*  vc_nhwc_convT1_fwd   : nn.ConvTranspose2d(Cin, 1, 4, 2, 1) forward (12:137) from a bf16 NHWC input [B,H,W,Cin]; y fp32
 *                         [B,1,2H,2W].
 *  vc_nhwc_tap_wgrad    : weight gradient of either single-channel end from a bf16 NHWC tensor a_nhwc [B,PH,PW,C] and the
 *                         one-channel fp32 tensor S [B,2PH,2PW]: dW[c][kh][kw] = sum a[b,i,j,c] * S[b,2i-1+kh,2j-1+kw].
 *                         KS = 3: Conv2d(1,C,3,2,1) (a = output gradient, S = x, 12:87); KS = 4: ConvTranspose2d(C,1,4,2,1)
 *                         (a = layer input, S = output gradient, 12:137).  C % 32 == 0, PW % 32 == 0. */
int vc_nhwc_convT1_fwd(const void* x_nhwc, const float* W, const float* bias, float* y, int B, int Cin, int H, int Wd,
                       void* stream);
size_t vc_nhwc_tap_wgrad_workspace_bytes(int C, int KS);
int vc_nhwc_tap_wgrad(int KS, const void* a_nhwc, const float* S, float* dW, int B, int C, int PH, int PW, void* workspace,
                      size_t workspace_bytes, void* stream);
/* Weight AND bias gradient of nn.Conv2d(1, Cout, 3, 2, 1) (12:87) from the bf16 NHWC gradient g_nhwc [B,H/2,W/2,Cout] of its
 * output and x [B,1,H,W]: vc_nhwc_tap_wgrad with KS = 3 plus db[c] = sum g (of the bf16 values).  For Cout % 32 == 0 the bias
 * gradient comes out of the same tensor-core pass (a constant-one tap). */
size_t vc_nhwc_conv1_wgrad_bias_workspace_bytes(int Cout);
int vc_nhwc_conv1_wgrad_bias(const void* g_nhwc, const float* x, float* dW, float* db, int B, int H, int Wd, int Cout,
                             void* workspace, size_t workspace_bytes, void* stream);
size_t vc_nhwc_bias_grad_workspace_bytes(int C);
int vc_nhwc_bias_grad(const void* g_nhwc, int64_t rows, int C, float* db, void* workspace, size_t workspace_bytes,
                      void* stream);
/* TMA-fed implicit-GEMM convolutions on bf16 NHWC activations (csrc/conv_nhwc.cu): the hidden conv layers of script 12
 * (nn.Conv2d(k3,s2,p1) 12:87-89 and nn.ConvTranspose2d(k4,s2,p1) 12:135-137), forward and data gradient, with the
 * elementwise neighbours fused into the epilogue.  op = VC_OP_CONV_FWD / CONVT_FWD / CONV_DGRAD / CONVT_DGRAD;
 * (B, Cin, H, W, Cout) is the LAYER's forward geometry (x is [B,Cin,H,W]) as in vc_dense_op.
 *   src_nhwc : bf16 [B][h][w][C] -- the layer input (forward) or the masked output gradient (data gradient)
 *   Wt       : fp32 weights in the reference's state_dict layout ([Cout,Cin,3,3] / [Cin,Cout,4,4])
 *   bias, relu : forward epilogue (bias may be NULL)
 *   mask_nhwc  : bf16 NHWC tensor of the OUTPUT's shape or NULL; out = acc * (mask > 0)  (ReLU backward of the layer below)
 *   out_nhwc / out_nchw : bf16 NHWC and / or fp32 NCHW result (either may be NULL, not both)
 * Supported: H, W even, output row width W/2 (conv) or W (convT) a multiple of 128, source channels 32 (Conv forward,
 * ConvT data gradient) or a multiple of 64, at most 256 GEMM columns (Cout, 4*Cout, 4*Cin, Cin respectively). */
int vc_conv_nhwc_supported(int op, int B, int Cin, int H, int W, int Cout);
/* profiling aid: clock64 stamps of CTA 0's producers / MMA issuer for the first 64 k-blocks (8 int64 each); NULL = off */
int vc_conv_nhwc_debug_trace(long long* buf);
size_t vc_conv_nhwc_workspace_bytes(int op, int B, int Cin, int H, int W, int Cout);
int vc_conv_nhwc(int op, const void* src_nhwc, const float* Wt, const float* bias, int relu, const void* mask_nhwc,
                 void* out_nhwc, float* out_nchw, int B, int Cin, int H, int W, int Cout, void* workspace,
                 size_t workspace_bytes, void* stream);
/* TMA-fed data gradients (bf16 tensor-core mode): the masked gradient g = dy * (y > 0) of a ReLU(conv) / ReLU(convT)
 * layer of 12:87-89,135-137 is produced ONCE, by the pass that also yields the bias gradient, as fp32 NCHW (for the
 * weight-gradient kernel) and as a bf16 NHWC copy; the data-gradient implicit GEMM then fetches its A tiles from
 * that copy with 4-D TMA boxes (no gather warps).  vc_relu_bwd_bias_nhwc: dy, y [B,C,H,W] fp32 -> g [B,C,H,W] fp32,
 * g_nhwc [B,H,W,C] bf16, db [C]; needs C % 8 == 0, C <= 256, W % 64 == 0.  vc_conv_dgrad_nhwc: op is
 * VC_OP_CONV_DGRAD (g_nhwc [B,OH,OW,Cout] -> dx [B,Cin,H,W]) or VC_OP_CONVT_DGRAD (g_nhwc [B,2H,2W,Cout] ->
 * dx [B,Cin,H,W]); geometry arguments as in vc_dense_op; ..._supported() says whether the shape qualifies (rows of 128
 * pixels, 32 / 64 / 128 channels), otherwise callers use vc_relu_bwd_bias + vc_dense_op. */
size_t vc_relu_bwd_bias_nhwc_workspace_bytes(int C);
int vc_relu_bwd_bias_nhwc(const float* dy, const float* y, int B, int C, int H, int W, float* g, void* g_nhwc, float* db,
                          void* workspace, size_t workspace_bytes, void* stream);
int vc_conv_dgrad_nhwc_supported(int op, int B, int Cin, int H, int W, int Cout);
size_t vc_conv_dgrad_nhwc_workspace_bytes(int op, int B, int Cin, int H, int W, int Cout);
int vc_conv_dgrad_nhwc(int op, const void* g_nhwc, const float* W, float* dx, int B, int Cin, int H, int Wd, int Cout,
                       void* workspace, size_t workspace_bytes, void* stream);

/* One entry point for every dense-layer op in either precision.  VC_PREC_FP32: the fp32 FFMA kernels above
 * (reference-precision path).  VC_PREC_BF16_TC: operands rounded to bf16, tcgen05 tensor-core MMAs with fp32
 * accumulation in tensor memory, fp32 outputs.  Argument meaning per op (in0, in1 -> out):
 *   LINEAR_FWD  x[B,Cin],  W[Cout,Cin]        -> y[B,Cout]          (H = W = 1; bias, relu used)
 *   LINEAR_DGRAD dy[B,Cout], W                -> dx[B,Cin]
 *   LINEAR_WGRAD dy[B,Cout], x[B,Cin]         -> dW[Cout,Cin]
 *   CONV_FWD    x[B,Cin,H,W], W[Cout,Cin,3,3] -> y[B,Cout,OH,OW]    (bias, relu used)
 *   CONV_DGRAD  dy, W -> dx[B,Cin,H,W]          CONV_WGRAD  dy, x -> dW[Cout,Cin,3,3]
 *   CONVT_FWD   x[B,Cin,H,W], W[Cin,Cout,4,4] -> y[B,Cout,2H,2W]    (bias, relu used)
 *   CONVT_DGRAD dy, W -> dx[B,Cin,H,W]          CONVT_WGRAD dy, x -> dW[Cin,Cout,4,4] */
#define VC_PREC_FP32 0
#define VC_PREC_BF16_TC 1
#define VC_OP_LINEAR_FWD 0
#define VC_OP_LINEAR_DGRAD 1
#define VC_OP_LINEAR_WGRAD 2
#define VC_OP_CONV_FWD 3
#define VC_OP_CONV_DGRAD 4
#define VC_OP_CONV_WGRAD 5
#define VC_OP_CONVT_FWD 6
#define VC_OP_CONVT_DGRAD 7
#define VC_OP_CONVT_WGRAD 8
size_t vc_dense_op_workspace_bytes(int op, int precision, int B, int Cin, int H, int W, int Cout);
int vc_dense_op(int op, int precision, const float* in0, const float* in1, const float* bias, float* out, int B,
                int Cin, int H, int W, int Cout, int relu, void* workspace, size_t workspace_bytes, void* stream);

/* Whole backward of one SMALL Linear(+ReLU) layer (latent head / lyrics branch / script-19 and script-06 MLPs,
 * 12:88-120,159-166; 19:79-95; 06:149-157) in one launch, bf16 operands / fp32 accumulate like VC_PREC_BF16_TC:
 * g = dy * (y > 0) (y = the layer's saved ReLU output, or NULL for no activation); dW[Cout,Cin] = g^T x;
 * db[Cout] = column sums of g (NULL: skip); dx[B,Cin] = g W (NULL: skip).  vc_linear_bwd_small_supported() says whether
 * a shape qualifies (B <= 256, Cin*Cout <= 2^20); other shapes go through vc_relu_bwd_bias + vc_dense_op. */
int vc_linear_bwd_small_supported(int B, int Cin, int Cout);
int vc_linear_bwd_small(const float* dy, const float* y, const float* x, const float* W, float* dW, float* db,
                        float* dx, int B, int Cin, int Cout, void* stream);

/* ------------------------------------------------------------------------------------------------- Adam
 * torch.optim.Adam defaults (12:233, 19:202, 06:310): one launch over a device-resident table of chunks. */
typedef struct VcAdamChunk {
  float* p;
  const float* g;
  float* m;
  float* v;
  int64_t n;
} VcAdamChunk;
int vc_adam_multi(const VcAdamChunk* table, int n_chunks, float lr, float beta1, float beta2, float eps,
                  float bias_corr1, float bias_corr2, void* stream);
/* The same update for a step captured in a CUDA graph (torch.optim.Adam(capturable=True) semantics): the step counter
 * and the bias corrections live in device memory; each call advances `step_state->step` by one on the device and
 * derives 1 - beta^step from it, so a replayed graph performs step t, t+1, ... without host arguments changing. */
typedef struct VcAdamDeviceStep {
  int32_t step;
  float bias_corr1;
  float bias_corr2;
  int32_t pad;
} VcAdamDeviceStep;
int vc_adam_multi_graph(const VcAdamChunk* table, int n_chunks, float lr, float beta1, float beta2, float eps,
                        VcAdamDeviceStep* step_state, void* stream);
/* The two halves of vc_adam_multi_graph on their own, for a step that updates its parameters in several launches
 * (sharded / bucketed updates that start as soon as a gradient bucket has been reduced): vc_adam_tick advances the
 * counter once per step, vc_adam_multi_at updates the chunks of `table` with the bias corrections of the current
 * device step. */
int vc_adam_tick(VcAdamDeviceStep* step_state, float beta1, float beta2, void* stream);
int vc_adam_multi_at(const VcAdamChunk* table, int n_chunks, float lr, float beta1, float beta2, float eps,
                     const VcAdamDeviceStep* step_state, void* stream);
/* Weight gradient of a wide nn.Linear (the two 67 M-parameter layers of 12:98-103,129-132) FUSED with its Adam update
 * (12:233,270): dW = dy^T x is formed in tensor memory and applied to W / exp_avg / exp_avg_sq in the epilogue; the
 * gradient is never written to memory.  Arithmetic order as vc_adam_multi, the quotient with sqrt.approx / div.approx
 * (this path exists in bf16 mode only; VAECLUST_FC_ADAM_IEEE=1 restores IEEE).  step_state (device, advanced by vc_adam_tick)
 * supplies the bias corrections when non-NULL, else bias_corr1 / bias_corr2 are used.  Single-process training only. */
int vc_linear_wgrad_adam_supported(int B, int Cin, int Cout);
size_t vc_linear_wgrad_adam_workspace_bytes(int B, int Cin, int Cout);
int vc_linear_wgrad_adam(const float* dy, const float* x, float* W, float* exp_avg, float* exp_avg_sq, void* w_shadow,
                         int B, int Cin, int Cout, float lr, float beta1, float beta2, float eps, float bias_corr1,
                         float bias_corr2, const VcAdamDeviceStep* step_state, void* workspace, size_t workspace_bytes,
                         void* stream);
/* The same wide layers read from a bf16 SHADOW of the weight ([Cout][Cin], kept in step with the fp32 master by
 * vc_linear_wgrad_adam's w_shadow argument, which may be NULL there): op = VC_OP_LINEAR_FWD (y = act(x W^T + b)) or
 * VC_OP_LINEAR_DGRAD (dx = dy W).  The weight tile is one TMA load straight into the MMA's operand layout (K-major boxes
 * forward, MN-major boxes for the data gradient) -- half the bytes of the fp32 stream and no conversion instructions;
 * the products are the same bf16 x bf16 ones as vc_dense_op in bf16 mode. */
int vc_linear_bf16w_supported(int op, int B, int Cin, int Cout);
size_t vc_linear_bf16w_workspace_bytes(int op, int B, int Cin, int Cout);
int vc_linear_bf16w(int op, const float* in, const void* w_shadow, const float* bias, float* out, int B, int Cin, int Cout,
                    int relu, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------ fused MLP-VAE train step (scripts 19 / 06 / 22)
 * The whole optimisation step of MLPVAE (19:64-121 model, 19:226-232 step), of script 06's MLPVAE (06:145-188,
 * 06:322-327) and of script 22's AE baseline (22:66-87,150-156) in four launches: pack -> [forward + loss + data
 * gradients per 128-row tile, tcgen05] -> [weight gradients with the batch as the contraction, tcgen05] -> [fold +
 * Adam + bf16 weight shadows].  hidden must be 256, latent in {8,16,32}, input_dim + cond_dim <= 512,
 * latent + cond_dim <= 64; anything else is VC_ERR_UNSUPPORTED (callers then use the per-layer ops above).
 *   variational = 1: heads mu | logvar, z = mu + eps*exp(logvar/2), loss = recon + beta*KL under `loss_mode`
 *   variational = 0: one head z (22:66-87), loss = mean squared error (VC_LOSS_MEAN_MEAN), no KL
 * Parameter order in VcMlpVaeParams: 0 enc.0 / enc1, 1 enc.2 / enc2, 2 mu (or the AE's enc.4), 3 logvar (unused for
 * the AE), 4 dec.0 / dec1, 5 dec.2 / dec2, 6 dec.4 / out; p = fp32 master, m / v = Adam moments (same shape).
 * The struct itself lives in HOST memory and holds device pointers.
 * The workspace belongs to ONE model instance and must persist between calls (it holds the bf16 weight shadows):
 * call vc_mlpvae_prepare once, and again whenever the fp32 masters were changed from outside (load_state_dict).
 * `phases` bit 0: pack + rows + weight-gradient partials; bit 1: fold the partials into the flat gradient block
 * (scaled by grad_weight = B_local / B_global) and write out3 = {loss, recon, kl} of THIS batch; bit 2: Adam from the
 * flat block + refresh the shadows.  One GPU: phases = 7.  Data parallel: 3, all-reduce(sum) the flat block
 * (vc_mlpvae_grad_layout), then 4.  step_state != NULL: bias corrections are read from (and, with bit 0, the step
 * counter advanced in) device memory, so the call sequence can be replayed from a CUDA graph. */
typedef struct VcMlpVaeDesc {
  int32_t input_dim, cond_dim, hidden, latent, variational, loss_mode;
  float beta;
  int32_t pad;
} VcMlpVaeDesc;
typedef struct VcMlpTensor {
  float* p;
  float* m;
  float* v;
} VcMlpTensor;
typedef struct VcMlpVaeParams {
  VcMlpTensor W[7];
  VcMlpTensor b[7];
} VcMlpVaeParams;
int vc_mlpvae_supported(const VcMlpVaeDesc* desc);
/* debug / profiling hook: a device buffer of >= 128 int64 receives clock64 stamps of the rows kernel's first tile
 * (NULL switches it off again); see profiles/trace_fused.py */
int vc_mlpvae_debug_trace(int64_t* device_buffer);
size_t vc_mlpvae_workspace_bytes(const VcMlpVaeDesc* desc, int B);
/* out[0] = byte offset (from the workspace pointer rounded up to 1024) of the flat fp32 gradient block, out[1] = its
 * element count; then for each of the 6 layers (heads = mu rows then logvar rows): weight offset in elements, row
 * pitch in elements, bias offset in elements.  Gradient of W_l[o][i] = block[woff + o*pitch + i]. */
int vc_mlpvae_grad_layout(const VcMlpVaeDesc* desc, int B, int64_t* out);
int vc_mlpvae_prepare(const VcMlpVaeDesc* desc, const VcMlpVaeParams* params, int B, void* workspace,
                      size_t workspace_bytes, void* stream);
int vc_mlpvae_step(const VcMlpVaeDesc* desc, const VcMlpVaeParams* params, const float* x, const float* c,
                   const float* eps, int B, float grad_weight, int phases, float lr, float beta1, float beta2,
                   float adam_eps, float bias_corr1, float bias_corr2, VcAdamDeviceStep* step_state, float* out3,
                   void* workspace, size_t workspace_bytes, void* stream);

/* --------------------------------------------------------------------------------------------- KMeans
 * One Lloyd iteration of sklearn:cluster/_k_means_lloyd.pyx:23-218 reached from 07:70-71, 13:120-121, 16:183,
 * 20:68-69: labels[i] = first argmin_j (||c_j||^2 - 2 x_i.c_j); per-cluster sums/counts of the points.
 * X [N,d] fp32 row-major (d % 4 == 0, d <= 256), C [k,d] (k <= 256).  `labels` holds the previous labels on
 * entry (any value on the first iteration) and the new ones on exit; *n_changed counts differences.
 * sums [k,d] and counts [k] are doubles reduced in a fixed order (deterministic).  update=0 skips the
 * M-step accumulation (final E-step, _kmeans.py:742-754). */
size_t vc_kmeans_workspace_bytes(int k, int d);
int vc_kmeans_assign_update(const float* X, int64_t N, int d, const float* C, int k, int32_t* labels,
                            double* sums, double* counts, int64_t* n_changed, int update, void* workspace,
                            size_t workspace_bytes, void* stream);
/* centers_new = sums/counts (count==0 keeps the old centre, _k_means_common.pyx:274-295);
 * *shift_sq = sum_j ||c_new_j - c_old_j||^2 (:298-311). */
int vc_kmeans_finalize(const double* sums, const double* counts, const float* C_old, int k, int d, float* C_new,
                       double* shift_sq, void* stream);
/* Sync-free form of the Lloyd loop (sklearn:cluster/_kmeans.py:705-738): the host enqueues
 *   begin ; { estep ; [all-reduce(stats[0 .. k*d+k+1)) when X is sharded] ; finalize } x m ; read ctl back ONCE
 * stats = [sums k*d | counts k | n_changed | shift_sq] (doubles, k*d+k+2); ctl = {state, n_iter, -, -} (int32 x 4):
 * state 0 running, 1 strict convergence (no label changed), 2 centre shift <= tol, 3 an empty cluster needs the host's
 * relocation path (_k_means_common.pyx:167-211; centres untouched, iteration not counted).  finalize replaces C by the
 * new centres in place (C_scratch: k*d floats); once state != 0 every later estep / finalize launch is a no-op, so
 * over-enqueued iterations change nothing.  Results are bit-identical to the stepwise entry points above.
 * tol < 0 disables both stopping rules (benchmarks of the steady loop run exactly the enqueued iterations). */
int vc_kmeans_lloyd_begin(int32_t* ctl, void* workspace, size_t workspace_bytes, int k, int d, void* stream);
int vc_kmeans_lloyd_estep(const float* X, int64_t N, int d, const float* C, int k, int32_t* labels, double* stats,
                          const int32_t* ctl, void* workspace, size_t workspace_bytes, void* stream);
int vc_kmeans_lloyd_finalize(double* stats, float* C, float* C_scratch, int k, int d, double tol, int32_t* ctl,
                             void* stream);
/* inertia = sum_i ||x_i - c_label(i)||^2 by direct differences (_k_means_common.pyx:94-124); optionally writes
 * per-point squared distances (used by the empty-cluster relocation, :167-211). */
int vc_kmeans_inertia(const float* X, int64_t N, int d, const float* C, int k, const int32_t* labels,
                      double* inertia, float* point_sqdist, void* workspace, size_t workspace_bytes, void* stream);
/* per-cluster sums/counts for given labels (M-step only; also first pass of Davies-Bouldin). */
int vc_cluster_sums(const float* X, int64_t N, int d, const int32_t* labels, int k, double* sums, double* counts,
                    void* workspace, size_t workspace_bytes, void* stream);
/* column mean and (population) variance in double: KMeans centering + tol (_kmeans.py:285-293,1487-1490),
 * StandardScaler (06:291-292, 07:67-68, 13:194, 16:178). */
size_t vc_col_stats_workspace_bytes(int d);
int vc_col_stats(const float* X, int64_t N, int d, double* mean, double* var, void* workspace,
                 size_t workspace_bytes, void* stream);
/* out = (X - shift) * scale  per column (shift/scale double [d]; scale may be NULL) */
int vc_col_affine(const float* X, int64_t N, int d, const double* shift, const double* scale, float* out,
                  void* stream);
/* squared distances from m query rows Q [m,d] to every row of X -> out [m,N] (k-means++ potentials,
 * _kmeans.py:180-278) */
int vc_sqdist_rows(const float* X, int64_t N, int d, const float* Q, int m, float* out, void* stream);

/* ----------------------------------------------------------------------------------------- silhouette
 * sklearn:metrics/cluster/_unsupervised.py:149-200,211-323 reached from 13:74,78, 16:68,73, 20:45, 22:61.
 * Xs is X permuted so that rows are grouped by cluster (cluster c occupies rows [off[c], off[c+1])); sqn are
 * the squared row norms; off_dev is a DEVICE array of k+1 int64 offsets.  For rows [row_begin,row_end) computes
 * cluster_sums[s, i-row_begin, c] = sum over the column tiles of cluster c owned by split s (s < nsplit; splits
 * only add parallelism for small N) of ||x_i - x_j|| (self-distance forced to 0, pairwise.py:2285).
 * cluster_sums is [nsplit, row_end-row_begin, k]. */
int vc_silhouette_gather(const float* X, int64_t N, int d, const int64_t* order, const double* shift, float* Xs,
                         float* sqn, void* stream);
int vc_silhouette_rows(const float* Xs, const float* sqn, int64_t N, int d, const int64_t* off_dev, int k,
                       int64_t row_begin, int64_t row_end, int nsplit, float* cluster_sums, void* stream);
/* Tensor-core variant (tcgen05, bf16 hi/lo split with fp32 accumulation in TMEM; d <= 128).  First split the sorted,
 * centred rows into two bf16 matrices H, L of shape [N, dp] (dp = d rounded up to a multiple of 64, caller-owned,
 * 128-byte aligned), then run row blocks like vc_silhouette_rows, except that cluster_sums holds P =
 * vc_silhouette_tc_planes(dp) partial planes per split (one per column part of a tile an epilogue thread owns):
 * [P*nsplit, row_end-row_begin, k] (pass P*nsplit to vc_silhouette_finalize).
 * sqn_split (may be NULL) receives the squared norms of the split rows (hi + lo); pass those as `sqn` to
 * vc_silhouette_rows_tc so the norm formula cancels consistently.  Xs/d are the fp32 rows (d % 4 == 0): pairs whose
 * squared distance falls below 1/256 of |x|^2 + |y|^2 (duplicates, near-duplicates) are recomputed from them as
 * sum (x - y)^2, the accuracy sklearn gets from its float64 upcast (pairwise.py:567-638). */
int vc_silhouette_tc_planes(int dp);
int vc_silhouette_split_bf16(const float* Xs, int64_t N, int d, int dp, void* H, void* L, float* sqn_split,
                             void* stream);
int vc_silhouette_rows_tc(const void* H, const void* L, const float* Xs, int d, const float* sqn, int64_t N, int dp,
                          const int64_t* off_dev, int k, int64_t row_begin, int64_t row_end, int nsplit,
                          float* cluster_sums, void* stream);
/* s_i = (b-a)/max(a,b) from cluster_sums (:191-200,315-323); writes s per row and *s_sum (double).
 * labels_sorted points at the label of row_begin. */
size_t vc_silhouette_finalize_workspace_bytes(void);
int vc_silhouette_finalize(const float* cluster_sums, int nsplit, const int32_t* labels_sorted,
                           const int64_t* counts_dev, int k, int64_t nrows, float* s_rows, double* s_sum,
                           void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------- Davies-Bouldin
 * sklearn:metrics/cluster/_unsupervised.py:457-487 reached from 13:94,98, 16:89,94.
 * intra_sum[c] = sum_{i in c} ||x_i - centroid_c|| (double). */
int vc_dbi_intra(const float* X, int64_t N, int d, const int32_t* labels, const double* centroids, int k,
                 double* intra_sum, void* workspace, size_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VAECLUST_H_ */
