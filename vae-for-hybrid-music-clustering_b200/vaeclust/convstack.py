"""The conv stacks of script 12 (12:87-89 encoder, 12:135-137 decoder) with bf16 NHWC activations between the layers.

In bf16 mode the per-layer autograd Functions of `ops` exchange fp32 NCHW tensors: every hidden activation is written and
re-read as fp32, the ReLU backward / bias gradient is a separate 12 B/element pass, and the implicit-GEMM kernels gather
and convert their operands with eight producer warps.  Here a whole stack is ONE autograd Function: hidden activations and
their gradients exist only as bf16 NHWC, every hidden-layer operator (forward, data gradient, weight gradient) is fed by
TMA (csrc/conv_nhwc.cu, csrc/conv_nhwc_wgrad.cu), bias + ReLU and the ReLU masks are fused into the producing epilogues.
The single-channel ends and the fully connected neighbours keep their fp32 NCHW interfaces, so the Functions drop into
`AudioEncoder.features` / `AudioDecoder.forward` without changing anything around them.  Arithmetic is the same as the
per-layer path (bf16 operands, fp32 accumulation); the only new rounding is that the bias gradients of the hidden layers
sum bf16-rounded gradients.
"""
from __future__ import annotations

import os

import torch

from . import ops
from ._lib import lib, ptr, stream_ptr

ENABLED = True      # bf16 mode only; `supported()` decides per shape, anything else takes the per-layer path
# The single-channel ends run on bf16 NHWC as well (csrc/nhwc_mma.cu: mma.sync kernels for the 32-channel layers), so no
# fp32 copy of the two widest activations exists at all: deconv2 writes one layout instead of two (0.145 -> 0.105 ms),
# deconv.4 forward reads half the bytes (0.115 -> 0.059 ms), both tap weight gradients likewise (0.069 / 0.077 ->
# 0.050 / 0.056 ms incl. the bias gradient).  VAECLUST_ENDS_NHWC=0 keeps fp32 NCHW at the two ends (the round-2 default
# before the tensor-core kernels existed: as CUDA-core kernels the NHWC ends were slower than the fp32 ones).
ENDS_NHWC = os.environ.get("VAECLUST_ENDS_NHWC", "1") == "1"

CONV_FWD, CONV_DGRAD, CONV_WGRAD, CONVT_FWD, CONVT_DGRAD, CONVT_WGRAD = 3, 4, 5, 6, 7, 8
_BF = torch.bfloat16


def _conv(op, src, w, bias, relu, mask, out_nhwc, out_nchw, dims):
    B, Cin, H, W, Cout = dims
    L = lib()
    ws = ops._workspace(L.size("vc_conv_nhwc_workspace_bytes", op, B, Cin, H, W, Cout), src.device)
    L.call("vc_conv_nhwc", op, ptr(src), ptr(w), ptr(bias), int(relu), ptr(mask), ptr(out_nhwc), ptr(out_nchw), B, Cin, H, W,
           Cout, ptr(ws), ws.numel(), stream_ptr())


def _wgrad(op, g, x, w, dims):
    B, Cin, H, W, Cout = dims
    L = lib()
    dw = torch.empty_like(w)
    ws = ops._workspace(L.size("vc_conv_nhwc_wgrad_workspace_bytes", op, B, Cin, H, W, Cout), g.device)
    L.call("vc_conv_nhwc_wgrad", op, ptr(g), ptr(x), ptr(dw), B, Cin, H, W, Cout, ptr(ws), ws.numel(), stream_ptr())
    return dw


def _tap_wgrad(KS, a, S, w):
    B, PH, PW, C = a.shape
    L = lib()
    dw = torch.empty_like(w)
    ws = ops._workspace(L.size("vc_nhwc_tap_wgrad_workspace_bytes", C, KS), a.device)
    L.call("vc_nhwc_tap_wgrad", KS, ptr(a), ptr(S), ptr(dw), B, C, PH, PW, ptr(ws), ws.numel(), stream_ptr())
    return dw


def _bias_nhwc(g):
    C = g.shape[-1]
    L = lib()
    db = torch.empty(C, dtype=torch.float32, device=g.device)
    ws = ops._workspace(L.size("vc_nhwc_bias_grad_workspace_bytes", C), g.device)
    L.call("vc_nhwc_bias_grad", ptr(g), g.numel() // C, C, ptr(db), ptr(ws), ws.numel(), stream_ptr())
    return db


def _bias_nchw(g):
    B, C = g.shape[0], g.shape[1]
    L = lib()
    db = torch.empty(C, dtype=torch.float32, device=g.device)
    ws = ops._workspace(L.size("vc_bias_grad_workspace_bytes", C), g.device)
    L.call("vc_bias_grad", ptr(g), B, C, g.shape[2] * g.shape[3], ptr(db), 0, ptr(ws), ws.numel(), stream_ptr())
    return db


def _empty_nhwc(B, H, W, C, dev):
    return torch.empty((B, H, W, C), dtype=_BF, device=dev)


_ok_cache: dict = {}


def _ok(*checks) -> bool:
    r = _ok_cache.get(checks)
    if r is None:
        L = lib()
        r = all(L.size(name, *args) == 1 for name, args in checks)
        _ok_cache[checks] = r
    return r


def encoder_supported(x, convs) -> bool:
    if not (ENABLED and ops.PRECISION == "bf16" and x.is_cuda and x.dim() == 4 and x.shape[1] == 1 and len(convs) == 3):
        return False
    B, _, H, W = x.shape
    c1, c2, c3 = (c.weight.shape[0] for c in convs)
    if H % 8 or W % 8 or c1 % 32 or convs[1].weight.shape[1] != c1 or convs[2].weight.shape[1] != c2:
        return False
    d2, d3 = (B, c1, H // 2, W // 2, c2), (B, c2, H // 4, W // 4, c3)
    return _ok(("vc_conv_nhwc_supported", (CONV_FWD, *d2)), ("vc_conv_nhwc_supported", (CONV_FWD, *d3)),
               ("vc_conv_nhwc_supported", (CONV_DGRAD, *d2)), ("vc_conv_nhwc_supported", (CONV_DGRAD, *d3)),
               ("vc_conv_nhwc_wgrad_supported", (CONV_WGRAD, *d2)), ("vc_conv_nhwc_wgrad_supported", (CONV_WGRAD, *d3))) \
        and (W // 8) % 64 == 0


def decoder_supported(h, deconvs) -> bool:
    if not (ENABLED and ops.PRECISION == "bf16" and h.is_cuda and h.dim() == 4 and len(deconvs) == 3):
        return False
    B, C0, H, W = h.shape
    c1, c2, c3 = (d.weight.shape[1] for d in deconvs)
    if c3 != 1 or C0 % 16 or W % 64 or deconvs[0].weight.shape[0] != C0 or c2 % 32:
        return False
    d1, d2 = (B, C0, H, W, c1), (B, c1, 2 * H, 2 * W, c2)
    return _ok(("vc_conv_nhwc_supported", (CONVT_FWD, *d1)), ("vc_conv_nhwc_supported", (CONVT_FWD, *d2)),
               ("vc_conv_nhwc_supported", (CONVT_DGRAD, *d1)), ("vc_conv_nhwc_supported", (CONVT_DGRAD, *d2)),
               ("vc_conv_nhwc_wgrad_supported", (CONVT_WGRAD, *d1)), ("vc_conv_nhwc_wgrad_supported", (CONVT_WGRAD, *d2)))


class EncoderConvFn(torch.autograd.Function):
    """x [B,1,H,W] fp32 -> ReLU(conv3(ReLU(conv2(ReLU(conv1(x)))))) as fp32 NCHW [B,C3,H/8,W/8] (12:87-89, 12:105)."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, w3, b3):
        x = ops._f32(x, "encoder.x")
        B, _, H, W = x.shape
        c1, c2, c3 = w1.shape[0], w2.shape[0], w3.shape[0]
        dev = x.device
        L = lib()
        a1 = _empty_nhwc(B, H // 2, W // 2, c1, dev)
        L.call("vc_nhwc_conv1_fwd", ptr(x), ptr(w1), ptr(b1), 1, ptr(a1), B, H, W, c1, stream_ptr())
        a2 = _empty_nhwc(B, H // 4, W // 4, c2, dev)
        _conv(CONV_FWD, a1, w2, b2, True, None, a2, None, (B, c1, H // 2, W // 2, c2))
        a3 = torch.empty((B, c3, H // 8, W // 8), dtype=torch.float32, device=dev)
        _conv(CONV_FWD, a2, w3, b3, True, None, None, a3, (B, c2, H // 4, W // 4, c3))
        ctx.save_for_backward(x, a1, a2, a3, w1, w2, w3)
        return a3

    @staticmethod
    def backward(ctx, da3):
        x, a1, a2, a3, w1, w2, w3 = ctx.saved_tensors
        da3 = ops._f32(da3, "encoder.da3")
        B, _, H, W = x.shape
        c1, c2, c3 = w1.shape[0], w2.shape[0], w3.shape[0]
        dev = x.device
        L = lib()
        s = stream_ptr()
        d2, d3 = (B, c1, H // 2, W // 2, c2), (B, c2, H // 4, W // 4, c3)
        # conv3: g3 = da3 * (a3 > 0) as bf16 NHWC (+ db3) from the fp32 NCHW pair the fully connected layer exchanges
        g3n = _empty_nhwc(B, H // 8, W // 8, c3, dev)
        db3 = torch.empty(c3, dtype=torch.float32, device=dev)
        ws = ops._workspace(L.size("vc_relu_bwd_bias_nhwc_workspace_bytes", c3), dev)
        L.call("vc_relu_bwd_bias_nhwc", ptr(da3), ptr(a3), B, c3, H // 8, W // 8, None, ptr(g3n), ptr(db3), ptr(ws),
               ws.numel(), s)
        dw3 = _wgrad(CONV_WGRAD, g3n, a2, w3, d3)
        g2n = torch.empty_like(a2)
        _conv(CONV_DGRAD, g3n, w3, None, False, a2, g2n, None, d3)          # ReLU mask of a2 in the epilogue
        db2 = _bias_nhwc(g2n)
        dw2 = _wgrad(CONV_WGRAD, g2n, a1, w2, d2)
        dx = None
        if ENDS_NHWC:
            g1n = torch.empty_like(a1)
            _conv(CONV_DGRAD, g2n, w2, None, False, a1, g1n, None, d2)
            dw1 = torch.empty_like(w1)          # weight and bias gradient from one tensor-core pass over g1n
            db1 = torch.empty(c1, dtype=torch.float32, device=dev)
            ws = ops._workspace(L.size("vc_nhwc_conv1_wgrad_bias_workspace_bytes", c1), dev)
            L.call("vc_nhwc_conv1_wgrad_bias", ptr(g1n), ptr(x), ptr(dw1), ptr(db1), B, H, W, c1, ptr(ws), ws.numel(), s)
            g1 = g1n.float().permute(0, 3, 1, 2).contiguous() if ctx.needs_input_grad[0] else None
        else:       # conv1's gradient leaves the NHWC world: fp32 NCHW for the single-channel weight-gradient kernel
            g1 = torch.empty((B, c1, H // 2, W // 2), dtype=torch.float32, device=dev)
            _conv(CONV_DGRAD, g2n, w2, None, False, a1, None, g1, d2)
            dw1 = torch.empty_like(w1)
            db1 = torch.empty(c1, dtype=torch.float32, device=dev)
            ws = ops._workspace(L.size("vc_conv1_wgrad_bias_workspace_bytes", c1), dev)
            L.call("vc_conv1_wgrad_bias", ptr(g1), ptr(x), ptr(dw1), ptr(db1), B, H, W, c1, 1, ptr(ws), ws.numel(), s)
        if ctx.needs_input_grad[0]:      # not on the training path (x is data)
            dx = torch.empty_like(x)
            ops._dense(CONV_DGRAD, 1, g1, w1, None, dx, B, 1, H, W, c1, False)
        return dx, dw1, db1, dw2, db2, dw3, db3


class DecoderConvFn(torch.autograd.Function):
    """h [B,C0,H,W] fp32 -> deconv3(ReLU(deconv2(ReLU(deconv1(h))))) as fp32 NCHW [B,1,8H,8W] (12:135-137, 12:141)."""

    @staticmethod
    def forward(ctx, h, w1, b1, w2, b2, w3, b3):
        h = ops._f32(h, "decoder.h").contiguous()
        B, C0, H, W = h.shape
        c1, c2 = w1.shape[1], w2.shape[1]
        dev = h.device
        L = lib()
        d0n = _empty_nhwc(B, H, W, C0, dev)
        L.call("vc_nhwc_from_nchw", ptr(h), ptr(d0n), B, C0, H, W, stream_ptr())
        d1n = _empty_nhwc(B, 2 * H, 2 * W, c1, dev)
        _conv(CONVT_FWD, d0n, w1, b1, True, None, d1n, None, (B, C0, H, W, c1))
        d2n = _empty_nhwc(B, 4 * H, 4 * W, c2, dev)
        xhat = torch.empty((B, 1, 8 * H, 8 * W), dtype=torch.float32, device=dev)
        if ENDS_NHWC:
            d2 = None
            _conv(CONVT_FWD, d1n, w2, b2, True, None, d2n, None, (B, c1, 2 * H, 2 * W, c2))
            L.call("vc_nhwc_convT1_fwd", ptr(d2n), ptr(w3), ptr(b3), ptr(xhat), B, c2, 4 * H, 4 * W, stream_ptr())
        else:       # the single-channel end reads fp32 NCHW: deconv2 writes both layouts
            d2 = torch.empty((B, c2, 4 * H, 4 * W), dtype=torch.float32, device=dev)
            _conv(CONVT_FWD, d1n, w2, b2, True, None, d2n, d2, (B, c1, 2 * H, 2 * W, c2))
            ops._dense(CONVT_FWD, 1, d2, w3, b3, xhat, B, c2, 4 * H, 4 * W, 1, False)
        ctx.save_for_backward(d0n, d1n, d2n, d2, w1, w2, w3)
        ctx.dims = (B, C0, H, W)
        return xhat

    @staticmethod
    def backward(ctx, dxhat):
        d0n, d1n, d2n, d2, w1, w2, w3 = ctx.saved_tensors
        dxhat = ops._f32(dxhat, "decoder.dxhat")
        B, C0, H, W = ctx.dims
        c1, c2 = w1.shape[1], w2.shape[1]
        dev = dxhat.device
        L = lib()
        e1, e2 = (B, C0, H, W, c1), (B, c1, 2 * H, 2 * W, c2)
        db3 = _bias_nchw(dxhat)
        if d2 is None:
            dw3 = _tap_wgrad(4, d2n, dxhat, w3)
        else:
            dw3 = torch.empty_like(w3)   # bf16-rounded operands like every other weight gradient of the stack
            ws = ops._workspace(L.size("vc_convT1_wgrad_workspace_bytes", c2), dev)
            L.call("vc_convT1_wgrad", ptr(dxhat), ptr(d2), ptr(dw3), B, c2, 4 * H, 4 * W, 1, ptr(ws), ws.numel(), stream_ptr())
        g2n = torch.empty_like(d2n)
        db2 = torch.empty(c2, dtype=torch.float32, device=dev)     # bias gradient of deconv.2 from the same pass
        ws = ops._workspace(L.size("vc_nhwc_convT1_dgrad_bias_workspace_bytes", c2), dev)
        L.call("vc_nhwc_convT1_dgrad_bias", ptr(dxhat), ptr(w3), ptr(d2n), ptr(g2n), ptr(db2), B, c2, 4 * H, 4 * W, ptr(ws),
               ws.numel(), stream_ptr())
        dw2 = _wgrad(CONVT_WGRAD, g2n, d1n, w2, e2)
        g1n = torch.empty_like(d1n)
        _conv(CONVT_DGRAD, g2n, w2, None, False, d1n, g1n, None, e2)
        db1 = _bias_nhwc(g1n)
        dw1 = _wgrad(CONVT_WGRAD, g1n, d0n, w1, e1)
        dh = None
        if ctx.needs_input_grad[0]:
            dh = torch.empty((B, C0, H, W), dtype=torch.float32, device=dev)
            _conv(CONVT_DGRAD, g1n, w1, None, False, None, None, dh, e1)
        return dh, dw1, db1, dw2, db2, dw3, db3


def encoder_convs(x, convs):
    return EncoderConvFn.apply(x, convs[0].weight, convs[0].bias, convs[1].weight, convs[1].bias, convs[2].weight,
                               convs[2].bias)


def decoder_deconvs(h, deconvs):
    return DecoderConvFn.apply(h, deconvs[0].weight, deconvs[0].bias, deconvs[1].weight, deconvs[1].bias,
                               deconvs[2].weight, deconvs[2].bias)
