"""torch.autograd.Functions over the C-ABI of libvaeclust.so.

Every op takes/returns fp32 CUDA tensors in torch's own layouts (NCHW activations, state_dict weight layouts) and
launches on torch's current stream.  CPU tensors are rejected: there is no fallback path.
"""
from __future__ import annotations

from typing import Optional

import torch

from ._lib import VC_LOSS_MEAN_MEAN, VC_LOSS_SUM_MEAN, VaeclustError, lib, ptr, stream_ptr

_ws_cache: dict = {}
_ws_retired: list = []   # superseded buffers stay alive: a captured CUDA graph may have baked their address in


def _workspace(nbytes: int, device) -> torch.Tensor:
    """Grow-only per-device scratch buffer (the library never allocates).  A buffer that is outgrown is RETIRED, not
    freed: GraphedVAEStep records the scratch address of its launches at capture time, so returning that memory to
    the caching allocator would let a later replay scribble over somebody else's tensor."""
    key = (device.type, device.index)
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        if buf is not None:
            _ws_retired.append(buf)
        buf = torch.empty(max(int(nbytes), 1 << 20), dtype=torch.uint8, device=device)
        _ws_cache[key] = buf
    return buf


def _f32(t: torch.Tensor, name: str) -> torch.Tensor:
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise VaeclustError(f"{name}: expected a CUDA tensor (vaeclust has no CPU path)")
    if t.dtype != torch.float32:
        raise VaeclustError(f"{name}: expected float32, got {t.dtype}")
    return t.contiguous()


# ----------------------------------------------------------------------------------------------- dense layers
# bf16 mode: conv / convT data gradients from a bf16 NHWC copy of the masked gradient, fetched by TMA
# (csrc/conv_igemm_tma.cu).  Measured on B200 at the script-12 shapes: the four data-gradient kernels drop from 0.494 to
# 0.335 ms/step, but emitting the NHWC copy from the fp32-NCHW ReLU-backward pass costs 0.33 ms more than the plain pass
# (step 3.90 -> 4.07 ms), so it stays OFF until activations flow in bf16 NHWC end to end.
USE_TMA_DGRAD = False
PRECISION = "fp32"  # "fp32": FFMA reference-precision kernels; "bf16": tcgen05 tensor cores (bf16 in, fp32 accumulate)
_PREC = {"fp32": 0, "bf16": 1}
_OPS = {"linear": (0, 1, 2), "conv": (3, 4, 5), "convT": (6, 7, 8)}  # (fwd, dgrad, wgrad) ids of vc_dense_op


def set_precision(p: str) -> None:
    """Select the arithmetic of the dense layers for ops created from now on."""
    global PRECISION
    if p not in _PREC:
        raise VaeclustError(f"precision must be one of {sorted(_PREC)}")
    PRECISION = p


def _dense(op: int, prec: int, in0, in1, bias, out, B, Cin, H, W, Cout, relu) -> None:
    nb = lib().size("vc_dense_op_workspace_bytes", op, prec, B, Cin, H, W, Cout)
    ws = _workspace(nb, out.device)
    lib().call("vc_dense_op", op, prec, ptr(in0), ptr(in1), ptr(bias), ptr(out), B, Cin, H, W, Cout, int(relu), ptr(ws),
               ws.numel(), stream_ptr())


class _DenseFn(torch.autograd.Function):
    """Shared forward/backward of Linear, Conv2d(k3,s2,p1) and ConvTranspose2d(k4,s2,p1) (+bias, +ReLU)."""

    @staticmethod
    def forward(ctx, kind: str, x, weight, bias, relu: bool):
        x, weight = _f32(x, kind + ".x"), _f32(weight, kind + ".weight")
        bias = _f32(bias, kind + ".bias") if bias is not None else None
        prec = _PREC[PRECISION]
        if kind == "linear":
            B, Cin = x.shape
            H = W = 1
            Cout = weight.shape[0]
            if weight.shape[1] != Cin:
                raise VaeclustError(f"linear: x is [*, {Cin}] but weight is {tuple(weight.shape)}")
            yshape = (B, Cout)
        elif kind == "conv":
            B, Cin, H, W = x.shape
            Cout = weight.shape[0]
            if tuple(weight.shape[1:]) != (Cin, 3, 3):
                raise VaeclustError(f"conv2d_s2k3: weight {tuple(weight.shape)} does not match Cin={Cin}, k=3")
            yshape = (B, Cout, (H - 1) // 2 + 1, (W - 1) // 2 + 1)
        else:
            B, Cin, H, W = x.shape
            Cout = weight.shape[1]
            if weight.shape[0] != Cin or tuple(weight.shape[2:]) != (4, 4):
                raise VaeclustError(f"convT2d_s2k4: weight {tuple(weight.shape)} does not match Cin={Cin}, k=4")
            yshape = (B, Cout, 2 * H, 2 * W)
        y = torch.empty(yshape, dtype=torch.float32, device=x.device)
        # wide Linear layers inside a fused training step: read the weight from its bf16 shadow (train.VAEStep keeps it)
        sh = getattr(weight, "_vc_shadow", None) if kind == "linear" and prec == _PREC["bf16"] else None
        if sh is not None and not lib().size("vc_linear_bf16w_supported", 0, B, Cin, Cout):
            sh = None
        if sh is not None:
            _linear_bf16w(0, x, sh, bias, y, B, Cin, Cout, relu)
        else:
            _dense(_OPS[kind][0], prec, x, weight, bias, y, B, Cin, H, W, Cout, relu)
        ctx.shadow = sh
        ctx.kind, ctx.prec, ctx.relu, ctx.has_bias = kind, prec, bool(relu), bias is not None
        # single-GPU training: the optimizer that wants this (wide Linear) weight updated straight from the
        # weight-gradient kernel's accumulator (train.VAEStep sets and clears the attribute around a step)
        ctx.adam = getattr(weight, "_vc_adam", None) if kind == "linear" else None
        ctx.wparam = weight if ctx.adam is not None else None
        ctx.dims = (B, Cin, H, W, Cout)
        ctx.save_for_backward(x, weight, y if relu else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight, y = ctx.saved_tensors
        dy = _f32(dy, ctx.kind + ".dy")
        B, Cin, H, W, Cout = ctx.dims
        _, op_d, op_w = _OPS[ctx.kind]
        s = stream_ptr()
        dx = dw = db = None
        want_db = ctx.has_bias and ctx.needs_input_grad[3]
        inner = 1 if ctx.kind == "linear" else dy.shape[2] * dy.shape[3]
        if (ctx.kind == "linear" and ctx.prec == _PREC["bf16"] and ctx.needs_input_grad[2]
                and lib().size("vc_linear_bwd_small_supported", B, Cin, Cout)):
            # latency-bound layer: ReLU mask, weight / bias / data gradients in one launch
            dw = torch.empty_like(weight)
            db = torch.empty(Cout, dtype=torch.float32, device=x.device) if want_db else None
            dx = torch.empty_like(x) if ctx.needs_input_grad[1] else None
            lib().call("vc_linear_bwd_small", ptr(dy), ptr(y) if ctx.relu else None, ptr(x), ptr(weight), ptr(dw),
                       ptr(db), ptr(dx), B, Cin, Cout, s)
            return None, dx, dw, db, None
        if (ctx.kind != "linear" and ctx.prec == _PREC["bf16"] and ctx.relu and want_db and ctx.needs_input_grad[1]
                and USE_TMA_DGRAD and dy.shape[1] % 8 == 0 and dy.shape[1] <= 256 and dy.shape[3] % 64 == 0
                and lib().size("vc_conv_dgrad_nhwc_supported", op_d, B, Cin, H, W, Cout)):
            # g = dy * (y > 0) once, as fp32 NCHW (weight gradient) and bf16 NHWC (TMA-fed data gradient), + db
            g = torch.empty_like(dy)
            gn = torch.empty((dy.shape[0], dy.shape[2], dy.shape[3], dy.shape[1]), dtype=torch.bfloat16, device=dy.device)
            db = torch.empty(Cout, dtype=torch.float32, device=x.device)
            ws = _workspace(max(lib().size("vc_relu_bwd_bias_nhwc_workspace_bytes", Cout),
                                lib().size("vc_conv_dgrad_nhwc_workspace_bytes", op_d, B, Cin, H, W, Cout)), x.device)
            lib().call("vc_relu_bwd_bias_nhwc", ptr(dy), ptr(y), B, Cout, dy.shape[2], dy.shape[3], ptr(g), ptr(gn),
                       ptr(db), ptr(ws), ws.numel(), s)
            dx = torch.empty_like(x)
            lib().call("vc_conv_dgrad_nhwc", op_d, ptr(gn), ptr(weight), ptr(dx), B, Cin, H, W, Cout, ptr(ws), ws.numel(), s)
            if ctx.needs_input_grad[2]:
                dw = torch.empty_like(weight)
                _dense(op_w, ctx.prec, g, x, None, dw, B, Cin, H, W, Cout, False)
            return None, dx, dw, db, None
        if ctx.relu and want_db:  # ReLU mask and bias gradient in one pass over dy / y
            g = torch.empty_like(dy)
            db = torch.empty(Cout, dtype=torch.float32, device=x.device)
            ws = _workspace(lib().size("vc_bias_grad_workspace_bytes", Cout), x.device)
            lib().call("vc_relu_bwd_bias", ptr(dy), ptr(y), B, Cout, inner, ptr(g), ptr(db), ptr(ws), ws.numel(), s)
            want_db = False
        elif ctx.relu:
            g = torch.empty_like(dy)
            lib().call("vc_relu_bwd", ptr(dy), ptr(y), ptr(g), dy.numel(), s)
        else:
            g = dy
        if ctx.needs_input_grad[1]:
            dx = torch.empty_like(x)
            if ctx.shadow is not None and lib().size("vc_linear_bf16w_supported", 1, B, Cin, Cout):
                _linear_bf16w(1, g, ctx.shadow, None, dx, B, Cin, Cout, False)
            else:
                _dense(op_d, ctx.prec, g, weight, None, dx, B, Cin, H, W, Cout, False)
        if ctx.needs_input_grad[2]:
            if (ctx.adam is not None and ctx.prec == _PREC["bf16"]
                    and lib().size("vc_linear_wgrad_adam_supported", B, Cin, Cout)):
                _linear_wgrad_adam(ctx.adam, ctx.wparam, g, x, B, Cin, Cout, ctx.shadow)   # no dW: the update happens here
            else:
                dw = torch.empty_like(weight)
                _dense(op_w, ctx.prec, g, x, None, dw, B, Cin, H, W, Cout, False)
        if want_db:
            db = torch.empty(Cout, dtype=torch.float32, device=x.device)
            ws = _workspace(lib().size("vc_bias_grad_workspace_bytes", Cout), x.device)
            lib().call("vc_bias_grad", ptr(g), B, Cout, inner, ptr(db), 0, ptr(ws), ws.numel(), s)
        return None, dx, dw, db, None


def _linear_bf16w(op, inp, shadow, bias, out, B, Cin, Cout, relu) -> None:
    ws = _workspace(lib().size("vc_linear_bf16w_workspace_bytes", op, B, Cin, Cout), out.device)
    lib().call("vc_linear_bf16w", op, ptr(inp), ptr(shadow), ptr(bias), ptr(out), B, Cin, Cout, int(relu), ptr(ws), ws.numel(),
               stream_ptr())


def _linear_wgrad_adam(opt, p, g, x, B, Cin, Cout, shadow=None) -> None:
    """dW = g^T x formed in tensor memory and applied to p / exp_avg / exp_avg_sq by the same kernel
    (vc_linear_wgrad_adam); the optimizer's later step() skips p because its .grad stays None."""
    st = opt._state(p)
    b1, b2 = opt.betas
    if opt.capturable:
        if not opt._ticked:
            raise VaeclustError("fused Linear weight update: Adam.begin_step() must run before backward")
        step_ptr, bc1, bc2 = ptr(opt.device_step()), 1.0, 1.0
    else:
        st["step"] += 1
        step_ptr, bc1, bc2 = None, 1.0 - b1 ** st["step"], 1.0 - b2 ** st["step"]
    ws = _workspace(lib().size("vc_linear_wgrad_adam_workspace_bytes", B, Cin, Cout), x.device)
    lib().call("vc_linear_wgrad_adam", ptr(g), ptr(x), ptr(p.data), ptr(st["exp_avg"]), ptr(st["exp_avg_sq"]), ptr(shadow), B,
               Cin, Cout,
               float(opt.param_groups[0]["lr"]), b1, b2, opt.eps, bc1, bc2, step_ptr, ptr(ws), ws.numel(), stream_ptr())


def linear(x, weight, bias=None, relu: bool = False):
    """y = act(x W^T + b): nn.Linear(+ReLU) of 12:98-103,114-117,129-132,159-164; 19:79-95; 06:149-157."""
    g = getattr(weight, "_vc_gather", None)
    if g is not None:      # sharded data parallelism: the other ranks' slices of this weight are still arriving
        g.wait()
        weight._vc_gather = None
    return _DenseFn.apply("linear", x, weight, bias, relu)


def conv2d_s2k3(x, weight, bias=None, relu: bool = False):
    """nn.Conv2d(Cin, Cout, 3, stride=2, padding=1) (+ReLU), 12:87-89."""
    return _DenseFn.apply("conv", x, weight, bias, relu)


def convT2d_s2k4(x, weight, bias=None, relu: bool = False):
    """nn.ConvTranspose2d(Cin, Cout, 4, stride=2, padding=1) (+ReLU), 12:135-137.  weight is [Cin, Cout, 4, 4]."""
    return _DenseFn.apply("convT", x, weight, bias, relu)


class CatGateFn(torch.autograd.Function):
    """[a, l*gate, gate] (12:178-180) or [a, l] (19:97-115) along dim 1."""

    @staticmethod
    def forward(ctx, a, l, gate, append_gate: bool):
        a, l = _f32(a, "cat.a"), _f32(l, "cat.l")
        gate = _f32(gate, "cat.gate") if gate is not None else None
        B, na, nl = a.shape[0], a.shape[1], l.shape[1]
        if l.shape[0] != B or (gate is not None and gate.numel() != B):
            raise VaeclustError("cat_gate: batch sizes differ")
        out = torch.empty((B, na + nl + (1 if append_gate else 0)), dtype=torch.float32, device=a.device)
        lib().call("vc_cat_gate_fwd", ptr(a), na, ptr(l), nl, ptr(gate), int(append_gate), B, ptr(out), stream_ptr())
        ctx.dims = (B, na, nl, bool(append_gate))
        ctx.save_for_backward(gate)
        return out

    @staticmethod
    def backward(ctx, dout):
        (gate,) = ctx.saved_tensors
        B, na, nl, app = ctx.dims
        dout = _f32(dout, "cat.dout")
        da = torch.empty((B, na), dtype=torch.float32, device=dout.device) if ctx.needs_input_grad[0] else None
        dl = torch.empty((B, nl), dtype=torch.float32, device=dout.device) if ctx.needs_input_grad[1] else None
        lib().call("vc_cat_gate_bwd", ptr(dout), na, nl, ptr(gate), int(app), B, ptr(da), ptr(dl), stream_ptr())
        return da, dl, None, None


def cat_gate(a, l, gate=None, append_gate: bool = False):
    return CatGateFn.apply(a, l, gate, append_gate)


# --------------------------------------------------------------------------------------------------- VAE loss
class ReparamKLFn(torch.autograd.Function):
    """(z, kl_rows) = (mu + eps*exp(lv/2), -0.5*sum_j(1+lv-mu^2-e^lv)); backward fused into one launch."""

    @staticmethod
    def forward(ctx, mu, lv, eps):
        mu, lv, eps = _f32(mu, "reparam.mu"), _f32(lv, "reparam.logvar"), _f32(eps, "reparam.eps")
        B, z = mu.shape
        zo = torch.empty_like(mu)
        kl = torch.empty(B, dtype=torch.float32, device=mu.device)
        lib().call("vc_reparam_kl_fwd", ptr(mu), ptr(lv), ptr(eps), B, z, ptr(zo), ptr(kl), stream_ptr())
        ctx.save_for_backward(mu, lv, eps)
        return zo, kl

    @staticmethod
    def backward(ctx, dz, dkl):
        mu, lv, eps = ctx.saved_tensors
        B, z = mu.shape
        dz = _f32(dz, "reparam.dz") if dz is not None else None
        dkl = _f32(dkl, "reparam.dkl") if dkl is not None else None
        dmu, dlv = torch.empty_like(mu), torch.empty_like(lv)
        lib().call("vc_reparam_kl_bwd", ptr(dz), ptr(dkl), ptr(mu), ptr(lv), ptr(eps), B, z, ptr(dmu), ptr(dlv),
                   stream_ptr())
        return dmu, dlv, None


def reparam_kl(mu, lv, eps):
    return ReparamKLFn.apply(mu, lv, eps)


class KLRowsFn(torch.autograd.Function):
    """kl_rows only (the reference's stand-alone kl_div / kl_divergence, 12:189-190, 19:124-126)."""

    @staticmethod
    def forward(ctx, mu, lv):
        mu, lv = _f32(mu, "kl.mu"), _f32(lv, "kl.logvar")
        B, z = mu.shape
        kl = torch.empty(B, dtype=torch.float32, device=mu.device)
        lib().call("vc_reparam_kl_fwd", ptr(mu), ptr(lv), None, B, z, None, ptr(kl), stream_ptr())
        ctx.save_for_backward(mu, lv)
        return kl

    @staticmethod
    def backward(ctx, dkl):
        mu, lv = ctx.saved_tensors
        B, z = mu.shape
        dkl = _f32(dkl, "kl.dkl")
        dmu, dlv = torch.empty_like(mu), torch.empty_like(lv)
        lib().call("vc_reparam_kl_bwd", None, ptr(dkl), ptr(mu), ptr(lv), None, B, z, ptr(dmu), ptr(dlv),
                   stream_ptr())
        return dmu, dlv


def kl_rows(mu, lv):
    return KLRowsFn.apply(mu, lv)


class SqErrSumFn(torch.autograd.Function):
    """sum (xhat[..., :H, :W] - x)^2 as a 0-d fp32 tensor; d/dxhat is produced in the same pass (crop of 12:260)."""

    @staticmethod
    def forward(ctx, xhat, x):
        xhat, x = _f32(xhat, "mse.xhat"), _f32(x, "mse.x")
        if xhat.dim() == 2:
            outer, Hh, Wh, H, W = xhat.shape[0], 1, xhat.shape[1], 1, x.shape[1]
            if x.shape[0] != outer:
                raise VaeclustError("sq_err_sum: batch sizes differ")
        else:
            if xhat.shape[:-2] != x.shape[:-2]:
                raise VaeclustError(f"sq_err_sum: leading dims differ {tuple(xhat.shape)} vs {tuple(x.shape)}")
            outer = 1
            for s_ in xhat.shape[:-2]:
                outer *= s_
            Hh, Wh, H, W = xhat.shape[-2], xhat.shape[-1], x.shape[-2], x.shape[-1]
        need_grad = ctx.needs_input_grad[0]
        dx = torch.empty_like(xhat) if need_grad else None
        sq = torch.empty((), dtype=torch.float64, device=xhat.device)
        nb = lib().size("vc_recon_loss_workspace_bytes")
        ws = _workspace(nb, xhat.device)
        lib().call("vc_recon_loss_fwd_bwd", ptr(xhat), ptr(x), outer, Hh, Wh, H, W, 1.0, ptr(dx), ptr(sq), ptr(ws),
                   ws.numel(), stream_ptr())
        ctx.save_for_backward(dx)
        return sq.to(torch.float32)

    @staticmethod
    def backward(ctx, g):
        (dx,) = ctx.saved_tensors
        return dx * g, None


def sq_err_sum(xhat, x):
    return SqErrSumFn.apply(xhat, x)


def mse_loss(xhat, x, reduction: str = "mean"):
    """Drop-in for nn.MSELoss()/F.mse_loss as used at 12:234,262 and 06:184 (crop-aware)."""
    s = sq_err_sum(xhat, x)
    if reduction == "mean":
        return s / x.numel()
    if reduction == "sum":
        return s
    raise VaeclustError("mse_loss: reduction must be 'mean' or 'sum'")


class VAELossFn(torch.autograd.Function):
    """Fused training loss: crop (12:260) + squared error + its gradient in ONE pass over xhat/x, then
    loss = recon + beta*kl under the script's reduction convention (VC_LOSS_MEAN_MEAN: 12:262-264, 06:182-188;
    VC_LOSS_SUM_MEAN: 19:226-228).  Returns (loss 0-d, stats [3] = loss, recon, kl); stats carries no gradient.
    `unit_grad=True` is the caller's promise that it will call loss.backward() directly (upstream gradient 1), which
    lets backward hand out the pre-scaled dxhat without another pass over it."""

    @staticmethod
    def forward(ctx, xhat, x, kl_rows_t, beta: float, mode: int, z_dim: int, unit_grad: bool):
        xhat, x, kl_rows_t = _f32(xhat, "vae_loss.xhat"), _f32(x, "vae_loss.x"), _f32(kl_rows_t, "vae_loss.kl_rows")
        B = kl_rows_t.numel()
        if xhat.dim() == 2:
            outer, Hh, Wh, H, W = xhat.shape[0], 1, xhat.shape[1], 1, x.shape[1]
        else:
            outer = 1
            for s_ in xhat.shape[:-2]:
                outer *= s_
            Hh, Wh, H, W = xhat.shape[-2], xhat.shape[-1], x.shape[-2], x.shape[-1]
        n = x.numel()
        gscale = 1.0 / n if mode == VC_LOSS_MEAN_MEAN else 1.0 / B
        dx = torch.empty_like(xhat)
        sq = torch.empty((), dtype=torch.float64, device=xhat.device)
        ws = _workspace(lib().size("vc_recon_loss_workspace_bytes"), xhat.device)
        s = stream_ptr()
        lib().call("vc_recon_loss_fwd_bwd", ptr(xhat), ptr(x), outer, Hh, Wh, H, W, gscale, ptr(dx), ptr(sq), ptr(ws),
                   ws.numel(), s)
        stats = torch.empty(3, dtype=torch.float32, device=xhat.device)
        lib().call("vc_loss_finalize", ptr(sq), ptr(kl_rows_t), B, z_dim, n, float(beta), mode, ptr(stats), s)
        ctx.save_for_backward(dx)
        ctx.kl_scale = float(beta) * (1.0 / (B * z_dim) if mode == VC_LOSS_MEAN_MEAN else 1.0 / B)
        ctx.B, ctx.unit_grad = B, bool(unit_grad)
        ctx.mark_non_differentiable(stats)
        return stats[0].clone(), stats

    @staticmethod
    def backward(ctx, g, _gstats):
        (dx,) = ctx.saved_tensors
        if ctx.unit_grad:
            dkl = torch.full((ctx.B,), ctx.kl_scale, dtype=torch.float32, device=dx.device)
            return dx, None, dkl, None, None, None, None
        return dx * g, None, (g * ctx.kl_scale).expand(ctx.B).contiguous(), None, None, None, None


def vae_loss(xhat, x, kl_rows_t, beta: float, mode: int, z_dim: int, unit_grad: bool = False):
    return VAELossFn.apply(xhat, x, kl_rows_t, beta, mode, z_dim, unit_grad)


__all__ = ["set_precision", "linear", "conv2d_s2k3", "convT2d_s2k4", "cat_gate", "reparam_kl", "kl_rows", "sq_err_sum", "mse_loss", "vae_loss",
           "VC_LOSS_MEAN_MEAN", "VC_LOSS_SUM_MEAN"]
