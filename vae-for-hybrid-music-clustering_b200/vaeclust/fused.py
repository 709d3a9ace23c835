"""The fused train step of the MLP VAEs (scripts 19 / 06) and of script 22's AE baseline: `vc_mlpvae_step`.

One object per model instance: it owns the persistent workspace (bf16 weight shadows, activation / gradient rows,
split-K partials, the flat gradient block) and the Adam moments, and replaces zero_grad / forward / loss / backward /
optimizer.step of 19:226-232, 06:322-327 and 22:150-156 by four kernel launches (csrc/mlpvae_step.cu).  The dense
arithmetic is the bf16 tensor-core mode (operands rounded to bf16, fp32 accumulation, fp32 master weights).
Under a process group the flat gradient block is all-reduced between the fold and the Adam launch.
"""
from __future__ import annotations

import ctypes
from typing import Dict, Optional

import torch

from . import dist as vdist
from ._lib import VC_LOSS_MEAN_MEAN, VC_LOSS_SUM_MEAN, VaeclustError, lib, ptr, stream_ptr
from .optim import Adam


class _Desc(ctypes.Structure):
    _fields_ = [("input_dim", ctypes.c_int32), ("cond_dim", ctypes.c_int32), ("hidden", ctypes.c_int32),
                ("latent", ctypes.c_int32), ("variational", ctypes.c_int32), ("loss_mode", ctypes.c_int32),
                ("beta", ctypes.c_float), ("pad", ctypes.c_int32)]


class _Tensor(ctypes.Structure):
    _fields_ = [("p", ctypes.c_void_p), ("m", ctypes.c_void_p), ("v", ctypes.c_void_p)]


class _Params(ctypes.Structure):
    _fields_ = [("W", _Tensor * 7), ("b", _Tensor * 7)]


def _layers(model, kind: str):
    """The seven Linear modules in VcMlpVaeParams order (None for the AE's missing logvar head)."""
    if kind == "mlp19":      # 19:79-95
        return [model.enc[0], model.enc[2], model.mu, model.logvar, model.dec[0], model.dec[2], model.dec[4]]
    if kind == "mlp06":      # 06:149-157
        return [model.enc1, model.enc2, model.mu, model.logvar, model.dec1, model.dec2, model.out]
    if kind == "ae22":       # 22:69-82
        return [model.enc[0], model.enc[2], model.enc[4], None, model.dec[0], model.dec[2], model.dec[4]]
    raise VaeclustError(f"FusedMLPStep: unknown model kind {kind!r}")


def describe(model, kind: str, beta: float) -> _Desc:
    L = _layers(model, kind)
    hidden, enc_in = L[0].weight.shape
    input_dim = L[6].weight.shape[0]
    latent = L[2].weight.shape[0]
    d = _Desc()
    d.input_dim, d.cond_dim, d.hidden, d.latent = input_dim, enc_in - input_dim, hidden, latent
    d.variational = 0 if kind == "ae22" else 1
    d.loss_mode = VC_LOSS_SUM_MEAN if kind == "mlp19" else VC_LOSS_MEAN_MEAN
    d.beta = float(beta)
    ok = (L[1].weight.shape == (hidden, hidden) and L[4].weight.shape == (hidden, latent + d.cond_dim)
          and L[5].weight.shape == (hidden, hidden) and L[6].weight.shape == (input_dim, hidden)
          and (L[3] is None or L[3].weight.shape == (latent, hidden)) and d.cond_dim >= 0)
    if not ok:
        d.hidden = -1    # not the topology the fused kernel implements
    return d


def supported(model, kind: str) -> bool:
    if kind not in ("mlp19", "mlp06", "ae22"):
        return False
    d = describe(model, kind, 1.0)
    return bool(lib().size("vc_mlpvae_supported", ctypes.addressof(d)))


class FusedMLPStep:
    """step(x[, c], eps=None, global_batch=None) -> device tensor [loss, recon, kl] of this rank's batch."""

    def __init__(self, model, kind: str, lr: float, beta: float, group=None, capturable: bool = False,
                 betas=(0.9, 0.999), adam_eps: float = 1e-8, max_batch: int = 256):
        self.model, self.kind, self.group = model, kind, group
        self.lr, self.betas, self.adam_eps = float(lr), (float(betas[0]), float(betas[1])), float(adam_eps)
        self.desc = describe(model, kind, beta)
        if not lib().size("vc_mlpvae_supported", ctypes.addressof(self.desc)):
            raise VaeclustError("FusedMLPStep: model shape not supported by vc_mlpvae_step (hidden 256, latent in "
                                "{8,16,32}, input_dim + cond_dim <= 512, latent + cond_dim <= 64)")
        self.layers = _layers(model, kind)
        self.multi = group is not None and vdist.world_size(group) > 1
        self.capturable = bool(capturable)
        # the Adam object keeps the moments (state_dict compatible with the unfused path); the update itself runs in C
        self.opt = Adam(model.parameters(), lr=lr, betas=betas, eps=adam_eps, capturable=capturable)
        self.params = _Params()
        self._names: Dict[int, str] = {id(p): n for n, p in model.named_parameters()}
        self._tensors = []
        for i, L in enumerate(self.layers):
            for field, t in (("W", None if L is None else L.weight), ("b", None if L is None else L.bias)):
                slot = getattr(self.params, field)[i]
                if t is None:
                    slot.p = slot.m = slot.v = None
                    continue
                if not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous():
                    raise VaeclustError("FusedMLPStep: parameters must be contiguous float32 CUDA tensors")
                st = self.opt._state(t)
                slot.p, slot.m, slot.v = t.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
                self._tensors.append(t)
        self.ws: Optional[torch.Tensor] = None
        self.ws_batch = 0
        self._versions = None
        self._t = 0
        self._need(max_batch)

    # ------------------------------------------------------------------------------------------------ workspace
    def _need(self, B: int) -> None:
        if self.ws is not None and B <= self.ws_batch:
            return
        nb = lib().size("vc_mlpvae_workspace_bytes", ctypes.addressof(self.desc), int(B))
        self.ws = torch.empty(nb, dtype=torch.uint8, device=self._tensors[0].device)
        self.ws_batch = int(B)
        self._versions = None      # the shadows live in the workspace: rebuild them

    def _sync_shadows(self, B: int) -> None:
        ver = tuple(t._version for t in self._tensors) + tuple(t.data_ptr() for t in self._tensors)
        if ver == self._versions:
            return
        lib().call("vc_mlpvae_prepare", ctypes.addressof(self.desc), ctypes.addressof(self.params), int(B), ptr(self.ws),
                   self.ws.numel(), stream_ptr())
        self._versions = ver

    def _flat(self, B: int):
        """(flat fp32 gradient block as a tensor view, layout int64[20]) for the plan of batch size B."""
        lay = (ctypes.c_int64 * 20)()
        lib().call("vc_mlpvae_grad_layout", ctypes.addressof(self.desc), int(B), ctypes.addressof(lay))
        base = (self.ws.data_ptr() + 1023) // 1024 * 1024 - self.ws.data_ptr()
        blk = self.ws[base + lay[0]: base + lay[0] + 4 * lay[1]].view(torch.float32)
        return blk, lay

    # ----------------------------------------------------------------------------------------------------- step
    def __call__(self, x, *extra, eps: Optional[torch.Tensor] = None, global_batch: Optional[int] = None,
                 local_rows: Optional[int] = None):
        d = self.desc
        B = int(x.shape[0])
        c = extra[0] if extra and extra[0] is not None else None
        if d.cond_dim > 0 and c is None:
            raise ValueError("Conditional model requires conditioning vector c.")
        for name, t, w in (("x", x, d.input_dim), ("c", c, d.cond_dim)):
            if t is None:
                continue
            if not t.is_cuda or t.dtype != torch.float32 or t.shape != (B, w):
                raise VaeclustError(f"FusedMLPStep: {name} must be a float32 CUDA tensor of shape ({B}, {w})")
        x = x.contiguous()
        c = c.contiguous() if c is not None else None
        if d.variational:
            if eps is None:
                eps = torch.randn(B, d.latent, device=x.device, dtype=torch.float32)
            eps = eps.contiguous()
        self._need(B)
        self._sync_shadows(B)
        out3 = torch.empty(3, dtype=torch.float32, device=x.device)
        b1, b2 = self.betas
        dev_step = ptr(self.opt.device_step()) if self.capturable else None   # created at the CURRENT step count
        self._t += 1
        for t in self._tensors:
            self.opt.state[id(t)]["step"] = self._t
        bc1, bc2 = (1.0, 1.0) if self.capturable else (1.0 - b1 ** self._t, 1.0 - b2 ** self._t)
        rows = B if local_rows is None else int(local_rows)
        weight = 1.0 if not self.multi else rows / float(global_batch or B * vdist.world_size(self.group))

        def call(phases):
            lib().call("vc_mlpvae_step", ctypes.addressof(d), ctypes.addressof(self.params), ptr(x), ptr(c), ptr(eps), B,
                       float(weight), phases, self.lr, b1, b2, self.adam_eps, bc1, bc2, dev_step, ptr(out3),
                       ptr(self.ws), self.ws.numel(), stream_ptr())

        if self.multi:
            call(3)
            blk, _ = self._flat(B)
            vdist.all_reduce_sum(blk, self.group)
            call(4)
        else:
            call(7)
        self._last_B = B
        return out3   # (the in-place update goes through raw pointers: torch's version counters do not move)

    def last_grads(self) -> Dict[str, torch.Tensor]:
        """Per-parameter gradients of the last step, copied out of the flat block (parity checks only)."""
        blk, lay = self._flat(self._last_B)
        out: Dict[str, torch.Tensor] = {n: None for n in self._names.values()}
        Z = self.desc.latent
        for l, idx in enumerate((0, 1, 2, 4, 5, 6)):
            woff, pitch, boff = lay[2 + 3 * l], lay[3 + 3 * l], lay[4 + 3 * l]
            mods = [(self.layers[idx], 0)]
            if idx == 2 and self.layers[3] is not None:
                mods.append((self.layers[3], Z))
            for mod, r0 in mods:
                o, i = mod.weight.shape
                rows = blk[woff + r0 * pitch: woff + (r0 + o) * pitch].view(o, pitch)
                out[self._names[id(mod.weight)]] = rows[:, :i].clone()
                out[self._names[id(mod.bias)]] = blk[boff + r0: boff + r0 + o].clone()
        return out
