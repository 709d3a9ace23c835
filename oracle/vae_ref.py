"""TEST INFRASTRUCTURE ONLY -- CPU restatement (plain torch fp32, functional form) of the
reference's VAE hot path.  Not imported by the product package; only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg use it.

Parity status: PINNED against the reference's own classes run in the build container
(oracle/make_golden.py imports /root/reference/scripts/{06,12,19}_*.py, loads one
state_dict into both, injects the same eps and asserts equality; the outputs are
committed under tests/golden/).  The reference itself ships no tests or golden vectors
(SURVEY.md section 4), so that import is the only pin available.

All functions take a ``params`` dict keyed exactly like the reference ``state_dict``.
Citations are into /root/reference/scripts/.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor


# ----------------------------------------------------------------------------- shapes
def conv_out(n: int) -> int:
    """Conv2d(k3, s2, p1) output length (12:87-89)."""
    return (n + 2 - 3) // 2 + 1


def conv_vae_shapes(feat_hw: Tuple[int, int]):
    """h_shape discovered by the dummy forward at 12:92-96."""
    f, t = feat_hw
    for _ in range(3):
        f, t = conv_out(f), conv_out(t)
    return (128, f, t)


# ------------------------------------------------------------------- parameter init
def _linear_init(g: torch.Generator, out_f: int, in_f: int):
    # nn.Linear.reset_parameters: kaiming_uniform_(a=sqrt(5)) -> U(+-1/sqrt(fan_in)) for both
    b = 1.0 / math.sqrt(in_f)
    w = (torch.rand(out_f, in_f, generator=g) * 2 - 1) * b
    bias = (torch.rand(out_f, generator=g) * 2 - 1) * b
    return w, bias


def init_conv_vae_params(feat_hw, z_dim=32, lyr_dim=384, in_ch=1, seed=0) -> Dict[str, Tensor]:
    """Random parameters with the reference's key names / shapes (12:83-165).
    (Distribution matches torch defaults; the RNG stream does not -- parity tests load
    identical tensors into both sides instead of relying on seeds.)"""
    g = torch.Generator().manual_seed(seed)
    C, H, W = conv_vae_shapes(feat_hw)
    flat = C * H * W
    p: Dict[str, Tensor] = {}

    def conv(name, co, ci, k, transposed=False):
        fan_in = (co if transposed else ci) * k * k  # ConvTranspose2d fan_in quirk: weight.size(1)*k*k
        b = 1.0 / math.sqrt(fan_in)
        shape = (ci, co, k, k) if transposed else (co, ci, k, k)
        p[name + ".weight"] = (torch.rand(*shape, generator=g) * 2 - 1) * b
        p[name + ".bias"] = (torch.rand(co, generator=g) * 2 - 1) * b

    def lin(name, out_f, in_f):
        p[name + ".weight"], p[name + ".bias"] = _linear_init(g, out_f, in_f)

    conv("audio_enc.conv.0", 32, in_ch, 3)
    conv("audio_enc.conv.2", 64, 32, 3)
    conv("audio_enc.conv.4", 128, 64, 3)
    lin("audio_enc.fc.0", 256, flat)
    lin("audio_enc.mu", z_dim, 256)
    lin("audio_enc.logvar", z_dim, 256)
    lin("lyr_proj.net.0", 256, lyr_dim)
    lin("lyr_proj.net.2", 128, 256)
    lin("fuse.0", 256, z_dim + 128 + 1)
    lin("mu", z_dim, 256)
    lin("logvar", z_dim, 256)
    lin("audio_dec.fc.0", 256, z_dim)
    lin("audio_dec.fc.2", flat, 256)
    conv("audio_dec.deconv.0", 64, 128, 4, transposed=True)
    conv("audio_dec.deconv.2", 32, 64, 4, transposed=True)
    conv("audio_dec.deconv.4", in_ch, 32, 4, transposed=True)
    return p


def init_mlp_vae_params(input_dim, latent_dim=16, hidden_dim=256, cond_dim=0, seed=0,
                        style="19") -> Dict[str, Tensor]:
    """Keys of script 19's MLPVAE (19:79-95) or script 06's (06:149-157)."""
    g = torch.Generator().manual_seed(seed)
    names19 = ["enc.0", "enc.2", "mu", "logvar", "dec.0", "dec.2", "dec.4"]
    names06 = ["enc1", "enc2", "mu", "logvar", "dec1", "dec2", "out"]
    names = names19 if style == "19" else names06
    dims = [(hidden_dim, input_dim + cond_dim), (hidden_dim, hidden_dim), (latent_dim, hidden_dim),
            (latent_dim, hidden_dim), (hidden_dim, latent_dim + cond_dim), (hidden_dim, hidden_dim),
            (input_dim, hidden_dim)]
    p = {}
    for n, (o, i) in zip(names, dims):
        p[n + ".weight"], p[n + ".bias"] = _linear_init(g, o, i)
    return p


# ------------------------------------------------- emulation of the bf16 tensor-core mode of the product
# The product's bf16 mode rounds BOTH operands of every tensor-core contraction to bf16 (8 mantissa bits) and
# accumulates in fp32: activations and weights in the forward and data-gradient ops, the incoming gradient and the saved
# input in the weight-gradient op.  The single-channel ends of script 12 (Conv2d 1->32, ConvTranspose2d 32->1) run on
# FFMA kernels that round their operands the same way.  With `emulate_bf16(True)` the functions below round at exactly those points
# (straight-through in the other direction), so a CUDA step can be checked against an oracle that computes in the SAME
# arithmetic with a tight tolerance; what remains is fp32 summation order.  Default: plain fp32 (the reference).
_EMU_BF16 = False


def emulate_bf16(on: bool) -> None:
    global _EMU_BF16
    _EMU_BF16 = bool(on)


class _RoundFwd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, t):
        return t.to(torch.bfloat16).to(torch.float32)

    @staticmethod
    def backward(ctx, g):
        return g


class _RoundBwd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, t):
        return t.view_as(t)

    @staticmethod
    def backward(ctx, g):
        return g.to(torch.bfloat16).to(torch.float32)


def _tc(op, x, w, b, rounded=True, **kw):
    """One tensor-core layer op(x, w, b): fp32, or operands rounded to bf16 in forward and backward."""
    if not (_EMU_BF16 and rounded):
        return op(x, w, b, **kw)
    # bias joins after the contraction: the gradient is rounded where the dgrad / wgrad kernels consume it
    y = _RoundBwd.apply(op(_RoundFwd.apply(x), _RoundFwd.apply(w), None, **kw))
    if b is None:
        return y
    return y + (b.view(1, -1, 1, 1) if y.dim() == 4 else b)


# ------------------------------------------------------------------------- forward
def reparam(mu: Tensor, logvar: Tensor, eps: Tensor) -> Tensor:
    """12:168-172 / 19:105-108 / 06:164-168 with eps injected instead of randn_like."""
    return mu + eps * torch.exp(0.5 * logvar)


def conv_vae_forward(p: Dict[str, Tensor], x: Tensor, lyr: Tensor, m: Tensor, eps: Tensor):
    """ConvMultiModalVAE.forward (12:174-186) -> (xhat_uncropped, mu, lv)."""
    lin = lambda t, k: _tc(F.linear, t, p[k + ".weight"], p[k + ".bias"])
    h = x
    for i in (0, 2, 4):  # AudioEncoder.conv 12:86-90
        h = F.relu(_tc(F.conv2d, h, p[f"audio_enc.conv.{i}.weight"], p[f"audio_enc.conv.{i}.bias"], stride=2, padding=1))
    hshape = h.shape[1:]
    h = F.relu(lin(h.flatten(1), "audio_enc.fc.0"))  # 12:98-101,106-107
    mu_a = lin(h, "audio_enc.mu")  # lv_a unused (12:175)
    l = F.relu(lin(lyr, "lyr_proj.net.0"))  # 12:114-117
    l = F.relu(lin(l, "lyr_proj.net.2"))
    l = l * m  # 12:178
    hf = F.relu(lin(torch.cat([mu_a, l, m], dim=1), "fuse.0"))  # 12:180
    mu = lin(hf, "mu")
    lv = lin(hf, "logvar")
    z = reparam(mu, lv, eps)
    d = F.relu(lin(z, "audio_dec.fc.0"))  # 12:129-132
    d = F.relu(lin(d, "audio_dec.fc.2"))
    d = d.view(z.size(0), *hshape)
    for i in (0, 2, 4):  # AudioDecoder.deconv 12:134-138 (no activation after the last one)
        d = _tc(F.conv_transpose2d, d, p[f"audio_dec.deconv.{i}.weight"], p[f"audio_dec.deconv.{i}.bias"], stride=2,
                padding=1)
        if i != 4:
            d = F.relu(d)
    return d, mu, lv


def conv_vae_loss(xhat: Tensor, x: Tensor, mu: Tensor, lv: Tensor, beta: float):
    """crop 12:260, MSELoss(mean) 12:262, kl_div 12:189-190, total 12:264."""
    xhat = xhat[:, :, : x.shape[2], : x.shape[3]]
    recon = torch.mean((xhat - x) ** 2)
    kl = -0.5 * torch.mean(1 + lv - mu.pow(2) - lv.exp())
    return recon + beta * kl, recon, kl


def mlp_vae_forward(p: Dict[str, Tensor], x: Tensor, eps: Tensor, c: Optional[Tensor] = None, style="19"):
    """MLPVAE.forward 19:117-121 (cond concatenated on both sides, 19:97-115) / 06:175-179."""
    n = (["enc.0", "enc.2", "mu", "logvar", "dec.0", "dec.2", "dec.4"] if style == "19"
         else ["enc1", "enc2", "mu", "logvar", "dec1", "dec2", "out"])
    lin = lambda t, k: _tc(F.linear, t, p[k + ".weight"], p[k + ".bias"])
    xi = x if c is None else torch.cat([x, c], dim=1)
    h = F.relu(lin(F.relu(lin(xi, n[0])), n[1]))
    mu, lv = lin(h, n[2]), lin(h, n[3])
    z = reparam(mu, lv, eps)
    zi = z if c is None else torch.cat([z, c], dim=1)
    xhat = lin(F.relu(lin(F.relu(lin(zi, n[4])), n[5])), n[6])
    return xhat, mu, lv


def init_ae_params(input_dim, z_dim=16, hidden=256, seed=0) -> Dict[str, Tensor]:
    """Keys of script 22's AE (22:66-82): enc.{0,2,4}, dec.{0,2,4}."""
    g = torch.Generator().manual_seed(seed)
    dims = {"enc.0": (hidden, input_dim), "enc.2": (hidden, hidden), "enc.4": (z_dim, hidden),
            "dec.0": (hidden, z_dim), "dec.2": (hidden, hidden), "dec.4": (input_dim, hidden)}
    p = {}
    for n, (o, i) in dims.items():
        p[n + ".weight"], p[n + ".bias"] = _linear_init(g, o, i)
    return p


def ae_forward(p: Dict[str, Tensor], x: Tensor):
    """AE.forward 22:84-87 -> (x_hat, z)."""
    lin = lambda t, k: _tc(F.linear, t, p[k + ".weight"], p[k + ".bias"])
    z = lin(F.relu(lin(F.relu(lin(x, "enc.0")), "enc.2")), "enc.4")
    return lin(F.relu(lin(F.relu(lin(z, "dec.0")), "dec.2")), "dec.4"), z


def mlp_vae_loss_sum(xhat, x, mu, lv, beta):
    """Script 19 convention 19:226-228 + 19:124-126: feature-sum, batch-mean."""
    rec = ((xhat - x) ** 2).sum(dim=1)
    kl = -0.5 * torch.sum(1 + lv - mu.pow(2) - lv.exp(), dim=1)
    return (rec + beta * kl).mean(), rec.mean(), kl.mean()


def mlp_vae_loss_mean(xhat, x, mu, lv, beta):
    """Script 06 convention vae_loss 06:182-188: mean / mean."""
    recon = torch.mean((xhat - x) ** 2)
    kl = -0.5 * torch.mean(1.0 + lv - mu.pow(2) - lv.exp())
    return recon + beta * kl, recon, kl


# ----------------------------------------------------------------------- optimiser
class AdamRef:
    """torch.optim.Adam defaults (12:233, 19:202, 06:310): betas (0.9,0.999), eps 1e-8, no wd.
    Params whose grad is None are skipped (audio_enc.logvar, 12:175)."""

    def __init__(self, params: Dict[str, Tensor], lr: float, b1=0.9, b2=0.999, eps=1e-8):
        self.p, self.lr, self.b1, self.b2, self.eps = params, lr, b1, b2, eps
        self.m = {k: torch.zeros_like(v) for k, v in params.items()}
        self.v = {k: torch.zeros_like(v) for k, v in params.items()}
        self.t = {k: 0 for k in params}

    @torch.no_grad()
    def step(self, grads: Dict[str, Optional[Tensor]]):
        for k, g in grads.items():
            if g is None:
                continue
            self.t[k] += 1
            t = self.t[k]
            self.m[k].mul_(self.b1).add_(g, alpha=1 - self.b1)
            self.v[k].mul_(self.b2).addcmul_(g, g, value=1 - self.b2)
            bc1, bc2 = 1 - self.b1 ** t, 1 - self.b2 ** t
            denom = (self.v[k].sqrt() / math.sqrt(bc2)).add_(self.eps)
            self.p[k].addcdiv_(self.m[k], denom, value=-self.lr / bc1)


def train_step(p: Dict[str, Tensor], opt: AdamRef, kind: str, batch: dict, beta: float):
    """One reference train step (12:256-267 / 19:226-232 / 06:322-327). kind in {conv, mlp19, mlp06}.
    Returns (loss, recon, kl) python floats and leaves grads in ``opt.last_grads``."""
    leaves = {k: v.detach().requires_grad_(True) for k, v in p.items()}
    if kind == "conv":
        xhat, mu, lv = conv_vae_forward(leaves, batch["x"], batch["lyr"], batch["m"], batch["eps"])
        loss, rec, kl = conv_vae_loss(xhat, batch["x"], mu, lv, beta)
    elif kind == "mlp19":
        xhat, mu, lv = mlp_vae_forward(leaves, batch["x"], batch["eps"], batch.get("c"), "19")
        loss, rec, kl = mlp_vae_loss_sum(xhat, batch["x"], mu, lv, beta)
    elif kind == "ae22":   # 22:150-156: loss = F.mse_loss(x_hat, xb)
        xhat, _z = ae_forward(leaves, batch["x"])
        loss = torch.mean((xhat - batch["x"]) ** 2)
        rec, kl = loss, torch.zeros(())
    else:
        xhat, mu, lv = mlp_vae_forward(leaves, batch["x"], batch["eps"], None, "06")
        loss, rec, kl = mlp_vae_loss_mean(xhat, batch["x"], mu, lv, beta)
    names = list(leaves)
    gs = torch.autograd.grad(loss, [leaves[k] for k in names], allow_unused=True)
    grads = dict(zip(names, gs))
    opt.last_grads = grads
    opt.step(grads)
    return float(loss.detach()), float(rec.detach()), float(kl.detach())


# --------------------------------------------------------------------------- loops
def train_loop(p: Dict[str, Tensor], kind: str, arrays, epochs: int, batch_size: int, lr: float, beta: float,
               eps_fn, seed: int):
    """The reference's epoch loop on CPU with the reference's own batch order: a real
    DataLoader(shuffle=True, drop_last=False, num_workers=0) over the sample indices under
    torch.manual_seed(seed) (12:226-227,246-276; 19:191-192,209-242; 06:300-339).
    arrays: (X, lyr, m) for kind "conv"; (X,) or (X, c) for "mlp19"; (X,) for "mlp06".
    eps_fn(epoch, step, rows) -> [rows, z] noise of that batch (stands in for randn_like).
    Epoch value = unweighted mean of batch values for conv / mlp06 (12:274-276, 06:330-331),
    sample-weighted mean for mlp19 (19:234-242).  Returns {"loss": [...], "recon": [...], "kl": [...]}."""
    from torch.utils.data import DataLoader
    if seed is not None:   # None: continue from the caller's generator state (model construction came first)
        torch.manual_seed(seed)
    opt = AdamRef(p, lr)
    arrays = [torch.as_tensor(a) for a in arrays]
    n = arrays[0].shape[0]
    dl = DataLoader(torch.arange(n), batch_size=batch_size, shuffle=True, drop_last=False, num_workers=0)
    log = {"loss": [], "recon": [], "kl": []}
    for epoch in range(1, epochs + 1):
        tot = [0.0, 0.0, 0.0]
        den = 0.0
        for s_i, idx in enumerate(dl):
            bs = idx.numel()
            b = {"x": arrays[0][idx], "eps": eps_fn(epoch, s_i, bs)}
            if kind == "conv":
                b["lyr"], b["m"] = arrays[1][idx], arrays[2][idx]
            elif kind == "mlp19" and len(arrays) > 1:
                b["c"] = arrays[1][idx]
            vals = train_step(p, opt, kind, b, beta)
            w = float(bs) if kind == "mlp19" else 1.0
            for j in range(3):
                tot[j] += vals[j] * w
            den += w
        for j, k in enumerate(("loss", "recon", "kl")):
            log[k].append(tot[j] / max(den, 1.0))
    return log


@torch.no_grad()
def encode_mu(p: Dict[str, Tensor], kind: str, arrays) -> Tensor:
    """Latent export: mu of every sample in the original order (12:289-309, 19:269-287, 06:354-358)."""
    arrays = [torch.as_tensor(a) for a in arrays]
    x = arrays[0]
    z_dim = p["mu.weight"].shape[0] if "mu.weight" in p else 0
    eps = torch.zeros(x.shape[0], z_dim)
    if kind == "ae22":
        return ae_forward(p, x)[1]
    if kind == "conv":
        _, mu, _ = conv_vae_forward(p, x, arrays[1], arrays[2], eps)
    elif kind == "mlp19":
        _, mu, _ = mlp_vae_forward(p, x, eps, arrays[1] if len(arrays) > 1 else None, "19")
    else:
        _, mu, _ = mlp_vae_forward(p, x, eps, None, "06")
    return mu
