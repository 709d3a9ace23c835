"""Drop-in model classes for the reference's VAE hot path, running on libvaeclust.so kernels.

Same class names, constructor signatures, attribute names, return tuples and state_dict keys as
  scripts/12_train_conv_multimodal_vae_medium.py:83-190   (AudioEncoder, LyricsProjector, AudioDecoder,
                                                            ConvMultiModalVAE, kl_div)
  scripts/19_train_beta_cvae_multimodal_hard.py:64-132     (MLPVAE, kl_divergence, make_onehot)
  scripts/06_train_basic_vae_easy.py:145-188               (MLPVAE -> EasyMLPVAE here / vaeclust.nn06.MLPVAE, vae_loss)
  scripts/22_compare_with_baselines_hard.py:66-87          (AE)
The torch.nn.Linear / Conv2d / ConvTranspose2d submodules are kept ONLY as parameter containers: constructing them
consumes the RNG exactly like the reference (identical initialisation under one seed) and yields the same
state_dict keys, so checkpoints interchange in both directions.  Their own forward() is never called; the layers
are evaluated by the CUDA ops in vaeclust.ops.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import numpy as np
import torch
import torch.nn as nn

from . import convstack, ops


def _conv_out(n: int) -> int:
    return (n + 2 - 3) // 2 + 1


def _lin(layer: nn.Linear, x: torch.Tensor, relu: bool = False) -> torch.Tensor:
    return ops.linear(x, layer.weight, layer.bias, relu)


# ------------------------------------------------------------------------------------------------ script 12
class AudioEncoder(nn.Module):
    """12:83-108.  3 x [Conv2d k3 s2 p1 + ReLU] -> flatten (c,h,w order) -> Linear(flat,256)+ReLU -> mu, logvar."""

    def __init__(self, in_ch: int, z_dim: int, feat_hw: Tuple[int, int]):
        super().__init__()
        self.conv = nn.Sequential(
            nn.Conv2d(in_ch, 32, 3, stride=2, padding=1), nn.ReLU(),
            nn.Conv2d(32, 64, 3, stride=2, padding=1), nn.ReLU(),
            nn.Conv2d(64, 128, 3, stride=2, padding=1), nn.ReLU(),
        )
        # the reference discovers this with a dummy forward (12:92-96); the arithmetic is closed-form
        h, w = int(feat_hw[0]), int(feat_hw[1])
        for _ in range(3):
            h, w = _conv_out(h), _conv_out(w)
        self.h_shape = torch.Size((128, h, w))
        flat = 128 * h * w
        self.fc = nn.Sequential(nn.Linear(flat, 256), nn.ReLU())
        self.mu = nn.Linear(256, z_dim)
        self.logvar = nn.Linear(256, z_dim)

    def features(self, x: torch.Tensor) -> torch.Tensor:
        convs = [self.conv[0], self.conv[2], self.conv[4]]
        if convstack.encoder_supported(x, convs):      # bf16 mode, T a multiple of 1024: bf16 NHWC between the layers
            h = convstack.encoder_convs(x, convs)
        else:
            h = x
            for c in convs:
                h = ops.conv2d_s2k3(h, c.weight, c.bias, relu=True)
        return _lin(self.fc[0], h.flatten(1), relu=True)

    def forward(self, x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        h = self.features(x)
        return _lin(self.mu, h), _lin(self.logvar, h)


class LyricsProjector(nn.Module):
    """12:111-120."""

    def __init__(self, in_dim: int, out_dim: int = 128):
        super().__init__()
        self.net = nn.Sequential(nn.Linear(in_dim, 256), nn.ReLU(), nn.Linear(256, out_dim), nn.ReLU())

    def forward(self, e: torch.Tensor) -> torch.Tensor:
        return _lin(self.net[2], _lin(self.net[0], e, relu=True), relu=True)


class AudioDecoder(nn.Module):
    """12:123-142.  Linear(z,256)+ReLU -> Linear(256,C*H*W)+ReLU -> view -> 3 x ConvTranspose2d k4 s2 p1."""

    def __init__(self, z_dim: int, out_ch: int, h_shape: Tuple[int, int, int]):
        super().__init__()
        C, H, W = (int(v) for v in h_shape)
        self.h_shape = (C, H, W)
        self.fc = nn.Sequential(nn.Linear(z_dim, 256), nn.ReLU(), nn.Linear(256, C * H * W), nn.ReLU())
        self.deconv = nn.Sequential(
            nn.ConvTranspose2d(C, 64, 4, stride=2, padding=1), nn.ReLU(),
            nn.ConvTranspose2d(64, 32, 4, stride=2, padding=1), nn.ReLU(),
            nn.ConvTranspose2d(32, out_ch, 4, stride=2, padding=1),
        )

    def forward(self, z: torch.Tensor) -> torch.Tensor:
        h = _lin(self.fc[2], _lin(self.fc[0], z, relu=True), relu=True)
        h = h.view(z.size(0), *self.h_shape)
        deconvs = [self.deconv[0], self.deconv[2], self.deconv[4]]
        if convstack.decoder_supported(h, deconvs):
            return convstack.decoder_deconvs(h, deconvs)
        h = ops.convT2d_s2k4(h, self.deconv[0].weight, self.deconv[0].bias, relu=True)
        h = ops.convT2d_s2k4(h, self.deconv[2].weight, self.deconv[2].bias, relu=True)
        return ops.convT2d_s2k4(h, self.deconv[4].weight, self.deconv[4].bias, relu=False)


class ConvMultiModalVAE(nn.Module):
    """12:145-186.  forward(x, lyr, m) -> (xhat, mu, lv); xhat is NOT cropped (the caller crops, 12:260)."""

    def __init__(self, in_ch: int, feat_hw: Tuple[int, int], z_dim: int, lyr_dim: int):
        super().__init__()
        self.audio_enc = AudioEncoder(in_ch, z_dim, feat_hw)
        self.lyr_proj = LyricsProjector(lyr_dim, out_dim=128)
        self.fuse = nn.Sequential(nn.Linear(z_dim + 128 + 1, 256), nn.ReLU())
        self.mu = nn.Linear(256, z_dim)
        self.logvar = nn.Linear(256, z_dim)
        self.audio_dec = AudioDecoder(z_dim, in_ch, self.audio_enc.h_shape)

    @staticmethod
    def reparam(mu: torch.Tensor, logvar: torch.Tensor) -> torch.Tensor:
        eps = torch.randn_like(mu)  # same draw as randn_like(std) at 12:170
        z, _ = ops.reparam_kl(mu, logvar, eps)
        return z

    def encode(self, x: torch.Tensor, lyr: torch.Tensor, m: torch.Tensor):
        """(mu, lv) only -- what the latent export at 12:289-309 needs (the reference also runs the decoder)."""
        # lv_a of the reference (12:175) is unused, so audio_enc.logvar is never evaluated and gets no gradient
        mu_a = _lin(self.audio_enc.mu, self.audio_enc.features(x))
        l = self.lyr_proj(lyr)
        h = _lin(self.fuse[0], ops.cat_gate(mu_a, l, m.reshape(-1), append_gate=True), relu=True)
        return _lin(self.mu, h), _lin(self.logvar, h)

    def forward(self, x: torch.Tensor, lyr: torch.Tensor, m: torch.Tensor):
        mu, lv = self.encode(x, lyr, m)
        z = self.reparam(mu, lv)
        return self.audio_dec(z), mu, lv


def kl_div(mu: torch.Tensor, logvar: torch.Tensor) -> torch.Tensor:
    """12:189-190: -0.5 * mean(1 + lv - mu^2 - exp(lv)) as a 0-d tensor."""
    return ops.kl_rows(mu, logvar).sum() / mu.numel()


class MSELoss(nn.Module):
    """nn.MSELoss() of 12:234 with the crop of 12:260 folded in (xhat may be larger than x)."""

    def forward(self, xhat: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
        return ops.mse_loss(xhat, x, "mean")


# ------------------------------------------------------------------------------------------------ script 19
class MLPVAE(nn.Module):
    """19:64-121 (Beta-VAE / CVAE).  cond is concatenated to BOTH encoder and decoder inputs."""

    def __init__(self, input_dim: int, latent_dim: int = 16, hidden_dim: int = 256, cond_dim: int = 0,
                 conditional: bool = False):
        super().__init__()
        self.conditional = conditional
        self.cond_dim = cond_dim
        enc_in = input_dim + (cond_dim if conditional else 0)
        self.enc = nn.Sequential(nn.Linear(enc_in, hidden_dim), nn.ReLU(), nn.Linear(hidden_dim, hidden_dim), nn.ReLU())
        self.mu = nn.Linear(hidden_dim, latent_dim)
        self.logvar = nn.Linear(hidden_dim, latent_dim)
        dec_in = latent_dim + (cond_dim if conditional else 0)
        self.dec = nn.Sequential(nn.Linear(dec_in, hidden_dim), nn.ReLU(), nn.Linear(hidden_dim, hidden_dim), nn.ReLU(),
                                 nn.Linear(hidden_dim, input_dim))

    def _with_cond(self, t: torch.Tensor, c: Optional[torch.Tensor]) -> torch.Tensor:
        if not self.conditional:
            return t
        if c is None:
            raise ValueError("Conditional model requires conditioning vector c.")
        return ops.cat_gate(t, c)

    def encode(self, x, c=None):
        h = _lin(self.enc[2], _lin(self.enc[0], self._with_cond(x, c), relu=True), relu=True)
        return _lin(self.mu, h), _lin(self.logvar, h)

    def reparameterize(self, mu, logvar):
        z, _ = ops.reparam_kl(mu, logvar, torch.randn_like(mu))
        return z

    def decode(self, z, c=None):
        h = _lin(self.dec[2], _lin(self.dec[0], self._with_cond(z, c), relu=True), relu=True)
        return _lin(self.dec[4], h)

    def forward(self, x, c=None):
        mu, logvar = self.encode(x, c)
        z = self.reparameterize(mu, logvar)
        return self.decode(z, c), mu, logvar


def kl_divergence(mu, logvar):
    """19:124-126: per-sample KL(q(z|x) || N(0,I)), shape [B]."""
    return ops.kl_rows(mu, logvar)


def make_onehot(idx: np.ndarray, num_classes: int) -> np.ndarray:
    """19:129-132."""
    out = np.zeros((idx.shape[0], num_classes), dtype=np.float32)
    out[np.arange(idx.shape[0]), idx] = 1.0
    return out


# ------------------------------------------------------------------------------------------------ script 06
class EasyMLPVAE(nn.Module):
    """06:145-179 (`MLPVAE(input_dim, hidden_dim, latent_dim)` there; note the different argument order)."""

    def __init__(self, input_dim: int, hidden_dim: int, latent_dim: int):
        super().__init__()
        self.enc1 = nn.Linear(input_dim, hidden_dim)
        self.enc2 = nn.Linear(hidden_dim, hidden_dim)
        self.mu = nn.Linear(hidden_dim, latent_dim)
        self.logvar = nn.Linear(hidden_dim, latent_dim)
        self.dec1 = nn.Linear(latent_dim, hidden_dim)
        self.dec2 = nn.Linear(hidden_dim, hidden_dim)
        self.out = nn.Linear(hidden_dim, input_dim)

    def encode(self, x):
        h = _lin(self.enc2, _lin(self.enc1, x, relu=True), relu=True)
        return _lin(self.mu, h), _lin(self.logvar, h)

    def reparameterize(self, mu, logvar):
        z, _ = ops.reparam_kl(mu, logvar, torch.randn_like(mu))
        return z

    def decode(self, z):
        return _lin(self.out, _lin(self.dec2, _lin(self.dec1, z, relu=True), relu=True))

    def forward(self, x):
        mu, logvar = self.encode(x)
        return self.decode(self.reparameterize(mu, logvar)), mu, logvar


# ------------------------------------------------------------------------------------------------ script 22
class AE(nn.Module):
    """22:66-87: the plain auto-encoder baseline.  forward(x) -> (x_hat, z)."""

    def __init__(self, in_dim: int, z_dim: int = 16, hidden: int = 256):
        super().__init__()
        self.enc = nn.Sequential(nn.Linear(in_dim, hidden), nn.ReLU(), nn.Linear(hidden, hidden), nn.ReLU(),
                                 nn.Linear(hidden, z_dim))
        self.dec = nn.Sequential(nn.Linear(z_dim, hidden), nn.ReLU(), nn.Linear(hidden, hidden), nn.ReLU(),
                                 nn.Linear(hidden, in_dim))

    def encode(self, x):
        """(z, None): the pair shape of the VAEs' encode(), so the latent export treats all models alike."""
        return _lin(self.enc[4], _lin(self.enc[2], _lin(self.enc[0], x, relu=True), relu=True)), None

    def forward(self, x):
        z, _ = self.encode(x)
        return _lin(self.dec[4], _lin(self.dec[2], _lin(self.dec[0], z, relu=True), relu=True)), z


def vae_loss(xhat, x, mu, logvar, beta: float) -> Tuple[torch.Tensor, Dict[str, float]]:
    """06:182-188: mean-MSE + beta * mean-KL; returns (loss, {"recon","kl","total"} as python floats)."""
    recon = ops.mse_loss(xhat, x, "mean")
    kl = ops.kl_rows(mu, logvar).sum() / mu.numel()
    loss = recon + beta * kl
    return loss, {"recon": float(recon.item()), "kl": float(kl.item()), "total": float(loss.item())}
