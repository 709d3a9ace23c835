"""Diagnostic (GPU box): fused MLP-VAE step vs the same-arithmetic oracle, per-parameter gradient error, for both
MN-major descriptor conventions of the weight-gradient kernel (variant 0 = canonical, 1 = LBO/SBO exchanged)."""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT), str(ROOT / "vae-for-hybrid-music-clustering_b200")]
from oracle import parity, vae_ref  # noqa: E402
from vaeclust import nn  # noqa: E402
from vaeclust.fused import FusedMLPStep  # noqa: E402

for variant in (0, 1):
    for B in (32, 300):
        D, Z, cond = 464, 16, 6
        model = nn.MLPVAE(D, Z, 256, cond_dim=cond, conditional=True)
        model.load_state_dict(vae_ref.init_mlp_vae_params(D, Z, 256, cond, seed=5, style="19"))
        model.cuda().train()
        g = torch.Generator().manual_seed(1)
        x = torch.randn(B, D, generator=g)
        x[:, :80] *= torch.logspace(1.0, 0.0, 80)
        x[:, 80:] *= 0.05
        batch = {"x": x, "eps": torch.randn(B, Z, generator=g),
                 "c": torch.nn.functional.one_hot(torch.randint(0, cond, (B,), generator=g), cond).float()}
        batch = {k: v.cuda() for k, v in batch.items()}
        step = FusedMLPStep(model, "mlp19", lr=1e-3, beta=4.0)
        step.variant = variant
        try:
            r = parity.vae_step_parity(model, step, "mlp19", batch, 4.0, 1e-3, "bf16")
            print(f"variant {variant} B={B}: ok={r['ok']} loss {r['loss']:.4f} ref {r['loss_ref']:.4f} loss_rel "
                  f"{r['loss_rel']:.2e} kl_rel {r['kl_rel']:.2e} grad_rel {r['grad_rel']}", flush=True)
        except Exception as e:   # noqa: BLE001
            print(f"variant {variant} B={B}: FAILED {e!r}", flush=True)
