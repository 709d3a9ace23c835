"""Timing (GPU box): fused MLP-VAE step, graph-replayed, CUDA events; per-kernel split from an eager pass."""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT), str(ROOT / "vae-for-hybrid-music-clustering_b200")]
from vaeclust import nn, ops  # noqa: E402
from vaeclust.train import GraphedVAEStep  # noqa: E402

ops.set_precision("bf16")
N, D, COND = 200_000, 464, 6
g = torch.Generator(device="cuda").manual_seed(11)
X = torch.cat([torch.randn(N, 80, device="cuda", generator=g) * torch.logspace(2, 0, 80, device="cuda"),
               torch.randn(N, 384, device="cuda", generator=g) * 0.05], dim=1).contiguous()
cond = torch.nn.functional.one_hot(torch.randint(0, COND, (N,), device="cuda", generator=g), COND).float()
for B in (256, 1024, 8192, 32768):
    for fused in ("always", "never"):
        torch.manual_seed(42)
        model = nn.MLPVAE(D, 16, 256, cond_dim=COND, conditional=True).cuda().train()
        step = GraphedVAEStep(model, "mlp19", lr=1e-3, beta=4.0, warmup=2, fused=fused)
        for s in range(5):
            lo = (s * B) % (N - B)
            step(X[lo: lo + B], cond[lo: lo + B])
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 50
        a.record()
        for s in range(iters):
            lo = ((5 + s) * B) % (N - B)
            out = step(X[lo: lo + B], cond[lo: lo + B])
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / iters
        print(f"B={B} fused={fused}: {ms * 1e3:.1f} us/step  {B / ms * 1e3 / 1e6:.2f} M clips/s  loss {float(out[0]):.2f}", flush=True)
