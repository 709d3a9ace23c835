"""A few EAGER fused MLP-VAE steps (one launch list under ncu shows the four kernels' durations)."""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT), str(ROOT / "vae-for-hybrid-music-clustering_b200")]
from vaeclust import nn  # noqa: E402
from vaeclust.fused import FusedMLPStep  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
torch.manual_seed(42)
model = nn.MLPVAE(464, 16, 256, cond_dim=6, conditional=True).cuda().train()
step = FusedMLPStep(model, "mlp19", lr=1e-3, beta=4.0, max_batch=B)
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn(B, 464, device="cuda", generator=g)
c = torch.nn.functional.one_hot(torch.randint(0, 6, (B,), device="cuda", generator=g), 6).float()
for _ in range(4):
    out = step(x, c)
torch.cuda.synchronize()
print("loss", float(out[0]))
