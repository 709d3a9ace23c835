"""Per-phase timeline of the rows kernel (CTA 0, first tile): clock64 stamps -> microseconds at 1.965 GHz."""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT), str(ROOT / "vae-for-hybrid-music-clustering_b200")]
from vaeclust import lib, nn  # noqa: E402
from vaeclust.fused import FusedMLPStep  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
torch.manual_seed(42)
model = nn.MLPVAE(464, 16, 256, cond_dim=6, conditional=True).cuda().train()
step = FusedMLPStep(model, "mlp19", lr=1e-3, beta=4.0, max_batch=B)
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn(B, 464, device="cuda", generator=g)
c = torch.nn.functional.one_hot(torch.randint(0, 6, (B,), device="cuda", generator=g), 6).float()
for _ in range(3):
    step(x, c)
buf = torch.zeros(128, dtype=torch.int64, device="cuda")
lib().call("vc_mlpvae_debug_trace", buf.data_ptr())
step(x, c)
torch.cuda.synchronize()
lib().call("vc_mlpvae_debug_trace", None)
t = buf.cpu().numpy()
t0 = t[0]
us = lambda v: (v - t0) / 1965.0
names = ["enc.0", "enc.2", "heads", "dec.0", "dec.2", "out a", "out b", "dd2 a", "dd2 b", "dd1", "dz", "dh2", "dh1"]
acc = t[64:77]
ai = 0
print("phase      reach   prev-epi-done  mmas-issued  stores-read  acc-ready  (us since kernel start of CTA 0)")
for p in range(13):
    r = t[p * 4: p * 4 + 4]
    has_epi = p != 5
    ready = us(acc[ai]) if has_epi else float("nan")
    if has_epi:
        ai += 1
    print(f"{names[p]:8s} {us(r[0]):8.2f} {us(r[1]):12.2f} {us(r[2]):12.2f} {us(r[3]):12.2f} {ready:10.2f}")
w = t[96:104]
print("worker thread 64, epilogue of enc.2 (us):", [round(float(us(v)), 2) for v in w], " acc-ready", round(float(us(acc[1])), 2))
print("issuer, heads phase: per k-block (A block taken, weights arrived):", [round(float(us(v)), 2) for v in t[104:112]])
