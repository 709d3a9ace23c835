"""Profiling driver: silhouette row blocks on N x d points, k clusters (tc or fp32 path)."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT), str(ROOT / "vae-for-hybrid-music-clustering_b200")]
import torch
from vaeclust import cluster

N = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 64
k = int(sys.argv[3]) if len(sys.argv) > 3 else 8
cluster.SILHOUETTE_PRECISION = sys.argv[4] if len(sys.argv) > 4 else "tc"
g = torch.Generator(device="cuda").manual_seed(3)
cent = torch.randn(k, d, device="cuda", generator=g) * 3
lab = torch.randint(0, k, (N,), device="cuda", generator=g)
X = cent[lab] + torch.randn(N, d, device="cuda", generator=g)
cluster.silhouette_samples(X[:8192].contiguous(), lab[:8192], _return_sum=True)
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
s, n = cluster.silhouette_samples(X, lab, _return_sum=True)
b.record()
torch.cuda.synchronize()
ms = a.elapsed_time(b)
print(f"N={N} d={d} k={k} {cluster.SILHOUETTE_PRECISION}: {ms:.2f} ms, {N * N / ms / 1e9:.2f} Gpairs/s, score={s / n:.7f}")
