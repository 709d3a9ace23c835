"""Profiling driver: a few fused Lloyd iterations on N x 32 points (used under ncu; see profiles/README.md)."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT), str(ROOT / "vae-for-hybrid-music-clustering_b200")]
import torch
from vaeclust import cluster

N = int(sys.argv[1]) if len(sys.argv) > 1 else 8_000_000
ks = [int(v) for v in sys.argv[2].split(",")] if len(sys.argv) > 2 else [8]   # one k or a comma list
d = int(sys.argv[3]) if len(sys.argv) > 3 else 32
g = torch.Generator(device="cuda").manual_seed(0)
X = torch.randn(N, d, device="cuda", generator=g)
for k in ks:
    C = X[:k].clone().contiguous()
    st = cluster._LloydState(N, d, k, X.device)
    st.labels.fill_(-1)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for it in range(4):
        if it == 1:
            a.record()
        cluster._estep(X, C, st, True)
        cluster._finalize(C, st)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 3
    print(f"N={N} k={k} d={d}: {ms:.3f} ms/iter, {(4 * d + 4) * N / ms / 1e6:.1f} GB/s algorithmic")
