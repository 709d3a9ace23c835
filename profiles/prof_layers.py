"""Profiling driver: one dense-layer entry point at a script-12 shape (used plain and under ncu; profiles/README.md).
usage: prof_layers.py <kind: conv|convT|linear> <which: fwd|dgrad|wgrad> B Cin H W Cout [reps]"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT), str(ROOT / "vae-for-hybrid-music-clustering_b200")]
import torch
from vaeclust import ops

kind, which = sys.argv[1], sys.argv[2]
B, Cin, H, W, Cout = (int(v) for v in sys.argv[3:8])
reps = int(sys.argv[8]) if len(sys.argv) > 8 else 5
ops.set_precision("bf16")
g = torch.Generator(device="cuda").manual_seed(0)
r = lambda *s: torch.randn(*s, device="cuda", generator=g)
if kind == "linear":
    x, w, b = r(B, Cin), r(Cout, Cin) * Cin ** -0.5, r(Cout)
    yshape = (B, Cout)
elif kind == "conv":
    x, w, b = r(B, Cin, H, W), r(Cout, Cin, 3, 3) * (Cin * 9) ** -0.5, r(Cout)
    yshape = (B, Cout, (H - 1) // 2 + 1, (W - 1) // 2 + 1)
else:
    x, w, b = r(B, Cin, H, W), r(Cin, Cout, 4, 4) * (Cin * 4) ** -0.5, r(Cout)
    yshape = (B, Cout, 2 * H, 2 * W)
dy = r(*yshape)
y = torch.empty(yshape, device="cuda")
dx, dw = torch.empty_like(x), torch.empty_like(w)
fw, dg, wg = ops._OPS[kind]
hw = (1, 1) if kind == "linear" else (H, W)


def run():
    if which == "fwd":
        ops._dense(fw, 1, x, w, b, y, B, Cin, hw[0], hw[1], Cout, True)
    elif which == "dgrad":
        ops._dense(dg, 1, dy, w, None, dx, B, Cin, hw[0], hw[1], Cout, False)
    else:
        ops._dense(wg, 1, dy, x, None, dw, B, Cin, hw[0], hw[1], Cout, False)


run()
torch.cuda.synchronize()
a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(reps):
    run()
e.record()
torch.cuda.synchronize()
print(f"{kind} {which} B={B} Cin={Cin} H={H} W={W} Cout={Cout}: {a.elapsed_time(e) / reps:.4f} ms")
