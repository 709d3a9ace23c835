"""The fused weight gradient + Adam update of the two 67M-parameter Linear layers (vc_linear_wgrad_adam) at the script-12
shape (B = 64, 262144 <-> 256): CUDA-event timings when run plainly, the target of `ncu --set full -k regex:fc_adam`."""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT / "vae-for-hybrid-music-clustering_b200")]
from vaeclust._lib import lib, ptr, stream_ptr  # noqa: E402

L = lib()
B, WIDE, HID = 64, 262144, 256
gen = torch.Generator(device="cuda").manual_seed(0)
rnd = lambda *s: torch.randn(*s, device="cuda", generator=gen)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
# argv[2] = "tlb": also touch one element per 64 KB of an 8 GB buffer between calls (4096 distinct 2 MB pages), so that the
# address-translation caches are cold at every launch, as they are inside a training step
thrash = torch.zeros(1 << 31, dtype=torch.float32, device="cuda") if len(sys.argv) > 2 and sys.argv[2] == "tlb" else None
for name, (cin, cout) in {"enc W [256][262144]": (WIDE, HID), "dec W [262144][256]": (HID, WIDE)}.items():
    dy, x = rnd(B, cout) * 1e-2, rnd(B, cin)
    if len(sys.argv) > 2 and sys.argv[2] == "sparse":
        # what training data looks like: ReLU outputs with dead units -> whole rows / columns of the gradient are exactly 0,
        # and exp_avg / exp_avg_sq stay exactly 0 there (IEEE sqrt and division take their special-case paths for 0)
        x = torch.relu(x) * (torch.rand(1, cin, device="cuda", generator=gen) > 0.3)
        dy = dy * (torch.rand(1, cout, device="cuda", generator=gen) > 0.3)
    W = rnd(cout, cin) * 0.01
    m, v = torch.zeros_like(W), torch.zeros_like(W)
    sh = W.to(torch.bfloat16)
    ws = torch.empty(L.size("vc_linear_wgrad_adam_workspace_bytes", B, cin, cout), dtype=torch.uint8, device="cuda")
    call = lambda: L.call("vc_linear_wgrad_adam", ptr(dy), ptr(x), ptr(W), ptr(m), ptr(v), ptr(sh), B, cin, cout, 1e-3, 0.9, 0.999,
                          1e-8, 0.1, 0.001, None, ptr(ws), ws.numel(), stream_ptr())
    call()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.zero_()
        if thrash is not None:
            thrash[::16384].add_(1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); call(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    gb = (24 + 2) * cin * cout / 1e9
    print(f"{name:22s} median {ts[len(ts) // 2]:7.1f} us  ({gb / ts[len(ts) // 2] * 1e3:5.2f} TB/s of p/m/v/shadow traffic)   min {ts[0]:7.1f} us")
    del W, m, v, sh
