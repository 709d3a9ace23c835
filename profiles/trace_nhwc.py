"""Timeline of CTA 0 of conv_nhwc_kernel for one layer: per k-block, when its producer started waiting for the ring slot,
got it, had issued the TMA; when the MMA issuer started waiting for the data, got it, had issued + committed."""
import sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT), str(ROOT / "vae-for-hybrid-music-clustering_b200")]
from vaeclust._lib import lib, ptr, stream_ptr  # noqa: E402
L = lib()
B = 64
name = sys.argv[1] if len(sys.argv) > 1 else "deconv2_dgrad"
cfg = {"deconv2_dgrad": (7, 64, 32, 256, 32), "conv2_fwd": (3, 32, 64, 512, 64), "deconv2_fwd": (6, 64, 32, 256, 32)}[name]
op, Cin, H, W, Cout = cfg
bf = torch.bfloat16
if op in (3, 4):
    xs, ys, w = (B, H, W, Cin), (B, H // 2, W // 2, Cout), torch.randn(Cout, Cin, 3, 3, device="cuda") * .05
else:
    xs, ys, w = (B, H, W, Cin), (B, 2 * H, 2 * W, Cout), torch.randn(Cin, Cout, 4, 4, device="cuda") * .05
x, y = torch.randn(xs, device="cuda").to(bf), torch.randn(ys, device="cuda").to(bf)
dims = (B, Cin, H, W, Cout)
ws = torch.empty(L.size("vc_conv_nhwc_workspace_bytes", op, *dims), dtype=torch.uint8, device="cuda")
fwd = op in (3, 6)
src, out = (x, torch.empty_like(y)) if fwd else (y, torch.empty_like(x))
def run():
    L.call("vc_conv_nhwc", op, ptr(src), ptr(w), None, 0, None, ptr(out), None, *dims, ptr(ws), ws.numel(), stream_ptr())
for _ in range(3):
    run()
buf = torch.zeros(64 * 8, dtype=torch.int64, device="cuda")
L.call("vc_conv_nhwc_debug_trace", buf.data_ptr())
run()
torch.cuda.synchronize()
L.call("vc_conv_nhwc_debug_trace", None)
t = buf.cpu().numpy().reshape(64, 8)
t0 = t[:, 0][t[:, 0] > 0].min()
print("kb   P:wait-start  P:slot-free  P:issued |  M:wait-start  M:data-ready  M:committed   (clocks since first stamp)")
for i in range(48):
    r = [int(v - t0) if v else -1 for v in t[i, :6]]
    print(f"{i:3d} {r[0]:12d} {r[1]:12d} {r[2]:10d} | {r[3]:13d} {r[4]:13d} {r[5]:12d}")
