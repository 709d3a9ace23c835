"""Calls every mma.sync kernel of the single-channel ends (csrc/nhwc_mma.cu, conv_direct.cu) at the script-12 shapes
(B = 64, 1 x 128 x 1024 input, 32 channels at 64 x 512), a few times each: CUDA-event timings when run plainly, the target of
`ncu --set full -k regex:hm_` when profiled."""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT / "vae-for-hybrid-music-clustering_b200")]
from vaeclust._lib import lib, ptr, stream_ptr  # noqa: E402

L = lib()
B, C, H, W = 64, 32, 64, 512
g = torch.Generator(device="cuda").manual_seed(0)
rnd = lambda *s: torch.randn(*s, device="cuda", generator=g)
x1 = rnd(B, 1, 2 * H, 2 * W)                       # one-channel fp32 tensor (input / output gradient)
an = rnd(B, H, W, C).to(torch.bfloat16)            # 32-channel bf16 NHWC tensor
gn = torch.empty_like(an)
w3, w4, bias = rnd(C, 1, 3, 3), rnd(C, 1, 4, 4), rnd(C)
b1 = rnd(1)
y1 = torch.empty_like(x1)
dw3, dw4, db = torch.empty_like(w3), torch.empty_like(w4), torch.empty(C, device="cuda")
ws = torch.empty(1 << 24, dtype=torch.uint8, device="cuda")
s = stream_ptr()
calls = {
    "conv1_fwd": lambda: L.call("vc_nhwc_conv1_fwd", ptr(x1), ptr(w3), ptr(bias), 1, ptr(gn), B, 2 * H, 2 * W, C, s),
    "convT1_fwd": lambda: L.call("vc_nhwc_convT1_fwd", ptr(an), ptr(w4), ptr(b1), ptr(y1), B, C, H, W, s),
    "convT1_dgrad_bias": lambda: L.call("vc_nhwc_convT1_dgrad_bias", ptr(x1), ptr(w4), ptr(an), ptr(gn), ptr(db), B, C, H, W,
                                        ptr(ws), ws.numel(), s),
    "tap_wgrad_k4": lambda: L.call("vc_nhwc_tap_wgrad", 4, ptr(an), ptr(x1), ptr(dw4), B, C, H, W, ptr(ws), ws.numel(), s),
    "conv1_wgrad_bias": lambda: L.call("vc_nhwc_conv1_wgrad_bias", ptr(an), ptr(x1), ptr(dw3), ptr(db), B, 2 * H, 2 * W, C,
                                       ptr(ws), ws.numel(), s),
}
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
for name, fn in calls.items():
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    print(f"{name:20s} median {ts[len(ts) // 2]:7.1f} us   min {ts[0]:7.1f} us")
