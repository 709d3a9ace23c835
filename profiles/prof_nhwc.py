"""Time the TMA-fed NHWC conv kernels at the script-12 shapes (CUDA events, 20 launches after 3 warm-ups).
usage: prof_nhwc.py [fwd|dgrad|wgrad|all]"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT), str(ROOT / "vae-for-hybrid-music-clustering_b200")]
from vaeclust._lib import lib, ptr, stream_ptr  # noqa: E402

L = lib()
B = 64
LAYERS = {"conv2": ("conv", 32, 64, 512, 64), "conv3": ("conv", 64, 32, 256, 128),
          "deconv1": ("convT", 128, 16, 128, 64), "deconv2": ("convT", 64, 32, 256, 32)}
which = sys.argv[1] if len(sys.argv) > 1 else "all"
bf = torch.bfloat16


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


for name, (kind, Cin, H, W, Cout) in LAYERS.items():
    if kind == "conv":
        xs, ys = (B, H, W, Cin), (B, H // 2, W // 2, Cout)
        w = torch.randn(Cout, Cin, 3, 3, device="cuda") * 0.05
        ops = {"fwd": 3, "dgrad": 4, "wgrad": 5}
    else:
        xs, ys = (B, H, W, Cin), (B, 2 * H, 2 * W, Cout)
        w = torch.randn(Cin, Cout, 4, 4, device="cuda") * 0.05
        ops = {"fwd": 6, "dgrad": 7, "wgrad": 8}
    x = torch.randn(xs, device="cuda").to(bf)
    y = torch.randn(ys, device="cuda").to(bf)
    bias = torch.randn(Cout, device="cuda")
    out_x, out_y = torch.empty_like(x), torch.empty_like(y)
    dims = (B, Cin, H, W, Cout)
    for pas in ("fwd", "dgrad", "wgrad"):
        if which not in ("all", pas):
            continue
        op = ops[pas]
        if pas == "wgrad":
            ws = torch.empty(L.size("vc_conv_nhwc_wgrad_workspace_bytes", op, *dims), dtype=torch.uint8, device="cuda")
            dw = torch.empty_like(w)
            us = timeit(lambda: L.call("vc_conv_nhwc_wgrad", op, ptr(y), ptr(x), ptr(dw), *dims, ptr(ws), ws.numel(), stream_ptr()))
            mb = (x.numel() + y.numel()) * 2 / 1e6
        else:
            ws = torch.empty(L.size("vc_conv_nhwc_workspace_bytes", op, *dims), dtype=torch.uint8, device="cuda")
            if pas == "fwd":
                us = timeit(lambda: L.call("vc_conv_nhwc", op, ptr(x), ptr(w), ptr(bias), 1, None, ptr(out_y), None, *dims,
                                           ptr(ws), ws.numel(), stream_ptr()))
                mb = (x.numel() + y.numel()) * 2 / 1e6
            else:
                us = timeit(lambda: L.call("vc_conv_nhwc", op, ptr(y), ptr(w), None, 0, ptr(x), ptr(out_x), None, *dims,
                                           ptr(ws), ws.numel(), stream_ptr()))
                mb = (2 * x.numel() + y.numel()) * 2 / 1e6
        print(f"{name:8s} {pas:6s} {us:8.1f} us  {mb:7.1f} MB algorithmic  {mb / us * 1e3 / 1e3:6.2f} TB/s")
