"""TMA-fed bf16-NHWC convolution kernels (csrc/conv_nhwc.cu and friends) against torch fp32 on bf16-rounded operands.

The kernels multiply bf16 operands exactly and accumulate in fp32, so against an fp32 torch convolution of the ROUNDED
inputs the only differences are summation order (1e-5 level) and, for bf16 outputs, the final rounding (2^-9 relative).
"""
import sys
from pathlib import Path

import pytest
import torch
import torch.nn.functional as F

ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT / "vae-for-hybrid-music-clustering_b200")]

pytestmark = pytest.mark.gpu

OPS = {"conv_fwd": 3, "conv_dgrad": 4, "convT_fwd": 6, "convT_dgrad": 7}


def _r(*shape, seed, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(*shape, device="cuda", generator=g) * scale


def _bf(t):
    return t.to(torch.bfloat16).float()


def _nhwc(t):
    return t.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)


def _run(op, src_nchw, w, bias, relu, mask_nchw, dims, want_nhwc=True, want_nchw=True):
    from vaeclust._lib import lib, ptr, stream_ptr
    L = lib()
    B, Cin, H, W, Cout = dims
    assert L.size("vc_conv_nhwc_supported", op, B, Cin, H, W, Cout) == 1
    if op == OPS["conv_fwd"]:
        oshape = (B, Cout, H // 2, W // 2)
    elif op == OPS["convT_fwd"]:
        oshape = (B, Cout, 2 * H, 2 * W)
    else:
        oshape = (B, Cin, H, W)
    src = _nhwc(src_nchw)
    mask = _nhwc(mask_nchw) if mask_nchw is not None else None
    o_nhwc = torch.full((oshape[0], oshape[2], oshape[3], oshape[1]), float("nan"), dtype=torch.bfloat16, device="cuda") \
        if want_nhwc else None
    o_nchw = torch.full(oshape, float("nan"), device="cuda") if want_nchw else None
    ws = torch.empty(L.size("vc_conv_nhwc_workspace_bytes", op, B, Cin, H, W, Cout), dtype=torch.uint8, device="cuda")
    L.call("vc_conv_nhwc", op, ptr(src), ptr(w), ptr(bias), int(relu), ptr(mask), ptr(o_nhwc), ptr(o_nchw), B, Cin, H, W,
           Cout, ptr(ws), ws.numel(), stream_ptr())
    torch.cuda.synchronize()
    return o_nhwc, o_nchw


def _check(o_nhwc, o_nchw, want):
    scale = float(want.abs().max())
    if o_nchw is not None:
        torch.testing.assert_close(o_nchw, want, rtol=2e-5, atol=2e-5 * scale)
    if o_nhwc is not None:
        got = o_nhwc.float().permute(0, 3, 1, 2)
        torch.testing.assert_close(got, want, rtol=2 ** -8, atol=2 ** -8 * scale * 0.05 + 1e-6)


@pytest.mark.parametrize("B,Cin,H,W,Cout", [(2, 32, 8, 512, 64),      # conv.2 of script 12: 32-channel source, pair view
                                            (3, 64, 6, 256, 128),     # conv.4: one 64-channel chunk per tap
                                            (1, 128, 4, 256, 64)])    # two chunks per tap
def test_conv_forward(B, Cin, H, W, Cout):
    x, w, b = _r(B, Cin, H, W, seed=1), _r(Cout, Cin, 3, 3, seed=2, scale=(Cin * 9) ** -0.5), _r(Cout, seed=3)
    want = F.relu(F.conv2d(_bf(x), _bf(w), b, stride=2, padding=1))
    _check(*_run(OPS["conv_fwd"], x, w, b, True, None, (B, Cin, H, W, Cout)), want)
    want = F.conv2d(_bf(x), _bf(w), None, stride=2, padding=1)
    _check(*_run(OPS["conv_fwd"], x, w, None, False, None, (B, Cin, H, W, Cout), want_nhwc=False), want)


@pytest.mark.parametrize("B,Cin,H,W,Cout", [(2, 128, 4, 128, 64),     # deconv.0: slab streams (N = 256, 18 k-blocks)
                                            (2, 64, 5, 256, 32)])     # deconv.2: resident slab, N = 128
def test_convT_forward(B, Cin, H, W, Cout):
    x, w, b = _r(B, Cin, H, W, seed=1), _r(Cin, Cout, 4, 4, seed=2, scale=(Cin * 4) ** -0.5), _r(Cout, seed=3)
    want = F.relu(F.conv_transpose2d(_bf(x), _bf(w), b, stride=2, padding=1))
    _check(*_run(OPS["convT_fwd"], x, w, b, True, None, (B, Cin, H, W, Cout)), want)


@pytest.mark.parametrize("B,Cin,H,W,Cout", [(2, 32, 8, 512, 64), (3, 64, 6, 256, 128)])
def test_conv_data_gradient_with_mask(B, Cin, H, W, Cout):
    g, w = _r(B, Cout, H // 2, W // 2, seed=4), _r(Cout, Cin, 3, 3, seed=2, scale=(Cin * 9) ** -0.5)
    yprev = _r(B, Cin, H, W, seed=5)
    dx = torch.nn.grad.conv2d_input((B, Cin, H, W), _bf(w), _bf(g), stride=2, padding=1)
    want = dx * (_bf(yprev) > 0)
    _check(*_run(OPS["conv_dgrad"], g, w, None, False, yprev, (B, Cin, H, W, Cout)), want)
    _check(*_run(OPS["conv_dgrad"], g, w, None, False, None, (B, Cin, H, W, Cout), want_nhwc=False), dx)


@pytest.mark.parametrize("B,Cin,H,W,Cout", [(2, 128, 4, 128, 64), (2, 64, 5, 256, 32), (1, 64, 3, 128, 32)])
def test_convT_data_gradient_with_mask(B, Cin, H, W, Cout):
    g, w = _r(B, Cout, 2 * H, 2 * W, seed=4), _r(Cin, Cout, 4, 4, seed=2, scale=(Cin * 4) ** -0.5)
    yprev = _r(B, Cin, H, W, seed=5)
    dx = F.conv2d(_bf(g), _bf(w), None, stride=2, padding=1)      # the data gradient of a transposed conv is a conv
    want = dx * (_bf(yprev) > 0)
    _check(*_run(OPS["convT_dgrad"], g, w, None, False, yprev, (B, Cin, H, W, Cout)), want)
    _check(*_run(OPS["convT_dgrad"], g, w, None, False, None, (B, Cin, H, W, Cout), want_nhwc=False), dx)


@pytest.mark.parametrize("B,C,H,W", [(2, 32, 16, 256), (3, 16, 5, 64), (1, 64, 7, 96), (2, 128, 3, 32),
                                     (2, 32, 4, 40), (1, 8, 6, 64)])   # the last two take the FFMA kernel
def test_single_channel_weight_gradients_tensor_core(B, C, H, W):
    """vc_conv1_wgrad_bias / vc_convT1_wgrad with round_bf16 = 1: C in {16, 32, 64, 128} and rows of 32-pixel segments run
    tap_wgrad_mma_kernel (mma.sync on bf16-rounded operands, fp32 accumulation), other shapes the FFMA kernel with the
    same rounding.  Against torch autograd on the rounded operands only the summation order differs."""
    from vaeclust._lib import lib, ptr, stream_ptr
    L = lib()
    a, S = _r(B, C, H, W, seed=21), _r(B, 1, 2 * H, 2 * W, seed=22)
    # Conv2d(1 -> C): a = gradient of the output [H, W], S = the input [2H, 2W]
    dw, db = torch.full((C, 1, 3, 3), float("nan"), device="cuda"), torch.full((C,), float("nan"), device="cuda")
    ws = torch.empty(L.size("vc_conv1_wgrad_bias_workspace_bytes", C), dtype=torch.uint8, device="cuda")
    for _ in range(2):      # second call: same bits (fixed-order reduction)
        prev = dw.clone()
        L.call("vc_conv1_wgrad_bias", ptr(a), ptr(S), ptr(dw), ptr(db), B, 2 * H, 2 * W, C, 1, ptr(ws), ws.numel(), stream_ptr())
    assert torch.equal(prev, dw)
    wr = torch.zeros(C, 1, 3, 3, device="cuda", dtype=torch.float64, requires_grad=True)
    F.conv2d(_bf(S).double(), wr, None, stride=2, padding=1).backward(_bf(a).double())
    torch.testing.assert_close(dw.double(), wr.grad, rtol=2e-5, atol=2e-5 * float(wr.grad.abs().max()))
    wantb = a.double().sum(dim=(0, 2, 3))
    torch.testing.assert_close(db.double(), wantb, rtol=1e-5, atol=1e-5 * float(wantb.abs().max()) + 1e-4)
    # ConvTranspose2d(C -> 1): a = the layer input [H, W], S = gradient of the output [2H, 2W]
    dwt = torch.full((C, 1, 4, 4), float("nan"), device="cuda")
    ws = torch.empty(L.size("vc_convT1_wgrad_workspace_bytes", C), dtype=torch.uint8, device="cuda")
    L.call("vc_convT1_wgrad", ptr(S), ptr(a), ptr(dwt), B, C, H, W, 1, ptr(ws), ws.numel(), stream_ptr())
    wr = torch.zeros(C, 1, 4, 4, device="cuda", dtype=torch.float64, requires_grad=True)
    F.conv_transpose2d(_bf(a).double(), wr, None, stride=2, padding=1).backward(_bf(S).double())
    torch.testing.assert_close(dwt.double(), wr.grad, rtol=2e-5, atol=2e-5 * float(wr.grad.abs().max()))
    # round_bf16 = 0 keeps the fp32 values (FFMA kernel)
    L.call("vc_convT1_wgrad", ptr(S), ptr(a), ptr(dwt), B, C, H, W, 0, ptr(ws), ws.numel(), stream_ptr())
    wr = torch.zeros(C, 1, 4, 4, device="cuda", dtype=torch.float64, requires_grad=True)
    F.conv_transpose2d(a.double(), wr, None, stride=2, padding=1).backward(S.double())
    torch.testing.assert_close(dwt.double(), wr.grad, rtol=2e-5, atol=2e-5 * float(wr.grad.abs().max()))


@pytest.mark.parametrize("B,H,W", [(2, 16, 256), (3, 5, 32), (1, 9, 96), (2, 1, 64)])
def test_single_channel_ends_tensor_core(B, H, W):
    """The mma.sync kernels of csrc/nhwc_mma.cu (32-channel layers): ConvTranspose2d(32 -> 1) forward from NHWC (tiles of 14
    pixels, so W = 32 / 96 / 256 all end in a partial tile), its data gradient with and without the ReLU mask, and both
    tap weight gradients incl. the bias gradient from the constant-one tap."""
    from vaeclust._lib import lib, ptr, stream_ptr
    L = lib()
    C = 32
    wt, bb = _r(C, 1, 4, 4, seed=5, scale=0.25), _r(1, seed=10)
    # forward
    xin = _r(B, C, H, W, seed=9)
    xn = _nhwc(xin)
    for bias in (bb, None):
        yo = torch.full((B, 1, 2 * H, 2 * W), float("nan"), device="cuda")
        L.call("vc_nhwc_convT1_fwd", ptr(xn), ptr(wt), ptr(bias), ptr(yo), B, C, H, W, stream_ptr())
        want = F.conv_transpose2d(_bf(xin).double(), _bf(wt).double(), None if bias is None else bias.double(), stride=2, padding=1)
        torch.testing.assert_close(yo.double(), want, rtol=2e-5, atol=2e-5 * float(want.abs().max()))
    # data gradient
    dy, act = _r(B, 1, 2 * H, 2 * W, seed=4), _r(B, C, H, W, seed=6)
    an = _nhwc(act)
    for mask in (an, None):
        g = torch.full((B, H, W, C), float("nan"), dtype=torch.bfloat16, device="cuda")
        L.call("vc_nhwc_convT1_dgrad", ptr(dy), ptr(wt), ptr(mask), ptr(g), B, C, H, W, stream_ptr())
        want = F.conv2d(_bf(dy), _bf(wt), None, stride=2, padding=1)
        if mask is not None:
            want = want * (_bf(act) > 0)
        _check(g, None, want)
    # data gradient + bias gradient of the layer below from the same pass
    g2 = torch.full((B, H, W, C), float("nan"), dtype=torch.bfloat16, device="cuda")
    db = torch.full((C,), float("nan"), device="cuda")
    ws = torch.empty(L.size("vc_nhwc_convT1_dgrad_bias_workspace_bytes", C), dtype=torch.uint8, device="cuda")
    L.call("vc_nhwc_convT1_dgrad_bias", ptr(dy), ptr(wt), ptr(an), ptr(g2), ptr(db), B, C, H, W, ptr(ws), ws.numel(), stream_ptr())
    L.call("vc_nhwc_convT1_dgrad", ptr(dy), ptr(wt), ptr(an), ptr(g), B, C, H, W, stream_ptr())
    assert torch.equal(g2, g)
    wantb = g.double().sum(dim=(0, 1, 2))
    torch.testing.assert_close(db.double(), wantb, rtol=1e-5, atol=1e-5 * float(wantb.abs().max()) + 1e-4)
    # Conv2d(1 -> 32) forward into NHWC (W a multiple of 64 takes the tensor-core kernel)
    xc, wc, bc = _r(B, 1, 2 * H, 2 * W, seed=31), _r(C, 1, 3, 3, seed=32, scale=1 / 3), _r(C, seed=33)
    for bias, relu in ((bc, 1), (None, 0)):
        yc = torch.full((B, H, W, C), float("nan"), dtype=torch.bfloat16, device="cuda")
        L.call("vc_nhwc_conv1_fwd", ptr(xc), ptr(wc), ptr(bias), relu, ptr(yc), B, 2 * H, 2 * W, C, stream_ptr())
        want = F.conv2d(_bf(xc), _bf(wc), bias, stride=2, padding=1)
        _check(yc, None, F.relu(want) if relu else want)
    # weight gradients (+ bias gradient for the k3 end)
    a_t, S = _r(B, C, H, W, seed=11), _r(B, 1, 2 * H, 2 * W, seed=12)
    an2 = _nhwc(a_t)
    dwt = torch.full((C, 1, 4, 4), float("nan"), device="cuda")
    ws = torch.empty(L.size("vc_nhwc_tap_wgrad_workspace_bytes", C, 4), dtype=torch.uint8, device="cuda")
    L.call("vc_nhwc_tap_wgrad", 4, ptr(an2), ptr(S), ptr(dwt), B, C, H, W, ptr(ws), ws.numel(), stream_ptr())
    wr = torch.zeros(C, 1, 4, 4, device="cuda", dtype=torch.float64, requires_grad=True)
    F.conv_transpose2d(_bf(a_t).double(), wr, None, stride=2, padding=1).backward(_bf(S).double())
    torch.testing.assert_close(dwt.double(), wr.grad, rtol=2e-5, atol=2e-5 * float(wr.grad.abs().max()))
    dw3, db3 = torch.full((C, 1, 3, 3), float("nan"), device="cuda"), torch.full((C,), float("nan"), device="cuda")
    ws = torch.empty(L.size("vc_nhwc_conv1_wgrad_bias_workspace_bytes", C), dtype=torch.uint8, device="cuda")
    for _ in range(2):
        prev = dw3.clone()
        L.call("vc_nhwc_conv1_wgrad_bias", ptr(an2), ptr(S), ptr(dw3), ptr(db3), B, 2 * H, 2 * W, C, ptr(ws), ws.numel(),
               stream_ptr())
    assert torch.equal(prev, dw3)
    wr = torch.zeros(C, 1, 3, 3, device="cuda", dtype=torch.float64, requires_grad=True)
    F.conv2d(_bf(S).double(), wr, None, stride=2, padding=1).backward(_bf(a_t).double())
    torch.testing.assert_close(dw3.double(), wr.grad, rtol=2e-5, atol=2e-5 * float(wr.grad.abs().max()))
    wantb = _bf(a_t).double().sum(dim=(0, 2, 3))
    torch.testing.assert_close(db3.double(), wantb, rtol=1e-5, atol=1e-5 * float(wantb.abs().max()) + 1e-4)


@pytest.mark.parametrize("B,C,H,W,with_g", [(2, 128, 3, 128, False), (3, 64, 2, 256, True), (1, 32, 5, 128, False),
                                            (2, 128, 2, 64, False)])     # the last shape takes the 64-pixel kernel
def test_relu_backward_to_nhwc(B, C, H, W, with_g):
    """vc_relu_bwd_bias_nhwc: g = dy * (y > 0) as bf16 NHWC (+ optionally fp32 NCHW) and its per-channel sums."""
    from vaeclust._lib import lib, ptr, stream_ptr
    L = lib()
    dy, y = _r(B, C, H, W, seed=41), _r(B, C, H, W, seed=42)
    gn = torch.full((B, H, W, C), float("nan"), dtype=torch.bfloat16, device="cuda")
    g = torch.full((B, C, H, W), float("nan"), device="cuda") if with_g else None
    db = torch.full((C,), float("nan"), device="cuda")
    ws = torch.empty(L.size("vc_relu_bwd_bias_nhwc_workspace_bytes", C), dtype=torch.uint8, device="cuda")
    L.call("vc_relu_bwd_bias_nhwc", ptr(dy), ptr(y), B, C, H, W, ptr(g), ptr(gn), ptr(db), ptr(ws), ws.numel(), stream_ptr())
    want = dy * (y > 0)
    assert torch.equal(gn, _nhwc(want))
    if with_g:
        assert torch.equal(g, want)
    wantb = want.double().sum(dim=(0, 2, 3))
    torch.testing.assert_close(db.double(), wantb, rtol=1e-5, atol=1e-5 * float(wantb.abs().max()) + 1e-5)


def test_single_channel_ends_and_helpers():
    from vaeclust._lib import lib, ptr, stream_ptr
    L = lib()
    B, H, W, C = 2, 16, 256, 32
    # Conv2d(1 -> 32) forward into NHWC
    x, w, b = _r(B, 1, H, W, seed=1), _r(C, 1, 3, 3, seed=2, scale=1 / 3), _r(C, seed=3)
    y = torch.empty(B, H // 2, W // 2, C, dtype=torch.bfloat16, device="cuda")
    L.call("vc_nhwc_conv1_fwd", ptr(x), ptr(w), ptr(b), 1, ptr(y), B, H, W, C, stream_ptr())
    want = F.relu(F.conv2d(_bf(x), _bf(w), b, stride=2, padding=1))
    _check(y, None, want)
    # ConvTranspose2d(32 -> 1) data gradient + mask into NHWC
    dy, wt, act = _r(B, 1, 2 * H, 2 * W, seed=4), _r(C, 1, 4, 4, seed=5, scale=0.25), _r(B, C, H, W, seed=6)
    g = torch.empty(B, H, W, C, dtype=torch.bfloat16, device="cuda")
    an = _nhwc(act)
    L.call("vc_nhwc_convT1_dgrad", ptr(dy), ptr(wt), ptr(an), ptr(g), B, C, H, W, stream_ptr())
    want = F.conv2d(_bf(dy), _bf(wt), None, stride=2, padding=1) * (_bf(act) > 0)
    _check(g, None, want)
    # ConvTranspose2d(32 -> 1) forward from NHWC
    xin, bb = _r(B, C, H, W, seed=9), _r(1, seed=10)
    xn = _nhwc(xin)
    yo = torch.empty(B, 1, 2 * H, 2 * W, device="cuda")
    L.call("vc_nhwc_convT1_fwd", ptr(xn), ptr(wt), ptr(bb), ptr(yo), B, C, H, W, stream_ptr())
    want = F.conv_transpose2d(_bf(xin), _bf(wt), bb, stride=2, padding=1)
    _check(None, yo, want)
    # weight gradients of both ends from NHWC
    for KS in (3, 4):
        a_t = _r(B, C, H, W, seed=11)
        S = _r(B, 1, 2 * H, 2 * W, seed=12)
        an2 = _nhwc(a_t)
        dwt = torch.full((C, 1, KS, KS), float("nan"), device="cuda")
        ws = torch.empty(L.size("vc_nhwc_tap_wgrad_workspace_bytes", C, KS), dtype=torch.uint8, device="cuda")
        L.call("vc_nhwc_tap_wgrad", KS, ptr(an2), ptr(S), ptr(dwt), B, C, H, W, ptr(ws), ws.numel(), stream_ptr())
        if KS == 3:      # Conv2d(1 -> C): a = output gradient on the H x W grid, S = input [2H, 2W]
            wr = torch.zeros(C, 1, 3, 3, device="cuda", requires_grad=True)
            F.conv2d(_bf(S), wr, None, stride=2, padding=1).backward(_bf(a_t))
        else:            # ConvTranspose2d(C -> 1): a = layer input, S = output gradient
            wr = torch.zeros(C, 1, 4, 4, device="cuda", requires_grad=True)
            F.conv_transpose2d(_bf(a_t), wr, None, stride=2, padding=1).backward(_bf(S))
        torch.testing.assert_close(dwt, wr.grad, rtol=2e-5, atol=2e-5 * float(wr.grad.abs().max()))
    # Conv2d(1 -> C) weight + bias gradient in one pass (fp32 NCHW dy)
    dyc, xc = _r(B, C, H, W, seed=13), _r(B, 1, 2 * H, 2 * W, seed=14)
    dwc, dbc = torch.full((C, 1, 3, 3), float("nan"), device="cuda"), torch.full((C,), float("nan"), device="cuda")
    ws = torch.empty(L.size("vc_conv1_wgrad_bias_workspace_bytes", C), dtype=torch.uint8, device="cuda")
    L.call("vc_conv1_wgrad_bias", ptr(dyc), ptr(xc), ptr(dwc), ptr(dbc), B, 2 * H, 2 * W, C, 1, ptr(ws), ws.numel(), stream_ptr())
    wr = torch.zeros(C, 1, 3, 3, device="cuda", requires_grad=True)
    F.conv2d(_bf(xc), wr, None, stride=2, padding=1).backward(_bf(dyc))
    torch.testing.assert_close(dwc, wr.grad, rtol=2e-5, atol=2e-5 * float(wr.grad.abs().max()))
    wantb = dyc.double().sum(dim=(0, 2, 3)).float()
    torch.testing.assert_close(dbc, wantb, rtol=1e-5, atol=1e-5 * float(wantb.abs().max()) + 1e-4)
    # layout conversion
    src = _r(B, 128, 4, 128, seed=7)
    dst = torch.empty(B, 4, 128, 128, dtype=torch.bfloat16, device="cuda")
    L.call("vc_nhwc_from_nchw", ptr(src), ptr(dst), B, 128, 4, 128, stream_ptr())
    torch.cuda.synchronize()
    assert torch.equal(dst, _nhwc(src))
    # bias gradient of an NHWC tensor
    for Cc, rows in ((32, 70001), (64, 4099), (128, 513)):
        gm = _r(rows, Cc, seed=8).to(torch.bfloat16)
        db = torch.empty(Cc, device="cuda")
        ws = torch.empty(L.size("vc_nhwc_bias_grad_workspace_bytes", Cc), dtype=torch.uint8, device="cuda")
        L.call("vc_nhwc_bias_grad", ptr(gm), rows, Cc, ptr(db), ptr(ws), ws.numel(), stream_ptr())
        wantb = gm.double().sum(0).float()
        torch.testing.assert_close(db, wantb, rtol=1e-5, atol=1e-5 * float(wantb.abs().max()) + 1e-4)


@pytest.mark.parametrize("kind,B,Cin,H,W,Cout", [
    ("conv", 2, 32, 8, 256, 64),      # conv.2: Q = 32-channel input through the pair view, 6 boxes, one CTA group
    ("conv", 3, 64, 6, 128, 128),     # conv.4: P = 128 channels (two boxes), tap rows split over three CTA groups
    ("convT", 2, 128, 4, 64, 64),     # deconv.0: 16 taps x 64 = 1024 columns -> two groups of 512
    ("convT", 2, 64, 5, 128, 32),     # deconv.2: Q = 32-channel gradient, pair view, 12 boxes -> two groups
    ("convT", 1, 64, 3, 192, 32),
])
def test_weight_gradients(kind, B, Cin, H, W, Cout):
    from vaeclust._lib import lib, ptr, stream_ptr
    L = lib()
    x = _r(B, Cin, H, W, seed=1)
    if kind == "conv":
        op, w = 5, _r(Cout, Cin, 3, 3, seed=2)
        g = _r(B, Cout, H // 2, W // 2, seed=3)
        wr = w.clone().requires_grad_(True)
        F.conv2d(_bf(x), wr, None, stride=2, padding=1).backward(_bf(g))
    else:
        op, w = 8, _r(Cin, Cout, 4, 4, seed=2)
        g = _r(B, Cout, 2 * H, 2 * W, seed=3)
        wr = w.clone().requires_grad_(True)
        F.conv_transpose2d(_bf(x), wr, None, stride=2, padding=1).backward(_bf(g))
    want = wr.grad
    assert L.size("vc_conv_nhwc_wgrad_supported", op, B, Cin, H, W, Cout) == 1
    ws = torch.empty(L.size("vc_conv_nhwc_wgrad_workspace_bytes", op, B, Cin, H, W, Cout), dtype=torch.uint8, device="cuda")
    gn, xn = _nhwc(g), _nhwc(x)      # named: a temporary would be freed (and its block reused) before the launch runs
    for _ in range(2):        # twice: the result must not depend on what the workspace held
        dw = torch.full_like(w, float("nan"))
        L.call("vc_conv_nhwc_wgrad", op, ptr(gn), ptr(xn), ptr(dw), B, Cin, H, W, Cout, ptr(ws), ws.numel(),
               stream_ptr())
        torch.cuda.synchronize()
        torch.testing.assert_close(dw, want, rtol=2e-5, atol=2e-5 * float(want.abs().max()))
        ws.fill_(255)


@pytest.mark.parametrize("ends_nhwc", [False, True])
def test_conv_stacks_match_the_per_layer_path(ends_nhwc):
    """The NHWC conv stacks (vaeclust.convstack) inside the script-12 model against the per-layer fp32-NCHW path: the same
    bf16 products and fp32 accumulation, so loss and gradients agree to summation order; only the hidden-layer bias
    gradients differ by the bf16 rounding of the gradients they sum."""
    from vaeclust import convstack, nn, ops
    ops.set_precision("bf16")
    try:
        torch.manual_seed(0)
        model = nn.ConvMultiModalVAE(1, (16, 1024), 32, 384).cuda()
        B = 2
        x, lyr = _r(B, 1, 16, 1024, seed=1), _r(B, 384, seed=2)
        m = torch.tensor([[1.0], [0.0]], device="cuda")
        eps = _r(B, 32, seed=3)
        out = {}
        ends_default = convstack.ENDS_NHWC
        convstack.ENDS_NHWC = ends_nhwc
        for flag in (False, True):
            convstack.ENABLED = flag
            used = []
            real = convstack.encoder_convs
            model.zero_grad()
            mu, lv = model.encode(x, lyr, m)
            z, klr = ops.reparam_kl(mu, lv, eps)
            xhat = model.audio_dec(z)
            loss = ((xhat - x) ** 2).mean() + 1e-3 * klr.mean()
            loss.backward()
            out[flag] = (loss.detach().clone(), xhat.detach().clone(),
                         {k: p.grad.detach().clone() for k, p in model.named_parameters() if p.grad is not None})
        assert convstack.encoder_supported(x, [model.audio_enc.conv[i] for i in (0, 2, 4)])
        torch.testing.assert_close(out[True][0], out[False][0], rtol=1e-5, atol=0)
        torch.testing.assert_close(out[True][1], out[False][1], rtol=1e-4, atol=1e-5)
        assert out[True][2].keys() == out[False][2].keys()
        for k, g in out[False][2].items():
            rel = float((out[True][2][k] - g).norm() / (g.norm() + 1e-30))
            assert rel < (2e-2 if k.endswith("bias") else 5e-3), (k, rel)
    finally:
        convstack.ENABLED = True
        convstack.ENDS_NHWC = ends_default
        ops.set_precision("fp32")


@pytest.mark.parametrize("graphed", [False, True])
def test_fused_fc_weight_gradient_adam_matches_plain_step(graphed, monkeypatch):
    """vc_linear_wgrad_adam (the wide Linear weights updated from the weight-gradient kernel's accumulator, gradient never
    written) against the ordinary wgrad + vc_adam_multi step: same arithmetic order, so parameters and moments after three
    steps agree to fp32 summation order of the GEMM."""
    from vaeclust import nn, ops, train
    monkeypatch.setenv("VAECLUST_FUSE_FC_ADAM_NUMEL", str(1 << 22))     # 64 x 1024 model: FC layers are 65536 x 256
    ops.set_precision("bf16")
    try:
        B = 64      # the skinny weight-gradient kernel wants the big operand to dominate: 65536 x 64 >= 2^22
        x, lyr = _r(B, 1, 64, 1024, seed=1), _r(B, 384, seed=2)
        m = torch.ones(B, 1, device="cuda")
        eps = _r(B, 32, seed=3)
        res = {}
        for fuse in (False, True):
            torch.manual_seed(0)
            model = nn.ConvMultiModalVAE(1, (64, 1024), 32, 384).cuda()
            if graphed:
                step = train.GraphedVAEStep(model, "conv12", lr=2e-3, beta=1e-3, warmup=1)
                step.step.fuse_fc_adam = fuse
            else:
                step = train.VAEStep(model, "conv12", lr=2e-3, beta=1e-3)
                step.fuse_fc_adam = fuse
            assert len((step.step if graphed else step)._wide) == 2
            losses = [float(step(x, lyr, m, eps=eps)[0])]
            torch.cuda.synchronize()
            opt = step.opt
            wide = (step.step if graphed else step)._wide
            # parameters and moments after ONE step (later steps start from last-bit-different weights and Adam's
            # m / (sqrt(v) + eps) amplifies that for near-zero gradients); the loss curve over three
            snap = ({k: p.detach().clone() for k, p in model.named_parameters()},
                    [opt.state[id(p)]["exp_avg_sq"].clone() for p in wide], [p.grad is None for p in wide])
            losses += [float(step(x, lyr, m, eps=eps)[0]) for _ in range(2)]
            res[fuse] = (losses,) + snap
        assert res[True][3] == [True, True] and res[False][3] == [False, False]
        torch.testing.assert_close(torch.tensor(res[True][0]), torch.tensor(res[False][0]), rtol=1e-4, atol=0)
        for k, p in res[False][1].items():
            # (lr = 2e-3: where |g| is at Adam's eps = 1e-8 scale, m / (sqrt(v) + eps) amplifies a last-bit difference of the
            #  contracted multiply-adds, so a handful of the 33 M elements differ by up to ~1e-3 of one update)
            torch.testing.assert_close(res[True][1][k], p, rtol=1e-4, atol=1e-5, msg=lambda s, k=k: f"{k}: {s}")
        for a, b in zip(res[True][2], res[False][2]):
            torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-12)
    finally:
        ops.set_precision("fp32")


@pytest.mark.parametrize("Cin,Cout", [(65536, 256), (256, 65536)])
def test_wide_linear_from_bf16_shadow(Cin, Cout):
    """vc_linear_bf16w (weight tile by TMA from a bf16 copy: K-major boxes forward, MN-major boxes for the data gradient)
    against vc_dense_op in bf16 mode on the fp32 weight: the same bf16 products, so only summation order differs."""
    from vaeclust._lib import lib, ptr, stream_ptr
    L = lib()
    B = 64
    w = _r(Cout, Cin, seed=1, scale=Cin ** -0.5)
    sh = w.to(torch.bfloat16)
    x, dy, b = _r(B, Cin, seed=2), _r(B, Cout, seed=3), _r(Cout, seed=4)
    for op, inp, bias, relu, oshape in ((0, x, b, 1, (B, Cout)), (1, dy, None, 0, (B, Cin))):
        assert L.size("vc_linear_bf16w_supported", op, B, Cin, Cout) == 1
        want = torch.full(oshape, float("nan"), device="cuda")
        ws = torch.empty(L.size("vc_dense_op_workspace_bytes", op, 1, B, Cin, 1, 1, Cout), dtype=torch.uint8, device="cuda")
        L.call("vc_dense_op", op, 1, ptr(inp), ptr(w), ptr(bias), ptr(want), B, Cin, 1, 1, Cout, relu, ptr(ws), ws.numel(),
               stream_ptr())
        got = torch.full(oshape, float("nan"), device="cuda")
        ws2 = torch.empty(L.size("vc_linear_bf16w_workspace_bytes", op, B, Cin, Cout), dtype=torch.uint8, device="cuda")
        L.call("vc_linear_bf16w", op, ptr(inp), ptr(sh), ptr(bias), ptr(got), B, Cin, Cout, relu, ptr(ws2), ws2.numel(),
               stream_ptr())
        torch.cuda.synchronize()
        torch.testing.assert_close(got, want, rtol=2e-5, atol=2e-5 * float(want.abs().max()))
        ref = _bf(inp) @ (_bf(w).t() if op == 0 else _bf(w))
        if op == 0:
            ref = torch.relu(ref + b)
        torch.testing.assert_close(got, ref, rtol=1e-4, atol=1e-4 * float(ref.abs().max()))


def test_shadows_follow_external_parameter_changes(monkeypatch):
    """A graph-replayed fused step must see a parameter change made between steps by something else (here an in-place
    scaling, as load_state_dict would do): the bf16 shadow is rebuilt before the replay."""
    from vaeclust import nn, ops, train
    monkeypatch.setenv("VAECLUST_FUSE_FC_ADAM_NUMEL", str(1 << 22))
    ops.set_precision("bf16")
    try:
        B = 64
        x, lyr = _r(B, 1, 64, 1024, seed=1), _r(B, 384, seed=2)
        m = torch.ones(B, 1, device="cuda")
        eps = _r(B, 32, seed=3)
        losses = {}
        for fuse in (False, True):
            torch.manual_seed(0)
            model = nn.ConvMultiModalVAE(1, (64, 1024), 32, 384).cuda()
            step = train.GraphedVAEStep(model, "conv12", lr=2e-3, beta=1e-3, warmup=1)
            step.step.fuse_fc_adam = fuse
            out = [float(step(x, lyr, m, eps=eps)[0]) for _ in range(3)]       # eager, capture, replay
            with torch.no_grad():
                model.audio_dec.fc[2].weight.add_(0.02)      # shifts every decoder activation: the loss moves visibly
                model.audio_enc.fc[0].weight.mul_(1.5)
            out.append(float(step(x, lyr, m, eps=eps)[0]))                       # replay after the change
            losses[fuse] = out
        assert abs(losses[True][3] - losses[True][2]) > 1e-3 * abs(losses[True][2])      # the change is visible
        torch.testing.assert_close(torch.tensor(losses[True]), torch.tensor(losses[False]), rtol=2e-4, atol=0)
    finally:
        ops.set_precision("fp32")
