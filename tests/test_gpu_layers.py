"""GPU: every CUDA op against the plain torch fp32 op it replaces (forward values and all gradients), on the
reference's layer shapes scaled down, including odd spatial sizes (T=646 -> 323 -> 162 -> 81 path of script 12)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

RTOL, ATOL = 2e-4, 2e-5  # fp32 FFMA in a different summation order than cuDNN/cuBLAS


def _dev():
    return torch.device("cuda")


def _rand(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(_dev())


def _check(fn_ours, fn_ref, inputs, names):
    ins_a = [t.clone().requires_grad_(t.dtype.is_floating_point and r) for t, r in inputs]
    ins_b = [t.clone().requires_grad_(t.dtype.is_floating_point and r) for t, r in inputs]
    ya, yb = fn_ours(*ins_a), fn_ref(*ins_b)
    torch.testing.assert_close(ya, yb, rtol=RTOL, atol=ATOL)
    g = torch.randn_like(yb)
    ya.backward(g)
    yb.backward(g)
    for a, b, n, (_, r) in zip(ins_a, ins_b, names, inputs):
        if r:
            scale = float(b.grad.abs().max().clamp_min(1e-6))
            torch.testing.assert_close(a.grad, b.grad, rtol=RTOL, atol=ATOL * max(1.0, scale), msg=lambda m: f"{n}: {m}")


@pytest.mark.parametrize("M,N,K,relu", [(64, 256, 470, True), (17, 16, 256, False), (3, 700, 33, True),
                                        (256, 464, 256, False), (5, 256, 4100, True)])
def test_linear(M, N, K, relu):
    from vaeclust import ops
    x, w, b = _rand(M, K, seed=1), _rand(N, K, seed=2, scale=K ** -0.5), _rand(N, seed=3)
    ref = lambda x, w, b: F.relu(F.linear(x, w, b)) if relu else F.linear(x, w, b)
    _check(lambda x, w, b: ops.linear(x, w, b, relu), ref, [(x, True), (w, True), (b, True)], ["x", "w", "b"])


@pytest.mark.parametrize("B,Cin,H,W,Cout,relu", [(2, 1, 16, 22, 32, True), (3, 32, 8, 11, 64, True),
                                                 (2, 64, 5, 7, 128, False), (1, 3, 9, 323, 8, True)])
def test_conv2d_s2k3(B, Cin, H, W, Cout, relu):
    from vaeclust import ops
    x, w, b = _rand(B, Cin, H, W, seed=1), _rand(Cout, Cin, 3, 3, seed=2, scale=(Cin * 9) ** -0.5), _rand(Cout, seed=3)
    def ref(x, w, b):
        y = F.conv2d(x, w, b, stride=2, padding=1)
        return F.relu(y) if relu else y
    _check(lambda x, w, b: ops.conv2d_s2k3(x, w, b, relu), ref, [(x, True), (w, True), (b, True)], ["x", "w", "b"])


@pytest.mark.parametrize("B,Cin,H,W,Cout,relu", [(2, 128, 2, 3, 64, True), (3, 64, 4, 6, 32, True),
                                                 (2, 32, 8, 12, 1, False), (1, 5, 3, 81, 7, True)])
def test_convT2d_s2k4(B, Cin, H, W, Cout, relu):
    from vaeclust import ops
    x, w, b = _rand(B, Cin, H, W, seed=1), _rand(Cin, Cout, 4, 4, seed=2, scale=(Cin * 4) ** -0.5), _rand(Cout, seed=3)
    def ref(x, w, b):
        y = F.conv_transpose2d(x, w, b, stride=2, padding=1)
        return F.relu(y) if relu else y
    _check(lambda x, w, b: ops.convT2d_s2k4(x, w, b, relu), ref, [(x, True), (w, True), (b, True)], ["x", "w", "b"])


def test_cat_gate():
    from vaeclust import ops
    a, l = _rand(9, 32, seed=1), _rand(9, 128, seed=2)
    m = (torch.rand(9, generator=torch.Generator().manual_seed(3)) < 0.6).float().to(_dev())
    _check(lambda a, l: ops.cat_gate(a, l, m, True), lambda a, l: torch.cat([a, l * m[:, None], m[:, None]], 1),
           [(a, True), (l, True)], ["a", "l"])
    _check(lambda a, l: ops.cat_gate(a, l), lambda a, l: torch.cat([a, l], 1), [(a, True), (l, True)], ["a", "l"])


@pytest.mark.parametrize("B,z", [(64, 32), (7, 16), (256, 16)])
def test_reparam_kl(B, z):
    from vaeclust import ops
    mu, lv, eps = _rand(B, z, seed=1), _rand(B, z, seed=2, scale=0.5), _rand(B, z, seed=3)
    def ours(mu, lv):
        zz, kl = ops.reparam_kl(mu, lv, eps)
        return torch.cat([zz, kl[:, None]], 1)
    def ref(mu, lv):
        zz = mu + eps * torch.exp(0.5 * lv)
        kl = -0.5 * torch.sum(1 + lv - mu.pow(2) - lv.exp(), dim=1)
        return torch.cat([zz, kl[:, None]], 1)
    _check(ours, ref, [(mu, True), (lv, True)], ["mu", "lv"])
    _check(lambda mu, lv: ops.kl_rows(mu, lv), lambda mu, lv: -0.5 * torch.sum(1 + lv - mu.pow(2) - lv.exp(), dim=1),
           [(mu, True), (lv, True)], ["mu", "lv"])


@pytest.mark.parametrize("shape_hat,shape_x", [((4, 1, 16, 24), (4, 1, 16, 22)), ((3, 1, 128, 1024), (3, 1, 128, 1024)),
                                               ((32, 464), (32, 464)), ((5, 1, 7, 9), (5, 1, 7, 9))])
def test_mse_and_fused_vae_loss(shape_hat, shape_x):
    from vaeclust import ops
    from vaeclust._lib import VC_LOSS_MEAN_MEAN, VC_LOSS_SUM_MEAN
    xhat, x = _rand(*shape_hat, seed=1), _rand(*shape_x, seed=2)
    crop = (lambda t: t[..., : shape_x[-2], : shape_x[-1]]) if len(shape_x) == 4 else (lambda t: t)
    _check(lambda xh: ops.mse_loss(xh, x), lambda xh: F.mse_loss(crop(xh), x), [(xhat, True)], ["xhat"])
    B, z = shape_x[0], 16
    mu, lv, eps = _rand(B, z, seed=3), _rand(B, z, seed=4, scale=0.3), _rand(B, z, seed=5)
    for mode, beta in ((VC_LOSS_MEAN_MEAN, 1.0), (VC_LOSS_SUM_MEAN, 4.0)):
        a = [t.clone().requires_grad_(True) for t in (xhat, mu, lv)]
        b = [t.clone().requires_grad_(True) for t in (xhat, mu, lv)]
        _, klr = ops.reparam_kl(a[1], a[2], eps)
        loss, stats = ops.vae_loss(a[0], x, klr, beta, mode, z, unit_grad=True)
        loss.backward()
        d = crop(b[0]) - x
        klt = 1 + b[2] - b[1].pow(2) - b[2].exp()
        if mode == VC_LOSS_MEAN_MEAN:
            rec, kl = d.pow(2).mean(), -0.5 * klt.mean()
        else:
            rec, kl = d.pow(2).flatten(1).sum(1).mean(), (-0.5 * klt.sum(1)).mean()
        ref = rec + beta * kl
        ref.backward()
        torch.testing.assert_close(stats, torch.stack([ref, rec, kl]).detach(), rtol=1e-5, atol=1e-6)
        for p, q in zip(a, b):
            torch.testing.assert_close(p.grad, q.grad, rtol=1e-4, atol=1e-7)


def test_adam_matches_torch():
    from vaeclust.optim import Adam
    torch.manual_seed(0)
    shapes = [(300, 70), (5,), (65536 + 3,), (128, 64, 4, 4)]
    pa = [torch.nn.Parameter(torch.randn(*s, device=_dev())) for s in shapes]
    pb = [torch.nn.Parameter(p.detach().clone()) for p in pa]
    oa, ob = Adam(pa, lr=2e-3), torch.optim.Adam(pb, lr=2e-3)
    for step in range(4):
        for i, (p, q) in enumerate(zip(pa, pb)):
            if i == 1 and step < 2:   # a parameter without gradient is skipped and its step does not advance
                p.grad = q.grad = None
                continue
            g = torch.randn_like(p) * (10.0 ** (step - 2))
            p.grad, q.grad = g.clone(), g.clone()
        oa.step()
        ob.step()
    for p, q in zip(pa, pb):
        torch.testing.assert_close(p, q, rtol=2e-6, atol=1e-7)


# ------------------------------------------------------------------------------------ bf16 tensor-core path
@pytest.fixture
def bf16_mode():
    from vaeclust import ops
    ops.set_precision("bf16")
    yield
    ops.set_precision("fp32")


def _r(t):  # what the tensor-core path sees: operands rounded to bf16
    return t.bfloat16().float()


def _check_bf16(fn_ours, fn_ref, x, w, b):
    """Forward against the fp32 op evaluated on bf16-rounded operands (tight: only the summation order differs);
    gradients against fp32 autograd with a bf16-sized tolerance (dy is rounded once more in the backward GEMMs)."""
    xa, wa, ba = (t.clone().requires_grad_(True) for t in (x, w, b))
    xb, wb, bb = (t.clone().requires_grad_(True) for t in (x, w, b))
    ya = fn_ours(xa, wa, ba)
    with torch.no_grad():
        yr, yf = fn_ref(_r(x), _r(w), b), fn_ref(x, w, b)
    # tiny Linear layers stay on the fp32 FFMA kernel in bf16 mode (launch-bound): accept either arithmetic, tightly
    err_r, err_f = float((ya - yr).abs().max()), float((ya - yf).abs().max())
    assert min(err_r, err_f) <= 2e-4 * (1.0 + float(yf.abs().max())), (err_r, err_f)
    rr = _r if err_r <= err_f else (lambda t: t)
    yb = fn_ref(rr(xb), rr(wb), bb)   # same operands -> same ReLU mask (rounding is straight-through)
    g = torch.randn_like(yb)
    ya.backward(g)
    yb.backward(g)
    for p, q, n in ((xa, xb, "x"), (wa, wb, "w"), (ba, bb, "b")):
        scale = float(q.grad.abs().max().clamp_min(1e-6))
        torch.testing.assert_close(p.grad, q.grad, rtol=3e-2, atol=2e-2 * scale, msg=lambda m: f"{n}: {m}")


@pytest.mark.parametrize("kind,B,C,H,W", [("conv", 2, 32, 16, 24), ("conv", 3, 8, 17, 32), ("conv", 2, 32, 128, 1024),
                                          ("conv", 1, 4, 1, 8), ("convT", 2, 32, 8, 12), ("convT", 3, 32, 7, 40),
                                          ("convT", 2, 32, 64, 512), ("convT", 1, 4, 1, 4)])
def test_single_channel_wgrad_tap_kernel(kind, B, C, H, W, bf16_mode):
    """Weight gradients of the single-channel ends (Conv2d(1 -> C) k3 with W % 8 == 0, ConvTranspose2d(C -> 1) k4 with
    W % 4 == 0) take the streaming tap-correlation kernel, which multiplies the fp32 values (no bf16 rounding of dy
    and x).  Against float64 autograd the only difference is fp32 summation: 2e-5 of the gradient's scale."""
    from vaeclust import ops
    if kind == "conv":
        x, w = _rand(B, 1, H, W, seed=1), _rand(C, 1, 3, 3, seed=2, scale=1 / 3)
        ours = lambda x, w: ops.conv2d_s2k3(x, w, None, False)
        ref = lambda x, w: F.conv2d(x, w, None, stride=2, padding=1)
    else:
        x, w = _rand(B, C, H, W, seed=1), _rand(C, 1, 4, 4, seed=2, scale=(C * 4) ** -0.5)
        ours = lambda x, w: ops.convT2d_s2k4(x, w, None, False)
        ref = lambda x, w: F.conv_transpose2d(x, w, None, stride=2, padding=1)
    wa = w.clone().requires_grad_(True)
    ya = ours(x, wa)
    g = torch.randn_like(ya)
    ya.backward(g)
    wb = w.double().requires_grad_(True)
    ref(x.double(), wb).backward(g.double())
    scale = float(wb.grad.abs().max())
    assert float((wa.grad.double() - wb.grad).abs().max()) <= 2e-5 * scale
    # deterministic: a second evaluation gives the same bits
    wa2 = w.clone().requires_grad_(True)
    ours(x, wa2).backward(g)
    assert torch.equal(wa.grad, wa2.grad)


@pytest.mark.parametrize("B,Cin,H,W,relu", [(2, 32, 8, 12, False), (3, 32, 7, 40, True), (2, 32, 64, 128, False),
                                             (1, 5, 1, 4, True), (2, 8, 33, 256, True)])
def test_convT_to1_quad_kernel_equals_scalar_kernel(B, Cin, H, W, relu, bf16_mode):
    """ConvTranspose2d(Cin -> 1) forward: the 2 x 4-pixel-per-thread kernel (W % 4 == 0, 16-byte aligned input) keeps
    the FMA order of the 2 x 1 kernel, so a misaligned copy of the same input (which takes the scalar kernel) must give
    identical bits; both match torch on the bf16-rounded operands (the direct kernels belong to the bf16 mode)."""
    from vaeclust import ops
    x, w, b = _rand(B, Cin, H, W, seed=4), _rand(Cin, 1, 4, 4, seed=5, scale=(Cin * 4) ** -0.5), _rand(1, seed=6)
    store = torch.empty(x.numel() + 1, device=x.device)
    x_off = store[1:].view_as(x)
    x_off.copy_(x)
    assert x.data_ptr() % 16 == 0 and x_off.data_ptr() % 16 == 4
    with torch.no_grad():
        y_quad, y_scalar = ops.convT2d_s2k4(x, w, b, relu), ops.convT2d_s2k4(x_off, w, b, relu)
        ref = F.conv_transpose2d(_r(x), _r(w), b, stride=2, padding=1)
        ref = F.relu(ref) if relu else ref
    assert torch.equal(y_quad, y_scalar)
    torch.testing.assert_close(y_quad, ref, rtol=2e-5, atol=2e-6)


@pytest.mark.parametrize("M,N,K,relu", [(64, 256, 470, True), (17, 16, 256, False), (3, 700, 33, True),
                                        (64, 256, 8200, True), (64, 4100, 256, True),
                                        # latent-head / lyrics-branch shapes: the one-launch mma.sync kernel
                                        (64, 256, 384, True), (64, 256, 161, True), (64, 32, 256, False),
                                        (64, 256, 32, True), (64, 128, 256, True), (7, 33, 50, False), (200, 9, 17, True),
                                        # big-matrix shapes take the skinny kernels (fp32 weight read in place)
                                        (64, 256, 16640, True), (64, 16500, 256, True), (50, 256, 16644, False)])
def test_linear_bf16(M, N, K, relu, bf16_mode):
    from vaeclust import ops
    x, w, b = _rand(M, K, seed=1), _rand(N, K, seed=2, scale=K ** -0.5), _rand(N, seed=3)
    ref = lambda x, w, b: F.relu(F.linear(x, w, b)) if relu else F.linear(x, w, b)
    _check_bf16(lambda x, w, b: ops.linear(x, w, b, relu), ref, x, w, b)


@pytest.mark.parametrize("B,Cin,H,W,Cout,relu", [(2, 1, 16, 22, 32, True), (3, 32, 8, 11, 64, True),
                                                 (2, 64, 5, 7, 128, False), (1, 3, 9, 323, 8, True),
                                                 (4, 32, 128, 256, 32, True),    # 256 row tiles: persistent loop
                                                 (2, 1, 128, 646, 32, True)])    # script-12 first layer at T=646
def test_conv2d_s2k3_bf16(B, Cin, H, W, Cout, relu, bf16_mode):
    from vaeclust import ops
    x, w, b = _rand(B, Cin, H, W, seed=1), _rand(Cout, Cin, 3, 3, seed=2, scale=(Cin * 9) ** -0.5), _rand(Cout, seed=3)
    def ref(x, w, b):
        y = F.conv2d(x, w, b, stride=2, padding=1)
        return F.relu(y) if relu else y
    _check_bf16(lambda x, w, b: ops.conv2d_s2k3(x, w, b, relu), ref, x, w, b)


@pytest.mark.parametrize("kind,B,Cin,H,W,Cout", [
    ("conv", 2, 32, 16, 128, 64),     # Cp 64, N 288: two 144-column tiles, one k-block per CTA
    ("conv", 8, 32, 64, 256, 64),     # 512 k-blocks: seven per CTA, the ring wraps
    ("conv", 1, 64, 5, 128, 128),     # odd height (bottom row group off the image), Cp 128, three 192-column tiles
    ("conv", 1, 64, 4, 256, 64),      # Cp 64 with 192-column tiles
    ("conv", 1, 16, 2, 128, 128),     # Cp 128, a single 144-column tile
    ("convT", 2, 64, 4, 64, 32),      # Cp 64, two 256-column tiles
    ("convT", 8, 64, 32, 128, 32),    # seven k-blocks per CTA
    ("convT", 1, 128, 3, 128, 64),    # Cp 128, four tiles
    ("convT", 1, 64, 1, 64, 16)])     # one input row: both outer row groups off the image
def test_wgrad_fast_producers_equal_general_kernel(kind, B, Cin, H, W, Cout, bf16_mode):
    """Model-shaped weight gradients (64-pixel k-blocks inside one row, Cp in {64, 128}) take conv_wgrad_fast_kernel.
    Operands that start 4 bytes off an 8-byte boundary take the general table-driven kernel: both sum the same bf16
    products in fp32 (different split-K order only), so they agree to 1e-4 of the gradient's scale, and with float64
    autograd on the bf16-rounded operands."""
    from vaeclust import ops
    if kind == "conv":
        x, w = _rand(B, Cin, H, W, seed=1), _rand(Cout, Cin, 3, 3, seed=2, scale=(Cin * 9) ** -0.5)
        ours = lambda x, w: ops.conv2d_s2k3(x, w, None, False)
        ref = lambda x, w: F.conv2d(x, w, None, stride=2, padding=1)
    else:
        x, w = _rand(B, Cin, H, W, seed=1), _rand(Cin, Cout, 4, 4, seed=2, scale=(Cin * 4) ** -0.5)
        ours = lambda x, w: ops.convT2d_s2k4(x, w, None, False)
        ref = lambda x, w: F.conv_transpose2d(x, w, None, stride=2, padding=1)

    def off4(t):   # same values, data pointer 4 bytes past a 16-byte boundary
        store = torch.empty(t.numel() + 1, device=t.device)
        v = store[1:].view_as(t)
        v.copy_(t)
        assert v.data_ptr() % 8 == 4
        return v

    wa = w.clone().requires_grad_(True)
    ya = ours(x, wa)
    g = torch.randn_like(ya)
    ya.backward(g)
    wb = w.clone().requires_grad_(True)
    ours(off4(x), wb).backward(off4(g))
    wr = _r(w).double().requires_grad_(True)
    ref(_r(x).double(), wr).backward(_r(g).double())
    scale = float(wr.grad.abs().max())
    assert float((wa.grad - wb.grad).abs().max()) <= 1e-4 * scale
    assert float((wa.grad.double() - wr.grad).abs().max()) <= 1e-4 * scale


@pytest.mark.parametrize("B,Cin,H,W,Cout,relu", [(2, 128, 2, 3, 64, True), (3, 64, 4, 6, 32, True),
                                                 (2, 32, 8, 12, 1, False), (1, 5, 3, 81, 7, True),
                                                 (2, 32, 64, 160, 32, True),     # 160 row tiles: persistent loop
                                                 (2, 128, 16, 81, 64, True),     # script-12 first deconv at T=646
                                                 (2, 32, 64, 162, 1, False),     # script-12 last deconv at T=646
                                                 (3, 32, 7, 40, 1, True)])       # odd height: half-empty row pair
def test_convT2d_s2k4_bf16(B, Cin, H, W, Cout, relu, bf16_mode):
    from vaeclust import ops
    x, w, b = _rand(B, Cin, H, W, seed=1), _rand(Cin, Cout, 4, 4, seed=2, scale=(Cin * 4) ** -0.5), _rand(Cout, seed=3)
    def ref(x, w, b):
        y = F.conv_transpose2d(x, w, b, stride=2, padding=1)
        return F.relu(y) if relu else y
    _check_bf16(lambda x, w, b: ops.convT2d_s2k4(x, w, b, relu), ref, x, w, b)


@pytest.mark.parametrize("kind,B,Cin,H,W,Cout", [
    ("conv", 2, 32, 8, 512, 64),      # script-12 conv.2: dy 64 channels, rows of 256 pixels, phase-merged N = 128
    ("conv", 3, 64, 6, 256, 128),     # conv.4: dy 128 channels = two 64-channel chunks per tap, N = 256
    ("convT", 2, 128, 4, 128, 64),    # deconv.0: stride-2 traversal of the tensor map, 16 taps
    ("convT", 2, 64, 5, 256, 32),     # deconv.2: 32-channel source through the pixel-pair view (12 k-blocks)
    ("convT", 1, 64, 3, 128, 32),
])
def test_tma_fed_data_gradients(kind, B, Cin, H, W, Cout, bf16_mode):
    """The TMA-fed data-gradient path (vc_relu_bwd_bias_nhwc + vc_conv_dgrad_nhwc: bf16 NHWC copy of g = dy * (y > 0),
    4-D tensor-map boxes with zero-filled padding and traversal stride) against (a) the gather kernel it replaces --
    the same bf16 products, so dx agrees to fp32 summation order and g / db / dw bit for bit or nearly -- and (b) fp32
    autograd on bf16-rounded operands."""
    from vaeclust import ops
    if kind == "conv":
        x, w, b = _rand(B, Cin, H, W, seed=1), _rand(Cout, Cin, 3, 3, seed=2, scale=(Cin * 9) ** -0.5), _rand(Cout, seed=3)
        ours = lambda x, w, b: ops.conv2d_s2k3(x, w, b, True)
        ref = lambda x, w, b: F.relu(F.conv2d(x, w, b, stride=2, padding=1))
    else:
        x, w, b = _rand(B, Cin, H, W, seed=1), _rand(Cin, Cout, 4, 4, seed=2, scale=(Cin * 4) ** -0.5), _rand(Cout, seed=3)
        ours = lambda x, w, b: ops.convT2d_s2k4(x, w, b, True)
        ref = lambda x, w, b: F.relu(F.conv_transpose2d(x, w, b, stride=2, padding=1))
    grads = {}
    calls = {}
    L = ops.lib()
    real_call = L.call
    for flag in (True, False):
        ops.USE_TMA_DGRAD = flag
        names = []
        L.call = lambda name, *a, _n=names: (_n.append(name), real_call(name, *a))[1]
        try:
            xa, wa, ba = (t.clone().requires_grad_(True) for t in (x, w, b))
            y = ours(xa, wa, ba)
            gsrc = torch.randn(y.shape, generator=torch.Generator().manual_seed(5)).cuda()
            y.backward(gsrc)
        finally:
            ops.USE_TMA_DGRAD = False
            L.call = real_call
        grads[flag] = (xa.grad.clone(), wa.grad.clone(), ba.grad.clone())
        calls[flag] = names
    assert "vc_conv_dgrad_nhwc" in calls[True] and "vc_conv_dgrad_nhwc" not in calls[False]
    sx = float(grads[False][0].abs().max())
    torch.testing.assert_close(grads[True][0], grads[False][0], rtol=1e-4, atol=1e-5 * sx)
    torch.testing.assert_close(grads[True][1], grads[False][1], rtol=1e-5, atol=1e-6 * float(grads[False][1].abs().max()))
    torch.testing.assert_close(grads[True][2], grads[False][2], rtol=1e-5, atol=1e-5 * float(grads[False][2].abs().max()))
    ops.USE_TMA_DGRAD = True
    try:
        _check_bf16(ours, ref, x, w, b)
    finally:
        ops.USE_TMA_DGRAD = False
