"""GPU: the fused MLP-VAE train step (vc_mlpvae_step: pack -> rows (tcgen05 fwd/bwd) -> weight gradients -> Adam) of
scripts 19 / 06 / 22 against (a) the golden fixtures of the UNMODIFIED reference classes, at the stated bf16 tolerance,
(b) the oracle computing in the same arithmetic (bf16 operands, fp32 accumulate): loss and EVERY parameter's gradient,
on ragged multi-tile batches too, (c) itself when replayed from a CUDA graph and when split for data-parallel use."""
import numpy as np
import pytest
import torch

from oracle import make_golden as mg
from oracle import parity, vae_ref

pytestmark = pytest.mark.gpu
BF16_LOSS_RTOL = 2e-2   # vs the fp32 reference run (operands rounded to 8 mantissa bits)


def _build(name):
    from vaeclust import nn
    c = mg.CASES[name]
    if c["kind"] == "mlp19":
        m = nn.MLPVAE(c["input_dim"], c["latent"], c["hidden"], cond_dim=c["cond"], conditional=bool(c["cond"]))
    elif c["kind"] == "mlp06":
        m = nn.EasyMLPVAE(c["input_dim"], c["hidden"], c["latent"])
    else:
        m = nn.AE(c["input_dim"], c["latent"], c["hidden"])
    m.load_state_dict(mg.case_params(name))
    return m.cuda().train(), c["kind"], c


def _args(b, c):
    return (b["x"], b["c"]) if c.get("cond") else (b["x"],)


@pytest.mark.parametrize("graphed", [False, True])
@pytest.mark.parametrize("name", ["cvae19", "beta19", "easy06", "ae22"])
def test_fused_steps_follow_reference_golden(name, graphed, golden_dir):
    from vaeclust.fused import FusedMLPStep
    from vaeclust.train import GraphedVAEStep
    gold = np.load(golden_dir / f"vae_{name}.npz")
    model, kind, c = _build(name)
    step = (GraphedVAEStep(model, kind, lr=c["lr"], beta=c["beta"], warmup=1, fused="always") if graphed
            else FusedMLPStep(model, kind, lr=c["lr"], beta=c["beta"]))
    for s in range(c["steps"]):
        b = {k: v.cuda() for k, v in mg.case_inputs(name, s).items()}
        stats = step(*_args(b, c), eps=b.get("eps")).cpu().numpy()
        assert stats[0] == pytest.approx(float(gold["loss"][s]), rel=BF16_LOSS_RTOL), s
        assert stats[1] == pytest.approx(float(gold["recon"][s]), rel=BF16_LOSS_RTOL), s
    # Adam moves every element by ~lr per step whatever the gradient's size, so an element whose bf16 gradient has the
    # other sign ends 2*lr*steps away: 1e-2 relative on the |.|-sums of the small tensors
    for n_, p in model.named_parameters():
        got = float(p.detach().double().abs().sum())
        assert got == pytest.approx(gold["param/" + n_][1], rel=1e-2, abs=1e-6), n_
    if graphed:
        assert step.graph is not None
        assert max(st["step"] for st in step.opt.state_dict()["state"].values()) == c["steps"]


@pytest.mark.parametrize("kind,B,D,cond,Z", [
    ("mlp19", 32, 464, 6, 16),      # one partial tile
    ("mlp19", 300, 464, 6, 16),     # three tiles, ragged tail
    ("mlp19", 1111, 464, 0, 16),    # beta-VAE (no conditioning), 9 tiles
    ("mlp19", 256, 464, 12, 32),    # genre+language conditioning, latent 32 (heads use both accumulator chunks)
    ("mlp06", 128, 80, 0, 16),      # script 06: one output half, mean/mean loss
    ("mlp06", 500, 80, 0, 8),
    ("ae22", 256, 464, 0, 16),      # script 22 AE: no logvar head, no KL
])
def test_fused_step_matches_same_arithmetic_oracle(kind, B, D, cond, Z):
    """Loss within 1e-4 and every parameter gradient within 2e-2 (relative Frobenius) of the oracle that rounds the
    operands of each contraction to bf16 at the same points; the fp32 oracle is reported alongside (oracle/parity.py)."""
    from vaeclust import nn
    from vaeclust.fused import FusedMLPStep
    g = torch.Generator().manual_seed(B + D + Z)
    if kind == "mlp19":
        model = nn.MLPVAE(D, Z, 256, cond_dim=cond, conditional=cond > 0)
        model.load_state_dict(vae_ref.init_mlp_vae_params(D, Z, 256, cond, seed=5, style="19"))
    elif kind == "mlp06":
        model = nn.EasyMLPVAE(D, 256, Z)
        model.load_state_dict(vae_ref.init_mlp_vae_params(D, Z, 256, 0, seed=5, style="06"))
    else:
        model = nn.AE(D, Z, 256)
        model.load_state_dict(vae_ref.init_ae_params(D, Z, 256, seed=5))
    model.cuda().train()
    x = torch.randn(B, D, generator=g)
    if kind != "mlp06":
        x[:, :80] *= torch.logspace(1.0, 0.0, 80)
        x[:, 80:] *= 0.05
    batch = {"x": x, "eps": torch.randn(B, Z, generator=g)}
    if cond:
        batch["c"] = torch.nn.functional.one_hot(torch.randint(0, cond, (B,), generator=g), cond).float()
    batch = {k: v.cuda() for k, v in batch.items()}
    beta = 4.0 if kind == "mlp19" else 1.0
    step = FusedMLPStep(model, kind, lr=1e-3, beta=beta)
    r = parity.vae_step_parity(model, step, kind, batch, beta, 1e-3, "bf16")
    assert r["ok"], {k: v for k, v in r.items() if k != "grad_rel"} | {"grad_rel": r["grad_rel"]}


def test_fused_step_split_for_data_parallel_equals_single_call():
    """phases 3 (+ all-reduce of the flat block, here the identity) + 4 must equal phases 7 bit for bit."""
    import ctypes
    from vaeclust import nn
    from vaeclust._lib import lib, ptr, stream_ptr
    from vaeclust.fused import FusedMLPStep
    outs = []
    for split in (False, True):
        model, kind, c = _build("cvae19")
        step = FusedMLPStep(model, kind, lr=c["lr"], beta=c["beta"])
        if split:
            step.multi = True
            import vaeclust.dist as vd
            real = vd.all_reduce_sum
            vd.all_reduce_sum = lambda t, group=None: t
        try:
            for s in range(2):
                b = {k: v.cuda() for k, v in mg.case_inputs("cvae19", s).items()}
                st = step(b["x"], b["c"], eps=b["eps"], global_batch=b["x"].shape[0])
        finally:
            if split:
                vd.all_reduce_sum = real
        torch.cuda.synchronize()
        outs.append((st.clone(), {n: p.detach().clone() for n, p in model.named_parameters()}))
    assert torch.equal(outs[0][0], outs[1][0])
    for n in outs[0][1]:
        assert torch.equal(outs[0][1][n], outs[1][1][n]), n


def test_external_weight_changes_are_picked_up():
    """load_state_dict between steps must refresh the bf16 shadows the kernels read (version counters)."""
    from vaeclust.fused import FusedMLPStep
    model, kind, c = _build("easy06")
    step = FusedMLPStep(model, kind, lr=c["lr"], beta=c["beta"])
    b = {k: v.cuda() for k, v in mg.case_inputs("easy06", 0).items()}
    first = float(step(b["x"], eps=b["eps"]).cpu()[0])
    model.load_state_dict(mg.case_params("easy06"))      # back to the initial weights
    again = float(step(b["x"], eps=b["eps"]).cpu()[0])
    assert again == pytest.approx(first, rel=1e-6)
