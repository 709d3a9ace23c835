"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by running the UNMODIFIED reference
classes (imported from /root/reference/scripts by oracle/ref_loader.py) and, in the same
pass, asserts that the restatement in oracle/vae_ref.py reproduces them.  Run in the build
container only:   python oracle/make_golden.py

Inputs are regenerated from seeds by ``case_inputs`` (so the fixtures stay small: they hold
the reference's OUTPUTS -- per-step losses, mu/logvar, per-parameter gradient and post-Adam
parameter checksums), parameters come from oracle.vae_ref.init_*_params(seed) loaded into
the reference module via load_state_dict, eps is injected by patching torch.randn_like.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from oracle import vae_ref  # noqa: E402

GOLDEN = ROOT / "tests" / "golden"
CASES = {
    # name: kind, dict of hyper-parameters
    "conv_16x22": dict(kind="conv", feat_hw=(16, 22), B=4, z=32, lyr=384, beta=1.0, lr=2e-3, steps=3),
    "conv_32x40": dict(kind="conv", feat_hw=(32, 40), B=3, z=16, lyr=384, beta=0.5, lr=2e-3, steps=2),
    "cvae19": dict(kind="mlp19", input_dim=464, cond=6, latent=16, hidden=256, B=32, beta=4.0, lr=1e-3, steps=3),
    "beta19": dict(kind="mlp19", input_dim=464, cond=0, latent=16, hidden=256, B=17, beta=4.0, lr=1e-3, steps=3),
    "easy06": dict(kind="mlp06", input_dim=80, cond=0, latent=16, hidden=256, B=16, beta=1.0, lr=1e-3, steps=3),
    "ae22": dict(kind="ae22", input_dim=464, cond=0, latent=16, hidden=256, B=24, beta=0.0, lr=1e-3, steps=3),
}


def case_inputs(name: str, step: int) -> dict:
    """Seeded synthetic batch following SURVEY.md 8(d) distributions (small shapes)."""
    c = CASES[name]
    rng = np.random.default_rng(1000 + 17 * step + sum(map(ord, name)))
    B = c["B"]
    if c["kind"] == "conv":
        F_, T_ = c["feat_hw"]
        x = rng.standard_normal((B, 1, F_, T_)).astype(np.float32)
        x = (x - x.mean(axis=(1, 2, 3), keepdims=True)) / x.std(axis=(1, 2, 3), keepdims=True)
        lyr = rng.standard_normal((B, c["lyr"])).astype(np.float32)
        lyr /= np.linalg.norm(lyr, axis=1, keepdims=True)
        m = (rng.random((B, 1)) < 0.593).astype(np.float32)
        lyr = lyr * m
        eps = rng.standard_normal((B, c["z"])).astype(np.float32)
        out = dict(x=x, lyr=lyr, m=m, eps=eps)
    else:
        D = c["input_dim"]
        x = rng.standard_normal((B, D)).astype(np.float32)
        if c["kind"] in ("mlp19", "ae22"):
            x[:, :80] *= np.logspace(1.0, 0.0, 80, dtype=np.float32)  # raw MFCC-stat scale (tamed for tests)
            x[:, 80:] *= 0.05
        eps = rng.standard_normal((B, c["latent"])).astype(np.float32)
        out = dict(x=x, eps=eps)
        if c["cond"]:
            g = rng.integers(0, c["cond"], size=B)
            oh = np.zeros((B, c["cond"]), dtype=np.float32)
            oh[np.arange(B), g] = 1.0
            out["c"] = oh
    return {k: torch.from_numpy(v) for k, v in out.items()}


def case_params(name: str) -> dict:
    c = CASES[name]
    if c["kind"] == "conv":
        return vae_ref.init_conv_vae_params(c["feat_hw"], z_dim=c["z"], lyr_dim=c["lyr"], seed=7)
    if c["kind"] == "ae22":
        return vae_ref.init_ae_params(c["input_dim"], c["latent"], c["hidden"], seed=7)
    style = "19" if c["kind"] == "mlp19" else "06"
    return vae_ref.init_mlp_vae_params(c["input_dim"], c["latent"], c["hidden"], c["cond"], seed=7, style=style)


def _checksums(d: dict) -> dict:
    out = {}
    for k, v in d.items():
        if v is None:
            out[k] = np.array([np.nan, np.nan])
        else:
            v = v.detach().double()
            out[k] = np.array([float(v.sum()), float(v.abs().sum())])
    return out


def run_reference(name: str):
    from oracle import ref_loader
    c = CASES[name]
    if c["kind"] == "conv":
        mod = ref_loader.load("12")
        model = mod.ConvMultiModalVAE(in_ch=1, feat_hw=c["feat_hw"], z_dim=c["z"], lyr_dim=c["lyr"])
    elif c["kind"] == "mlp19":
        mod = ref_loader.load("19")
        model = mod.MLPVAE(c["input_dim"], c["latent"], c["hidden"], cond_dim=c["cond"], conditional=bool(c["cond"]))
    elif c["kind"] == "ae22":
        mod = ref_loader.load("22")
        model = mod.AE(in_dim=c["input_dim"], z_dim=c["latent"])
    else:
        mod = ref_loader.load("06")
        model = mod.MLPVAE(c["input_dim"], c["hidden"], c["latent"])
    model.load_state_dict(case_params(name))
    model.train()
    opt = torch.optim.Adam(model.parameters(), lr=c["lr"])
    rec = dict(loss=[], recon=[], kl=[])
    real_randn_like = torch.randn_like
    first = {}
    for s in range(c["steps"]):
        b = case_inputs(name, s)
        torch.randn_like = lambda t, _e=b["eps"]: _e.clone()
        try:
            opt.zero_grad()
            if c["kind"] == "conv":
                xhat, mu, lv = model(b["x"], b["lyr"], b["m"])
                xc = xhat[:, :, : b["x"].shape[2], : b["x"].shape[3]]
                recon = torch.nn.MSELoss()(xc, b["x"])
                kl = mod.kl_div(mu, lv)
                loss = recon + c["beta"] * kl
            elif c["kind"] == "mlp19":
                xhat, mu, lv = model(b["x"], b.get("c")) if c["cond"] else model(b["x"])
                r = torch.nn.functional.mse_loss(xhat, b["x"], reduction="none").sum(dim=1)
                k = mod.kl_divergence(mu, lv)
                loss = (r + c["beta"] * k).mean()
                recon, kl = r.mean(), k.mean()
            elif c["kind"] == "ae22":
                xhat, mu = model(b["x"])          # (x_hat, z): z stands where mu does in the fixture
                lv = torch.zeros_like(mu)
                loss = torch.nn.functional.mse_loss(xhat, b["x"])
                recon, kl = loss, torch.zeros(())
            else:
                xhat, mu, lv = model(b["x"])
                loss, parts = mod.vae_loss(xhat, b["x"], mu, lv, c["beta"])
                recon, kl = torch.tensor(parts["recon"]), torch.tensor(parts["kl"])
            loss.backward()
        finally:
            torch.randn_like = real_randn_like
        if s == 0:
            first = dict(mu=mu.detach().numpy().copy(), lv=lv.detach().numpy().copy(),
                         xhat_sum=np.array([float(xhat.double().sum()), float(xhat.double().abs().sum())]),
                         xhat_shape=np.array(xhat.shape),
                         grads=_checksums({k: p.grad for k, p in model.named_parameters()}))
        opt.step()
        rec["loss"].append(float(loss)); rec["recon"].append(float(recon)); rec["kl"].append(float(kl))
    final = _checksums(dict(model.named_parameters()))
    return rec, first, final, model


def run_restatement(name: str):
    c = CASES[name]
    p = {k: v.clone() for k, v in case_params(name).items()}
    opt = vae_ref.AdamRef(p, c["lr"])
    losses = []
    g0 = None
    for s in range(c["steps"]):
        losses.append(vae_ref.train_step(p, opt, c["kind"], case_inputs(name, s), c["beta"]))
        if s == 0:
            g0 = _checksums(opt.last_grads)
    return losses, g0, _checksums(p)


def main():
    GOLDEN.mkdir(parents=True, exist_ok=True)
    for name in CASES:
        rec, first, final, _ = run_reference(name)
        losses, g0, pf = run_restatement(name)
        for s, (l, r, k) in enumerate(losses):
            assert abs(l - rec["loss"][s]) <= 2e-5 * abs(rec["loss"][s]) + 1e-7, (name, s, l, rec["loss"][s])
            assert abs(r - rec["recon"][s]) <= 2e-5 * abs(rec["recon"][s]) + 1e-7
            assert abs(k - rec["kl"][s]) <= 2e-5 * abs(rec["kl"][s]) + 1e-7
        for k_, v in first["grads"].items():
            if np.isnan(v[0]):
                assert np.isnan(g0[k_][0]), k_
            else:
                assert abs(v[1] - g0[k_][1]) <= 1e-4 * v[1] + 1e-7, (name, k_, v, g0[k_])
        for k_, v in final.items():
            assert abs(v[1] - pf[k_][1]) <= 1e-5 * v[1] + 1e-7, (name, k_, v, pf[k_])
        flat = dict(loss=np.array(rec["loss"]), recon=np.array(rec["recon"]), kl=np.array(rec["kl"]),
                    mu=first["mu"], lv=first["lv"], xhat_sum=first["xhat_sum"], xhat_shape=first["xhat_shape"])
        for k_, v in first["grads"].items():
            flat["grad/" + k_] = v
        for k_, v in final.items():
            flat["param/" + k_] = v
        np.savez_compressed(GOLDEN / f"vae_{name}.npz", **flat)
        print(f"{name}: reference == restatement; losses {rec['loss']}")


if __name__ == "__main__":
    main()
