"""TEST INFRASTRUCTURE ONLY -- loop-level golden fixtures (tests/golden/loop_*.npz) from the UNMODIFIED reference.

  loop12  the reference's script 12 `main()` itself (12:193-315), run on CPU with synthetic .npy files and a patched
          sys.argv: train_log.csv rows (12:278-279), exported latents + ids (12:289-309), checkpoint keys (12:281-285),
          and align_lyrics_to_audio's output (12:35-60).
  loop19  the hot loop of script 19 (19:209-245) and its export (19:269-287) re-driven with the reference's own
          MLPVAE / kl_divergence / NpyDataset / set_seed (its main() needs matplotlib and fixed data paths).
  loop06  the hot loop of script 06 (06:314-344, 354-358) re-driven with the reference's MLPVAE / vae_loss /
          NumpyDataset / set_seed (its main() needs librosa and audio files).

In the same pass the restated loop of oracle/vae_ref.py (train_loop, encode_mu) is asserted equal to the reference
run, which PINS it.  The reparameterisation noise is injected on every side by `eps_for(call_index, rows, z)`:
torch.randn_like is patched while the reference runs (call k of the run gets generator seed 5000 + k).

Run in the build container only:   python oracle/make_golden_loops.py
"""
from __future__ import annotations

import os
import sys
import tempfile
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from oracle import vae_ref  # noqa: E402

GOLDEN = ROOT / "tests" / "golden"

LOOPS = {
    "loop12": dict(kind="conv", N=150, feat_hw=(16, 22), z=32, lyr=384, n_lyr=100, B=64, epochs=2, lr=2e-3, beta=1.0,
                   seed=123),
    "loop19": dict(kind="mlp19", N=300, input_dim=464, cond=6, latent=16, hidden=256, B=128, epochs=2, lr=1e-3,
                   beta=4.0, seed=42),
    "loop06": dict(kind="mlp06", N=300, input_dim=80, latent=16, hidden=256, B=128, epochs=3, lr=1e-3, beta=1.0,
                   seed=42),
}


def eps_for(call: int, rows: int, z: int) -> torch.Tensor:
    return torch.randn(rows, z, generator=torch.Generator().manual_seed(5000 + call))


def steps_per_epoch(name: str) -> int:
    c = LOOPS[name]
    return (c["N"] + c["B"] - 1) // c["B"]


def eps_fn(name: str):
    """eps_fn(epoch, step, rows) of vaeclust.train.fit / oracle.vae_ref.train_loop for this case."""
    c = LOOPS[name]
    z, spe = c.get("z", c.get("latent")), steps_per_epoch(name)
    return lambda epoch, step, rows: eps_for((epoch - 1) * spe + step, rows, z)


def loop_inputs(name: str) -> dict:
    """Seeded synthetic dataset of the case (numpy arrays, as the scripts np.load them)."""
    c = LOOPS[name]
    rng = np.random.default_rng(77 + sum(map(ord, name)))
    N = c["N"]
    if c["kind"] == "conv":
        F_, T_ = c["feat_hw"]
        X = rng.standard_normal((N, 1, F_, T_)).astype(np.float32)
        X = (X - X.mean(axis=(1, 2, 3), keepdims=True)) / X.std(axis=(1, 2, 3), keepdims=True)
        ids = (rng.permutation(10 * N)[:N] + 1000).astype(np.int64)
        lyr_ids = np.concatenate([rng.permutation(ids)[: c["n_lyr"] - 10], np.arange(10, dtype=np.int64)])  # 10 strangers
        lyr_emb = rng.standard_normal((c["n_lyr"], c["lyr"])).astype(np.float32)
        lyr_emb /= np.linalg.norm(lyr_emb, axis=1, keepdims=True)
        return dict(X=X, ids=ids, lyr_ids=lyr_ids, lyr_emb=lyr_emb)
    D = c["input_dim"]
    X = rng.standard_normal((N, D)).astype(np.float32)
    out = dict(X=X)
    if c["kind"] == "mlp19":
        X[:, :80] *= np.logspace(1.0, 0.0, 80, dtype=np.float32)
        X[:, 80:] *= 0.05
        g = rng.integers(0, c["cond"], size=N)
        oh = np.zeros((N, c["cond"]), dtype=np.float32)
        oh[np.arange(N), g] = 1.0
        out["cond"] = oh
    return out


class _PatchedRandn:
    """torch.randn_like -> eps_for(call index, ...) while the reference runs."""

    def __enter__(self):
        self.real, self.calls = torch.randn_like, 0

        def fake(t, *a, **k):
            e = eps_for(self.calls, t.shape[0], t.shape[1])
            self.calls += 1
            return e

        torch.randn_like = fake
        return self

    def __exit__(self, *exc):
        torch.randn_like = self.real


def run_loop12():
    from oracle import ref_loader
    name = "loop12"
    c, d = LOOPS[name], loop_inputs(name)
    mod = ref_loader.load("12")
    old_argv, old_cwd = sys.argv, os.getcwd()
    with tempfile.TemporaryDirectory(prefix="vaeclust_loop12_") as tmp:
        for k in ("X", "ids", "lyr_ids", "lyr_emb"):
            np.save(os.path.join(tmp, k + ".npy"), d[k])
        sys.argv = ["12", "--x", f"{tmp}/X.npy", "--ids", f"{tmp}/ids.npy", "--lyr_emb", f"{tmp}/lyr_emb.npy",
                    "--lyr_ids", f"{tmp}/lyr_ids.npy", "--z_dim", str(c["z"]), "--beta", str(c["beta"]), "--epochs",
                    str(c["epochs"]), "--batch", str(c["B"]), "--lr", str(c["lr"]), "--out_latents", f"{tmp}/Z.npy",
                    "--out_latent_ids", f"{tmp}/Zids.npy", "--out_dir", f"{tmp}/out"]
        try:
            torch.manual_seed(c["seed"])     # script 12 does not seed itself
            with _PatchedRandn():
                mod.main()
        finally:
            sys.argv = old_argv
            os.chdir(old_cwd)
        rows = [l.strip().split(",") for l in open(f"{tmp}/out/train_log.csv").read().splitlines()]
        assert rows[0] == ["epoch", "loss", "recon", "kl"]
        log = np.array([[float(v) for v in r[1:]] for r in rows[1:]])
        Z, Zids = np.load(f"{tmp}/Z.npy"), np.load(f"{tmp}/Zids.npy")
        ck = torch.load(f"{tmp}/out/ckpt_epoch_{c['epochs']:03d}.pt", weights_only=False)
        assert sorted(ck) == ["cfg", "feat_hw", "lyr_dim", "model"]
    lyr_al, mask = mod.align_lyrics_to_audio(d["ids"], d["lyr_ids"], d["lyr_emb"])
    # initial parameters the reference drew under this seed (the drop-in must draw the same ones)
    torch.manual_seed(c["seed"])
    m0 = mod.ConvMultiModalVAE(in_ch=1, feat_hw=c["feat_hw"], z_dim=c["z"], lyr_dim=c["lyr"])
    p0 = {k: v.detach().clone() for k, v in m0.state_dict().items()}
    # restated loop: same init, same DataLoader order (the generator state after model construction), same eps
    p = {k: v.clone() for k, v in p0.items()}
    torch.manual_seed(c["seed"])
    mod.ConvMultiModalVAE(in_ch=1, feat_hw=c["feat_hw"], z_dim=c["z"], lyr_dim=c["lyr"])   # consume the init draws
    got = vae_ref.train_loop(p, "conv", (d["X"], lyr_al, mask), c["epochs"], c["B"], c["lr"], c["beta"], eps_fn(name),
                             seed=None)
    for j, k in enumerate(("loss", "recon", "kl")):
        np.testing.assert_allclose(got[k], log[:, j], rtol=2e-5, atol=2e-6)
    mu = vae_ref.encode_mu(p, "conv", (d["X"], lyr_al, mask)).numpy()
    np.testing.assert_allclose(mu, Z, rtol=1e-4, atol=1e-5)
    assert np.array_equal(Zids, d["ids"])
    return dict(log=log, Z=Z, Zids=Zids, lyr_aligned=lyr_al, mask=mask,
                init_abs_sum=np.array([float(v.double().abs().sum()) for v in p0.values()]),
                final_abs_sum=np.array([float(v.double().abs().sum()) for v in ck["model"].values()]),
                param_names=np.array(list(p0)))


def _redrive_mlp(name: str):
    """19:194-245,269-287 / 06:300-344,354-358 with the reference's own classes."""
    from oracle import ref_loader
    c, d = LOOPS[name], loop_inputs(name)
    from torch.utils.data import DataLoader
    if c["kind"] == "mlp19":
        mod = ref_loader.load("19")
        mod.set_seed(c["seed"])
        ds = mod.NpyDataset(d["X"], cond=d["cond"])
        dl = DataLoader(ds, batch_size=c["B"], shuffle=True, drop_last=False)
        model = mod.MLPVAE(input_dim=c["input_dim"], latent_dim=c["latent"], hidden_dim=c["hidden"], cond_dim=c["cond"],
                           conditional=True)
    else:
        mod = ref_loader.load("06")
        mod.set_seed(c["seed"])
        ds = mod.NumpyDataset(d["X"])
        dl = DataLoader(ds, batch_size=c["B"], shuffle=True, num_workers=0, pin_memory=False, drop_last=False)
        model = mod.MLPVAE(input_dim=c["input_dim"], hidden_dim=c["hidden"], latent_dim=c["latent"])
    p0 = {k: v.detach().clone() for k, v in model.state_dict().items()}
    opt = torch.optim.Adam(model.parameters(), lr=c["lr"])
    model.train()
    log = []
    with _PatchedRandn():
        for _epoch in range(c["epochs"]):
            tot, den = np.zeros(3), 0.0
            for batch in dl:
                if c["kind"] == "mlp19":
                    xb, cb = batch
                    x_hat, mu, logvar = model(xb, cb)
                    rec = torch.nn.functional.mse_loss(x_hat, xb, reduction="none").sum(dim=1)
                    kl = mod.kl_divergence(mu, logvar)
                    loss = (rec + c["beta"] * kl).mean()
                    opt.zero_grad(); loss.backward(); opt.step()
                    bs = xb.size(0)
                    tot += np.array([float(loss.item()), float(rec.mean().item()), float(kl.mean().item())]) * bs
                    den += bs
                else:
                    opt.zero_grad(set_to_none=True)
                    xhat, mu, logvar = model(batch)
                    loss, parts = mod.vae_loss(xhat, batch, mu, logvar, beta=c["beta"])
                    loss.backward(); opt.step()
                    tot += np.array([parts["total"], parts["recon"], parts["kl"]])
                    den += 1
            log.append(tot / den)
    log = np.array(log)
    model.eval()
    with torch.no_grad():
        if c["kind"] == "mlp19":
            mus = [model.encode(xb, cb)[0].numpy() for xb, cb in DataLoader(ds, batch_size=512, shuffle=False)]
            Z = np.concatenate(mus, axis=0).astype(np.float32)
        else:
            Z = model.encode(torch.from_numpy(d["X"]))[0].numpy().astype(np.float32)
    # the restated loop, pinned here
    p = {k: v.clone() for k, v in p0.items()}
    kind = c["kind"]
    if kind == "mlp19":
        mod.set_seed(c["seed"])
        mod.MLPVAE(input_dim=c["input_dim"], latent_dim=c["latent"], hidden_dim=c["hidden"], cond_dim=c["cond"],
                   conditional=True)
        arrays = (d["X"], d["cond"])
    else:
        mod.set_seed(c["seed"])
        mod.MLPVAE(input_dim=c["input_dim"], hidden_dim=c["hidden"], latent_dim=c["latent"])
        arrays = (d["X"],)
    got = vae_ref.train_loop(p, kind, arrays, c["epochs"], c["B"], c["lr"], c["beta"], eps_fn(name), seed=None)
    for j, k in enumerate(("loss", "recon", "kl")):
        np.testing.assert_allclose(got[k], log[:, j], rtol=2e-5, atol=2e-6)
    np.testing.assert_allclose(vae_ref.encode_mu(p, kind, arrays).numpy(), Z, rtol=1e-4, atol=1e-5)
    return dict(log=log, Z=Z, init_abs_sum=np.array([float(v.double().abs().sum()) for v in p0.values()]),
                final_abs_sum=np.array([float(v.detach().double().abs().sum()) for v in model.state_dict().values()]),
                param_names=np.array(list(p0)))


def main():
    GOLDEN.mkdir(parents=True, exist_ok=True)
    for name in LOOPS:
        out = run_loop12() if name == "loop12" else _redrive_mlp(name)
        np.savez_compressed(GOLDEN / f"{name}.npz", **out)
        print(f"{name}: reference loop == restated loop; epoch log\n{out['log']}")


if __name__ == "__main__":
    main()
