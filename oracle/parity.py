"""TEST INFRASTRUCTURE ONLY -- parity checkers that put ONE step of the CUDA path next to the CPU oracle on the same
seeded inputs, at whatever size the caller runs (tests/, __graft_entry__.smoke() and the cpu_baseline leg of bench.py
call these with the very shapes the benchmark times, so the tcgen05 instantiations of the headline meet the oracle).

The oracle side is oracle/vae_ref.py (pinned against the reference classes, see its header) and scikit-learn itself
for the clustering calls (the reference's clustering arithmetic IS scikit-learn).
"""
from __future__ import annotations

import time
from typing import Dict

import numpy as np
import torch

from . import vae_ref

# Stated tolerances.  fp32 mode must meet the north star's 1e-3 on the loss against the reference arithmetic.  The bf16
# tensor-core mode (operands rounded to 8 mantissa bits, fp32 accumulation) is checked twice:
#   * against the oracle computing in the SAME arithmetic (vae_ref.emulate_bf16: operands rounded at the same points,
#     fp32 accumulate) -- only summation order differs (and, through it, the handful of ReLU units whose pre-activation
#     sits within an fp32 rounding of zero): measured on B200 <= 5e-3 on the small shapes, 2.3e-2 on the 67 M-element
#     FC weight gradient of the full-size script-12 step; bound 4e-2.  This is the correctness gate;
#   * against the plain fp32 oracle -- the loss agrees to ~1e-5, but gradients of layers below a ReLU differ by a few
#     per cent because pre-activations within rounding distance of 0 flip their mask (a fraction p of flipped units
#     moves the gradient by ~sqrt(p)); measured on B200 at the benchmark shapes: <= 0.09.  Reported, bounded at 0.15.
TOL = {
    "bf16": dict(loss_rel=1e-4, grad_rel=4e-2, loss_rel_vs_fp32=5e-3, grad_rel_vs_fp32=0.15),
    "fp32": dict(loss_rel=1e-3, grad_rel=2e-3),
}


def _rel(a: torch.Tensor, b: torch.Tensor) -> float:
    """||a - b|| / ||b|| in float64."""
    a, b = a.detach().double().cpu().reshape(-1), b.detach().double().cpu().reshape(-1)
    den = float(b.norm())
    return float((a - b).norm()) / den if den > 0 else float((a - b).norm())


def _oracle_step(p0, kind, b_cpu, beta, lr, emulate):
    p_ref = {n: v.clone() for n, v in p0.items()}
    opt = vae_ref.AdamRef(p_ref, lr)
    vae_ref.emulate_bf16(emulate)
    try:
        t0 = time.perf_counter()
        ref = vae_ref.train_step(p_ref, opt, kind, b_cpu, beta)
        dt = time.perf_counter() - t0
    finally:
        vae_ref.emulate_bf16(False)
    return ref, opt.last_grads, p_ref, dt


def _compare(stats, grads, after, ref, g_ref, p_ref, names):
    out = {"loss_ref": ref[0], "loss_rel": abs(float(stats[0]) - ref[0]) / abs(ref[0]),
           "recon_rel": abs(float(stats[1]) - ref[1]) / abs(ref[1]),
           "kl_rel": abs(float(stats[2]) - ref[2]) / max(abs(ref[2]), 1e-12)}
    grad_rel, none_mismatch = {}, []
    for n in names:
        if (g_ref[n] is None) != (grads[n] is None):
            none_mismatch.append(n)
        elif grads[n] is not None:
            grad_rel[n] = _rel(grads[n], g_ref[n])
    worst = max(grad_rel, key=grad_rel.get)
    out["grad_rel_max"], out["grad_rel_max_param"] = grad_rel[worst], worst
    out["grad_rel"] = {n: float(f"{v:.3g}") for n, v in grad_rel.items()}
    out["grad_none_mismatch"] = none_mismatch
    # parameters after the first Adam step: |delta| ~ lr wherever |g| >> eps, so a sign flip of a near-zero gradient
    # element moves a parameter by 2 lr; relative to the parameter norm
    out["param_rel_after_adam"] = max(_rel(after[n], p_ref[n]) for n in names)
    return out


def vae_step_parity(model, step, kind: str, batch: Dict[str, torch.Tensor], beta: float, lr: float,
                    precision: str) -> dict:
    """Run ONE optimisation step of `step` (a vaeclust.train.VAEStep or a fused-step object over `model`, first step of
    its optimizer) and of the oracle from the same parameters, batch and eps.  kind: "conv" | "mlp19" | "mlp06";
    batch holds CUDA tensors x, eps and (conv) lyr, m or (mlp19) c.  Returns the comparison and whether it is within
    TOL[precision]."""
    names = [n for n, _ in model.named_parameters()]
    p0 = {n: p.detach().cpu().clone() for n, p in model.named_parameters()}
    b_cpu = {k: v.detach().cpu() for k, v in batch.items()}
    if kind == "conv":
        args = (batch["x"], batch["lyr"], batch["m"])
    elif kind == "mlp19" and "c" in batch:
        args = (batch["x"], batch["c"])
    else:
        args = (batch["x"],)
    # the per-parameter comparison needs every gradient: switch off the fused "weight gradient -> Adam" kernel of the wide
    # Linear layers for this step (tests/test_gpu_models.py compares the fused update with the plain one separately)
    for obj in (step, getattr(step, "step", None)):
        if obj is not None and hasattr(obj, "fuse_fc_adam"):
            obj.fuse_fc_adam = False
    stats = step(*args, eps=batch["eps"]).detach().cpu().numpy()
    torch.cuda.synchronize()
    if hasattr(step, "last_grads"):      # fused steps never materialise .grad; they expose a debug copy
        grads = {n: (g.detach().clone() if g is not None else None) for n, g in step.last_grads().items()}
    else:
        grads = {n: (p.grad.detach().clone() if p.grad is not None else None) for n, p in model.named_parameters()}
    after = {n: p.detach().cpu().clone() for n, p in model.named_parameters()}
    tol = TOL[precision]
    ref, g_ref, p_ref, dt = _oracle_step(p0, kind, b_cpu, beta, lr, emulate=False)
    out = {"loss": float(stats[0]), "oracle_step_s": dt, "precision": precision, "tolerance": tol}
    if precision == "fp32":
        out.update(_compare(stats, grads, after, ref, g_ref, p_ref, names))
        out["ok"] = bool(out["loss_rel"] <= tol["loss_rel"] and out["grad_rel_max"] <= tol["grad_rel"]
                         and not out["grad_none_mismatch"])
        return out
    vs32 = _compare(stats, grads, after, ref, g_ref, p_ref, names)
    ref_e, g_e, p_e, _ = _oracle_step(p0, kind, b_cpu, beta, lr, emulate=True)
    out.update(_compare(stats, grads, after, ref_e, g_e, p_e, names))   # headline fields: same-arithmetic oracle
    out["vs_fp32_oracle"] = {k: vs32[k] for k in ("loss_ref", "loss_rel", "recon_rel", "kl_rel", "grad_rel_max",
                                                   "grad_rel_max_param", "param_rel_after_adam")}
    out["ok"] = bool(out["loss_rel"] <= tol["loss_rel"] and out["grad_rel_max"] <= tol["grad_rel"]
                     and not out["grad_none_mismatch"] and vs32["loss_rel"] <= tol["loss_rel_vs_fp32"]
                     and vs32["grad_rel_max"] <= tol["grad_rel_vs_fp32"])
    return out


def kmeans_estep_parity(X: np.ndarray, centers: np.ndarray) -> dict:
    """ONE E-step from the same float32 array and centres through vaeclust's kernel and through scikit-learn's own
    lloyd_iter_chunked_dense (BLAS sgemm).  Labels must be equal except at exact TIES: a differing point counts as a
    tie when the float64 distances to the two candidate centroids differ by less than the fp32 rounding error of
    ||c||^2 - 2 x.c (2^-23 of |x|^2 + |c|^2); there the sgemm's summation order decides, not the arithmetic."""
    import os
    from sklearn.cluster._kmeans import _labels_inertia_threadpool_limit
    from vaeclust import cluster
    n, d = X.shape
    k = centers.shape[0]
    Xd, Cd = torch.from_numpy(X).cuda(), torch.from_numpy(centers).cuda()
    st = cluster._LloydState(n, d, k, Xd.device)
    st.labels.fill_(-1)
    cluster._estep(Xd, Cd, st, update=False)
    ours = st.labels.cpu().numpy()
    sk, _ = _labels_inertia_threadpool_limit(X, np.ones(n, dtype=np.float32), centers, n_threads=os.cpu_count())
    idx = np.where(ours != sk)[0]
    beyond = 0
    if idx.size:
        xi, C64 = X[idx].astype(np.float64), centers.astype(np.float64)
        da, db = ((xi - C64[ours[idx]]) ** 2).sum(1), ((xi - C64[sk[idx]]) ** 2).sum(1)
        scale = (xi ** 2).sum(1) + (C64[sk[idx]] ** 2).sum(1)
        beyond = int((np.abs(da - db) > scale * 2.0 ** -23).sum())
    return {"n": int(n), "k": int(k), "label_mismatches": int(idx.size), "mismatches_beyond_fp32_ties": beyond,
            "ok": beyond == 0}


def kmeans_fit_parity(X: np.ndarray, init: np.ndarray, max_iter: int = 300, tol: float = 1e-4) -> dict:
    """vaeclust.cluster.KMeans(init=array, n_init=1).fit(X) against sklearn.cluster.KMeans on the same array.
    Gate: n_iter_ equal, inertia within 1e-5 relative, and the FIRST E-step (same centres on both sides) equal up to
    exact fp32 ties (kmeans_estep_parity).  Labels of the whole fit are bit-equal whenever no tie occurs (every test
    case in tests/test_gpu_cluster.py); on millions of points a handful of ties per iteration is certain, the two
    runs then continue from centroids that differ in the last bits and boundary points may settle differently, so
    the whole-fit agreement is REPORTED (label_mismatches, agreement) next to the gate.  scikit-learn itself is not
    bit-reproducible across thread counts for the same reason (per-thread fp32 centroid partials)."""
    from sklearn.cluster import KMeans as SkKMeans
    from vaeclust import cluster
    k = init.shape[0]
    t0 = time.perf_counter()
    ours = cluster.KMeans(k, init=init, n_init=1, max_iter=max_iter, tol=tol).fit(X)
    torch.cuda.synchronize()
    t_ours = time.perf_counter() - t0
    t0 = time.perf_counter()
    ref = SkKMeans(k, init=init, n_init=1, max_iter=max_iter, tol=tol).fit(X)
    t_ref = time.perf_counter() - t0
    mism = int((ours.labels_ != ref.labels_).sum())
    Xc = X - X.mean(axis=0, dtype=np.float64).astype(np.float32)
    first = kmeans_estep_parity(Xc, init - X.mean(axis=0, dtype=np.float64).astype(np.float32))
    out = {"n": int(X.shape[0]), "d": int(X.shape[1]), "k": int(k), "label_mismatches": mism,
           "label_agreement": 1.0 - mism / X.shape[0], "first_estep": first,
           "n_iter": int(ours.n_iter_), "n_iter_ref": int(ref.n_iter_),
           "inertia_rel": abs(ours.inertia_ - ref.inertia_) / abs(ref.inertia_),
           "centers_max_abs_diff": float(np.abs(ours.cluster_centers_ - ref.cluster_centers_).max()),
           "fit_s": t_ours, "sklearn_fit_s": t_ref}
    out["ok"] = bool(first["ok"] and out["n_iter"] == out["n_iter_ref"] and out["inertia_rel"] <= 1e-5
                     and out["label_agreement"] >= 0.999)
    return out


def silhouette_dbi_parity(X: np.ndarray, labels: np.ndarray) -> dict:
    """vaeclust silhouette_score / davies_bouldin_score against sklearn on the same array and labels (1e-5 relative)."""
    from sklearn.metrics import davies_bouldin_score as sk_dbi
    from sklearn.metrics import silhouette_score as sk_sil
    from vaeclust import cluster
    t0 = time.perf_counter()
    s = cluster.silhouette_score(X, labels)
    d = cluster.davies_bouldin_score(X, labels)
    t_ours = time.perf_counter() - t0
    t0 = time.perf_counter()
    s_ref = float(sk_sil(X, labels))
    t_sil = time.perf_counter() - t0
    d_ref = float(sk_dbi(X, labels))
    out = {"n": int(X.shape[0]), "d": int(X.shape[1]), "k": int(len(np.unique(labels))), "silhouette": s,
           "silhouette_ref": s_ref, "silhouette_rel": abs(s - s_ref) / abs(s_ref), "dbi": d, "dbi_ref": d_ref,
           "dbi_rel": abs(d - d_ref) / abs(d_ref), "ours_s": t_ours, "sklearn_silhouette_s": t_sil}
    out["ok"] = bool(out["silhouette_rel"] <= 1e-5 and out["dbi_rel"] <= 1e-5)
    return out
