"""Clustering-sweep orchestration of scripts 13 / 16 / 20 / 22 on top of `vaeclust.cluster`.

What the reference's sweep bodies do, with the O(N d) / O(N^2 d) parts on the GPU:

  * `run_cluster_suite`  (13:116-151)  KMeans + Ward + a DBSCAN eps list -> one metrics row per labelling
  * `sweep_rows`         (16:171-244)  the k-list / eps x min_samples sweep body of script 16 with `conservative_score`
  * `hard_metrics`       (20:63-78)    KMeans(n_init=20) + silhouette / NMI / ARI / purity
  * `conservative_score`, `load_repr`, `encode_str_labels`, `safe_ari`, `cluster_purity` (16:99-119, 13:56-111, 20:29-37)

KMeans, silhouette and Davies-Bouldin go through libvaeclust.so.  Agglomerative (Ward) and DBSCAN are different
algorithms outside the north star and stay on scikit-learn, as SURVEY.md 8(a) a18 says; they are imported lazily so the
module loads without them.  The label statistics (ARI / NMI / purity, 8(f) rank 3) are computed here from ONE k x g
contingency table: built with a device histogram when the labels already live on the GPU (no host round trip of the
label vectors), on the host otherwise -- they are O(N) integer counts, not a kernel.  Calinski-Harabasz (09:57) reuses
the cluster-sum and inertia kernels.
"""
from __future__ import annotations

import math
from typing import Dict, Iterable, List, Optional, Sequence

import numpy as np
import torch

from . import cluster as vcl
from ._lib import lib, ptr, stream_ptr
from .ops import _workspace


# ------------------------------------------------------------------------------------------ label statistics
def contingency_table(y_true, y_pred) -> np.ndarray:
    """[n_true_classes, n_pred_clusters] int64 co-occurrence counts (sklearn:metrics/cluster/_supervised.py
    `contingency_matrix`); labels are re-encoded to 0..g-1 / 0..k-1 in sorted order first."""
    on_dev = isinstance(y_true, torch.Tensor) and y_true.is_cuda or isinstance(y_pred, torch.Tensor) and y_pred.is_cuda
    if on_dev:
        dev = y_pred.device if isinstance(y_pred, torch.Tensor) and y_pred.is_cuda else y_true.device
        a = (y_true if isinstance(y_true, torch.Tensor) else torch.from_numpy(np.asarray(y_true))).to(dev).reshape(-1)
        b = (y_pred if isinstance(y_pred, torch.Tensor) else torch.from_numpy(np.asarray(y_pred))).to(dev).reshape(-1)
        if a.numel() != b.numel():
            raise ValueError(f"labels_true and labels_pred must have same size, got {a.numel()} and {b.numel()}")
        ua, ia = torch.unique(a, return_inverse=True)
        ub, ib = torch.unique(b, return_inverse=True)
        g, k = int(ua.numel()), int(ub.numel())
        return torch.bincount(ia * k + ib, minlength=g * k).view(g, k).cpu().numpy().astype(np.int64)
    a = np.asarray(y_true.cpu() if isinstance(y_true, torch.Tensor) else y_true).reshape(-1)
    b = np.asarray(y_pred.cpu() if isinstance(y_pred, torch.Tensor) else y_pred).reshape(-1)
    if a.shape[0] != b.shape[0]:
        raise ValueError(f"labels_true and labels_pred must have same size, got {a.shape[0]} and {b.shape[0]}")
    ua, ia = np.unique(a, return_inverse=True)
    ub, ib = np.unique(b, return_inverse=True)
    g, k = len(ua), len(ub)
    return np.bincount(ia.astype(np.int64) * k + ib, minlength=g * k).reshape(g, k).astype(np.int64)


def _ari_from_table(T: np.ndarray) -> float:
    """adjusted_rand_score from the pair-confusion counts (sklearn `_supervised.py` pair_confusion_matrix + ARI);
    exact Python integers, so 50 M points cannot overflow."""
    n = int(T.sum())
    rows = [int(v) for v in T.sum(axis=1)]
    cols = [int(v) for v in T.sum(axis=0)]
    sum_sq = sum(int(v) * int(v) for v in T.reshape(-1))
    same_both = sum_sq - n
    pred_only = sum(c * c for c in cols) - sum_sq      # same predicted cluster, different class
    true_only = sum(r * r for r in rows) - sum_sq      # same class, different predicted cluster
    neither = n * n - pred_only - true_only - sum_sq
    if pred_only == 0 and true_only == 0:
        return 1.0
    tn, fp, fn, tp = float(neither), float(pred_only), float(true_only), float(same_both)
    return 2.0 * (tp * tn - fn * fp) / ((tp + fn) * (fn + tn) + (tp + fp) * (fp + tn))


def _entropy(counts: np.ndarray) -> float:
    c = counts[counts > 0].astype(np.float64)
    if c.size <= 1:
        return 0.0
    tot = c.sum()
    return float(-np.sum((c / tot) * (np.log(c) - math.log(tot))))


def _nmi_from_table(T: np.ndarray) -> float:
    """normalized_mutual_info_score, average_method="arithmetic" (sklearn default; 20:75)."""
    g, k = T.shape
    if (g == 1 and k == 1) or (g == 0 and k == 0):
        return 1.0
    n = float(T.sum())
    ri, ci = np.nonzero(T)
    nij = T[ri, ci].astype(np.float64)
    a = T.sum(axis=1).astype(np.float64)
    b = T.sum(axis=0).astype(np.float64)
    log_outer = -np.log(a[ri]) - np.log(b[ci]) + 2.0 * math.log(n)
    mi = float(np.sum(nij / n * (np.log(nij) - math.log(n)) + nij / n * log_outer))
    mi = min(max(mi, 0.0), float("inf"))
    if math.isclose(mi, 0.0, abs_tol=float(np.finfo(np.float64).eps)):
        return 0.0
    h_true, h_pred = _entropy(a), _entropy(b)
    return mi / max(0.5 * (h_true + h_pred), float(np.finfo(np.float64).eps))


def adjusted_rand_score(y_true, y_pred) -> float:
    return _ari_from_table(contingency_table(y_true, y_pred))


def normalized_mutual_info_score(y_true, y_pred) -> float:
    return _nmi_from_table(contingency_table(y_true, y_pred))


def cluster_purity(y_true, y_pred) -> float:
    """20:29-37: share of points whose class is the majority class of their predicted cluster; 0.0 when empty."""
    T = contingency_table(y_true, y_pred)
    n = int(T.sum())
    if n == 0:
        return 0.0
    return float(T.max(axis=0).sum()) / float(n)


def label_scores(y_true, y_pred) -> Dict[str, float]:
    """ARI, NMI and purity from one contingency table."""
    T = contingency_table(y_true, y_pred)
    n = int(T.sum())
    return {"ari": _ari_from_table(T), "nmi": _nmi_from_table(T),
            "purity": (float(T.max(axis=0).sum()) / float(n)) if n else 0.0}


def calinski_harabasz_score(X, labels) -> float:
    """sklearn.metrics.calinski_harabasz_score (`_unsupervised.py:374-403`; 09:57): between- over within-cluster
    dispersion.  Cluster sums and the within-cluster squared distances are the KMeans kernels (`vc_cluster_sums`,
    `vc_kmeans_inertia`: direct differences, double accumulation)."""
    Xd = vcl._pad_cols(vcl._as_dev_f32(X))
    N, d = Xd.shape
    lab, k = vcl._encode_labels(labels, Xd.device)
    vcl._check_number_of_labels(k, N)
    lab32 = lab.to(torch.int32).contiguous()
    stats = vcl._cluster_sums(Xd, lab32, k)
    counts = stats[k * d:]
    cent64 = stats[: k * d].view(k, d) / counts.unsqueeze(1)
    mean64 = stats[: k * d].view(k, d).sum(dim=0) / float(N)
    extra = float((counts * ((cent64 - mean64) ** 2).sum(dim=1)).sum().item())
    cent32 = cent64.to(torch.float32).contiguous()
    inertia = torch.zeros(1, dtype=torch.float64, device=Xd.device)
    ws = _workspace(lib().size("vc_kmeans_workspace_bytes", k, d), Xd.device)
    lib().call("vc_kmeans_inertia", ptr(Xd), N, d, ptr(cent32), k, ptr(lab32), ptr(inertia), None, ptr(ws), ws.numel(),
               stream_ptr())
    intra = float(inertia.item())
    return 1.0 if intra == 0.0 else extra * (N - k) / (intra * (k - 1.0))


# --------------------------------------------------------------------------- the reference's wrapper helpers
def encode_str_labels(y_str: np.ndarray) -> np.ndarray:
    """13:56-58: string labels -> ints in sorted order of the distinct strings."""
    table = {v: i for i, v in enumerate(sorted(set(np.asarray(y_str).tolist())))}
    return np.array([table[v] for v in np.asarray(y_str).tolist()], dtype=int)


def safe_ari(y_true_str, yhat) -> Optional[float]:
    """13:103-111 / 16:99-106: None without ground truth or on any error; DBSCAN noise (-1) stays a label."""
    if y_true_str is None:
        return None
    try:
        return float(adjusted_rand_score(encode_str_labels(y_true_str), yhat))
    except Exception:
        return None


def conservative_score(sil, dbi, ari, noise_frac) -> float:
    """16:109-119: silhouette + ARI - 0.2 DBI - 0.8 noise share; missing metrics count as -1 / 10 / 0."""
    s = -1.0 if sil is None else float(sil)
    d = 10.0 if dbi is None else float(dbi)
    a = 0.0 if ari is None else float(ari)
    return s + a - 0.2 * d - 0.8 * float(noise_frac)


def load_repr(path: str) -> np.ndarray:
    """16:122-126: .npy -> float32 [N, features] (trailing axes flattened)."""
    X = np.load(path).astype(np.float32)
    return X.reshape(X.shape[0], -1) if X.ndim > 2 else X


def _host_labels(y) -> np.ndarray:
    return np.asarray(y.cpu() if isinstance(y, torch.Tensor) else y)


def _noise_stats(yhat: np.ndarray):
    uniq = np.unique(yhat)
    n_noise = int(np.sum(yhat == -1)) if -1 in uniq else 0
    n_found = int(sum(1 for u in uniq.tolist() if u != -1))
    return n_found, n_noise


def _sk_host(X) -> np.ndarray:
    return X.detach().cpu().numpy() if isinstance(X, torch.Tensor) else np.asarray(X)


def _ward(X, k: int) -> np.ndarray:
    from sklearn.cluster import AgglomerativeClustering  # outside the north star: stays on scikit-learn
    return AgglomerativeClustering(n_clusters=k, linkage="ward").fit_predict(_sk_host(X))


def _dbscan(X, eps: float, min_samples: int) -> np.ndarray:
    from sklearn.cluster import DBSCAN  # outside the north star: stays on scikit-learn
    return DBSCAN(eps=eps, min_samples=min_samples).fit_predict(_sk_host(X))


def _resident(X):
    """One H2D copy of the representation for the whole suite (every KMeans / silhouette / DBI call reuses it)."""
    return vcl._as_dev_f32(X)


def run_cluster_suite(X, y_true_str, n_clusters: int, tag: str, *, dbscan_eps: Sequence[float] = (0.4, 0.6, 0.8, 1.0, 1.2),
                      other_algorithms: bool = True) -> List[dict]:
    """13:116-151.  Rows (same keys, same order): KMeans(n_init=10, random_state=42), Ward, DBSCAN(min_samples=5) for
    eps in 0.4..1.2.  `other_algorithms=False` keeps only the KMeans row (no scikit-learn needed)."""
    Xd = _resident(X)
    labelled = [("kmeans", f"k={n_clusters}",
                 vcl.KMeans(n_clusters=n_clusters, n_init=10, random_state=42).fit_predict(Xd))]
    if other_algorithms:
        labelled.append(("agglomerative", f"k={n_clusters},ward", _ward(X, n_clusters)))
        for eps in dbscan_eps:
            labelled.append(("dbscan", f"eps={eps},min=5", _dbscan(X, eps, 5)))
    out = []
    for algo, params, yhat in labelled:
        yhat = _host_labels(yhat)
        n_found, n_noise = _noise_stats(yhat)
        out.append({"representation": tag, "algo": algo, "params": params, "n_clusters_found": n_found,
                    "n_noise": n_noise, "silhouette": vcl.safe_silhouette(Xd, yhat),
                    "davies_bouldin": vcl.safe_db_index(Xd, yhat), "ari": safe_ari(y_true_str, yhat)})
    return out


def sweep_rows(X, y_true_str, rep_name: str, k_list: Iterable[int] = (4, 5, 6, 7, 8),
               eps_list: Iterable[float] = (0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.9, 1.0),
               min_samples_list: Iterable[int] = (3, 5, 8), *, other_algorithms: bool = True) -> List[dict]:
    """The per-representation body of script 16's sweep (16:181-244): for every k a KMeans(n_init=10,
    random_state=42) row and a Ward row, then DBSCAN over eps x min_samples; each row carries `conservative_score`."""
    Xd = _resident(X)

    def row(algo, params, yhat, n_found, n_noise, noise_frac):
        sil, dbi, ari = vcl.safe_silhouette(Xd, yhat), vcl.safe_dbi(Xd, yhat), safe_ari(y_true_str, yhat)
        return {"representation": rep_name, "algo": algo, "params": params, "n_clusters_found": n_found,
                "n_noise": n_noise, "noise_frac": noise_frac, "silhouette": sil, "davies_bouldin": dbi, "ari": ari,
                "score": conservative_score(sil, dbi, ari, noise_frac)}

    rows = []
    for k in k_list:
        yhat = _host_labels(vcl.KMeans(n_clusters=k, n_init=10, random_state=42).fit_predict(Xd))
        rows.append(row("kmeans", f"k={k}", yhat, len(np.unique(yhat)), 0, 0.0))
        if other_algorithms:
            yhat = _ward(X, k)
            rows.append(row("agglomerative", f"k={k},ward", yhat, len(np.unique(yhat)), 0, 0.0))
    if other_algorithms:
        for eps in eps_list:
            for ms in min_samples_list:
                yhat = _dbscan(X, eps, ms)
                n_found, n_noise = _noise_stats(yhat)
                rows.append(row("dbscan", f"eps={eps},min={ms}", yhat, n_found, n_noise,
                                float(n_noise) / float(len(yhat))))
    return rows


def hard_metrics(Z, y_genre, k: Optional[int] = None, seed: int = 42, feature_space: str = "latents") -> dict:
    """20:63-78: KMeans(k, random_state=seed, n_init=20) on the latents -> silhouette (nan-flavoured wrapper),
    NMI, ARI and purity against the genre indices.  Also returns the labels and the cluster x genre table
    (`pd.crosstab(pred, true)`, 20:86-87)."""
    y_genre = _host_labels(y_genre)
    k = int(y_genre.max() + 1) if k is None else int(k)
    Zd = _resident(Z)
    y_pred = _host_labels(vcl.KMeans(n_clusters=k, random_state=seed, n_init=20).fit_predict(Zd))
    sc = label_scores(y_genre, y_pred)
    return {"feature_space": feature_space, "k": k, "silhouette": vcl.safe_silhouette_nan(Zd, y_pred),
            "nmi": sc["nmi"], "ari": sc["ari"], "purity": sc["purity"], "labels": y_pred,
            "composition": contingency_table(y_pred, y_genre)}


__all__ = ["contingency_table", "adjusted_rand_score", "normalized_mutual_info_score", "cluster_purity",
           "label_scores", "calinski_harabasz_score", "encode_str_labels", "safe_ari", "conservative_score",
           "load_repr", "run_cluster_suite", "sweep_rows", "hard_metrics"]
