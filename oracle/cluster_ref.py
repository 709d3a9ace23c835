"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the clustering arithmetic the reference
reaches through scikit-learn (un-vendored dependency; requirements.txt pins
scikit-learn==1.8.0, this image has 1.9.0).  Reference call sites:
  KMeans(...).fit_predict      scripts/07:70-71, 13:120-121, 16:183, 20:68-69, 22:91-92
  silhouette_score             scripts/13:74,78, 16:68,73, 20:45, 22:61
  davies_bouldin_score         scripts/13:94,98, 16:89,94
Algorithm sources restated here (sklearn = site-packages/sklearn):
  sklearn/cluster/_kmeans.py:285-293 (tol), :705-758 (Lloyd driver), :1484-1545 (centering)
  sklearn/cluster/_k_means_lloyd.pyx:23-218 (E+M step), _k_means_common.pyx:94-124 (inertia),
  :167-211 (empty-cluster relocation), :274-311 (average, shift)
  sklearn/metrics/cluster/_unsupervised.py:149-200,211-323 (silhouette), :413-487 (DBI)

Parity status: PINNED against sklearn itself (tests/test_oracle_cluster.py runs both on the
same seeded inputs, plus sklearn's docstring known-answer values).  The reference ships no
tests or golden vectors of its own for this path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this.
"""
from __future__ import annotations

import numpy as np

CHUNK = 256  # _k_means_common.pyx:13


def kmeans_tolerance(X: np.ndarray, tol: float) -> float:
    return float(np.mean(np.var(X, axis=0)) * tol) if tol != 0 else 0.0


def lloyd_estep(X: np.ndarray, centers: np.ndarray) -> np.ndarray:
    """argmin_j (||c_j||^2 - 2 x.c_j), first minimum, strict '<' (_k_means_lloyd.pyx:196-213)."""
    c2 = (centers * centers).sum(axis=1)
    labels = np.empty(X.shape[0], dtype=np.int32)
    for s in range(0, X.shape[0], CHUNK * 64):
        pd = c2[None, :] - 2.0 * (X[s:s + CHUNK * 64] @ centers.T)
        labels[s:s + CHUNK * 64] = np.argmin(pd, axis=1)
    return labels


def lloyd_mstep(X, labels, centers_old):
    """centers_new, weights, shift for one iteration incl. relocation of empty clusters."""
    k, d = centers_old.shape
    w = np.bincount(labels, minlength=k).astype(X.dtype)
    sums = np.zeros((k, d), dtype=np.float64)
    np.add.at(sums, labels, X.astype(np.float64))
    sums = sums.astype(X.dtype)
    empty = np.where(w == 0)[0]
    if empty.size:  # _k_means_common.pyx:167-211
        dist = ((X - centers_old[labels]) ** 2).sum(axis=1)
        if dist.max() != 0:
            far = np.argpartition(dist, -empty.size)[:-empty.size - 1:-1]
            for new_id, fi in zip(empty, far):
                old_id = labels[fi]
                sums[old_id] -= X[fi]
                sums[new_id] = X[fi]
                w[new_id] = 1
                w[old_id] -= 1
    centers_new = sums.copy()
    nz = w > 0
    centers_new[nz] *= (1.0 / w[nz]).astype(X.dtype)[:, None]
    if (~nz).any():  # :286-295 an empty cluster is placed on the biggest cluster's centre
        centers_new[~nz] = centers_new[int(np.argmax(w))]
    shift = np.sqrt(((centers_new - centers_old) ** 2).sum(axis=1))
    return centers_new, w, shift


def inertia(X, centers, labels) -> float:
    """Direct differences, not the GEMM form (_k_means_common.pyx:94-124)."""
    diff = X.astype(np.float64) - centers.astype(np.float64)[labels]
    return float((diff * diff).sum())


def kmeans_lloyd(X: np.ndarray, init: np.ndarray, max_iter: int = 300, tol: float = 1e-4):
    """KMeans(init=<array>, n_init=1).fit(X): returns labels, inertia, centers, n_iter."""
    X = np.ascontiguousarray(X).copy()
    tol_ = kmeans_tolerance(X, tol)
    mean = X.mean(axis=0)
    X -= mean
    centers = (np.asarray(init, dtype=X.dtype).copy() - mean)
    labels_old = np.full(X.shape[0], -1, dtype=np.int32)
    strict = False
    n_iter = 0
    for i in range(max_iter):
        labels = lloyd_estep(X, centers)
        centers, _, shift = lloyd_mstep(X, labels, centers)
        n_iter = i + 1
        if np.array_equal(labels, labels_old):
            strict = True
            break
        if float((shift ** 2).sum()) <= tol_:
            break
        labels_old = labels
    if not strict:
        labels = lloyd_estep(X, centers)
    return labels, inertia(X, centers, labels), centers + mean, n_iter


def silhouette_samples(X: np.ndarray, labels: np.ndarray) -> np.ndarray:
    """_unsupervised.py:211-323; distances in float64 from ||x||^2-2x.y+||y||^2 (pairwise.py:567-638),
    self-distance forced to 0 (pairwise.py:2285)."""
    _, lab = np.unique(labels, return_inverse=True)
    k = lab.max() + 1
    freq = np.bincount(lab, minlength=k)
    n = X.shape[0]
    Xd = X.astype(np.float64)
    sq = (Xd * Xd).sum(axis=1)
    onehot = np.zeros((n, k))
    onehot[np.arange(n), lab] = 1.0
    a = np.empty(n)
    b = np.empty(n)
    step = max(1, (1 << 24) // max(n, 1))
    for s in range(0, n, step):
        e = min(n, s + step)
        d2 = sq[s:e, None] - 2.0 * (Xd[s:e] @ Xd.T) + sq[None, :]
        np.maximum(d2, 0, out=d2)
        D = np.sqrt(d2).astype(X.dtype).astype(np.float64) if X.dtype == np.float32 else np.sqrt(d2)
        D[np.arange(e - s), np.arange(s, e)] = 0.0
        cd = D @ onehot  # per-cluster distance sums (_unsupervised.py:184-189)
        rows = np.arange(e - s)
        a[s:e] = cd[rows, lab[s:e]]
        cd[rows, lab[s:e]] = np.inf
        b[s:e] = (cd / freq[None, :]).min(axis=1)
    denom = (freq - 1).take(lab, mode="clip")
    with np.errstate(divide="ignore", invalid="ignore"):
        a = a / denom
        s_ = (b - a) / np.maximum(a, b)
    return np.nan_to_num(s_)


def silhouette_score(X, labels) -> float:
    return float(np.mean(silhouette_samples(X, labels)))


def davies_bouldin_score(X: np.ndarray, labels: np.ndarray) -> float:
    """_unsupervised.py:457-487 (float64 throughout)."""
    _, lab = np.unique(labels, return_inverse=True)
    k = lab.max() + 1
    Xd = X.astype(np.float64)
    cent = np.zeros((k, X.shape[1]))
    intra = np.zeros(k)
    for j in range(k):
        pts = Xd[lab == j]
        cent[j] = pts.mean(axis=0)
        intra[j] = np.mean(np.sqrt(((pts - cent[j]) ** 2).sum(axis=1)))
    M = np.sqrt(np.maximum(((cent[:, None, :] - cent[None, :, :]) ** 2).sum(-1), 0))
    if np.allclose(intra, 0) or np.allclose(M, 0):
        return 0.0
    M[M == 0] = np.inf
    comb = intra[:, None] + intra[None, :]
    return float(np.mean(np.max(comb / M, axis=1)))
