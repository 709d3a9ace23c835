"""GPU: the clustering-sweep bodies of scripts 13 / 16 / 20 (`vaeclust.sweep`) against the same
bodies run on scikit-learn with identical inputs and seeds.  KMeans / silhouette / Davies-Bouldin go through the C-ABI;
Ward and DBSCAN are scikit-learn on both sides, so those rows differ only through the metric kernels."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
REL = 1e-5


def blobs(n, d, k, seed, spread=3.0):
    rng = np.random.default_rng(seed)
    cent = rng.normal(0, spread, size=(k, d))
    g = rng.integers(0, k, size=n)
    return (cent[g] + rng.normal(size=(n, d))).astype(np.float32), g


def _sk_safe(fn, X, yhat):
    uniq = np.unique(yhat)
    if len([u for u in uniq.tolist() if u != -1]) < 2:
        return None
    if -1 in uniq:
        m = yhat != -1
        if m.sum() < 3:
            return None
        try:
            return float(fn(X[m], yhat[m]))
        except Exception:
            return None
    try:
        return float(fn(X, yhat))
    except Exception:
        return None


def _close(a, b):
    if a is None or b is None:
        return a is None and b is None
    return a == pytest.approx(b, rel=REL, abs=1e-7)


def test_run_cluster_suite_matches_sklearn_rows():
    """13:116-151 on 1200 x 16 standardised blobs: every row (KMeans, Ward, 5 DBSCAN settings)."""
    from sklearn import cluster as skc, metrics as skm
    from sklearn.preprocessing import StandardScaler
    from vaeclust import sweep
    X, g = blobs(1200, 16, 6, seed=5)
    X = StandardScaler().fit_transform(X).astype(np.float32)
    genres = np.array(["g%d" % v for v in g], dtype=object)
    rows = sweep.run_cluster_suite(X, genres, 6, tag="vae_mm_latents", dbscan_eps=(0.4, 1.2, 2.0, 2.6, 3.2))
    assert [r["algo"] for r in rows] == ["kmeans", "agglomerative"] + ["dbscan"] * 5
    assert list(rows[0].keys()) == ["representation", "algo", "params", "n_clusters_found", "n_noise", "silhouette",
                                    "davies_bouldin", "ari"]
    want = [skc.KMeans(n_clusters=6, n_init=10, random_state=42).fit_predict(X),
            skc.AgglomerativeClustering(n_clusters=6, linkage="ward").fit_predict(X)]
    want += [skc.DBSCAN(eps=e, min_samples=5).fit_predict(X) for e in (0.4, 1.2, 2.0, 2.6, 3.2)]
    yt = sweep.encode_str_labels(genres)
    n_metric_rows = 0
    for r, y in zip(rows, want):
        uniq = set(y.tolist())
        assert r["n_clusters_found"] == len(uniq - {-1}) and r["n_noise"] == int(np.sum(y == -1))
        assert _close(r["silhouette"], _sk_safe(skm.silhouette_score, X, y)), r
        assert _close(r["davies_bouldin"], _sk_safe(skm.davies_bouldin_score, X, y)), r
        assert r["ari"] == pytest.approx(skm.adjusted_rand_score(yt, y), abs=1e-12)
        n_metric_rows += r["silhouette"] is not None
    assert n_metric_rows >= 3 and rows[2]["silhouette"] is None  # eps=0.4 finds nothing in 16-d: wrapper -> None


def test_sweep_rows_kmeans_part_and_score():
    """16:181-198: KMeans rows for k = 4..8 with the conservative score; device-resident input."""
    from sklearn import cluster as skc, metrics as skm
    from vaeclust import sweep
    X, g = blobs(3000, 32, 6, seed=9)
    rows = sweep.sweep_rows(torch.from_numpy(X).cuda(), None, "vae_mm_latents", other_algorithms=False)
    assert [r["params"] for r in rows] == [f"k={k}" for k in (4, 5, 6, 7, 8)]
    for r, k in zip(rows, (4, 5, 6, 7, 8)):
        y = skc.KMeans(n_clusters=k, n_init=10, random_state=42).fit_predict(X)
        sil, dbi = skm.silhouette_score(X, y), skm.davies_bouldin_score(X, y)
        assert r["silhouette"] == pytest.approx(sil, rel=REL) and r["davies_bouldin"] == pytest.approx(dbi, rel=REL)
        assert r["ari"] is None and r["n_clusters_found"] == k and r["noise_frac"] == 0.0
        assert r["score"] == pytest.approx(sil - 0.2 * dbi, rel=1e-4)


def test_hard_metrics_matches_script_20_on_sklearn():
    from sklearn import cluster as skc, metrics as skm
    from vaeclust import sweep
    Z, g = blobs(4000, 16, 6, seed=20, spread=1.5)
    m = sweep.hard_metrics(Z, g, seed=42)
    y = skc.KMeans(n_clusters=6, random_state=42, n_init=20).fit_predict(Z)
    assert m["k"] == 6 and np.array_equal(m["labels"], y)
    assert m["silhouette"] == pytest.approx(skm.silhouette_score(Z, y), rel=REL)
    assert m["nmi"] == pytest.approx(skm.normalized_mutual_info_score(g, y), rel=1e-10)
    assert m["ari"] == pytest.approx(skm.adjusted_rand_score(g, y), rel=1e-10)
    T = skm.cluster.contingency_matrix(g, y)
    assert m["purity"] == pytest.approx(T.max(axis=0).sum() / len(g))
    assert np.array_equal(m["composition"], T.T)


@pytest.mark.parametrize("n,d,k", [(3000, 16, 6), (5000, 64, 8), (777, 10, 3)])
def test_calinski_harabasz_matches_sklearn(n, d, k):
    from sklearn import metrics as skm
    from vaeclust import sweep
    X, g = blobs(n, d, k, seed=n)
    assert sweep.calinski_harabasz_score(X, g) == pytest.approx(skm.calinski_harabasz_score(X, g), rel=REL)
    with pytest.raises(ValueError):
        sweep.calinski_harabasz_score(X, np.zeros(n, dtype=int))


def test_label_scores_from_device_labels():
    from sklearn import metrics as skm
    from vaeclust import sweep
    rng = np.random.default_rng(3)
    a, b = rng.integers(0, 6, 200_000), rng.integers(-1, 9, 200_000)
    sc = sweep.label_scores(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda())
    assert sc["ari"] == pytest.approx(skm.adjusted_rand_score(a, b), abs=1e-12)
    assert sc["nmi"] == pytest.approx(skm.normalized_mutual_info_score(a, b), rel=1e-9, abs=1e-12)
    assert np.array_equal(sweep.contingency_table(torch.from_numpy(a).cuda(), b), skm.cluster.contingency_matrix(a, b))


def test_pca_matches_sklearn():
    """scripts 09:77-80 / 22:122-124: PCA(n_components, random_state).fit_transform on MFCC-statistics-sized inputs."""
    from sklearn.decomposition import PCA as SkPCA
    from vaeclust.cluster import PCA
    rng = np.random.default_rng(4)
    for n, d, nc in ((3000, 80, 16), (5000, 464, 32)):
        A = rng.normal(size=(d, d)) / np.sqrt(d)
        X = (rng.normal(size=(n, d)) * np.logspace(1, -1, d) @ A + rng.normal(size=d)).astype(np.float32)
        ref = SkPCA(n_components=nc, random_state=42).fit(X)
        ours = PCA(n_components=nc, random_state=42)
        Z = ours.fit_transform(X)
        np.testing.assert_allclose(ours.explained_variance_, ref.explained_variance_, rtol=2e-4)
        np.testing.assert_allclose(ours.explained_variance_ratio_, ref.explained_variance_ratio_, rtol=2e-4)
        np.testing.assert_allclose(ours.mean_, ref.mean_, rtol=1e-5, atol=1e-5)
        # well-separated leading components agree element-wise (same svd_flip sign rule); the projection too
        lead = 8
        np.testing.assert_allclose(ours.components_[:lead], ref.components_[:lead], atol=2e-3)
        Zr = ref.transform(X)
        assert Z.shape == Zr.shape and Z.dtype == np.float32
        np.testing.assert_allclose(Z[:, :lead], Zr[:, :lead], atol=2e-3 * np.abs(Zr[:, :lead]).max())
        np.testing.assert_allclose(ours.transform(X[:100]), Z[:100], rtol=1e-5, atol=1e-5)
