"""CPU: the numpy restatement in oracle/cluster_ref.py is pinned against scikit-learn itself (the un-vendored
dependency that executes the reference's clustering arithmetic) and against sklearn's docstring known answers."""
import numpy as np
import pytest

from oracle import cluster_ref as cr

sk_cluster = pytest.importorskip("sklearn.cluster")
sk_metrics = pytest.importorskip("sklearn.metrics")


def blobs(n, d, k, seed, spread=4.0):
    rng = np.random.default_rng(seed)
    cent = rng.normal(0, spread, size=(k, d))
    g = rng.integers(0, k, size=n)
    return (cent[g] + rng.normal(size=(n, d))).astype(np.float32), g


@pytest.mark.parametrize("n,d,k", [(500, 8, 3), (2000, 16, 6), (3000, 32, 8)])
def test_lloyd_matches_sklearn(n, d, k):
    X, _ = blobs(n, d, k, seed=n + d)
    init = X[np.random.default_rng(1).permutation(n)[:k]].copy()
    km = sk_cluster.KMeans(n_clusters=k, init=init, n_init=1, tol=1e-4, max_iter=300).fit(X)
    labels, inertia, centers, n_iter = cr.kmeans_lloyd(X, init)
    assert np.array_equal(labels, km.labels_)
    assert n_iter == km.n_iter_
    assert inertia == pytest.approx(km.inertia_, rel=1e-5)
    np.testing.assert_allclose(centers, km.cluster_centers_, rtol=1e-4, atol=1e-5)


def test_silhouette_and_dbi_match_sklearn():
    X, _ = blobs(1500, 16, 5, seed=3)
    labels = sk_cluster.KMeans(n_clusters=5, n_init=2, random_state=0).fit_predict(X)
    assert cr.silhouette_score(X, labels) == pytest.approx(sk_metrics.silhouette_score(X, labels), rel=1e-5)
    np.testing.assert_allclose(cr.silhouette_samples(X, labels), sk_metrics.silhouette_samples(X, labels),
                               rtol=1e-4, atol=1e-5)
    assert cr.davies_bouldin_score(X, labels) == pytest.approx(sk_metrics.davies_bouldin_score(X, labels), rel=1e-5)


def test_known_answers():
    # sklearn docstring: davies_bouldin_score([[0,1],[1,1],[3,4]], [0,0,1]) = 0.12...
    X = np.array([[0, 1], [1, 1], [3, 4]], dtype=np.float64)
    assert cr.davies_bouldin_score(X, np.array([0, 0, 1])) == pytest.approx(0.128036879, abs=1e-8)
    # singleton cluster -> s = 0 for that sample (_unsupervised.py:322-323)
    Xs = np.array([[0.0, 0.0], [0.1, 0.0], [5.0, 5.0]], dtype=np.float32)
    s = cr.silhouette_samples(Xs, np.array([0, 0, 1]))
    assert s[2] == 0.0 and s[0] > 0.9


def test_empty_cluster_relocation_matches_sklearn():
    rng = np.random.default_rng(5)
    X = rng.normal(size=(200, 4)).astype(np.float32)
    init = X[:4].copy()
    init[3] = 100.0  # far away: empty after the first E-step
    km = sk_cluster.KMeans(n_clusters=4, init=init, n_init=1).fit(X)
    labels, inertia, centers, n_iter = cr.kmeans_lloyd(X, init)
    assert np.array_equal(labels, km.labels_)
    assert inertia == pytest.approx(km.inertia_, rel=1e-5)
