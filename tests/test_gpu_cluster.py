"""GPU: KMeans / silhouette / Davies-Bouldin / StandardScaler through the C-ABI against the oracle restatement and
scikit-learn on identical seeded inputs.  Bars (north star): labels bit-exact given identical init (fp32),
inertia / silhouette / DBI within 1e-5 relative."""
import numpy as np
import pytest
import torch

from oracle import cluster_ref as cr

pytestmark = pytest.mark.gpu
REL = 1e-5


def blobs(n, d, k, seed, spread=4.0):
    rng = np.random.default_rng(seed)
    cent = rng.normal(0, spread, size=(k, d))
    g = rng.integers(0, k, size=n)
    return (cent[g] + rng.normal(size=(n, d))).astype(np.float32), g


@pytest.mark.parametrize("n,d,k", [(3000, 16, 6), (3000, 32, 6), (5000, 64, 8), (20000, 32, 32), (777, 8, 3),
                                   (1000, 20, 5), (900, 6, 4), (4000, 128, 7)])
def test_kmeans_explicit_init_matches_oracle(n, d, k):
    from vaeclust.cluster import KMeans
    X, _ = blobs(n, d, k, seed=n + d + k)
    init = X[np.random.default_rng(2).permutation(n)[:k]].copy()
    km = KMeans(k, init=init, n_init=1).fit(X)
    labels, inertia, centers, n_iter = cr.kmeans_lloyd(X, init)
    assert km.labels_.dtype == np.int32 and km.cluster_centers_.dtype == np.float32
    assert np.array_equal(km.labels_, labels)
    assert km.n_iter_ == n_iter
    assert km.inertia_ == pytest.approx(inertia, rel=REL)
    np.testing.assert_allclose(km.cluster_centers_, centers, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("n,k,scale", [(30001, 16, 1.0), (30001, 17, 1.0), (12345, 31, 1e3), (50000, 32, 1e-3),
                                       (255, 24, 1.0), (400003, 32, 1.0), (20000, 5, 1.0), (77777, 9, 1.0)])
def test_kmeans_tensor_core_estep_d32(n, k, scale, monkeypatch):
    """d == 32, 5 <= k <= 32 runs the tcgen05 E-step (tf32 hi/lo split + exact-recompute guard).  Its labels must be
    bit-identical to the FFMA kernel's (VAECLUST_KMEANS_TC=0) at every Lloyd iteration from the same centres -- ragged
    tile counts, several tiles per CTA, large / small magnitudes, points on bisectors -- and the first E-step must
    equal the oracle's up to rounding-level ties."""
    from vaeclust import cluster
    X, _ = blobs(n, 32, max(k // 2, 2), seed=n + k)        # fewer blobs than centres: many near-bisector points
    X = (X * scale).astype(np.float32)
    X -= X.mean(0)
    C0 = X[np.random.default_rng(3).permutation(n)[:k]].copy()
    Xt, C = torch.from_numpy(X).cuda(), torch.from_numpy(C0).cuda().contiguous()
    for it in range(6):
        out = {}
        for tc in ("1", "0"):
            monkeypatch.setenv("VAECLUST_KMEANS_TC", tc)
            st = cluster._LloydState(n, 32, k, Xt.device)
            st.labels.fill_(-1)
            cluster._estep(Xt, C, st, True)
            out[tc] = st
        assert torch.equal(out["1"].labels, out["0"].labels), it
        assert torch.equal(out["1"].counts, out["0"].counts)
        torch.testing.assert_close(out["1"].sums, out["0"].sums, rtol=1e-6, atol=1e-6 * scale * n / k)
        if it == 0 and n <= 60000:
            # vs the BLAS-ordered oracle only rounding-level bisector ties may differ (see the near-ties test below)
            assert (out["1"].labels.cpu().numpy() != cr.lloyd_estep(X, C0)).sum() <= 3
        cluster._finalize(C, out["1"])
        C = out["1"].centers_new.clone().contiguous()


def test_kmeans_estep_on_split_blobs_differs_only_at_near_ties():
    """k larger than the number of blobs puts many points near a bisector.  From identical centres, one E-step may
    differ from the BLAS-ordered oracle only where the float64 margin between the two candidates is a rounding
    error; whole fits on such data still agree in inertia."""
    from vaeclust import cluster
    X, _ = blobs(20000, 32, 16, seed=20064)
    init = X[np.random.default_rng(2).permutation(20000)[:32]].copy()
    Xc = X - X.mean(0)
    C = init - X.mean(0)
    want = cr.lloyd_estep(Xc, C)
    Xt, Ct = torch.from_numpy(Xc).cuda(), torch.from_numpy(C).cuda().contiguous()
    st = cluster._LloydState(20000, 32, 32, Xt.device)
    cluster._estep(Xt, Ct, st, False)
    got = st.labels.cpu().numpy()
    bad = np.where(got != want)[0]
    d2 = ((Xc[bad, None, :].astype(np.float64) - C[None].astype(np.float64)) ** 2).sum(-1)
    for i, r in enumerate(bad):
        margin = abs(d2[i, got[r]] - d2[i, want[r]])
        assert margin <= 1e-4 * d2[i].min(), (r, margin)
    assert bad.size <= 5, bad.size
    km = cluster.KMeans(32, init=init, n_init=1).fit(X)
    _, inertia, _, _ = cr.kmeans_lloyd(X, init)
    assert km.inertia_ == pytest.approx(inertia, rel=1e-3)


def test_kmeans_matches_sklearn_explicit_init():
    sk = pytest.importorskip("sklearn.cluster")
    from vaeclust.cluster import KMeans
    X, _ = blobs(6000, 32, 8, seed=9)
    init = X[:8].copy()
    ours = KMeans(8, init=init, n_init=1).fit(X)
    ref = sk.KMeans(8, init=init, n_init=1).fit(X)
    assert np.array_equal(ours.labels_, ref.labels_) and ours.n_iter_ == ref.n_iter_
    assert ours.inertia_ == pytest.approx(ref.inertia_, rel=REL)
    np.testing.assert_allclose(ours.cluster_centers_, ref.cluster_centers_, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("k,n_init", [(6, 10), (5, 20)])
def test_kmeans_plusplus_random_state_matches_sklearn(k, n_init):
    """The reference's actual call: KMeans(n_clusters=k, random_state=42, n_init=...) (07:70, 13:120, 20:68)."""
    sk = pytest.importorskip("sklearn.cluster")
    from vaeclust.cluster import KMeans
    X, _ = blobs(2924, 16, 6, seed=4, spread=2.0)
    ours = KMeans(n_clusters=k, random_state=42, n_init=n_init).fit(X)
    ref = sk.KMeans(n_clusters=k, random_state=42, n_init=n_init).fit(X)
    assert ours.inertia_ == pytest.approx(ref.inertia_, rel=1e-4)
    agree = (ours.labels_ == ref.labels_).mean()
    assert agree == 1.0, f"label agreement {agree}"


def test_kmeans_empty_cluster_and_duplicates():
    from vaeclust.cluster import KMeans
    rng = np.random.default_rng(5)
    X = rng.normal(size=(200, 4)).astype(np.float32)
    init = X[:4].copy()
    init[3] = 100.0
    km = KMeans(4, init=init, n_init=1).fit(X)
    labels, inertia, _, _ = cr.kmeans_lloyd(X, init)
    assert np.array_equal(km.labels_, labels) and km.inertia_ == pytest.approx(inertia, rel=REL)
    Xd = np.repeat(rng.normal(size=(3, 4)).astype(np.float32), 10, axis=0)   # 3 distinct points, k=3
    km = KMeans(3, init=Xd[[0, 10, 20]].copy(), n_init=1).fit(Xd)
    assert km.inertia_ == pytest.approx(0.0, abs=1e-9) and len(set(km.labels_.tolist())) == 3


def test_kmeans_accepts_device_tensors_and_predict():
    from vaeclust.cluster import KMeans
    X, _ = blobs(4096, 32, 4, seed=1)
    Xt = torch.from_numpy(X).cuda()
    km = KMeans(4, init=Xt[:4].clone(), n_init=1).fit(Xt)
    assert km.labels_.is_cuda and km.cluster_centers_.is_cuda
    assert torch.equal(km.predict(Xt), km.labels_)


@pytest.mark.parametrize("mode", ["tc", "fp32"])
@pytest.mark.parametrize("n,d,k", [(1500, 16, 5), (3000, 32, 6), (5000, 64, 8), (2924, 16, 2), (700, 10, 4),
                                   (2000, 100, 3)])
def test_silhouette_matches_oracle(n, d, k, mode, monkeypatch):
    from vaeclust import cluster
    from vaeclust.cluster import silhouette_samples, silhouette_score
    monkeypatch.setattr(cluster, "SILHOUETTE_PRECISION", mode)
    X, g = blobs(n, d, k, seed=n + k)
    labels = cr.kmeans_lloyd(X, X[:k].copy())[0]
    want = cr.silhouette_samples(X, labels)
    got = silhouette_samples(X, labels)
    np.testing.assert_allclose(got, want, rtol=2e-4, atol=2e-5)
    assert silhouette_score(X, labels) == pytest.approx(float(np.mean(want)), rel=REL, abs=1e-7)
    assert isinstance(silhouette_score(X, labels), float)


def test_silhouette_and_dbi_match_sklearn_and_known_answers():
    skm = pytest.importorskip("sklearn.metrics")
    from vaeclust.cluster import davies_bouldin_score, silhouette_samples, silhouette_score
    X, _ = blobs(4000, 64, 6, seed=21)
    labels = cr.kmeans_lloyd(X, X[:6].copy())[0]
    assert silhouette_score(X, labels) == pytest.approx(skm.silhouette_score(X, labels), rel=REL)
    assert davies_bouldin_score(X, labels) == pytest.approx(skm.davies_bouldin_score(X, labels), rel=REL)
    # sklearn docstring known answer (_unsupervised.py:451-455) and the singleton rule (:322-323)
    Xk = np.array([[0, 1], [1, 1], [3, 4]], dtype=np.float32)
    assert davies_bouldin_score(Xk, [0, 0, 1]) == pytest.approx(0.128036879, abs=1e-6)
    s = silhouette_samples(np.array([[0, 0], [0.1, 0], [5, 5]], dtype=np.float32), [0, 0, 1])
    assert s[2] == 0.0 and s[0] > 0.9
    # non-contiguous label values are label-encoded; duplicates give zero distances without NaNs
    Xd = np.repeat(X[:50], 4, axis=0)
    lab = np.tile([3, 3, 9, 9], 50)
    assert silhouette_score(Xd, lab) == pytest.approx(skm.silhouette_score(Xd, lab), rel=1e-4, abs=1e-6)


def test_dbi_matches_oracle_many_shapes():
    from vaeclust.cluster import davies_bouldin_score
    for n, d, k in [(3000, 16, 6), (5000, 64, 8), (999, 7, 3)]:
        X, _ = blobs(n, d, k, seed=n)
        labels = cr.kmeans_lloyd(X, X[:k].copy())[0]
        assert davies_bouldin_score(X, labels) == pytest.approx(cr.davies_bouldin_score(X, labels), rel=REL)


def test_standard_scaler_matches_sklearn():
    skp = pytest.importorskip("sklearn.preprocessing")
    from vaeclust.cluster import StandardScaler
    rng = np.random.default_rng(3)
    X = (rng.normal(size=(2924, 80)) * np.logspace(2, 0, 80) + 5).astype(np.float32)
    X[:, 7] = 3.0  # constant column
    ours, ref = StandardScaler().fit(X), skp.StandardScaler().fit(X)
    np.testing.assert_allclose(ours.mean_, ref.mean_, rtol=1e-6)
    np.testing.assert_allclose(ours.scale_, ref.scale_, rtol=1e-5)
    np.testing.assert_allclose(ours.transform(X), ref.transform(X), rtol=1e-5, atol=1e-6)


def test_logical_shards_equal_one_shard():
    """the sharded algorithm with logical shards on one GPU (partials summed in a fixed order)
    gives identical labels and counts and the same sums up to fp32 accumulation order."""
    from vaeclust import cluster
    X, _ = blobs(10000, 32, 8, seed=2)
    Xt = torch.from_numpy(X).cuda()
    C = Xt[:8].clone().contiguous()
    st = cluster._LloydState(10000, 32, 8, Xt.device)
    cluster._estep(Xt, C, st, True)
    full_counts = st.counts.clone()
    full_sums, full_labels = st.sums.clone(), st.labels.clone()
    acc = torch.zeros_like(st.stats)
    labs = []
    for lo, hi in [(0, 2500), (2500, 6001), (6001, 10000)]:
        s2 = cluster._LloydState(hi - lo, 32, 8, Xt.device)
        cluster._estep(Xt[lo:hi].contiguous(), C, s2, True)
        acc += s2.stats
        labs.append(s2.labels.clone())
    assert torch.equal(torch.cat(labs), full_labels)
    assert torch.equal(acc[8 * 32: 8 * 32 + 8], full_counts)
    # sums are accumulated in fp32 per warp before the double fold, so a different tiling moves the last bits
    torch.testing.assert_close(acc[: 8 * 32], full_sums, rtol=1e-5, atol=1e-3)


def test_full_size_properties():
    """Size-independent properties at a size the CPU oracle cannot reach quickly: inertia is non-increasing over
    Lloyd iterations, counts sum to N, sums/counts reproduce the centres, a converged fit is a fixed point."""
    from vaeclust import cluster
    N, d, k = 4_000_000, 32, 16
    g = torch.Generator(device="cuda").manual_seed(1)
    cent = torch.randn(16, d, device="cuda", generator=g) * 3
    X = cent[torch.randint(0, 16, (N,), device="cuda", generator=g)] + torch.randn(N, d, device="cuda", generator=g)
    C = X[:k].clone().contiguous()
    st = cluster._LloydState(N, d, k, X.device)
    st.labels.fill_(-1)
    last = None
    for it in range(6):
        cluster._estep(X, C, st, True)
        assert float(st.counts.sum()) == N
        inertia, _ = cluster._point_sqdist(X, C, st)
        if last is not None:
            assert float(inertia) <= last * (1 + 1e-6)
        last = float(inertia)
        cluster._finalize(C, st)
        ref_c = torch.zeros(k, d, dtype=torch.float64, device="cuda").index_add_(0, st.labels.long(), X.double())
        torch.testing.assert_close(st.sums.view(k, d), ref_c, rtol=1e-5, atol=1.0)
        C = st.centers_new.clone()
    km = cluster.KMeans(k, init=C, n_init=1, tol=0.0, max_iter=100).fit(X)
    again = cluster.KMeans(k, init=km.cluster_centers_, n_init=1, tol=0.0, max_iter=5).fit(X)
    # cluster_centers_ had the mean added back, so the refit starts within an ulp of the fixed point
    if km.n_iter_ < 100:  # converged: a refit moves at most a handful of borderline points
        assert (again.labels_ != km.labels_).float().mean().item() <= 1e-4
    assert again.inertia_ <= km.inertia_ * (1 + 1e-6)
