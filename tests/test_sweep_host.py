"""CPU: label statistics and wrapper helpers of `vaeclust.sweep` (scripts 13 / 16 / 20) against scikit-learn and the
reference's own helper semantics (13:56-58,103-111; 16:109-126; 20:29-37)."""
import numpy as np
import pytest

from vaeclust import sweep


def _pairs(seed, n, g, k, noise=False):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, g, size=n)
    b = (a + (rng.random(n) < 0.3) * rng.integers(0, k, size=n)) % k
    if noise:
        b = np.where(rng.random(n) < 0.1, -1, b)
    return a, b


@pytest.mark.parametrize("n,g,k,noise", [(3000, 6, 6, False), (2924, 6, 4, False), (500, 3, 8, True), (50, 2, 2, False),
                                          (1000, 1, 5, False), (1000, 5, 1, False)])
def test_label_scores_match_sklearn(n, g, k, noise):
    from sklearn import metrics as skm
    a, b = _pairs(n + g + k, n, g, k, noise)
    sc = sweep.label_scores(a, b)
    assert sc["ari"] == pytest.approx(skm.adjusted_rand_score(a, b), rel=1e-12, abs=1e-15)
    assert sc["nmi"] == pytest.approx(skm.normalized_mutual_info_score(a, b), rel=1e-10, abs=1e-14)
    T = skm.cluster.contingency_matrix(a, b)
    assert np.array_equal(sweep.contingency_table(a, b), T)
    assert sc["purity"] == pytest.approx(T.max(axis=0).sum() / n)


def test_label_scores_degenerate_cases():
    one = np.zeros(10, dtype=int)
    assert sweep.adjusted_rand_score(one, one) == 1.0
    assert sweep.normalized_mutual_info_score(one, one) == 1.0
    assert sweep.adjusted_rand_score([0, 0, 1, 1], [1, 1, 0, 0]) == 1.0
    assert sweep.adjusted_rand_score([0, 0, 1, 1], [0, 1, 2, 3]) == 0.0
    assert sweep.normalized_mutual_info_score([0, 0, 1, 1], [0, 0, 0, 0]) == 0.0
    assert sweep.cluster_purity([], []) == 0.0
    with pytest.raises(ValueError):
        sweep.contingency_table([0, 1], [0, 1, 2])


def test_ari_exact_integers_at_50M_scale():
    """counts of the size of config C4 (50 M points): the pair counts exceed 2^53 and must not lose the sign."""
    T = np.array([[20_000_000, 5_000_000], [5_000_000, 20_000_000]], dtype=np.int64)
    v = sweep._ari_from_table(T)
    assert 0.0 < v < 1.0
    assert v == pytest.approx(sweep._ari_from_table(T // 1000), rel=1e-3)  # ARI has a 1/n finite-size term


def test_reference_helper_semantics(tmp_path):
    y = np.array(["rock", "folk", "rock", "__MISSING__", "pop"], dtype=object)
    assert sweep.encode_str_labels(y).tolist() == [3, 1, 3, 0, 2]
    assert sweep.safe_ari(None, np.zeros(5, dtype=int)) is None
    assert sweep.safe_ari(y, np.array([0, 1, 0, 2, 3])) == 1.0
    assert sweep.safe_ari(y, np.array([0, 1])) is None          # size mismatch is swallowed, as in the reference
    assert sweep.conservative_score(None, None, None, 0.0) == pytest.approx(-1.0 - 2.0)
    assert sweep.conservative_score(0.5, 1.0, 0.25, 0.5) == pytest.approx(0.5 + 0.25 - 0.2 - 0.4)
    x = np.arange(2 * 1 * 3 * 4, dtype=np.float64).reshape(2, 1, 3, 4)
    np.save(tmp_path / "x.npy", x)
    r = sweep.load_repr(str(tmp_path / "x.npy"))
    assert r.shape == (2, 12) and r.dtype == np.float32


def test_reference_helpers_agree_with_reference_scripts():
    """Same inputs through the reference's own script-16 helpers (imported from /root/reference when it is there)."""
    from oracle import ref_loader
    if not ref_loader.available():  # the reference tree does not travel to the GPU box
        pytest.skip("reference not mounted here")
    ref16 = ref_loader.load("16")
    rng = np.random.default_rng(0)
    for _ in range(20):
        args = [None if rng.random() < 0.3 else float(rng.normal()) for _ in range(3)] + [float(rng.random())]
        assert sweep.conservative_score(*args) == pytest.approx(ref16.conservative_score(*args))
    y = np.array(list("abcabdzz"), dtype=object)
    assert np.array_equal(sweep.encode_str_labels(y), ref16.encode_str_labels(y))
    yh = rng.integers(-1, 3, size=8)
    assert sweep.safe_ari(y, yh) == pytest.approx(ref16.safe_ari(y, yh))
