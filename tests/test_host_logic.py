"""CPU: host-side logic of the drop-in layer -- batch order replay, sharding arithmetic, the Lloyd driver's sklearn
semantics (through oracle-backed stand-ins for the CUDA calls) and state_dict compatibility."""
import numpy as np
import pytest
import torch

from oracle import cluster_ref as cr
from oracle import make_golden as mg
from vaeclust import cluster, dist, nn, train

import _cpu_stubs


def test_shard_bounds_cover_exactly():
    for n in (0, 1, 7, 64, 1000, 50_000_001):
        for w in (1, 2, 3, 8):
            b = [dist.shard_bounds(n, r, w) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1


def test_epoch_permutation_replays_dataloader():
    ds = torch.utils.data.TensorDataset(torch.arange(103))
    torch.manual_seed(42)
    dl = torch.utils.data.DataLoader(ds, batch_size=16, shuffle=True)
    want = [b[0].clone() for _ in range(2) for b in dl]
    torch.manual_seed(42)
    got = [idx for _ in range(2) for idx in train.batches(103, 16)]
    assert len(want) == len(got)
    for a, b in zip(want, got):
        assert torch.equal(a, b)


@pytest.mark.parametrize("name", ["conv_16x22", "cvae19", "easy06"])
def test_state_dict_keys_and_shapes_match_reference(name, golden_dir):
    c = mg.CASES[name]
    if c["kind"] == "conv":
        m = nn.ConvMultiModalVAE(1, c["feat_hw"], c["z"], c["lyr"])
    elif c["kind"] == "mlp19":
        m = nn.MLPVAE(c["input_dim"], c["latent"], c["hidden"], cond_dim=c["cond"], conditional=bool(c["cond"]))
    else:
        m = nn.EasyMLPVAE(c["input_dim"], c["hidden"], c["latent"])
    ref = mg.case_params(name)  # keys/shapes verified against the reference classes by make_golden.py
    sd = m.state_dict()
    assert list(sd.keys()) == list(ref.keys())
    for k_, v in ref.items():
        assert tuple(sd[k_].shape) == tuple(v.shape), k_
    m.load_state_dict(ref)  # checkpoints interchange
    gold = np.load(golden_dir / f"vae_{name}.npz")
    assert {k_[6:] for k_ in gold.files if k_.startswith("param/")} == set(sd.keys())


def test_conditional_model_requires_cond():
    m = nn.MLPVAE(8, 2, 16, cond_dim=3, conditional=True)
    with pytest.raises(ValueError, match="conditioning vector"):
        m.encode(torch.zeros(2, 8))


def test_lloyd_driver_matches_oracle_and_sklearn(monkeypatch):
    _cpu_stubs.install(monkeypatch, cluster)
    rng = np.random.default_rng(0)
    cent = rng.normal(0, 4, size=(5, 12))
    X = (cent[rng.integers(0, 5, 1500)] + rng.normal(size=(1500, 12))).astype(np.float32)
    init = X[rng.permutation(1500)[:5]].copy()
    km = cluster.KMeans(5, init=init, n_init=1).fit(X)
    labels, inertia, centers, n_iter = cr.kmeans_lloyd(X, init)
    assert np.array_equal(km.labels_, labels) and km.n_iter_ == n_iter
    assert km.inertia_ == pytest.approx(inertia, rel=1e-5)
    np.testing.assert_allclose(km.cluster_centers_, centers, rtol=1e-4, atol=1e-5)
    sk = pytest.importorskip("sklearn.cluster")
    ref = sk.KMeans(5, init=init, n_init=1).fit(X)
    assert np.array_equal(km.labels_, ref.labels_) and km.n_iter_ == ref.n_iter_


def test_lloyd_driver_relocates_empty_cluster(monkeypatch):
    _cpu_stubs.install(monkeypatch, cluster)
    rng = np.random.default_rng(5)
    X = rng.normal(size=(200, 4)).astype(np.float32)
    init = X[:4].copy()
    init[3] = 100.0
    km = cluster.KMeans(4, init=init, n_init=1).fit(X)
    labels, inertia, _, _ = cr.kmeans_lloyd(X, init)
    assert np.array_equal(km.labels_, labels)
    assert km.inertia_ == pytest.approx(inertia, rel=1e-5)


def test_kmeans_argument_errors(monkeypatch):
    _cpu_stubs.install(monkeypatch, cluster)
    X = np.zeros((5, 4), dtype=np.float32)
    with pytest.raises(ValueError, match="n_samples=5 should be >= n_clusters=6"):
        cluster.KMeans(6, n_init=1, init=np.zeros((6, 4), dtype=np.float32)).fit(X)
    with pytest.raises(ValueError, match="does not match"):
        cluster.KMeans(2, n_init=1, init=np.zeros((3, 4), dtype=np.float32)).fit(X)
    with pytest.raises(ValueError, match="Number of labels is 1"):
        cluster.silhouette_score(X, np.zeros(5, dtype=np.int64))


def test_wrappers_guard_like_the_reference(monkeypatch):
    _cpu_stubs.install(monkeypatch, cluster)
    X = np.random.default_rng(0).normal(size=(30, 4)).astype(np.float32)
    assert cluster.safe_silhouette(X, np.zeros(30, dtype=int)) is None          # 13:64-67
    assert cluster.safe_db_index(X, np.zeros(30, dtype=int)) is None
    assert np.isnan(cluster.safe_silhouette_nan(X, np.zeros(30, dtype=int)))    # 20:42-43
    lab = np.r_[np.full(28, -1), [0, 1]]
    assert cluster.safe_silhouette(X, lab) is None                              # < 3 non-noise points
    lab = np.r_[np.full(10, -1), np.zeros(10, int), np.ones(10, int)]
    v = cluster.safe_silhouette(X, lab)
    m = lab != -1
    assert v == pytest.approx(cr.silhouette_score(X[m], lab[m]), rel=1e-5)


def test_is_same_clustering():
    a = np.array([0, 0, 1, 2, 2])
    assert cluster._is_same_clustering(a, np.array([2, 2, 0, 1, 1]), 3)
    assert not cluster._is_same_clustering(a, np.array([2, 1, 0, 1, 1]), 3)


def test_bench_reference_arm_prints_one_json_line():
    """bench.py --impl reference: stdout is exactly one JSON line (library chatter on fd 1 goes to stderr) carrying
    the reference-arm keys of the bench contract."""
    import json
    import subprocess
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parents[1]
    out = subprocess.run([sys.executable, str(root / "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=300, cwd=root)
    assert out.returncode == 0, out.stderr[-500:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "VAE train clips/s" and d["value"] > 0
    assert d["cpu_baseline"]["kind"] in ("port", "reference") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
