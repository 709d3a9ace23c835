"""CPU, world_size=2 over gloo: the sharded KMeans / silhouette / Davies-Bouldin / moments / gradient-bucket logic
gives the single-process answer.  The CUDA entry points are replaced by oracle-backed stand-ins (tests/_cpu_stubs)."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as td
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parents[1]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _data():
    rng = np.random.default_rng(11)
    cent = rng.normal(0, 4, size=(4, 8))
    X = (cent[rng.integers(0, 4, 901)] + rng.normal(size=(901, 8))).astype(np.float32)
    init = X[rng.permutation(901)[:4]].copy()
    return X, init


def _worker(rank, world, port, out_dir):
    for p in (str(ROOT), str(ROOT / "vae-for-hybrid-music-clustering_b200"), str(ROOT / "tests")):
        sys.path.insert(0, p)
    td.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    import _cpu_stubs
    from vaeclust import cluster, dist

    _cpu_stubs.install(None, cluster)
    X, init = _data()
    lo, hi = dist.shard_bounds(X.shape[0], rank, world)
    res = {}
    km = cluster.KMeans(4, init=init, n_init=1, group=td.group.WORLD).fit(X[lo:hi])
    res["labels"], res["centers"], res["inertia"], res["n_iter"] = km.labels_, km.cluster_centers_, km.inertia_, km.n_iter_
    full = np.zeros(X.shape[0], dtype=np.int64)
    full[lo:hi] = km.labels_
    t = torch.from_numpy(full)
    td.all_reduce(t)
    res["sil"] = cluster.silhouette_score(X, t.numpy(), group=td.group.WORLD)
    res["sil_samples"] = cluster.silhouette_samples(X, t.numpy(), group=td.group.WORLD)
    res["dbi"] = cluster.davies_bouldin_score(X[lo:hi], km.labels_, group=td.group.WORLD)
    sc = cluster.StandardScaler().fit(X[lo:hi], group=td.group.WORLD)
    res["mean"], res["var"] = sc.mean_, sc.var_
    # gradient bucket: uneven local batches 5 and 3 of a global batch of 8
    p = torch.nn.Parameter(torch.zeros(6))
    bucket = dist.GradBucket([p], td.group.WORLD)
    nloc = 5 if rank == 0 else 3
    bucket.flat.copy_(torch.full((6,), float(rank + 1)))  # local batch-mean gradient
    bucket.all_reduce_mean(nloc, 8)
    res["grad"] = p.grad.numpy().copy()
    # overlapped exchange: one "big" parameter reduced from its autograd hook, two small ones through the flat buffer
    big = torch.nn.Parameter(torch.ones(40))
    s1, s2 = torch.nn.Parameter(torch.ones(3)), torch.nn.Parameter(torch.ones(2, 2))
    unused = torch.nn.Parameter(torch.ones(5))
    sync = dist.OverlappedGradSync([big, s1, s2, unused], td.group.WORLD, big_numel=16)
    for _ in range(2):  # twice: the second step must not see stale gradients
        sync.begin(nloc, 8)
        ((rank + 1.0) * (big.sum() + 2.0 * s1.sum() + 3.0 * s2.sum())).backward()
        sync.finish()
    res["ov_big"], res["ov_s1"], res["ov_s2"] = big.grad.numpy().copy(), s1.grad.numpy().copy(), s2.grad.numpy().copy()
    res["ov_unused_none"] = np.array(unused.grad is None)
    # empty-cluster relocation across ranks
    init2 = init.copy()
    init2[3] = 1000.0
    km2 = cluster.KMeans(4, init=init2, n_init=1, group=td.group.WORLD).fit(X[lo:hi])
    res["labels2"], res["inertia2"] = km2.labels_, km2.inertia_
    # sharded optimizer: reduce-scatter -> update of this rank's slice -> all-gather (here: SGD on CPU tensors)
    bigp = torch.nn.Parameter(torch.arange(40, dtype=torch.float32))
    smallp = torch.nn.Parameter(torch.ones(3))

    class _Opt:   # minimal stand-in with the two methods ShardedGradSync uses
        def __init__(self):
            self.sh = {}

        def shard(self, p, lo_, hi_, gs):
            self.sh[id(p)] = (lo_, hi_, gs)

        def update_now(self, p):
            lo_, hi_, gs = self.sh[id(p)]
            p.data.view(-1)[lo_:hi_] -= 0.1 * gs

    sopt = _Opt()
    ssync = dist.ShardedGradSync([bigp, smallp], td.group.WORLD, sopt, big_numel=16)
    for _ in range(2):
        ssync.begin(nloc, 8)
        ((rank + 1.0) * ((bigp * torch.arange(40.0)).sum() + smallp.sum())).backward()
        ssync.finish()
        smallp.data -= 0.1 * smallp.grad
        ssync.join()
    res["sh_big"], res["sh_small"] = bigp.detach().numpy().copy(), smallp.detach().numpy().copy()
    # k-means++ over row shards (device-prefix-sum path) and best-of-n_init restarts dealt to the ranks
    cluster._HOST_CUMSUM_MAX = 0
    km3 = cluster.KMeans(4, n_init=1, random_state=7, group=td.group.WORLD).fit(X[lo:hi])
    res["pp_centers"], res["pp_labels"] = km3.cluster_centers_, km3.labels_
    cluster._HOST_CUMSUM_MAX = 1 << 21
    km4 = cluster.KMeans(4, n_init=5, random_state=3, group=td.group.WORLD, parallel="restarts").fit(X)
    res["rs_labels"], res["rs_inertia"], res["rs_centers"] = km4.labels_, km4.inertia_, km4.cluster_centers_
    res["rs_n_iter"] = km4.n_iter_
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), **res)
    td.destroy_process_group()


@pytest.mark.timeout(300)
def test_world2_matches_single_process(tmp_path):
    from oracle import cluster_ref as cr
    from vaeclust import dist

    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    X, init = _data()
    labels, inertia, centers, n_iter = cr.kmeans_lloyd(X, init)
    r = [np.load(tmp_path / f"r{i}.npz") for i in range(2)]
    got = np.concatenate([r[0]["labels"], r[1]["labels"]])
    assert np.array_equal(got, labels)
    for i in range(2):
        assert int(r[i]["n_iter"]) == n_iter
        assert float(r[i]["inertia"]) == pytest.approx(inertia, rel=1e-6)
        np.testing.assert_allclose(r[i]["centers"], centers, rtol=1e-5, atol=1e-6)
        assert float(r[i]["sil"]) == pytest.approx(cr.silhouette_score(X, labels), rel=1e-5)
        np.testing.assert_allclose(r[i]["sil_samples"], cr.silhouette_samples(X, labels), rtol=1e-4, atol=1e-5)
        assert float(r[i]["dbi"]) == pytest.approx(cr.davies_bouldin_score(X, labels), rel=1e-6)
        np.testing.assert_allclose(r[i]["mean"], X.astype(np.float64).mean(0), rtol=1e-9, atol=1e-12)
        np.testing.assert_allclose(r[i]["var"], X.astype(np.float64).var(0), rtol=1e-9)
        np.testing.assert_allclose(r[i]["grad"], np.full(6, (5 * 1.0 + 3 * 2.0) / 8))
        w = (5 * 1.0 + 3 * 2.0) / 8
        np.testing.assert_allclose(r[i]["ov_big"], np.full(40, w), rtol=1e-6)
        np.testing.assert_allclose(r[i]["ov_s1"], np.full(3, 2 * w), rtol=1e-6)
        np.testing.assert_allclose(r[i]["ov_s2"], np.full((2, 2), 3 * w), rtol=1e-6)
        assert bool(r[i]["ov_unused_none"])
    np.testing.assert_array_equal(r[0]["centers"], r[1]["centers"])  # bit-identical replicas
    # sharded optimizer: both ranks hold the same, fully updated parameter = 2 SGD steps on the weighted mean gradient
    w = (5 * 1.0 + 3 * 2.0) / 8
    np.testing.assert_allclose(r[0]["sh_big"], np.arange(40.0) - 2 * 0.1 * w * np.arange(40.0), rtol=1e-6)
    np.testing.assert_array_equal(r[0]["sh_big"], r[1]["sh_big"])
    np.testing.assert_allclose(r[0]["sh_small"], np.ones(3) - 2 * 0.1 * w, rtol=1e-6)
    init2 = init.copy()
    init2[3] = 1000.0
    labels2, inertia2, _, _ = cr.kmeans_lloyd(X, init2)
    assert np.array_equal(np.concatenate([r[0]["labels2"], r[1]["labels2"]]), labels2)
    assert float(r[0]["inertia2"]) == pytest.approx(inertia2, rel=1e-6)
    # sharded k-means++ == the same (float64 prefix-sum) path on one process; restarts over ranks == sequential n_init
    import _cpu_stubs
    from vaeclust import cluster
    _cpu_stubs.install(None, cluster)
    cluster._HOST_CUMSUM_MAX = 0
    one = cluster.KMeans(4, n_init=1, random_state=7).fit(X)
    cluster._HOST_CUMSUM_MAX = 1 << 21
    np.testing.assert_allclose(r[0]["pp_centers"], one.cluster_centers_, rtol=1e-6, atol=1e-6)
    assert np.array_equal(np.concatenate([r[0]["pp_labels"], r[1]["pp_labels"]]), one.labels_)
    seq = cluster.KMeans(4, n_init=5, random_state=3).fit(X)
    sk = pytest.importorskip("sklearn.cluster").KMeans(4, n_init=5, random_state=3).fit(X)
    for i in range(2):
        assert np.array_equal(r[i]["rs_labels"], seq.labels_) and int(r[i]["rs_n_iter"]) == seq.n_iter_
        assert float(r[i]["rs_inertia"]) == pytest.approx(seq.inertia_, rel=1e-9)
        np.testing.assert_array_equal(r[i]["rs_centers"], seq.cluster_centers_)
    # ... and the seeds are sklearn's own draws: same winner as scikit-learn's sequential best-of-n_init
    assert float(seq.inertia_) == pytest.approx(float(sk.inertia_), rel=1e-5)
