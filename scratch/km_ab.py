import os, sys, numpy as np, torch
sys.path[:0] = [".", "vae-for-hybrid-music-clustering_b200"]
from vaeclust import cluster
from oracle import cluster_ref as cr
def blobs(n, d, k, seed, spread=4.0):
    rng = np.random.default_rng(seed)
    cent = rng.normal(0, spread, size=(k, d))
    g = rng.integers(0, k, size=n)
    return (cent[g] + rng.normal(size=(n, d))).astype(np.float32), g
for (n, k, scale) in [(30001, 16, 1.0), (30001, 17, 1.0), (12345, 31, 1e3), (50000, 32, 1e-3), (255, 24, 1.0), (1000003, 32, 1.0), (1000003, 16, 1.0)]:
    X, _ = blobs(n, 32, k, seed=n + k); X = (X * scale).astype(np.float32)
    init = X[np.random.default_rng(3).permutation(n)[:k]].copy()
    Xt, Ct = torch.from_numpy(X).cuda(), torch.from_numpy(init).cuda().contiguous()
    res = {}
    for tc in ("1", "0"):
        os.environ["VAECLUST_KMEANS_TC"] = tc
        st = cluster._LloydState(n, 32, k, Xt.device)
        cluster._estep(Xt, Ct, st, True)
        torch.cuda.synchronize()
        res[tc] = (st.labels.cpu().numpy().copy(), st.sums.cpu().numpy().copy(), st.counts.cpu().numpy().copy())
    nd = int((res["1"][0] != res["0"][0]).sum())
    want = cr.lloyd_estep(X, init) if n < 100000 else res["0"][0]
    print(n, k, scale, "tc-vs-simt label diffs", nd, "simt-vs-oracle", int((res["0"][0] != want).sum()),
          "tc-vs-oracle", int((res["1"][0] != want).sum()), "counts eq", np.array_equal(res["1"][2], res["0"][2]),
          "sums maxrel", float(np.abs(res["1"][1] - res["0"][1]).max() / (np.abs(res["0"][1]).max() + 1e-30)), flush=True)
