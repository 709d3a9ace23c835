import os, sys, numpy as np, torch
sys.path[:0] = [".", "vae-for-hybrid-music-clustering_b200"]
from vaeclust import cluster
def blobs(n, d, k, seed, spread=4.0):
    rng = np.random.default_rng(seed)
    cent = rng.normal(0, spread, size=(k, d))
    g = rng.integers(0, k, size=n)
    return (cent[g] + rng.normal(size=(n, d))).astype(np.float32), g
for (n, k, scale) in [(30001, 16, 1.0), (50000, 32, 1e-3), (2000003, 32, 1.0)]:
    X, _ = blobs(n, 32, k, seed=n + k); X = (X * scale).astype(np.float32)
    X = X - X.mean(0)
    init = X[np.random.default_rng(3).permutation(n)[:k]].copy()
    Xt, C = torch.from_numpy(X).cuda(), torch.from_numpy(init).cuda().contiguous()
    for it in range(12):
        res = {}
        for tc in ("1", "0"):
            os.environ["VAECLUST_KMEANS_TC"] = tc
            st = cluster._LloydState(n, 32, k, Xt.device)
            st.labels.fill_(-1)
            cluster._estep(Xt, C, st, True)
            torch.cuda.synchronize()
            res[tc] = (st.labels.cpu().numpy().copy(), st)
        bad = np.where(res["1"][0] != res["0"][0])[0]
        msg = ""
        if bad.size:
            Cn = C.cpu().numpy().astype(np.float64)
            for r in bad[:5]:
                x = X[r].astype(np.float64)
                d2 = ((x[None] - Cn) ** 2).sum(-1)
                a_, b_ = res["1"][0][r], res["0"][0][r]
                cm = np.sqrt((Cn ** 2).sum(-1).max())
                msg += f" [r={r} tc={a_} simt={b_} margin/(|x|cmax)={abs(d2[a_]-d2[b_])/(np.linalg.norm(x)*cm):.3e} sorted={np.sort(d2)[:3]}]"
        print(n, k, "iter", it, "diffs", bad.size, msg, flush=True)
        st = res["0"][1]
        cluster._finalize(C, st)
        C = st.centers_new.clone().contiguous()
