"""Diagnostic (GPU box): where do label differences between vaeclust's E-step and scikit-learn's come from at k=32?
One E-step from the SAME centred float32 array and the SAME centres through (a) the tcgen05 kernel, (b) the FFMA
kernel, (c) scikit-learn's lloyd_iter_chunked_dense (BLAS sgemm); differing points are then examined in float64:
a point whose two candidate centroids are closer than the fp32 rounding error of ||c||^2 - 2 x.c is a TIE that the
summation order of the sgemm decides, not an arithmetic error."""
import os
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT), str(ROOT / "vae-for-hybrid-music-clustering_b200")]
from vaeclust import cluster  # noqa: E402


def main():
    from sklearn.cluster._kmeans import _labels_inertia_threadpool_limit
    rng = np.random.default_rng(1)
    n, d = 4_000_000, 32
    cent = rng.normal(0, 3, size=(16, d)).astype(np.float32)
    X = cent[rng.integers(0, 16, n)] + rng.standard_normal((n, d), dtype=np.float32)
    X -= X.mean(axis=0)
    perm = np.random.default_rng(2).permutation(n)
    Xd = torch.from_numpy(X).cuda()
    for k in (8, 16, 32):
        C = X[perm[:k]].copy()
        Cd = torch.from_numpy(C).cuda()
        labs = {}
        for name, env in (("tc", "1"), ("ffma", "0")):
            os.environ["VAECLUST_KMEANS_TC"] = env
            st = cluster._LloydState(n, d, k, Xd.device)
            st.labels.fill_(-1)
            cluster._estep(Xd, Cd, st, update=False)
            labs[name] = st.labels.cpu().numpy()
        os.environ.pop("VAECLUST_KMEANS_TC")
        sk, _ = _labels_inertia_threadpool_limit(X, np.ones(n, dtype=np.float32), C, n_threads=os.cpu_count())
        # float64 ground truth
        X64, C64 = X.astype(np.float64), C.astype(np.float64)
        print(f"k={k}: tc!=ffma {int((labs['tc'] != labs['ffma']).sum())}  ffma!=sklearn "
              f"{int((labs['ffma'] != sk).sum())}  tc!=sklearn {int((labs['tc'] != sk).sum())}")
        idx = np.where(labs["ffma"] != sk)[0]
        if idx.size:
            xi = X64[idx]
            da = ((xi - C64[labs["ffma"][idx]]) ** 2).sum(1)
            db = ((xi - C64[sk[idx]]) ** 2).sum(1)
            scale = (xi ** 2).sum(1) + (C64[sk[idx]] ** 2).sum(1)
            rel = np.abs(da - db) / scale
            ours_better = int((da < db).sum())
            print(f"   {idx.size} differing points: exact margin / (|x|^2+|c|^2): max {rel.max():.3e} median "
                  f"{np.median(rel):.3e} (fp32 eps 6e-8); ours is the exact nearest for {ours_better}, sklearn for "
                  f"{int((db < da).sum())}")
        # which fp32 evaluation order reproduces sklearn?  sequential FMA chain vs numpy's own sgemm
        pd = (C * C).sum(1)[None, :] - 2.0 * (X[:200000] @ C.T)
        print(f"   numpy matmul argmin != sklearn on first 200k: {int((pd.argmin(1) != sk[:200000]).sum())};"
              f" != ffma: {int((pd.argmin(1) != labs['ffma'][:200000]).sum())}")


if __name__ == "__main__":
    main()
