"""Test-only stand-ins for the CUDA entry points of vaeclust.cluster, built on the oracle, so that the HOST logic
(Lloyd driver, sharding, collectives) can run under gloo on CPU tensors.  Never imported by the product."""
import numpy as np
import torch

from oracle import cluster_ref as cr


def install(monkeypatch_or_module, cluster):
    def setattr_(name, fn):
        if hasattr(monkeypatch_or_module, "setattr"):
            monkeypatch_or_module.setattr(cluster, name, fn)
        else:
            setattr(cluster, name, fn)

    def as_dev(X):
        t = X if isinstance(X, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(np.asarray(X, dtype=np.float32)))
        return t.to(torch.float32).contiguous()

    def col_stats(X, group=None):
        from vaeclust import dist as vdist
        x = X.double()
        mean, var, n = x.mean(0), x.var(0, unbiased=False), X.shape[0]
        if group is not None and vdist.world_size(group) > 1:
            mean, var, n = vdist.merge_moments(mean, var, n, group)
        return mean, var, n

    def col_affine(X, shift, scale):
        out = (X.double() - shift).float()
        if scale is not None:
            out = (out.double() / scale).float()
        return out

    class State:
        def __init__(self, N, d, k, device):
            self.labels = torch.empty(N, dtype=torch.int32)
            self.stats = torch.zeros(k * d + k + 2, dtype=torch.float64)
            self.nchg = torch.zeros(1, dtype=torch.int64)
            self.ctl = torch.zeros(4, dtype=torch.int32)
            self.centers_new = torch.empty((k, d), dtype=torch.float32)
            self.k, self.d = k, d
        sums = property(lambda s: s.stats[: s.k * s.d])
        counts = property(lambda s: s.stats[s.k * s.d: s.k * s.d + s.k])
        shift = property(lambda s: s.stats[s.k * s.d + s.k + 1:])

    def estep(X, centers, st, update):
        Xn, Cn = X.numpy(), centers.numpy()
        new = torch.from_numpy(cr.lloyd_estep(Xn, Cn))
        st.nchg[0] = int((new != st.labels).sum())
        st.labels.copy_(new)
        if update:
            sums = np.zeros((st.k, st.d))
            np.add.at(sums, new.numpy(), Xn.astype(np.float64))
            st.sums.copy_(torch.from_numpy(sums.reshape(-1)))
            st.counts.copy_(torch.from_numpy(np.bincount(new.numpy(), minlength=st.k).astype(np.float64)))

    def finalize(centers, st):
        cnt = st.counts.numpy()
        sums = st.sums.numpy().reshape(st.k, st.d)
        new = centers.numpy().copy()
        nz = cnt > 0
        new[nz] = (sums[nz] / cnt[nz, None]).astype(np.float32)
        if (~nz).any():
            new[~nz] = new[int(np.argmax(cnt))]
        st.centers_new.copy_(torch.from_numpy(new))
        st.shift[0] = float(((new.astype(np.float64) - centers.numpy().astype(np.float64)) ** 2).sum())

    # the sync-free loop's entry points (vc_kmeans_lloyd_*): same device-side state machine, on CPU tensors
    def lloyd_begin(st):
        st.ctl.zero_()

    def lloyd_estep(X, centers, st):
        if int(st.ctl[0]) != 0:
            return
        estep(X, centers, st, True)
        st.stats[st.k * st.d + st.k] = float(st.nchg[0])

    def lloyd_finalize(centers, st, tol):
        if int(st.ctl[0]) != 0:
            return
        if (st.counts.numpy() == 0).any():
            st.ctl[0] = 3
            return
        finalize(centers, st)
        centers.copy_(st.centers_new)
        st.ctl[1] += 1
        if float(st.stats[st.k * st.d + st.k]) == 0.0:
            st.ctl[0] = 1
        elif float(st.shift[0]) <= tol:
            st.ctl[0] = 2

    def sqdist_rows(X, Q):
        diff = Q.float().unsqueeze(1) - X.float().unsqueeze(0)
        return (diff * diff).sum(-1)

    def point_sqdist(X, centers, st):
        diff = X.numpy() - centers.numpy()[st.labels.numpy()]
        d2 = (diff * diff).sum(1)
        return torch.tensor([float(d2.astype(np.float64).sum())], dtype=torch.float64), torch.from_numpy(d2)

    def sil_prepare(Xd, order, mean64):
        Xs = (Xd.double() - mean64).float()[order].contiguous()
        return Xs, (Xs * Xs).sum(1)

    def sil_block(Xs, sqn, off, k, r0, r1, lab_sorted, counts, s_sorted, s_sum):
        s = cr.silhouette_samples(Xs.numpy(), lab_sorted.numpy())[r0:r1]
        s_sorted[r0:r1] = torch.from_numpy(s.astype(np.float32))
        s_sum[0] = float(s.sum())

    def cluster_sums(Xd, lab32, k):
        N, d = Xd.shape
        sums = np.zeros((k, d))
        np.add.at(sums, lab32.numpy(), Xd.numpy().astype(np.float64))
        cnt = np.bincount(lab32.numpy(), minlength=k).astype(np.float64)
        return torch.from_numpy(np.concatenate([sums.reshape(-1), cnt]))

    def dbi_intra(Xd, lab32, centroids, k):
        diff = Xd.numpy().astype(np.float64) - centroids.numpy()[lab32.numpy()]
        dist = np.sqrt((diff * diff).sum(1))
        return torch.from_numpy(np.bincount(lab32.numpy(), weights=dist, minlength=k))

    for name, fn in dict(_cluster_sums=cluster_sums, _dbi_intra=dbi_intra).items():
        setattr_(name, fn)
    for name, fn in dict(_as_dev_f32=as_dev, _col_stats=col_stats, _col_affine=col_affine, _LloydState=State,
                         _estep=estep, _finalize=finalize, _lloyd_begin=lloyd_begin, _lloyd_estep=lloyd_estep,
                         _lloyd_finalize=lloyd_finalize, _sqdist_rows=sqdist_rows, _point_sqdist=point_sqdist, _sil_prepare=sil_prepare,
                         _sil_block=sil_block).items():
        setattr_(name, fn)
