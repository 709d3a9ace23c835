"""Latent-space clustering evaluation on B200: KMeans (Lloyd), silhouette, Davies-Bouldin, StandardScaler.

sklearn-shaped call signatures as the reference uses them (07:67-71, 13:62-121, 16:57-96,183, 20:40-69, 22:55-99), so
`from vaeclust.cluster import KMeans, silhouette_score, davies_bouldin_score, StandardScaler` replaces the sklearn
imports of those scripts.  Inputs may be numpy arrays (copied to the GPU once) or CUDA tensors (zero-copy).
All O(N) / O(N^2) arithmetic runs in libvaeclust.so; the host only drives the Lloyd loop, draws the k-means++ random
numbers with numpy's RandomState exactly as sklearn does, and finishes O(k^2 d) scalars.  No CPU fallback.

Multi-GPU: pass `group=<torch.distributed process group>`; X is then this rank's row shard (KMeans, DBI) or the
replicated matrix of which this rank evaluates its row block (silhouette).  See DESIGN.md (e).
"""
from __future__ import annotations

import warnings
from typing import Optional, Tuple

import numpy as np
import torch

from . import dist as vdist
from ._lib import VaeclustError, lib, ptr, stream_ptr
from .ops import _workspace


def _device() -> torch.device:
    if not torch.cuda.is_available():
        raise VaeclustError("vaeclust.cluster needs a CUDA device (there is no CPU path)")
    return torch.device("cuda", torch.cuda.current_device())


def _as_dev_f32(X) -> torch.Tensor:
    """[N,d] float32 contiguous CUDA tensor from numpy / torch input."""
    if isinstance(X, torch.Tensor):
        t = X if X.is_cuda else X.to(_device())
    else:
        t = torch.from_numpy(np.ascontiguousarray(np.asarray(X, dtype=np.float32))).to(_device())
    if t.dim() != 2:
        raise ValueError(f"Expected 2D array, got {t.dim()}D array instead")
    return t.to(torch.float32).contiguous()


def _pad_cols(t: torch.Tensor) -> torch.Tensor:
    """Kernels need d % 4 == 0 (16-byte rows); zero columns change no distance."""
    d = t.shape[1]
    if d % 4 == 0:
        return t
    out = torch.zeros((t.shape[0], (d + 3) // 4 * 4), dtype=t.dtype, device=t.device)
    out[:, :d] = t
    return out


def _col_stats(X: torch.Tensor, group=None) -> Tuple[torch.Tensor, torch.Tensor, int]:
    """(mean, population variance) per column in float64 on device, and the global row count."""
    N, d = X.shape
    mean = torch.empty(d, dtype=torch.float64, device=X.device)
    var = torch.empty(d, dtype=torch.float64, device=X.device)
    ws = _workspace(lib().size("vc_col_stats_workspace_bytes", d), X.device)
    lib().call("vc_col_stats", ptr(X), N, d, ptr(mean), ptr(var), ptr(ws), ws.numel(), stream_ptr())
    if group is not None and vdist.world_size(group) > 1:
        mean, var, N = vdist.merge_moments(mean, var, N, group)
    return mean, var, N


def _col_affine(X: torch.Tensor, shift: torch.Tensor, scale: Optional[torch.Tensor]) -> torch.Tensor:
    out = torch.empty_like(X)
    lib().call("vc_col_affine", ptr(X), X.shape[0], X.shape[1], ptr(shift), ptr(scale), ptr(out), stream_ptr())
    return out


class StandardScaler:
    """sklearn.preprocessing.StandardScaler as used at 06:291-292, 07:67-68, 13:194, 16:178 (with_mean, with_std)."""

    def fit(self, X, group=None):
        Xd = _as_dev_f32(X)
        mean, var, n = _col_stats(Xd, group)
        scale = torch.sqrt(var)
        # sklearn's _handle_zeros_in_scale: (near-)constant columns are not scaled
        eps = 10 * torch.finfo(torch.float64).eps
        scale = torch.where(scale < eps * torch.abs(mean).clamp_min(1.0), torch.ones_like(scale), scale)
        self._mean_d, self._scale_d = mean, scale
        self.mean_, self.var_, self.scale_ = mean.cpu().numpy(), var.cpu().numpy(), scale.cpu().numpy()
        self.n_samples_seen_ = int(n)
        return self

    def transform(self, X):
        is_np = not isinstance(X, torch.Tensor)
        out = _col_affine(_as_dev_f32(X), self._mean_d, self._scale_d)
        return out.cpu().numpy() if is_np else out

    def fit_transform(self, X, y=None, group=None):
        return self.fit(X, group).transform(X)


class PCA:
    """sklearn.decomposition.PCA as scripts 09 and 22 use it (09:77-80, 22:122-124: PCA(n_components, random_state)
    .fit_transform(X) on a few hundred features): the `covariance_eigh` solver sklearn selects for such shapes --
    column means and the d x d covariance X^T X on the GPU (fp32 FFMA GEMM with a fixed-order split over the rows), the
    eigendecomposition of the small symmetric matrix on the host in float64, the projection on the GPU again.
    Component signs follow sklearn's svd_flip (largest-magnitude entry of each component positive)."""

    def __init__(self, n_components: Optional[int] = None, *, random_state=None, whiten: bool = False):
        if whiten:
            raise VaeclustError("vaeclust PCA: whiten is not used by the reference and not supported")
        self.n_components, self.random_state = n_components, random_state

    def fit(self, X, y=None):
        self._fit(X)
        return self

    def _fit(self, X):
        from .ops import _dense
        Xd = _as_dev_f32(X)
        N, d = Xd.shape
        nc = min(N, d) if self.n_components is None else int(self.n_components)
        if not 1 <= nc <= min(N, d):
            raise ValueError(f"n_components={nc} must be between 1 and min(n_samples, n_features)={min(N, d)}")
        mean64, _, _ = _col_stats(Xd)
        Xc = _col_affine(Xd, mean64, None)
        cov = torch.empty((d, d), dtype=torch.float32, device=Xd.device)
        _dense(2, 0, Xc, Xc, None, cov, N, d, 1, 1, d, False)          # LINEAR_WGRAD, fp32: Xc^T Xc
        C = cov.cpu().numpy().astype(np.float64) / (N - 1)
        C = 0.5 * (C + C.T)
        evals, evecs = np.linalg.eigh(C)
        evals, evecs = evals[::-1], evecs[:, ::-1]
        evals = np.clip(evals, 0.0, None)
        comps = evecs.T
        signs = np.sign(comps[np.arange(comps.shape[0]), np.argmax(np.abs(comps), axis=1)])
        signs[signs == 0] = 1.0
        comps = comps * signs[:, None]
        total = evals.sum()
        self.mean_ = mean64.cpu().numpy().astype(np.float32)
        self.components_ = comps[:nc].astype(np.float32)
        self.explained_variance_ = evals[:nc].astype(np.float32)
        self.explained_variance_ratio_ = (evals[:nc] / total).astype(np.float32)
        self.singular_values_ = np.sqrt(evals[:nc] * (N - 1)).astype(np.float32)
        self.n_components_, self.n_features_in_, self.n_samples_ = nc, d, N
        self._comp_d = torch.from_numpy(self.components_).to(Xd.device).contiguous()
        self._mean_d = mean64
        return Xc

    def _project(self, Xc: torch.Tensor) -> torch.Tensor:
        from .ops import _dense
        out = torch.empty((Xc.shape[0], self.n_components_), dtype=torch.float32, device=Xc.device)
        _dense(0, 0, Xc, self._comp_d, None, out, Xc.shape[0], Xc.shape[1], 1, 1, self.n_components_, False)
        return out

    def transform(self, X):
        is_np = not isinstance(X, torch.Tensor)
        out = self._project(_col_affine(_as_dev_f32(X), self._mean_d, None))
        return out.cpu().numpy() if is_np else out

    def fit_transform(self, X, y=None):
        is_np = not isinstance(X, torch.Tensor)
        out = self._project(self._fit(X))
        return out.cpu().numpy() if is_np else out


# ------------------------------------------------------------------------------------------------- KMeans
def _is_same_clustering(a: np.ndarray, b: np.ndarray, k: int) -> bool:
    """sklearn:cluster/_k_means_common.pyx:_is_same_clustering -- equal up to a permutation of labels."""
    mapping = np.full(k, -1, dtype=np.int64)
    for x, y in zip(a.tolist(), b.tolist()):
        if mapping[x] == -1:
            mapping[x] = y
        elif mapping[x] != y:
            return False
    return True


class _LloydState:
    """Device buffers of one Lloyd run (allocated once per fit, reused across n_init restarts)."""

    def __init__(self, N: int, d: int, k: int, device):
        self.labels = torch.empty(N, dtype=torch.int32, device=device)
        # stats = [sums k*d | counts k | n_changed 1 | shift_sq 1]  (float64): one buffer, one all-reduce per iteration
        self.stats = torch.zeros(k * d + k + 2, dtype=torch.float64, device=device)
        self.nchg = torch.zeros(1, dtype=torch.int64, device=device)
        self.ctl = torch.zeros(4, dtype=torch.int32, device=device)      # {state, n_iter, -, -} of the sync-free loop
        self.centers_new = torch.empty((k, d), dtype=torch.float32, device=device)
        self.ws = torch.empty(max(lib().size("vc_kmeans_workspace_bytes", k, d), 1 << 16), dtype=torch.uint8, device=device)
        self.k, self.d = k, d

    @property
    def sums(self):
        return self.stats[: self.k * self.d]

    @property
    def counts(self):
        return self.stats[self.k * self.d: self.k * self.d + self.k]

    @property
    def shift(self):
        return self.stats[self.k * self.d + self.k + 1:]


def _estep(X, centers, st: _LloydState, update: bool):
    lib().call("vc_kmeans_assign_update", ptr(X), X.shape[0], X.shape[1], ptr(centers), st.k, ptr(st.labels),
               ptr(st.sums), ptr(st.counts), ptr(st.nchg), int(update), ptr(st.ws), st.ws.numel(), stream_ptr())


def _finalize(centers, st: _LloydState):
    lib().call("vc_kmeans_finalize", ptr(st.sums), ptr(st.counts), ptr(centers), st.k, st.d, ptr(st.centers_new),
               ptr(st.shift), stream_ptr())


def _point_sqdist(X, centers, st: _LloydState) -> Tuple[torch.Tensor, torch.Tensor]:
    out = torch.empty(X.shape[0], dtype=torch.float32, device=X.device)
    inertia = torch.empty(1, dtype=torch.float64, device=X.device)
    lib().call("vc_kmeans_inertia", ptr(X), X.shape[0], X.shape[1], ptr(centers), st.k, ptr(st.labels), ptr(inertia),
               ptr(out), ptr(st.ws), st.ws.numel(), stream_ptr())
    return inertia, out


def _relocate_empty(X, centers, st: _LloydState, counts_host: np.ndarray, group) -> None:
    """sklearn:cluster/_k_means_common.pyx:167-211 on the (all-reduced) sums/counts."""
    empty = np.where(counts_host == 0)[0]
    _, dist = _point_sqdist(X, centers, st)
    n_e = int(min(empty.size, dist.numel()))
    top_d, top_i = torch.topk(dist, n_e)
    rows = X[top_i]
    labs = st.labels[top_i].to(torch.int64)
    if group is not None and vdist.world_size(group) > 1:
        top_d, rows, labs = vdist.gather_top(top_d, rows, labs, n_e, group)
    if float(top_d.max()) == 0.0:
        return
    sums = st.sums.view(st.k, st.d)
    for idx in range(n_e):
        new_id, old_id = int(empty[idx]), int(labs[idx])
        x = rows[idx].to(torch.float64)
        sums[old_id] -= x
        sums[new_id] = x
        st.counts[new_id] = 1.0
        st.counts[old_id] -= 1.0


def _lloyd_begin(st: _LloydState) -> None:
    lib().call("vc_kmeans_lloyd_begin", ptr(st.ctl), ptr(st.ws), st.ws.numel(), st.k, st.d, stream_ptr())


def _lloyd_estep(X, centers, st: _LloydState) -> None:
    lib().call("vc_kmeans_lloyd_estep", ptr(X), X.shape[0], X.shape[1], ptr(centers), st.k, ptr(st.labels),
               ptr(st.stats), ptr(st.ctl), ptr(st.ws), st.ws.numel(), stream_ptr())


def _lloyd_finalize(centers, st: _LloydState, tol: float) -> None:
    lib().call("vc_kmeans_lloyd_finalize", ptr(st.stats), ptr(centers), ptr(st.centers_new), st.k, st.d, float(tol),
               ptr(st.ctl), stream_ptr())


LLOYD_SYNC_EVERY = 4   # iterations enqueued between two read-backs of the device-side convergence state


def lloyd(X: torch.Tensor, centers_init: torch.Tensor, max_iter: int = 300, tol: float = 0.0, group=None,
          state: Optional[_LloydState] = None, sync_every: Optional[int] = None):
    """_kmeans_single_lloyd (sklearn:cluster/_kmeans.py:630-758) on device.  X is this rank's (centred) shard.
    Returns (labels int32 [N_local], inertia float, centers float32 [k,d], n_iter).

    The loop is sync-free: per iteration the host enqueues the fused E+M pass, (sharded X) ONE all-reduce of
    [sums | counts | n_changed], and the M-step / convergence kernel; the convergence state lives on the device and is
    read back every `sync_every` iterations.  Iterations enqueued after convergence are no-ops on the device, so labels,
    centres and n_iter_ are the same as for a loop that checks after every iteration."""
    N, d = X.shape
    k = centers_init.shape[0]
    st = state or _LloydState(N, d, k, X.device)
    st.labels.fill_(-1)
    centers = centers_init.to(torch.float32).contiguous().clone()
    multi = group is not None and vdist.world_size(group) > 1
    every = int(sync_every or LLOYD_SYNC_EVERY)
    kd = k * d
    _lloyd_begin(st)
    n_iter, code = 0, 0
    while n_iter < max_iter and code == 0:
        for _ in range(min(every, max_iter - n_iter)):
            _lloyd_estep(X, centers, st)
            if multi:
                vdist.all_reduce_sum(st.stats[: kd + k + 1], group)
            _lloyd_finalize(centers, st, tol)
        ctl = st.ctl.cpu()   # the one host sync of these iterations
        code, n_iter = int(ctl[0]), int(ctl[1])
        if code == 3:
            # an empty cluster: this iteration goes through the host's relocation path (identical on every rank: the
            # statistics are all-reduced), then the device loop resumes
            host = st.stats[kd:].cpu().numpy()
            _relocate_empty(X, centers, st, host[:k], group)
            _finalize(centers, st)
            shift_sq = float(st.shift.item())
            centers.copy_(st.centers_new)
            n_iter += 1
            code = 1 if int(host[k]) == 0 else (2 if shift_sq <= tol else 0)
            st.ctl[:2] = torch.tensor([code, n_iter], dtype=torch.int32, device=X.device)
    if code != 1:   # not strict convergence: one more E-step with the final centres (_kmeans.py:742-754)
        _estep(X, centers, st, update=False)
    inertia, _ = _point_sqdist(X, centers, st)
    if multi:
        vdist.all_reduce_sum(inertia, group)
    return st.labels, float(inertia.item()), centers, n_iter


_HOST_CUMSUM_MAX = 1 << 21  # below this (and unsharded), the k-means++ prefix sum replays numpy's sequential float32 cumsum


def _sqdist_rows(X: torch.Tensor, Q: torch.Tensor) -> torch.Tensor:
    out = torch.empty((Q.shape[0], X.shape[0]), dtype=torch.float32, device=X.device)
    lib().call("vc_sqdist_rows", ptr(X), X.shape[0], X.shape[1], ptr(Q.contiguous()), Q.shape[0], ptr(out),
               stream_ptr())
    return out


def _fetch_rows(X: torch.Tensor, gids: np.ndarray, lo: int, group) -> torch.Tensor:
    """Rows with GLOBAL indices `gids` of a row-sharded matrix (this rank holds [lo, lo + N_local)): the owner of each
    row contributes it, everybody else zeros, one all-reduce (k-means++ candidates: a handful of rows)."""
    loc = torch.from_numpy(np.asarray(gids, dtype=np.int64) - lo).to(X.device)
    mine = (loc >= 0) & (loc < X.shape[0])
    rows = torch.zeros((loc.numel(), X.shape[1]), dtype=torch.float32, device=X.device)
    if bool(mine.any()):
        rows[mine] = X[loc[mine]]
    return vdist.all_reduce_sum(rows, group)


def kmeans_plusplus(X: torch.Tensor, k: int, random_state: np.random.RandomState, group=None):
    """_kmeans_plusplus (sklearn:cluster/_kmeans.py:180-278): greedy k-means++ with 2+floor(ln k) local trials.
    The RandomState draws happen on the host in sklearn's order (every rank holds the same generator state); the
    distance scans run on the GPU.

    One GPU and N <= 2^21: the potentials and the prefix sum are evaluated exactly as numpy does (float32 sequential
    cumsum, float32 dot), so the seeds are sklearn's own -- every real-size call site of the reference.  Larger N, or X
    sharded by rows over `group`: float64 prefix sums on the device (per-rank cumsum + exclusive prefix of the rank
    totals) and `randint` for the first seed; the draw sequence is still sklearn's but a candidate may differ where
    a uniform draw falls within rounding distance of a prefix-sum step, so parity there is statistical
    (SURVEY.md 8(f)-1), and so is the choice of seeds."""
    N, d = X.shape
    multi = group is not None and vdist.world_size(group) > 1
    n_trials = 2 + int(np.log(k))
    centers = torch.empty((k, d), dtype=torch.float32, device=X.device)
    indices = np.full(k, -1, dtype=np.int64)
    if not multi and N <= _HOST_CUMSUM_MAX:  # same call, same float32 weights as sklearn (sample_weight / sample_weight.sum())
        w = np.ones(N, dtype=np.float32)
        center_id = int(random_state.choice(N, p=w / w.sum()))
        centers[0] = X[center_id]
        indices[0] = center_id
        closest = _sqdist_rows(X, centers[0:1])[0]
        pot = float(np.float32(closest.cpu().numpy() @ np.ones(N, dtype=np.float32)))
        for c in range(1, k):
            rand_vals = random_state.uniform(size=n_trials) * pot
            cs = np.cumsum(closest.cpu().numpy())  # float32 sequential, like np.cumsum(sample_weight*closest)
            cand = np.searchsorted(cs, rand_vals)
            np.clip(cand, None, N - 1, out=cand)
            cand_t = torch.from_numpy(cand.astype(np.int64)).to(X.device)
            dist = _sqdist_rows(X, X[cand_t])
            torch.minimum(dist, closest.unsqueeze(0), out=dist)
            pots = dist.cpu().numpy() @ np.ones(N, dtype=np.float32)
            best = int(np.argmin(pots))
            pot = float(pots[best])
            closest = dist[best].clone()
            centers[c] = X[int(cand[best])]
            indices[c] = int(cand[best])
        return centers, indices
    # ---- large or sharded X: global row index = rank offset + local index
    world, rank = (vdist.world_size(group), vdist.rank(group)) if multi else (1, 0)
    sizes = torch.zeros(world, dtype=torch.int64, device=X.device)
    sizes[rank] = N
    vdist.all_reduce_sum(sizes, group)
    sizes_h = sizes.cpu().numpy()
    lo = int(sizes_h[:rank].sum())
    n_total = int(sizes_h.sum())
    center_id = int(random_state.randint(n_total))
    centers[0] = _fetch_rows(X, np.array([center_id]), lo, group)[0]
    indices[0] = center_id
    closest = _sqdist_rows(X, centers[0:1])[0]

    def rank_totals(v: torch.Tensor) -> np.ndarray:
        t = torch.zeros(world, dtype=torch.float64, device=X.device)
        t[rank] = v.sum(dtype=torch.float64)
        return vdist.all_reduce_sum(t, group).cpu().numpy()

    tot = rank_totals(closest)
    for c in range(1, k):
        pot = float(tot.sum())
        rand_vals = random_state.uniform(size=n_trials) * pot
        prefix = float(tot[:rank].sum())
        cs = torch.cumsum(closest.to(torch.float64), 0) + prefix
        rv = torch.from_numpy(rand_vals).to(X.device)
        loc = torch.searchsorted(cs, rv)
        # the rank whose prefix range holds the value owns the candidate; values beyond the last step clip to N-1
        upper = prefix + float(tot[rank])
        mine = (rv > prefix if rank > 0 else torch.ones_like(rv, dtype=torch.bool)) & \
               ((rv <= upper) if rank < world - 1 else torch.ones_like(rv, dtype=torch.bool))
        gid = torch.where(mine, loc.clamp_max(N - 1) + lo, torch.zeros_like(loc))
        cand = vdist.all_reduce_sum(gid, group).cpu().numpy()
        np.clip(cand, None, n_total - 1, out=cand)
        dist = _sqdist_rows(X, _fetch_rows(X, cand, lo, group))
        torch.minimum(dist, closest.unsqueeze(0), out=dist)
        per_rank = torch.zeros((world, n_trials), dtype=torch.float64, device=X.device)
        per_rank[rank] = dist.sum(dim=1, dtype=torch.float64)
        per_rank = vdist.all_reduce_sum(per_rank, group).cpu().numpy()
        best = int(np.argmin(per_rank.sum(axis=0)))
        tot = per_rank[:, best].copy()
        closest = dist[best].clone()
        centers[c] = _fetch_rows(X, cand[best: best + 1], lo, group)[0]
        indices[c] = int(cand[best])
    return centers, indices


def _skip_kmeans_plusplus_draws(n: int, k: int, random_state: np.random.RandomState, exact: bool) -> None:
    """Advance `random_state` by exactly what one k-means++ initialisation consumes (restarts handled by another rank)."""
    if exact:
        w = np.ones(n, dtype=np.float32)
        random_state.choice(n, p=w / w.sum())
    else:
        random_state.randint(n)
    for _ in range(1, k):
        random_state.uniform(size=2 + int(np.log(k)))


class KMeans:
    """sklearn.cluster.KMeans as the reference uses it: KMeans(n_clusters, n_init, random_state[, init, max_iter,
    tol]).fit / fit_predict; attributes labels_ (int32), cluster_centers_ (float32), inertia_, n_iter_.
    Lloyd only (sklearn's default algorithm).  `init` may be "k-means++" (default), "random" or an ndarray.

    Multi-GPU (`group` = a torch.distributed process group), two ways to use the ranks:
      parallel="points"    X is THIS RANK'S ROW SHARD of the data; every Lloyd iteration all-reduces the per-centroid
                           statistics, k-means++ scans are sharded too.  labels_ are the local rows' labels.
      parallel="restarts"  X is the full (small) matrix on every rank; the n_init restarts are dealt round-robin to
                           the ranks (each advances the shared RandomState past the restarts it skips, so restart i
                           starts from the seeds sklearn would draw for it), then all-reduce(min) on the inertia and a
                           broadcast of the winner (sklearn:cluster/_kmeans.py:1506-1541; ties -> lowest restart)."""

    def __init__(self, n_clusters: int = 8, *, init="k-means++", n_init="auto", max_iter: int = 300,
                 tol: float = 1e-4, verbose: int = 0, random_state=None, copy_x: bool = True, algorithm: str = "lloyd",
                 group=None, parallel: str = "points"):
        if algorithm not in ("lloyd", "auto", "full"):
            raise VaeclustError("vaeclust KMeans implements Lloyd only")
        if parallel not in ("points", "restarts"):
            raise ValueError("parallel must be 'points' or 'restarts'")
        self.n_clusters, self.init, self.n_init = int(n_clusters), init, n_init
        self.max_iter, self.tol, self.verbose = int(max_iter), float(tol), verbose
        self.random_state, self.copy_x, self.algorithm, self.group = random_state, copy_x, algorithm, group
        self.parallel = parallel

    def _check_rs(self):
        rs = self.random_state
        if rs is None or isinstance(rs, (int, np.integer)):
            return np.random.RandomState(rs)
        return rs

    def fit(self, X, y=None, sample_weight=None):
        if sample_weight is not None:
            raise VaeclustError("vaeclust KMeans: sample_weight is not used by the reference and not supported")
        return_numpy = not isinstance(X, torch.Tensor)
        Xd = _as_dev_f32(X)
        N, d_true = Xd.shape
        k = self.n_clusters
        world = vdist.world_size(self.group) if self.group is not None else 1
        restarts = world > 1 and self.parallel == "restarts"
        multi = world > 1 and not restarts                      # rows sharded
        grp = self.group if multi else None
        n_total = N if not multi else int(vdist.all_reduce_sum(torch.tensor([N], dtype=torch.int64, device=Xd.device),
                                                               self.group).item())
        if n_total < k:
            raise ValueError(f"n_samples={n_total} should be >= n_clusters={k}.")
        init = self.init
        init_is_array = not isinstance(init, str)
        if init_is_array:
            init_t = _as_dev_f32(init)
            if tuple(init_t.shape) != (k, d_true):
                raise ValueError(f"The shape of the initial centers {tuple(init_t.shape)} does not match "
                                 f"(n_clusters, n_features) = ({k}, {d_true}).")
        elif init not in ("k-means++", "random"):
            raise ValueError(f"init should be 'k-means++', 'random' or an array, got {init!r}")
        n_init = self.n_init
        if n_init == "auto":
            n_init = 1 if (init_is_array or init == "k-means++") else 10
        if init_is_array and n_init != 1:
            warnings.warn("Explicit initial center position passed: performing only one init in KMeans instead of "
                          f"n_init={n_init}.", RuntimeWarning, stacklevel=2)
            n_init = 1
        # tolerance and centring (_kmeans.py:285-293, 1487-1490)
        mean64, var64, _ = _col_stats(Xd, grp)
        tol_abs = float(var64.mean().item()) * self.tol if self.tol != 0 else 0.0
        mean32 = mean64.to(torch.float32)
        Xc = _pad_cols(_col_affine(Xd, mean32.to(torch.float64), None))
        d = Xc.shape[1]
        rs = self._check_rs()
        st = _LloydState(N, d, k, Xc.device)
        rank = vdist.rank(self.group) if world > 1 else 0
        exact_draws = not multi and N <= _HOST_CUMSUM_MAX
        best = None   # (labels, inertia, centers, n_iter, restart index)
        for i in range(int(n_init)):
            if restarts and i % world != rank:
                # another rank runs this restart: consume its share of the random stream
                if init == "k-means++":
                    _skip_kmeans_plusplus_draws(N, k, rs, exact_draws)
                elif init == "random":
                    rs.choice(N, size=k, replace=False, p=np.ones(N) / N)
                continue
            if init_is_array:
                c0 = init_t - mean32
            elif init == "k-means++":
                c0, _idx = kmeans_plusplus(Xc, k, rs, grp)
            else:
                # sklearn passes p=sample_weight/sum (a different numpy algorithm than p=None): same call here
                seeds = rs.choice(n_total, size=k, replace=False, p=np.ones(n_total) / n_total)
                if multi:
                    sizes = torch.zeros(world, dtype=torch.int64, device=Xc.device)
                    sizes[rank] = N
                    lo = int(vdist.all_reduce_sum(sizes, grp).cpu().numpy()[:rank].sum())
                    c0 = _fetch_rows(Xc, seeds, lo, grp)
                else:
                    c0 = Xc[torch.from_numpy(seeds).to(Xc.device)]
            c0 = _pad_cols(c0.contiguous())
            labels, inertia, centers, n_iter = lloyd(Xc, c0, self.max_iter, tol_abs, grp, st)
            take = best is None
            if not take and inertia < best[1]:
                take = multi or not _is_same_clustering(labels.cpu().numpy(), best[0].cpu().numpy(), k)
            if take:
                best = (labels.clone(), inertia, centers.clone(), n_iter, i)
        if restarts:
            # all-reduce(min) on (inertia, restart index), then the winner broadcasts its result
            mine = torch.tensor([best[1] if best is not None else float("inf"),
                                 float(best[4]) if best is not None else float("inf")], dtype=torch.float64,
                                device=Xc.device)
            allv = [torch.empty_like(mine) for _ in range(world)]
            torch.distributed.all_gather(allv, mine, group=self.group)
            table = torch.stack(allv).cpu().numpy()
            order = np.lexsort((table[:, 1], table[:, 0]))           # lowest inertia, then lowest restart index
            owner = int(order[0])
            src = torch.distributed.get_global_rank(self.group, owner)
            lab = best[0] if rank == owner else torch.empty(N, dtype=torch.int32, device=Xc.device)
            cen = best[2] if rank == owner else torch.empty((k, d), dtype=torch.float32, device=Xc.device)
            torch.distributed.broadcast(lab, src=src, group=self.group)
            torch.distributed.broadcast(cen, src=src, group=self.group)
            meta = torch.tensor([float(best[3]) if rank == owner else 0.0], dtype=torch.float64, device=Xc.device)
            torch.distributed.broadcast(meta, src=src, group=self.group)
            best = (lab, float(table[owner, 0]), cen, int(meta.item()), int(table[owner, 1]))
        labels, inertia, centers, n_iter = best[:4]
        centers = centers[:, :d_true] + mean32
        self._labels_d, self._centers_d = labels, centers
        self.labels_ = labels.cpu().numpy() if return_numpy else labels
        self.cluster_centers_ = centers.cpu().numpy() if return_numpy else centers
        self.inertia_, self.n_iter_ = inertia, n_iter
        self.n_features_in_ = d_true
        if not multi:
            distinct = int(torch.unique(labels).numel())
            if distinct < k:
                warnings.warn(f"Number of distinct clusters ({distinct}) found smaller than n_clusters ({k}). "
                              "Possibly due to duplicate points in X.", RuntimeWarning, stacklevel=2)
        return self

    def fit_predict(self, X, y=None, sample_weight=None):
        return self.fit(X, sample_weight=sample_weight).labels_

    def predict(self, X):
        return_numpy = not isinstance(X, torch.Tensor)
        Xd = _pad_cols(_as_dev_f32(X))
        C = _pad_cols(self._centers_d.contiguous())
        st = _LloydState(Xd.shape[0], Xd.shape[1], self.n_clusters, Xd.device)
        _estep(Xd, C, st, update=False)
        return st.labels.cpu().numpy() if return_numpy else st.labels


# --------------------------------------------------------------------------------------------- silhouette
def _encode_labels(labels, device) -> Tuple[torch.Tensor, int]:
    """LabelEncoder (sklearn:_unsupervised.py:300-303): labels -> 0..k-1 (int64 on device)."""
    lt = labels if isinstance(labels, torch.Tensor) else torch.from_numpy(np.asarray(labels))
    lt = lt.to(device).reshape(-1).to(torch.int64)
    uniq, inv = torch.unique(lt, return_inverse=True)
    return inv, int(uniq.numel())


def _check_number_of_labels(k: int, n: int) -> None:
    if not 1 < k < n:
        raise ValueError(f"Number of labels is {k}. Valid values are 2 to n_samples - 1 (inclusive)")


def _sil_prepare(Xd: torch.Tensor, order: torch.Tensor, mean64: torch.Tensor):
    """Rows permuted into cluster order and centred, plus their squared norms."""
    N, d = Xd.shape
    Xs = torch.empty_like(Xd)
    sqn = torch.empty(N, dtype=torch.float32, device=Xd.device)
    lib().call("vc_silhouette_gather", ptr(Xd), N, d, ptr(order), ptr(mean64), ptr(Xs), ptr(sqn), stream_ptr())
    return Xs, sqn


import os as _os

# bf16 hi/lo split of both operands: hi.hi + hi.lo + lo.hi (+ lo.lo, 2^-16 of the product, with VAECLUST_SIL_MMA=4)
SILHOUETTE_MMA_PER_PRODUCT = 4 if _os.environ.get("VAECLUST_SIL_MMA", "3").startswith("4") else 3
SILHOUETTE_PRECISION = "tc"  # "tc": tcgen05 bf16 hi/lo split (d <= 128); "fp32": FFMA tile kernel


def _sil_block(Xs, sqn, off, k: int, r0: int, r1: int, lab_sorted, counts, s_sorted, s_sum) -> None:
    """Rows [r0, r1) against all columns: per-row silhouette into s_sorted[r0:r1] and their sum into s_sum."""
    N, d = Xs.shape
    nrows = r1 - r0
    tiles = (nrows + 127) // 128
    sms = lib().size("vc_sm_count")
    nsplit = int(max(1, min(32, (2 * sms) // max(tiles, 1))))
    dp = (d + 63) // 64 * 64
    if SILHOUETTE_PRECISION == "tc" and dp <= 128:
        # the tensor-core kernel writes several partial planes per split (one per column part of a tile)
        planes = lib().size("vc_silhouette_tc_planes", dp)
        sums = torch.empty((planes * nsplit, nrows, k), dtype=torch.float32, device=Xs.device)
        H = torch.empty((N, dp), dtype=torch.bfloat16, device=Xs.device)
        L = torch.empty((N, dp), dtype=torch.bfloat16, device=Xs.device)
        sqn_hl = torch.empty_like(sqn)  # norms of the split rows: the formula must cancel against the same operands
        lib().call("vc_silhouette_split_bf16", ptr(Xs), N, d, dp, ptr(H), ptr(L), ptr(sqn_hl), stream_ptr())
        lib().call("vc_silhouette_rows_tc", ptr(H), ptr(L), ptr(Xs), d, ptr(sqn_hl), N, dp, ptr(off), k, r0, r1, nsplit,
                   ptr(sums), stream_ptr())
        nsplit *= planes
    else:
        sums = torch.empty((nsplit, nrows, k), dtype=torch.float32, device=Xs.device)
        lib().call("vc_silhouette_rows", ptr(Xs), ptr(sqn), N, d, ptr(off), k, r0, r1, nsplit, ptr(sums),
                   stream_ptr())
    ws = _workspace(lib().size("vc_silhouette_finalize_workspace_bytes"), Xs.device)
    lib().call("vc_silhouette_finalize", ptr(sums), nsplit, ptr(lab_sorted[r0:]), ptr(counts), k, nrows,
               ptr(s_sorted[r0:]), ptr(s_sum), ptr(ws), ws.numel(), stream_ptr())


def silhouette_samples(X, labels, *, metric: str = "euclidean", group=None, _return_sum: bool = False):
    """sklearn.metrics.silhouette_samples (euclidean only).  With `group`, every rank holds the full X and
    evaluates rows [r*N/G, (r+1)*N/G); per-row values are all-gathered."""
    if metric != "euclidean":
        raise VaeclustError("vaeclust silhouette supports metric='euclidean' only (what the reference uses)")
    return_numpy = not isinstance(X, torch.Tensor)
    Xd = _pad_cols(_as_dev_f32(X))
    N, d = Xd.shape
    lab, k = _encode_labels(labels, Xd.device)
    if lab.numel() != N:
        raise ValueError(f"Found input variables with inconsistent numbers of samples: [{N}, {lab.numel()}]")
    _check_number_of_labels(k, N)
    # group rows by cluster (stable), centre for conditioning (distances are translation invariant)
    order = torch.argsort(lab, stable=True)
    lab_sorted = lab[order].to(torch.int32).contiguous()
    counts = torch.bincount(lab, minlength=k).to(torch.int64)
    off = torch.zeros(k + 1, dtype=torch.int64, device=Xd.device)
    off[1:] = torch.cumsum(counts, 0)
    mean64, _, _ = _col_stats(Xd)
    Xs, sqn = _sil_prepare(Xd, order, mean64)
    world, rank = (vdist.world_size(group), vdist.rank(group)) if group is not None else (1, 0)
    r0, r1 = vdist.shard_bounds(N, rank, world)
    s_sorted = torch.zeros(N, dtype=torch.float32, device=Xd.device)
    s_sum = torch.zeros(1, dtype=torch.float64, device=Xd.device)
    if r1 > r0:
        _sil_block(Xs, sqn, off, k, r0, r1, lab_sorted, counts, s_sorted, s_sum)
    if world > 1:
        vdist.all_reduce_sum(s_sum, group)
        if not _return_sum:
            vdist.all_reduce_sum(s_sorted, group)  # disjoint row blocks, zeros elsewhere
    if _return_sum:
        return float(s_sum.item()), N
    s = torch.empty_like(s_sorted)
    s[order] = s_sorted
    return s.cpu().numpy() if return_numpy else s


def silhouette_score(X, labels, *, metric: str = "euclidean", sample_size=None, random_state=None, group=None):
    """sklearn.metrics.silhouette_score -> python float (13:74,78; 16:68,73; 20:45; 22:61)."""
    if sample_size is not None:
        raise VaeclustError("vaeclust silhouette_score: sample_size is not used by the reference and not supported")
    total, n = silhouette_samples(X, labels, metric=metric, group=group, _return_sum=True)
    return float(np.float32(total / n)) if not isinstance(X, torch.Tensor) or X.dtype == torch.float32 else total / n


# ----------------------------------------------------------------------------------------- Davies-Bouldin
def _cluster_sums(Xd: torch.Tensor, lab32: torch.Tensor, k: int) -> torch.Tensor:
    """[k*d sums | k counts] in float64 for the local rows."""
    N, d = Xd.shape
    stats = torch.zeros(k * d + k, dtype=torch.float64, device=Xd.device)
    ws = _workspace(lib().size("vc_kmeans_workspace_bytes", k, d), Xd.device)
    lib().call("vc_cluster_sums", ptr(Xd), N, d, ptr(lab32), k, ptr(stats), ptr(stats[k * d:]), ptr(ws), ws.numel(),
               stream_ptr())
    return stats


def _dbi_intra(Xd: torch.Tensor, lab32: torch.Tensor, centroids: torch.Tensor, k: int) -> torch.Tensor:
    """sum over local rows of ||x - centroid(label)|| per cluster (float64)."""
    N, d = Xd.shape
    intra = torch.zeros(k, dtype=torch.float64, device=Xd.device)
    ws = _workspace(lib().size("vc_kmeans_workspace_bytes", k, d), Xd.device)
    lib().call("vc_dbi_intra", ptr(Xd), N, d, ptr(lab32), ptr(centroids), k, ptr(intra), ptr(ws), ws.numel(),
               stream_ptr())
    return intra


def davies_bouldin_score(X, labels, *, group=None) -> float:
    """sklearn.metrics.davies_bouldin_score (sklearn:_unsupervised.py:413-487; 13:94,98; 16:89,94).
    With `group`, X/labels are this rank's row shard; two tiny all-reduces (C5 of SURVEY.md 2.1)."""
    Xd = _pad_cols(_as_dev_f32(X))
    N, d = Xd.shape
    multi = group is not None and vdist.world_size(group) > 1
    if multi:
        lt = labels if isinstance(labels, torch.Tensor) else torch.from_numpy(np.asarray(labels))
        lab = lt.to(Xd.device).reshape(-1).to(torch.int64)
        kk = torch.tensor([int(lab.max().item()) + 1], dtype=torch.int64, device=Xd.device)
        k = int(vdist.all_reduce_max(kk, group).item())  # sharded callers pass labels already in 0..k-1
        n_total = int(vdist.all_reduce_sum(torch.tensor([N], dtype=torch.int64, device=Xd.device), group).item())
    else:
        lab, k = _encode_labels(labels, Xd.device)
        n_total = N
    _check_number_of_labels(k, n_total)
    lab32 = lab.to(torch.int32).contiguous()
    stats = _cluster_sums(Xd, lab32, k)
    if multi:
        vdist.all_reduce_sum(stats, group)
    counts = stats[k * d:]
    centroids = (stats[: k * d].view(k, d) / counts.clamp_min(1.0).unsqueeze(1)).contiguous()
    intra = _dbi_intra(Xd, lab32, centroids, k)
    if multi:
        vdist.all_reduce_sum(intra, group)
    # O(k^2 d) tail on the host in float64, as sklearn does
    intra_h = (intra / counts.clamp_min(1.0)).cpu().numpy()
    cent_h = centroids.cpu().numpy()
    diff = cent_h[:, None, :] - cent_h[None, :, :]
    M = np.sqrt(np.maximum((diff * diff).sum(-1), 0.0))
    if np.allclose(intra_h, 0) or np.allclose(M, 0):
        return 0.0
    M[M == 0] = np.inf
    comb = intra_h[:, None] + intra_h[None, :]
    return float(np.mean(np.max(comb / M, axis=1)))


# -------------------------------------------------------------------- the reference's own wrapper helpers
def safe_silhouette(X, labels):
    """13:62-80 / 16:57-75: None if < 2 clusters; DBSCAN noise (-1) dropped, needs >= 3 points; errors -> None."""
    labels = np.asarray(labels.cpu() if isinstance(labels, torch.Tensor) else labels)
    uniq = set(labels.tolist())
    if len(uniq) < 2:
        return None
    try:
        if -1 in uniq:
            mask = labels != -1
            if mask.sum() < 3 or len(set(labels[mask].tolist())) < 2:
                return None
            Xm = X[torch.from_numpy(mask).to(X.device)] if isinstance(X, torch.Tensor) else X[mask]
            return float(silhouette_score(Xm, labels[mask]))
        return float(silhouette_score(X, labels))
    except VaeclustError:
        raise   # a failed CUDA launch / missing library is not a "metric undefined" case
    except Exception:
        return None


def safe_db_index(X, labels):
    """13:83-100 / 16:78-96 (`safe_dbi`): same guards as safe_silhouette."""
    labels = np.asarray(labels.cpu() if isinstance(labels, torch.Tensor) else labels)
    uniq = set(labels.tolist())
    if len(uniq) < 2:
        return None
    try:
        if -1 in uniq:
            mask = labels != -1
            if mask.sum() < 3 or len(set(labels[mask].tolist())) < 2:
                return None
            Xm = X[torch.from_numpy(mask).to(X.device)] if isinstance(X, torch.Tensor) else X[mask]
            return float(davies_bouldin_score(Xm, labels[mask]))
        return float(davies_bouldin_score(X, labels))
    except VaeclustError:
        raise
    except Exception:
        return None


safe_dbi = safe_db_index


def safe_silhouette_nan(X, labels) -> float:
    """20:40-47 / 22:55-63 flavour: nan instead of None."""
    v = safe_silhouette(X, labels)
    return float("nan") if v is None else v


def eval_kmeans(X, k: int, seed: int, n_init: int = 20):
    """22:90-99: KMeans(k, random_state=seed, n_init=20) labels + silhouette."""
    km = KMeans(n_clusters=k, random_state=seed, n_init=n_init)
    labels = km.fit_predict(X)
    return labels, safe_silhouette_nan(X, labels)


__all__ = ["KMeans", "StandardScaler", "PCA", "silhouette_score", "silhouette_samples", "davies_bouldin_score",
           "safe_silhouette", "safe_db_index", "safe_dbi", "safe_silhouette_nan", "eval_kmeans", "lloyd",
           "kmeans_plusplus"]
