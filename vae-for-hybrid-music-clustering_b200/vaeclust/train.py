"""Training steps and loops of the reference's VAE scripts on the CUDA ops (12:246-309, 19:209-287, 06:314-360).

What is kept from the reference: batch order (the DataLoader's RandomSampler recipe is replayed, so one seed gives
the same batches), `drop_last=False`, the two different epoch-average conventions, the log/checkpoint formats and
the latent export.  What changes: the dataset lives in HBM and batches are gathered on the device; the loss,
its gradient and the three logged scalars come from one fused kernel; scalars are read back once per epoch
instead of 3 `.item()` syncs per step; the latent export skips the decoder (the reference runs it for nothing).
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import Callable, List, Optional

import numpy as np
import torch

from . import dist as vdist
from . import ops
from ._lib import VC_LOSS_MEAN_MEAN, VC_LOSS_SUM_MEAN
from .optim import Adam


def epoch_permutation(n: int) -> torch.Tensor:
    """The index order DataLoader(shuffle=True, num_workers=0) would produce for one epoch: the loader iterator
    draws its _base_seed from the global CPU generator, RandomSampler then draws its own seed from the same
    generator and calls torch.randperm(n, generator=fresh generator)."""
    torch.empty((), dtype=torch.int64).random_()  # the DataLoader iterator's _base_seed draw comes first
    seed = int(torch.empty((), dtype=torch.int64).random_().item())
    g = torch.Generator()
    g.manual_seed(seed)
    return torch.randperm(n, generator=g)


def batches(n: int, batch_size: int, shuffle: bool = True):
    order = epoch_permutation(n) if shuffle else torch.arange(n)
    for s in range(0, n, batch_size):  # drop_last=False (12:227, 19:192)
        yield order[s: s + batch_size]


def make_step(model, kind: str, lr: float, beta: float, group=None, capturable: bool = False, fused: str = "auto"):
    """The optimisation step for `model`: the fused four-launch step of vaeclust.fused (vc_mlpvae_step) for the MLP
    models when the dense precision is "bf16" and the shape qualifies, else the per-layer VAEStep.
    fused: "auto" | "never" | "always" (raise if the fused kernel cannot run this model)."""
    from . import fused as vfused
    want = fused != "never" and kind in ("mlp19", "mlp06", "ae22") and (ops.PRECISION == "bf16" or fused == "always")
    if want and next(model.parameters()).is_cuda and vfused.supported(model, kind):
        return vfused.FusedMLPStep(model, kind, lr, beta, group=group, capturable=capturable)
    if fused == "always":
        raise ValueError(f"make_step: the fused step does not support this {kind} model")
    return VAEStep(model, kind, lr, beta, group=group, capturable=capturable)


class VAEStep:
    """One optimisation step for any of the models.  kind: "conv12" | "mlp19" | "mlp06" | "ae22".
    Under a process group each rank passes ITS slice of the global batch; gradients are combined with weights
    B_local/B_global so the result equals the reference's single-process batch mean."""

    def __init__(self, model, kind: str, lr: float, beta: float, group=None, capturable: bool = False):
        self.model, self.kind, self.beta, self.group = model, kind, float(beta), group
        # single GPU, bf16 mode: the wide Linear weights (>= 2^24 elements: the two 67 M-parameter layers of script 12) are
        # updated by their weight-gradient kernel itself (ops._linear_wgrad_adam); their .grad is never materialised.
        # Set `fuse_fc_adam = False` (or VAECLUST_FUSE_FC_ADAM=0) to get ordinary gradients, e.g. to inspect them.
        self.fuse_fc_adam = os.environ.get("VAECLUST_FUSE_FC_ADAM", "1") != "0"
        wide = int(os.environ.get("VAECLUST_FUSE_FC_ADAM_NUMEL", str(1 << 24)))
        self._wide = [m.weight for m in model.modules() if isinstance(m, torch.nn.Linear) and m.weight.numel() >= wide]
        # bf16 shadows of those weights (forward / data gradient read them by TMA; the fused update refreshes them):
        # id(p) -> (tensor, p._version, p.data_ptr()) -- rebuilt when the parameter was changed by anything else
        self.bf16_shadows = os.environ.get("VAECLUST_FC_SHADOW", "1") != "0"
        self._shadow = {}
        self._dp_shadows = None      # data parallel: number of weights whose slices are gathered as bf16 (dist.ShardedGradSync)
        self.mode = VC_LOSS_SUM_MEAN if kind == "mlp19" else VC_LOSS_MEAN_MEAN
        self.multi = group is not None and vdist.world_size(group) > 1
        self.opt = Adam(model.parameters(), lr=lr, capturable=capturable)
        # data parallel: the big tensors go reduce-scatter -> Adam on the slice -> all-gather (sharded optimizer state);
        # VAECLUST_DP=allreduce selects the plain overlapped all-reduce + replicated Adam of round 1
        self.sync = None
        if self.multi:
            big = int(os.environ.get("VAECLUST_DP_BIG_NUMEL", str(1 << 20)))   # tensors at least this large are sharded
            if os.environ.get("VAECLUST_DP", "sharded") == "sharded":
                self.sync = vdist.ShardedGradSync(model.parameters(), group, self.opt, big_numel=big)
            else:
                self.sync = vdist.OverlappedGradSync(model.parameters(), group, big_numel=big)

    def __call__(self, x, *extra, eps: Optional[torch.Tensor] = None, global_batch: Optional[int] = None,
                 local_rows: Optional[int] = None):
        """`local_rows` overrides the weight of this rank's contribution (default: x.shape[0]); a rank whose shard of
        the last batch is empty passes one dummy row with local_rows=0 so that it still takes part in the collectives."""
        m = self.model
        fused = []
        if self.sync is not None:
            if (self._dp_shadows is None and self.bf16_shadows and ops.PRECISION == "bf16"
                    and hasattr(self.sync, "enable_bf16_gather") and os.environ.get("VAECLUST_DP_BF16_GATHER", "1") != "0"):
                # the wide weights travel between the ranks as bf16 (what forward / data gradient read); sync_params()
                # brings the fp32 masters up to date
                self._dp_shadows = self.sync.enable_bf16_gather(self._wide)
            rows = x.shape[0] if local_rows is None else int(local_rows)
            self.sync.begin(rows, global_batch or x.shape[0] * vdist.world_size(self.group))
        else:
            self.opt.zero_grad()
            if self.fuse_fc_adam and self._wide and ops.PRECISION == "bf16":
                fused = self._wide
                self.opt.begin_step()            # capturable: the device step counter advances here, once
                if self.bf16_shadows:
                    self.refresh_shadows()
                for p in fused:
                    p._vc_adam = self.opt
                    if self.bf16_shadows:
                        p._vc_shadow = self._shadow[id(p)][0]
        try:
            return self._run(m, x, extra, eps, fused)
        finally:
            for p in fused:
                p._vc_adam = None
                p._vc_shadow = None

    @torch.no_grad()
    def refresh_shadows(self) -> None:
        """Bring the bf16 shadows of the wide weights up to date with parameters that something other than the fused
        update changed (load_state_dict, an ordinary optimizer step): torch bumps `_version` on every in-place write, the
        C-ABI kernels do not.  Called at the start of every eager step and before every graph replay."""
        for p in self._wide:
            sh = self._shadow.get(id(p))
            if sh is None or sh[1] != p._version or sh[2] != p.data_ptr():
                if sh is None:
                    sh = (torch.empty(p.shape, dtype=torch.bfloat16, device=p.device), 0, 0)
                sh[0].copy_(p.detach())          # in place: a captured graph keeps reading this buffer
                self._shadow[id(p)] = (sh[0], p._version, p.data_ptr())

    def _run(self, m, x, extra, eps, fused):
        if self.kind == "conv12":
            lyr, mask = extra
            mu, lv = m.encode(x, lyr, mask)
        elif self.kind == "mlp19":
            c = extra[0] if extra else None
            mu, lv = m.encode(x, c)
        elif self.kind == "ae22":   # 22:150-156: plain MSE, no sampling, no KL
            xhat, _z = m(x)
            klr = torch.zeros(x.shape[0], dtype=torch.float32, device=x.device)
            loss, stats = ops.vae_loss(xhat, x, klr, 0.0, self.mode, 1, unit_grad=True)
            loss.backward()
            if self.multi:
                self.sync.finish()
            self.opt.step()
            if self.multi and hasattr(self.sync, "join"):
                self.sync.join()
            return stats
        else:
            mu, lv = m.encode(x)
        if eps is None:
            eps = torch.randn_like(mu)
        z, klr = ops.reparam_kl(mu, lv, eps)
        if self.kind == "conv12":
            xhat = m.audio_dec(z)
        elif self.kind == "mlp19":
            xhat = m.decode(z, extra[0] if extra else None)
        else:
            xhat = m.decode(z)
        loss, stats = ops.vae_loss(xhat, x, klr, self.beta, self.mode, mu.shape[1], unit_grad=True)
        loss.backward()
        if self.multi:
            self.sync.finish()
        self.opt.step()
        if self.multi and hasattr(self.sync, "join"):
            self.sync.join()
        return stats  # device tensor [loss, recon, kl] of the local batch


    def sync_params(self, force: bool = False) -> None:
        """Data parallel with sharded optimizer state: bring every rank's copy of the parameters up to date (the
        all-gather of the last step's slices is otherwise deferred into the next step).  Call before reading them."""
        if self.sync is not None and hasattr(self.sync, "flush"):
            self.sync.flush(force)


class GraphedVAEStep:
    """The same optimisation step recorded once into a CUDA graph and replayed: the ~125 kernel launches
    of a step are submitted by one cudaGraphLaunch, which removes the host launch gaps between the short kernels of
    the latent head and of the backward pass.  The first `warmup` calls run eagerly (they allocate optimizer state and
    warm the allocator), the next call captures; inputs are copied into static buffers, `eps` (optional) likewise, else
    it is drawn inside the graph from torch's graph-safe generator.  Returns the static [loss, recon, kl] tensor --
    read it before the next call.  Shapes must not change after capture (pass full batches).  Under a process group
    the overlapped NCCL gradient exchange is captured with the step (set TORCH_NCCL_ASYNC_ERROR_HANDLING=0 before
    init_process_group, as for any NCCL-in-graph capture); `global_batch` is fixed at capture."""

    def __init__(self, model, kind: str, lr: float, beta: float, warmup: int = 2, group=None, fused: str = "auto"):
        self.step = make_step(model, kind, lr, beta, group=group, capturable=True, fused=fused)
        self.warmup, self.calls = int(warmup), 0
        self.graph = None
        self.static_in = None
        self.static_eps = None
        self.out = None

    @property
    def opt(self):
        return self.step.opt

    def sync_params(self) -> None:
        if hasattr(self.step, "sync_params"):
            self.step.sync_params(force=self.graph is not None)

    def __call__(self, x, *extra, eps: Optional[torch.Tensor] = None, global_batch: Optional[int] = None):
        self.calls += 1
        if self.graph is None and self.calls <= self.warmup:
            return self.step(x, *extra, eps=eps, global_batch=global_batch)
        if self.graph is None:
            self.static_in = [t.clone() if torch.is_tensor(t) else t for t in (x,) + tuple(extra)]
            self.static_eps = eps.clone() if eps is not None else None
            if hasattr(self.step.opt, "zero_grad"):
                self.step.opt.zero_grad()   # gradients must be (re)allocated from the graph's private pool
            torch.cuda.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self.out = self.step(*self.static_in, eps=self.static_eps, global_batch=global_batch)
        else:
            if (eps is None) != (self.static_eps is None):
                raise ValueError("GraphedVAEStep: pass eps on every call or on none (fixed at capture)")
            for s, t in zip(self.static_in, (x,) + tuple(extra)):
                if torch.is_tensor(s):
                    if s.shape != t.shape:
                        raise ValueError(f"GraphedVAEStep: input shape changed after capture ({tuple(t.shape)} vs "
                                         f"{tuple(s.shape)})")
                    s.copy_(t, non_blocking=True)
            if eps is not None:
                self.static_eps.copy_(eps, non_blocking=True)
        if getattr(self.step, "_shadow", None):      # parameters changed from outside since the last step?
            self.step.refresh_shadows()
        self.graph.replay()
        return self.out


class HostBatchFeeder:
    """Feeds batches from a HOST-resident (pinned) dataset, the situation of the reference's DataLoader (12:226-227,
    19:191-192), without stalling the step: the rows of batch s+1 are copied host->device on a side stream into the
    other half of a double buffer while batch s trains.  Large per-sample arrays (the spectrograms) go row by row
    straight out of the pinned dataset (no host-side gather pass); small ones are gathered on the host first."""

    def __init__(self, arrays, batch_size: int, device, small_bytes: int = 1 << 16):
        self.device = device
        self.host = [a if a.is_pinned() else a.pin_memory() for a in (t.contiguous() for t in arrays)]
        self.copy_stream = torch.cuda.Stream(device=device)
        self.slots = []
        for _ in range(2):
            dev = [torch.empty((batch_size,) + tuple(a.shape[1:]), dtype=a.dtype, device=device) for a in self.host]
            stage = [torch.empty((batch_size,) + tuple(a.shape[1:]), dtype=a.dtype).pin_memory()
                     if a[0].numel() * a.element_size() <= small_bytes else None for a in self.host]
            self.slots.append(dict(dev=dev, stage=stage, ready=torch.cuda.Event(), free=torch.cuda.Event(),
                                   copied=torch.cuda.Event(), used=False, n=0))
        self.bytes_per_sample = sum(a[0].numel() * a.element_size() for a in self.host)
        self._next = 0
        self._queue = []

    def prefetch(self, idx) -> None:
        """Enqueue the copy of the batch with sample indices `idx` (1-D CPU int64 tensor or list)."""
        slot = self.slots[self._next]
        self._next ^= 1
        idx_t = idx if isinstance(idx, torch.Tensor) else torch.as_tensor(idx)
        ids = idx_t.tolist()
        n = len(ids)
        if slot["used"]:
            # the pinned staging rows are rewritten by the HOST below: the H2D copies that last read them must have
            # completed (a stream-side wait is not enough when the host runs ahead of the GPU)
            slot["copied"].synchronize()
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(slot["free"])  # the step that last used this slot has been enqueued and finished
            for h, d, st in zip(self.host, slot["dev"], slot["stage"]):
                if st is not None:
                    torch.index_select(h, 0, idx_t, out=st[:n])
                    d[:n].copy_(st[:n], non_blocking=True)
                else:
                    for i, j in enumerate(ids):
                        d[i].copy_(h[j], non_blocking=True)
            slot["copied"].record(self.copy_stream)
            slot["ready"].record(self.copy_stream)
        slot["n"], slot["used"] = n, True
        self._queue.append(slot)

    def get(self):
        """Device tensors of the oldest prefetched batch; the current stream waits for its copy."""
        slot = self._queue.pop(0)
        torch.cuda.current_stream(self.device).wait_event(slot["ready"])
        self._last = slot
        return [d[: slot["n"]] for d in slot["dev"]]

    def release(self) -> None:
        """Call after the step that consumed the batch has been enqueued: its slot may be refilled once that work is done."""
        self._last["free"].record(torch.cuda.current_stream(self.device))


class StepResultReader:
    """Per-step device->host read of a step's small result tensor (loss, recon, kl -- the reference's three `.item()` calls,
    12:268-270) that does not idle the GPU: the copy into pinned host memory is enqueued right behind the step and the host
    consumes it ONE STEP LATER, while the next step already runs.  `push(st)` returns the previous step's values (None on
    the first call); `flush()` returns the last step's values (waits for its copy)."""

    def __init__(self, numel: int, device, dtype=torch.float32, depth: int = 2):
        self.buf = [torch.empty(numel, dtype=dtype).pin_memory() for _ in range(depth)]
        self.ev = [torch.cuda.Event() for _ in range(depth)]
        self.device = device
        self.n = 0

    def _take(self, k: int):
        self.ev[k].synchronize()
        return self.buf[k].clone()

    def push(self, st: torch.Tensor):
        k = self.n % len(self.buf)
        self.buf[k].copy_(st.detach().reshape(-1), non_blocking=True)
        self.ev[k].record(torch.cuda.current_stream(self.device))
        self.n += 1
        return self._take((self.n - 2) % len(self.buf)) if self.n >= 2 else None

    def flush(self):
        return self._take((self.n - 1) % len(self.buf)) if self.n >= 1 else None


@dataclass
class TrainLog:
    loss: List[float] = field(default_factory=list)
    recon: List[float] = field(default_factory=list)
    kl: List[float] = field(default_factory=list)


def _to_dev(a, device):
    t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))
    return t.to(device=device, dtype=torch.float32).contiguous()


def fit(model, kind: str, arrays, epochs: int, batch_size: int, lr: float, beta: float, device=None,
        on_epoch: Optional[Callable] = None, group=None, eps_fn: Optional[Callable] = None) -> TrainLog:
    """The hot loop of 12:246-287 / 19:209-245 / 06:314-344 with a device-resident dataset.
    arrays: (X, lyr, mask) for conv12; (X,) or (X, cond) for mlp19; (X,) for mlp06.
    Epoch averages follow the scripts: unweighted mean of batch means for 06/12 (12:274-276), sample-weighted
    for 19 (19:234-242).  `eps_fn(epoch, step, rows)` (optional) supplies the reparameterisation noise of a GLOBAL
    batch as a CPU tensor (parity tests inject the oracle's stream); default: torch.randn on the device.
    Under a process group every rank draws the same permutation and trains on its slice of each global batch; all
    ranks take part in every step (a rank with an empty slice contributes a zero-weight dummy row) and the logged
    statistics are the global-batch values, all-reduced once per epoch."""
    device = device or torch.device("cuda", torch.cuda.current_device())
    data = [_to_dev(a, device) for a in arrays]
    n = data[0].shape[0]
    model.to(device).train()
    step = make_step(model, kind, lr, beta, group)
    world, rank = (vdist.world_size(group), vdist.rank(group)) if group is not None else (1, 0)
    log = TrainLog()
    for epoch in range(1, epochs + 1):
        acc = torch.zeros(3, dtype=torch.float64, device=device)
        denom = 0.0
        for s_i, idx in enumerate(batches(n, batch_size)):  # every rank draws the same permutation (same seed)
            gb = idx.numel()
            lo, hi = vdist.shard_bounds(gb, rank, world)
            rows = hi - lo
            if rows == 0:
                lo, hi = 0, 1           # dummy row, weight 0: keeps this rank inside the step's collectives
            sel = idx[lo:hi].to(device)
            eps = None
            if eps_fn is not None:
                eps = eps_fn(epoch, s_i, gb)[lo:hi].to(device)
            stats = step(*[t.index_select(0, sel) for t in data], eps=eps, global_batch=gb, local_rows=rows)
            w = float(gb) if kind in ("mlp19", "ae22") else 1.0   # 19:234-237, 22:157-158: sample-weighted
            acc += stats.to(torch.float64) * (w * rows / gb)   # this rank's share of the global-batch mean
            denom += w
        if world > 1:
            vdist.all_reduce_sum(acc, group)
            if hasattr(step, "sync_params"):
                step.sync_params()      # replicas up to date before the epoch callback (checkpoints) reads them
        tot = acc.cpu().numpy() / max(denom, 1.0)  # the one host sync of the epoch
        log.loss.append(float(tot[0])); log.recon.append(float(tot[1])); log.kl.append(float(tot[2]))
        if on_epoch is not None:
            on_epoch(epoch, log, model)
    return log


# ------------------------------------------------------------------- host helpers around the loop (drop-in formats)
def align_lyrics_to_audio(audio_ids: np.ndarray, lyr_ids: np.ndarray, lyr_emb: np.ndarray, verbose: bool = True):
    """12:35-60: (lyr_aligned [N, D] float32, mask [N, 1] float32); rows of tracks without lyrics stay zero.  Same
    semantics as the reference's dict lookup (a track id listed twice keeps its LAST embedding), vectorised."""
    audio_ids = np.asarray(audio_ids).astype(np.int64)
    lyr_ids = np.asarray(lyr_ids).astype(np.int64)
    D = lyr_emb.shape[1]
    lyr_aligned = np.zeros((len(audio_ids), D), dtype=np.float32)
    mask = np.zeros((len(audio_ids), 1), dtype=np.float32)
    if len(lyr_ids):
        # last occurrence of every id: reverse, unique (first of the reversed = last of the original)
        uniq, first_rev = np.unique(lyr_ids[::-1], return_index=True)
        last = len(lyr_ids) - 1 - first_rev
        pos = np.searchsorted(uniq, audio_ids)
        pos_c = np.clip(pos, 0, len(uniq) - 1)
        hit = uniq[pos_c] == audio_ids
        lyr_aligned[hit] = lyr_emb[last[pos_c[hit]]]
        mask[hit, 0] = 1.0
    if verbose:
        print(f"Lyrics aligned: {int(mask.sum())}/{len(audio_ids)} tracks have embeddings.")
    return lyr_aligned, mask


class Script12Writer:
    """Per-epoch outputs of script 12 (12:236-240,278-285): `train_log.csv` with header `epoch,loss,recon,kl` and
    `%.6f` rows, and `ckpt_epoch_XXX.pt` = {"model": state_dict, "cfg": dict, "feat_hw", "lyr_dim"}.  Pass an instance
    as fit(on_epoch=...)."""

    def __init__(self, out_dir, cfg: dict, feat_hw, lyr_dim: int, save_ckpt: bool = True):
        from pathlib import Path
        self.out_dir = Path(out_dir)
        self.out_dir.mkdir(parents=True, exist_ok=True)
        self.cfg, self.feat_hw, self.lyr_dim, self.save_ckpt = dict(cfg), tuple(feat_hw), int(lyr_dim), save_ckpt
        self.log_path = self.out_dir / "train_log.csv"
        with self.log_path.open("w", encoding="utf-8") as f:
            f.write("epoch,loss,recon,kl\n")

    def __call__(self, epoch: int, log: TrainLog, model) -> None:
        with self.log_path.open("a", encoding="utf-8") as f:
            f.write(f"{epoch},{log.loss[-1]:.6f},{log.recon[-1]:.6f},{log.kl[-1]:.6f}\n")
        if self.save_ckpt:
            torch.save({"model": model.state_dict(), "cfg": self.cfg, "feat_hw": self.feat_hw, "lyr_dim": self.lyr_dim},
                       self.out_dir / f"ckpt_epoch_{epoch:03d}.pt")


def save_script19_model(path, model, *, input_dim: int, latent_dim: int, hidden_dim: int, beta: float, use_cvae: bool,
                        cond_dim: int, include_genre_in_input: bool = False, include_lang_in_input: bool = False,
                        cond_on: str = "genre_lang", seed: int = 42) -> None:
    """The model dict of 19:248-264 (keys and value types as the reference writes them)."""
    torch.save({"state_dict": model.state_dict(), "input_dim": int(input_dim), "latent_dim": int(latent_dim),
                "hidden_dim": int(hidden_dim), "beta": float(beta), "use_cvae": bool(use_cvae), "cond_dim": int(cond_dim),
                "include_genre_in_input": bool(include_genre_in_input),
                "include_lang_in_input": bool(include_lang_in_input), "cond_on": cond_on, "seed": int(seed)}, path)


def save_script06_outputs(out_dir, model, cfg: dict, log: TrainLog) -> None:
    """06:347-351: bare state_dict `vae_basic.pt`, `train_config.json`, `history.json` {"epoch","recon","kl","total"}."""
    import json
    from pathlib import Path
    out = Path(out_dir)
    out.mkdir(parents=True, exist_ok=True)
    torch.save(model.state_dict(), out / "vae_basic.pt")
    with open(out / "train_config.json", "w", encoding="utf-8") as f:
        json.dump(cfg, f, indent=2)
    hist = {"epoch": list(range(1, len(log.loss) + 1)), "recon": log.recon, "kl": log.kl, "total": log.loss}
    with open(out / "history.json", "w", encoding="utf-8") as f:
        json.dump(hist, f, indent=2)


@torch.no_grad()
def export_latents(model, kind: str, arrays, batch_size: int = 512, device=None) -> np.ndarray:
    """mu for every sample in the original order, float32 (12:289-309, 19:269-287, 06:354-360)."""
    device = device or torch.device("cuda", torch.cuda.current_device())
    data = [_to_dev(a, device) for a in arrays]
    model.to(device).eval()
    outs = []
    for s in range(0, data[0].shape[0], batch_size):
        part = [t[s: s + batch_size] for t in data]
        if kind == "conv12":
            mu, _ = model.encode(*part)
        elif kind == "mlp19":
            mu, _ = model.encode(part[0], part[1] if len(part) > 1 else None)
        elif kind == "ae22":
            mu, _ = model.encode(part[0])     # z of 22:84-87
        else:
            mu, _ = model.encode(part[0])
        outs.append(mu)
    return torch.cat(outs, 0).cpu().numpy().astype(np.float32)
