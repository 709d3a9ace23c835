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

_OVERLAP_MIN_NUMEL = 1 << 22   # parameters at least this large are updated early, on a side stream (Adam.overlap)


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


class VAEStep:
    """One optimisation step for any of the three models.  kind: "conv12" | "mlp19" | "mlp06".
    Under a process group each rank passes ITS slice of the global batch; gradients are combined with weights
    B_local/B_global so the result equals the reference's single-process batch mean."""

    def __init__(self, model, kind: str, lr: float, beta: float, group=None, capturable: bool = False):
        self.model, self.kind, self.beta, self.group = model, kind, float(beta), group
        self.mode = VC_LOSS_SUM_MEAN if kind == "mlp19" else VC_LOSS_MEAN_MEAN
        self.multi = group is not None and vdist.world_size(group) > 1
        self.opt = Adam(model.parameters(), lr=lr, capturable=capturable)
        self.sync = vdist.OverlappedGradSync(model.parameters(), group) if self.multi else None
        # opt-in (VAECLUST_ADAM_OVERLAP=1), one GPU: the Adam sweep over a large weight (28 B/parameter) starts on a
        # side stream as soon as backward has produced its gradient and runs under the remaining backward kernels.
        # Bit-identical, but measured SLOWER on script 12 (4.86 -> 4.95-5.00 ms/step, B200): the co-resident Adam CTAs
        # take issue slots and L2 bandwidth from the instruction-bound implicit-GEMM producers, which costs more than
        # the 0.6 ms sweep that gets hidden.  Kept off by default; DESIGN.md 5 records the experiment.
        big = [p for p in model.parameters() if p.numel() >= _OVERLAP_MIN_NUMEL]
        self.overlap = bool(big) and not self.multi and os.environ.get("VAECLUST_ADAM_OVERLAP", "0") == "1"
        if self.overlap:
            self.opt.overlap(big)

    def __call__(self, x, *extra, eps: Optional[torch.Tensor] = None, global_batch: Optional[int] = None):
        m = self.model
        if self.sync is not None:
            self.sync.begin(x.shape[0], global_batch or x.shape[0] * vdist.world_size(self.group))
        else:
            self.opt.zero_grad()
        if self.overlap:
            self.opt.begin_step()
        if self.kind == "conv12":
            lyr, mask = extra
            mu, lv = m.encode(x, lyr, mask)
        elif self.kind == "mlp19":
            c = extra[0] if extra else None
            mu, lv = m.encode(x, c)
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
        return stats  # device tensor [loss, recon, kl] of the local batch


class GraphedVAEStep:
    """The same optimisation step recorded once into a CUDA graph and replayed: the ~125 kernel launches
    of a step are submitted by one cudaGraphLaunch, which removes the host launch gaps between the short kernels of
    the latent head and of the backward pass.  The first `warmup` calls run eagerly (they allocate optimizer state and
    warm the allocator), the next call captures; inputs are copied into static buffers, `eps` (optional) likewise, else
    it is drawn inside the graph from torch's graph-safe generator.  Returns the static [loss, recon, kl] tensor --
    read it before the next call.  Shapes must not change after capture (pass full batches).  Under a process group
    the overlapped NCCL gradient exchange is captured with the step (set TORCH_NCCL_ASYNC_ERROR_HANDLING=0 before
    init_process_group, as for any NCCL-in-graph capture); `global_batch` is fixed at capture."""

    def __init__(self, model, kind: str, lr: float, beta: float, warmup: int = 2, group=None):
        self.step = VAEStep(model, kind, lr, beta, group=group, capturable=True)
        self.warmup, self.calls = int(warmup), 0
        self.graph = None
        self.static_in = None
        self.static_eps = None
        self.out = None

    @property
    def opt(self):
        return self.step.opt

    def __call__(self, x, *extra, eps: Optional[torch.Tensor] = None, global_batch: Optional[int] = None):
        self.calls += 1
        if self.graph is None and self.calls <= self.warmup:
            return self.step(x, *extra, eps=eps, global_batch=global_batch)
        if self.graph is None:
            self.static_in = [t.clone() if torch.is_tensor(t) else t for t in (x,) + tuple(extra)]
            self.static_eps = eps.clone() if eps is not None else None
            self.step.opt.zero_grad()       # gradients must be (re)allocated from the graph's private pool
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
            self.slots.append(dict(dev=dev, stage=stage, ready=torch.cuda.Event(), free=torch.cuda.Event(), n=0))
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
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(slot["free"])  # the step that last used this slot has been enqueued and finished
            for h, d, st in zip(self.host, slot["dev"], slot["stage"]):
                if st is not None:
                    torch.index_select(h, 0, idx_t, out=st[:n])
                    d[:n].copy_(st[:n], non_blocking=True)
                else:
                    for i, j in enumerate(ids):
                        d[i].copy_(h[j], non_blocking=True)
            slot["ready"].record(self.copy_stream)
        slot["n"] = n
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


@dataclass
class TrainLog:
    loss: List[float] = field(default_factory=list)
    recon: List[float] = field(default_factory=list)
    kl: List[float] = field(default_factory=list)


def _to_dev(a, device):
    t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))
    return t.to(device=device, dtype=torch.float32).contiguous()


def fit(model, kind: str, arrays, epochs: int, batch_size: int, lr: float, beta: float, device=None,
        on_epoch: Optional[Callable] = None, group=None) -> TrainLog:
    """The hot loop of 12:246-287 / 19:209-245 / 06:314-344 with a device-resident dataset.
    arrays: (X, lyr, mask) for conv12; (X,) or (X, cond) for mlp19; (X,) for mlp06.
    Epoch averages follow the scripts: unweighted mean of batch means for 06/12 (12:274-276), sample-weighted
    for 19 (19:234-242)."""
    device = device or torch.device("cuda", torch.cuda.current_device())
    data = [_to_dev(a, device) for a in arrays]
    n = data[0].shape[0]
    model.to(device).train()
    step = VAEStep(model, kind, lr, beta, group)
    world, rank = (vdist.world_size(group), vdist.rank(group)) if group is not None else (1, 0)
    log = TrainLog()
    for epoch in range(1, epochs + 1):
        acc = torch.zeros(3, dtype=torch.float64, device=device)
        denom = 0.0
        for idx in batches(n, batch_size):  # every rank draws the same permutation (same seed)
            gb = idx.numel()
            lo, hi = vdist.shard_bounds(gb, rank, world)
            if hi <= lo:
                continue
            sel = idx[lo:hi].to(device)
            stats = step(*[t.index_select(0, sel) for t in data], global_batch=gb)
            w = float(hi - lo) if kind == "mlp19" else 1.0
            acc += stats.to(torch.float64) * w
            denom += w
        tot = acc.cpu().numpy() / max(denom, 1.0)  # the one host sync of the epoch
        log.loss.append(float(tot[0])); log.recon.append(float(tot[1])); log.kl.append(float(tot[2]))
        if on_epoch is not None:
            on_epoch(epoch, log, model)
    return log


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
        else:
            mu, _ = model.encode(part[0])
        outs.append(mu)
    return torch.cat(outs, 0).cpu().numpy().astype(np.float32)
