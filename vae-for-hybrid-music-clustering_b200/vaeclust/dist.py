"""Multi-GPU plumbing: one process per GPU, torch.distributed (NCCL over NVLink 5 / NVSwitch; gloo in CPU tests).

The reference is single-process; everything here is new.  Only the exchanges the hot path really
has are implemented:
  * VAE data-parallel step: one flat-bucket gradient all-reduce per step (weighted by local batch size so an
    uneven last batch still equals the reference's global batch mean);
  * KMeans: all-reduce of [k*d sums | k counts] (+ label-change counter) per Lloyd iteration, float64 so that
    every rank derives bit-identical centroids; all-gather of farthest-point candidates for empty clusters;
  * silhouette: row blocks are independent; one all-reduce of the score sum (or of the disjoint per-row values);
  * Davies-Bouldin / StandardScaler: all-reduce of per-cluster or per-column moments.
Functions take whatever device the tensors live on, so the same code runs under gloo on CPU tensors in tests.
"""
from __future__ import annotations

import os
from typing import List, Optional, Tuple

import torch
import torch.distributed as td


def is_initialized() -> bool:
    return td.is_available() and td.is_initialized()


def world_size(group=None) -> int:
    return td.get_world_size(group) if is_initialized() else 1


def rank(group=None) -> int:
    return td.get_rank(group) if is_initialized() else 0


def init_from_env(backend: Optional[str] = None):
    """Initialise torch.distributed from torchrun's environment (RANK/LOCAL_RANK/WORLD_SIZE/MASTER_*).
    Returns (rank, world, local_rank).  Single-process runs return (0, 1, 0) without initialising anything."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world <= 1:
        return 0, 1, 0
    r = int(os.environ["RANK"])
    local = int(os.environ.get("LOCAL_RANK", r))
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        torch.cuda.set_device(local)
    if not td.is_initialized():
        kw = {"device_id": torch.device("cuda", local)} if backend == "nccl" else {}
        td.init_process_group(backend=backend, rank=r, world_size=world, **kw)
    return r, world, local


def shard_bounds(n: int, r: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced row range [lo, hi) of rank r (first n % world ranks get one extra row)."""
    base, rem = divmod(int(n), int(world))
    lo = r * base + min(r, rem)
    return lo, lo + base + (1 if r < rem else 0)


def all_reduce_sum(t: torch.Tensor, group=None) -> torch.Tensor:
    if world_size(group) > 1:
        td.all_reduce(t, op=td.ReduceOp.SUM, group=group)
    return t


def all_reduce_max(t: torch.Tensor, group=None) -> torch.Tensor:
    if world_size(group) > 1:
        td.all_reduce(t, op=td.ReduceOp.MAX, group=group)
    return t


def merge_moments(mean: torch.Tensor, var: torch.Tensor, n: int, group=None):
    """Combine per-rank (mean, population variance, count) into global ones (Chan et al. pairwise update written
    as one all-reduce of [n, n*mean, n*(var + mean^2)] about a shared pivot = rank 0's mean)."""
    if world_size(group) <= 1:
        return mean, var, n
    pivot = mean.clone()
    td.broadcast(pivot, src=td.get_global_rank(group, 0) if group is not None else 0, group=group)
    dm = mean - pivot
    nn = torch.tensor([float(n)], dtype=torch.float64, device=mean.device)
    buf = torch.cat([nn, dm * n, (var + dm * dm) * n])
    td.all_reduce(buf, op=td.ReduceOp.SUM, group=group)
    d = mean.numel()
    ntot = buf[0]
    m1 = buf[1: 1 + d] / ntot
    m2 = buf[1 + d:] / ntot
    return pivot + m1, (m2 - m1 * m1).clamp_min(0.0), int(ntot.item())


def gather_top(dist: torch.Tensor, rows: torch.Tensor, labs: torch.Tensor, n: int, group=None):
    """All-gather each rank's `n` farthest-point candidates and keep the global top n (empty-cluster relocation)."""
    w = world_size(group)
    if w <= 1:
        return dist, rows, labs
    dev = dist.device
    pad = n - dist.numel()
    if pad > 0:  # a rank with fewer than n points pads with -1 distances
        dist = torch.cat([dist, torch.full((pad,), -1.0, dtype=dist.dtype, device=dev)])
        rows = torch.cat([rows, torch.zeros((pad, rows.shape[1]), dtype=rows.dtype, device=dev)])
        labs = torch.cat([labs, torch.zeros(pad, dtype=labs.dtype, device=dev)])
    packed = torch.cat([dist.to(torch.float64).unsqueeze(1), labs.to(torch.float64).unsqueeze(1),
                        rows.to(torch.float64)], dim=1).contiguous()
    out: List[torch.Tensor] = [torch.empty_like(packed) for _ in range(w)]
    td.all_gather(out, packed, group=group)
    allp = torch.cat(out, dim=0)
    top = torch.topk(allp[:, 0], n).indices
    sel = allp[top]
    return sel[:, 0].to(torch.float32), sel[:, 2:].to(torch.float32), sel[:, 1].to(torch.int64)


class GradBucket:
    """Flat fp32 gradient bucket for the data-parallel VAE step: parameters' .grad become views into one buffer so
    the exchange is a single NCCL all-reduce (1.5 MB for script 19, 540 MB for script 12)."""

    def __init__(self, params, group=None):
        self.params = [p for p in params if p.requires_grad]
        self.group = group
        n = sum(p.numel() for p in self.params)
        dev = self.params[0].device
        self.flat = torch.zeros(n, dtype=torch.float32, device=dev)
        o = 0
        for p in self.params:
            p.grad = self.flat[o: o + p.numel()].view_as(p)
            o += p.numel()

    def zero(self) -> None:
        self.flat.zero_()

    def all_reduce_mean(self, local_batch: int, global_batch: int) -> None:
        """grad <- sum_r (B_r / B) grad_r : equals the gradient of the reference's global batch mean."""
        w = world_size(self.group)
        if w <= 1:
            return
        self.flat.mul_(float(local_batch) / float(global_batch))
        td.all_reduce(self.flat, op=td.ReduceOp.SUM, group=self.group)


class OverlappedGradSync:
    """Gradient exchange of the data-parallel VAE step, overlapped with backward.

    Script 12 has 135 M parameters, 134 M of them in two Linear weights (12:98, 12:131).  Each big gradient is
    all-reduced IN PLACE, asynchronously, from a post-accumulate hook the moment autograd has produced it: the decoder
    FC gradient travels while the encoder is still back-propagating, the encoder FC gradient while the convolutions
    are.  The few small gradients are packed into one flat buffer and reduced once at the end (latency-bound).
    Contributions are weighted B_local / B_global so an uneven last batch still gives the gradient of the reference's
    global batch mean (12:262-264, 19:228).  NCCL applies the weight inside the reduction (premul-sum); other backends
    scale first."""

    def __init__(self, params, group=None, big_numel: int = 1 << 20):
        self.params = [p for p in params if p.requires_grad]
        self.group = group
        self.big = [p for p in self.params if p.numel() >= big_numel]
        self.small = [p for p in self.params if p.numel() < big_numel]
        self.factor = 1.0
        self.works = []
        self._hooks = [p.register_post_accumulate_grad_hook(self._on_big_grad) for p in self.big]
        self._flat = None

    def _reduce(self, t: torch.Tensor, async_op: bool):
        if td.get_backend(self.group) == "nccl" and hasattr(td, "_make_nccl_premul_sum"):
            return td.all_reduce(t, op=td._make_nccl_premul_sum(self.factor), group=self.group, async_op=async_op)
        t.mul_(self.factor)
        return td.all_reduce(t, op=td.ReduceOp.SUM, group=self.group, async_op=async_op)

    def _on_big_grad(self, p) -> None:
        if world_size(self.group) > 1 and p.grad is not None:
            self.works.append(self._reduce(p.grad, True))

    def begin(self, local_batch: int, global_batch: int) -> None:
        """Call before backward: clears the gradients and fixes this step's weight."""
        self.factor = float(local_batch) / float(global_batch)
        self.works = []
        for p in self.params:
            p.grad = None

    def finish(self) -> None:
        """Call after backward: reduces the small gradients and waits for everything in flight."""
        if world_size(self.group) <= 1:
            return
        live = [p for p in self.small if p.grad is not None]
        if live:
            n = sum(p.numel() for p in live)
            if self._flat is None or self._flat.numel() != n or self._flat.device != live[0].device:
                self._flat = torch.empty(n, dtype=torch.float32, device=live[0].device)
            flat = torch.cat([p.grad.reshape(-1) for p in live], out=self._flat)   # packed into the cached buffer
            self._reduce(flat, False)
            o = 0
            for p in live:
                n = p.numel()
                p.grad = flat[o: o + n].view_as(p)
                o += n
        for w in self.works:
            w.wait()
        self.works = []


class ShardedGradSync(OverlappedGradSync):
    """Gradient exchange + optimizer for the data-parallel script-12 step with the optimizer state SHARDED over the ranks
    (SURVEY.md 8(e) C1): for each large parameter (the two 67 M-element FC weights, 99.5 % of the model)
        reduce-scatter(gradient)  ->  Adam on this rank's 1/world slice  ->  all-gather(parameter)
    The reduce-scatter and the slice update are issued from the parameter's post-accumulate hook and run under the rest
    of backward (NCCL's stream, then a side stream).  The all-gather of the updated slices is DEFERRED to the start of
    the next step, where it runs under the forward convolutions; the layer that reads the parameter waits for it
    (vaeclust.ops.linear looks for `weight._vc_gather`).  So per step the wire carries the same bytes as the all-reduce
    it replaces, but half of them move while the forward computes instead of after backward, and the 28 B/parameter
    Adam sweep and the moments shrink by the world size.  Small parameters keep the flat all-reduce + replicated update
    (bit-identical replicas).
    Protocol per step: begin() -> forward/backward -> finish() -> opt.step() -> join();  flush() brings every replica
    up to date (call it before reading parameters: checkpoints, evaluation, the end of training)."""

    def __init__(self, params, group, opt, big_numel: int = 1 << 20, update=None):
        super().__init__(params, group, big_numel=big_numel)
        self.opt = opt
        self.update = update or opt.update_now        # (tests on CPU tensors pass their own slice update)
        w, r = world_size(group), rank(group)
        self.sharded = {}
        keep_big = []
        for p in self.big:
            if w > 1 and p.numel() % w == 0:
                n = p.numel() // w
                gs = torch.empty(n, dtype=torch.float32, device=p.device)
                self.sharded[id(p)] = (r * n, (r + 1) * n, gs)
                if hasattr(opt, "shard"):
                    opt.shard(p, r * n, (r + 1) * n, gs)
            else:
                keep_big.append(p)          # not divisible: stays on the in-place all-reduce path
        self._plain_big = {id(p) for p in keep_big}
        self.side = torch.cuda.Stream(device=self.params[0].device) if self.params[0].is_cuda else None
        self.gathers = []
        self.dirty = []      # sharded parameters whose other slices are stale (updated by their owners, not gathered yet)
        self._premul = None if hasattr(td, "_make_nccl_premul_sum") else False
        # bf16 mode (enable_bf16_gather): forward and data gradient of the wide Linear layers read a bf16 copy of the weight
        # anyway, so the owners' updated slices travel as bf16 (half the all-gather bytes, and it is the all-gather the
        # forward waits for); the fp32 master keeps only this rank's slice current until flush() gathers it in fp32.
        self.shadow = {}     # id(p) -> bf16 tensor of p's shape
        self._stale_fp32 = {}

    def enable_bf16_gather(self, params) -> int:
        """Give every sharded CUDA parameter in `params` a bf16 shadow (`p._vc_shadow`, read by vaeclust.ops.linear in
        bf16 mode) and gather updated slices into it instead of into the fp32 parameter.  Returns how many took it."""
        n = 0
        if self.side is None or world_size(self.group) <= 1 or td.get_backend(self.group) != "nccl":
            return 0
        for p in params:
            if id(p) in self.sharded and id(p) not in self.shadow and p.is_cuda and p.dim() == 2:
                sh = torch.empty(p.shape, dtype=torch.bfloat16, device=p.device)
                sh.copy_(p.detach())
                self.shadow[id(p)] = sh
                p._vc_shadow = sh
                n += 1
        return n

    def _on_big_grad(self, p) -> None:
        if world_size(self.group) <= 1 or p.grad is None:
            return
        if id(p) in self._plain_big:
            return super()._on_big_grad(p)
        lo, hi, gs = self.sharded[id(p)]
        flat_g = p.grad.reshape(-1)
        if td.get_backend(self.group) == "nccl":
            cur = torch.cuda.current_stream(p.device)
            self.side.wait_stream(cur)
            with torch.cuda.stream(self.side):
                # weight B_local / B_global inside the reduction when NCCL's pre-multiplied sum is available for
                # reduce-scatter (saves a pass over the gradient), else as a separate scaling pass
                if self._premul is None:
                    try:
                        td.reduce_scatter_tensor(gs, flat_g, op=td._make_nccl_premul_sum(self.factor), group=self.group)
                        self._premul = True
                    except Exception:   # noqa: BLE001  (older NCCL / torch builds)
                        self._premul = False
                        flat_g.mul_(self.factor)
                        td.reduce_scatter_tensor(gs, flat_g, op=td.ReduceOp.SUM, group=self.group)
                elif self._premul:
                    td.reduce_scatter_tensor(gs, flat_g, op=td._make_nccl_premul_sum(self.factor), group=self.group)
                else:
                    flat_g.mul_(self.factor)
                    td.reduce_scatter_tensor(gs, flat_g, op=td.ReduceOp.SUM, group=self.group)   # side waits for the result
                with torch.no_grad():
                    self.update(p)
                    sh = self.shadow.get(id(p))
                    if sh is not None:      # this rank's slice of the bf16 copy follows its fp32 master
                        sh.view(-1)[lo:hi].copy_(p.data.view(-1)[lo:hi])
            self.dirty.append(p)
        else:   # gloo (CPU tests): no reduce-scatter there; same data movement through an all-reduce
            flat_g.mul_(self.factor)
            td.all_reduce(flat_g, op=td.ReduceOp.SUM, group=self.group)
            gs.copy_(flat_g[lo:hi])
            with torch.no_grad():
                self.update(p)
                flat_p = p.data.view(-1)
                parts = [torch.empty(hi - lo, dtype=flat_p.dtype) for _ in range(world_size(self.group))]
                td.all_gather(parts, flat_p[lo:hi].clone(), group=self.group)
                flat_p.copy_(torch.cat(parts))

    def _start_gathers(self) -> None:
        """All-gather the slices updated last step; the consumer of each parameter waits for `p._vc_gather`."""
        if not self.dirty:
            return
        cur = torch.cuda.current_stream(self.params[0].device)
        self.side.wait_stream(cur)
        with torch.cuda.stream(self.side), torch.no_grad():
            # backward produced the gradients in reverse layer order, the forward needs the weights in layer order: gather
            # in the reverse of the order the hooks fired, so the first layer's weight arrives first
            for p in reversed(self.dirty):
                lo, hi, _ = self.sharded[id(p)]
                sh = self.shadow.get(id(p))
                flat_p = p.data.view(-1) if sh is None else sh.view(-1)
                p._vc_gather = td.all_gather_into_tensor(flat_p, flat_p[lo:hi], group=self.group, async_op=True)
                if sh is not None:
                    self._stale_fp32[id(p)] = p
        self.dirty = []

    def begin(self, local_batch: int, global_batch: int) -> None:
        super().begin(local_batch, global_batch)
        if hasattr(self.opt, "begin_step"):
            self.opt.begin_step()
        if self.side is not None:
            self._start_gathers()

    def join(self) -> None:
        """Call after opt.step(): the side stream (slice updates) joins the step's stream."""
        if self.side is not None:
            torch.cuda.current_stream(self.params[0].device).wait_stream(self.side)
        for p in self.big:          # a gather nobody waited for (layer not evaluated this step) must not dangle
            w = getattr(p, "_vc_gather", None)
            if w is not None:
                w.wait()
                p._vc_gather = None

    def flush(self, force: bool = False) -> None:
        """Bring every replica up to date now (before parameters are read outside the step).  force: gather every
        sharded parameter whatever the bookkeeping says (graph replays update slices without running the hooks)."""
        if force:
            self.dirty = [p for p in self.big if id(p) in self.sharded]
            self._stale_fp32.update({id(p): p for p in self.dirty if id(p) in self.shadow})
        if self.side is None:
            return
        if self.dirty:
            self._start_gathers()
            self.join()
        if self._stale_fp32:        # the fp32 masters: every rank holds only its own slice up to date
            with torch.no_grad():
                for p in self._stale_fp32.values():
                    lo, hi, _ = self.sharded[id(p)]
                    flat_p = p.data.view(-1)
                    td.all_gather_into_tensor(flat_p, flat_p[lo:hi], group=self.group)
            self._stale_fp32 = {}
