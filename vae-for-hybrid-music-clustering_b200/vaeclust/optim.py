"""torch.optim.Adam replacement (12:233, 19:202, 06:310; defaults betas=(0.9,0.999), eps=1e-8, no weight decay).

One vc_adam_multi launch per step over a device-resident chunk table instead of torch's ~4 foreach kernels per
state tensor.  Parameters whose .grad is None are skipped and their step counter does not advance (this is what
keeps `audio_enc.logvar` of script 12 frozen).

Sharded mode (`shard(rank, world)`, used by the data-parallel script-12 step): each rank keeps the
Adam moments of, and updates, only its 1/world slice of every LARGE parameter; the gradient slice comes from a
reduce-scatter and the updated slices are all-gathered by the caller (vaeclust.dist.ShardedGradSync).  Small
parameters are updated in full on every rank from all-reduced gradients, which yields bit-identical replicas.
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional

import numpy as np
import torch

from ._lib import VaeclustError, lib, ptr, stream_ptr

_CHUNK = 1 << 16


class Adam:
    def __init__(self, params: Iterable[torch.nn.Parameter], lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, capturable: bool = False):
        if weight_decay != 0.0:
            raise VaeclustError("vaeclust.optim.Adam: weight_decay is not used by the reference and not supported")
        self.params: List[torch.nn.Parameter] = [p for p in params]
        if not self.params:
            raise ValueError("optimizer got an empty parameter list")
        for p in self.params:
            if not p.is_cuda or p.dtype != torch.float32:
                raise VaeclustError("vaeclust.optim.Adam: parameters must be float32 CUDA tensors (no CPU path)")
        self.lr, self.betas, self.eps = float(lr), (float(betas[0]), float(betas[1])), float(eps)
        self.state: Dict[int, dict] = {}
        self._tables: Dict[object, Optional[dict]] = {}   # chunk tables on the device, one per launch slot
        self._retired: list = []               # superseded tables: a captured graph may still read them
        self._spares: Dict[object, Optional[dict]] = {}   # pre-allocated (pinned, device) pair per slot
        # capturable (torch.optim.Adam(capturable=True)): step counter and bias corrections live on the device, so a
        # step recorded in a CUDA graph advances correctly on every replay (vc_adam_multi_graph)
        self.capturable = bool(capturable)
        self._dev_step = None
        self._ticked = False                   # capturable: the device step counter was advanced by begin_step()
        # id(param) -> (lo, hi, grad_slice_fn): this rank owns elements [lo, hi) of the flattened parameter
        self._shards: Dict[int, tuple] = {}
        self.skip: set = set()                 # ids of parameters some fused kernel updates itself this step
        self.param_groups = [{"params": self.params, "lr": self.lr, "betas": self.betas, "eps": self.eps}]

    def zero_grad(self, set_to_none: bool = True) -> None:
        for p in self.params:
            if p.grad is None:
                continue
            if set_to_none:
                p.grad = None
            else:
                p.grad.detach_()
                p.grad.zero_()

    def shard(self, p: torch.nn.Parameter, lo: int, hi: int, grad_slice) -> None:
        """From now on update only elements [lo, hi) of `p` (flattened), reading their gradient from the 1-D fp32
        tensor `grad_slice` (hi - lo elements; the output of a reduce-scatter).  Must be called before the first step
        that touches `p`: the moments are allocated for the slice only."""
        if id(p) in self.state:
            raise VaeclustError("Adam.shard: parameter already has full-size optimizer state")
        if grad_slice.numel() != hi - lo or grad_slice.dtype != torch.float32 or not grad_slice.is_contiguous():
            raise VaeclustError("Adam.shard: grad_slice must be a contiguous float32 tensor of hi - lo elements")
        self._shards[id(p)] = (int(lo), int(hi), grad_slice)

    def _state(self, p) -> dict:
        st = self.state.get(id(p))
        if st is None:
            sh = self._shards.get(id(p))
            n = p.numel() if sh is None else sh[1] - sh[0]
            shape = p.shape if sh is None else (n,)
            st = {"step": 0, "exp_avg": torch.zeros(shape, dtype=torch.float32, device=p.device),
                  "exp_avg_sq": torch.zeros(shape, dtype=torch.float32, device=p.device)}
            self.state[id(p)] = st
        return st

    # ----------------------------------------------------------------------------------------------- the update
    def _segments(self, plist):
        """(param base ptr, grad base ptr, m ptr, v ptr, n) of what this rank updates for each parameter."""
        segs = []
        for p in plist:
            if not p.data.is_contiguous():
                raise VaeclustError("vaeclust.optim.Adam: non-contiguous parameter")
            st = self._state(p)
            st["step"] += 1
            sh = self._shards.get(id(p))
            if sh is None:
                g = p.grad
                if not g.is_contiguous():
                    g = g.contiguous()
                    p.grad = g
                segs.append((p.data_ptr(), g.data_ptr(), st, p.numel()))
            else:
                lo, hi, gs = sh
                segs.append((p.data_ptr() + 4 * lo, gs.data_ptr(), st, hi - lo))
        return segs

    @torch.no_grad()
    def begin_step(self) -> None:
        """Call once at the start of a step whose parameters are updated by SEVERAL launches (update_now for sharded
        tensors as soon as their gradient slice has been reduced, step() for the rest): in capturable mode the device
        step counter advances here, once, and every launch of the step reads the bias corrections of that step."""
        self._done = set()
        if self.capturable:
            lib().call("vc_adam_tick", ptr(self.device_step()), self.betas[0], self.betas[1], stream_ptr())
            self._ticked = True

    @torch.no_grad()
    def update_now(self, p) -> None:
        """Adam update of ONE parameter (its shard, if sharded) on the current stream, now; step() then skips it."""
        self._update([p], slot=("now", id(p)))
        self._done.add(id(p))

    def _update(self, plist, slot="rest") -> None:
        """One launch over the chunk table of `plist` on the current stream (one launch per distinct step count)."""
        by_step: Dict[int, list] = {}
        for p, seg in zip(plist, self._segments(plist)):
            by_step.setdefault(seg[2]["step"], []).append(seg)
        if self.capturable and len(by_step) > 1:
            raise VaeclustError("vaeclust.optim.Adam(capturable=True): every stepped parameter must share one step count")
        for t, segs in by_step.items():
            rows = []
            # one block per chunk: few elements (everything but the wide layers) get smaller chunks so that the launch
            # still fills the GPU (34 blocks of 64 Ki elements took 0.076 ms for 2 M parameters)
            total = sum(n for _pp, _gp, _st, n in segs)
            chunk = _CHUNK if total >= (1 << 25) else (1 << 12)
            for pp, gp, st, n in segs:
                if n == 0:
                    continue
                base = np.arange(0, n, chunk, dtype=np.int64)
                tab = np.empty((base.size, 5), dtype=np.int64)
                tab[:, 0] = pp + 4 * base
                tab[:, 1] = gp + 4 * base
                tab[:, 2] = st["exp_avg"].data_ptr() + 4 * base
                tab[:, 3] = st["exp_avg_sq"].data_ptr() + 4 * base
                tab[:, 4] = np.minimum(chunk, n - base)
                rows.append(tab)
            if not rows:
                continue
            table = np.concatenate(rows, axis=0)
            key = table.tobytes()
            dev = plist[0].device
            c = self._tables.get(slot)
            if c is None or c["key"] != key or c["dev"].device != dev:
                if not self.capturable:
                    c = {"key": key, "dev": torch.from_numpy(table).to(dev, non_blocking=False)}
                else:
                    # never rewrite a table in place: a captured graph's memcpy node re-reads the pinned buffer and
                    # its kernel node the device buffer on every replay, so a changed table gets FRESH buffers and
                    # the old pair is retired (kept alive).  Nothing may be allocated while a stream is capturing,
                    # so one spare pair of the right shape is always kept ready.
                    if c is not None:
                        self._retired.append(c)
                        # tables filled during a capture belong to a graph and live as long as the optimizer; of the
                        # others only the most recent few can still be in flight
                        keep = [r for r in self._retired if r.get("captured")]
                        self._retired = keep + [r for r in self._retired if not r.get("captured")][-4:]
                    sp = self._spares.get(slot)
                    if sp is None or tuple(sp["pinned"].shape) != table.shape or sp["dev"].device != dev:
                        sp = self._new_pair(table.shape, dev)
                    self._spares[slot] = None
                    sp["pinned"].numpy()[...] = table
                    sp["dev"].copy_(sp["pinned"], non_blocking=True)   # async + pinned: legal inside a capture
                    sp["key"] = key
                    sp["captured"] = torch.cuda.is_current_stream_capturing()
                    c = sp
                self._tables[slot] = c
            if self.capturable and self._spares.get(slot) is None and not torch.cuda.is_current_stream_capturing():
                self._spares[slot] = self._new_pair(table.shape, dev)
            b1, b2 = self.betas
            if self.capturable:
                if self._dev_step is None:
                    self._dev_step = torch.zeros(4, dtype=torch.int32, device=dev)
                    self._dev_step[0] = t - 1
                lib().call("vc_adam_multi_at" if self._ticked else "vc_adam_multi_graph", ptr(c["dev"]),
                           int(table.shape[0]), self.lr, b1, b2, self.eps, ptr(self._dev_step), stream_ptr())
            else:
                lib().call("vc_adam_multi", ptr(c["dev"]), int(table.shape[0]), self.lr, b1, b2, self.eps,
                           1.0 - b1 ** t, 1.0 - b2 ** t, stream_ptr())

    @staticmethod
    def _new_pair(shape, dev) -> dict:
        return {"key": None, "pinned": torch.empty(shape, dtype=torch.int64).pin_memory(),
                "dev": torch.empty(shape, dtype=torch.int64, device=dev)}

    @torch.no_grad()
    def step(self) -> None:
        self.lr = float(self.param_groups[0]["lr"])
        done = getattr(self, "_done", None) or ()
        todo = [p for p in self.params if id(p) not in self.skip and id(p) not in done
                and (p.grad is not None or id(p) in self._shards)]
        if todo:
            self._update(todo)
        self._done = set()
        self._ticked = False

    def device_step(self) -> torch.Tensor:
        """The device-resident VcAdamDeviceStep of capturable mode (allocated on first use)."""
        if self._dev_step is None:
            t0 = max((st["step"] for st in self.state.values()), default=0)
            self._dev_step = torch.zeros(4, dtype=torch.int32, device=self.params[0].device)
            self._dev_step[0] = t0
        return self._dev_step

    def state_dict(self) -> dict:
        if self.capturable and self._dev_step is not None:   # replays advance the device counter only
            t = int(self._dev_step[0])
            for st in self.state.values():
                st["step"] = t
        return {"state": {i: {k: (v.clone() if torch.is_tensor(v) else v) for k, v in self._state(p).items()}
                          for i, p in enumerate(self.params)},
                "param_groups": [{"lr": self.lr, "betas": self.betas, "eps": self.eps}]}
