"""torch.optim.Adam replacement (12:233, 19:202, 06:310; defaults betas=(0.9,0.999), eps=1e-8, no weight decay).

One vc_adam_multi launch per step over a device-resident chunk table instead of torch's ~4 foreach kernels per
state tensor.  Parameters whose .grad is None are skipped and their step counter does not advance (this is what
keeps `audio_enc.logvar` of script 12 frozen).
"""
from __future__ import annotations

import math
from typing import Dict, Iterable, List

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
        self._tables: Dict[tuple, dict] = {}   # chunk tables on the device, one per launch group
        # capturable (torch.optim.Adam(capturable=True)): step counter and bias corrections live on the device, so a
        # step recorded in a CUDA graph advances correctly on every replay (vc_adam_multi_graph)
        self.capturable = bool(capturable)
        self._dev_step = None
        self._early: Dict[int, int] = {}       # id(param) -> slot of the parameters updated from a grad hook
        self._early_stream = None
        self._early_done = None                # ids updated early in the current step (None: not in overlap mode)
        self._early_pending = False
        self._ticked = False
        self._hooks: list = []
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

    def _state(self, p) -> dict:
        st = self.state.get(id(p))
        if st is None:
            st = {"step": 0, "exp_avg": torch.zeros_like(p.data), "exp_avg_sq": torch.zeros_like(p.data)}
            self.state[id(p)] = st
        return st

    # ------------------------------------------------------------------ overlap of large updates with backward
    def overlap(self, params: Iterable[torch.nn.Parameter], stream: "torch.cuda.Stream | None" = None) -> None:
        """Update `params` EARLY: a post-accumulate-grad hook launches the Adam update of each of them on `stream`
        (a side stream) the moment backward has written its gradient, so the 28 B/parameter sweep over a large
        tensor runs under the remaining (instruction-bound, not HBM-bound) backward kernels instead of after them.
        `step()` then updates only the other parameters and joins the side stream.  The caller must call
        `begin_step()` before backward of every step.  Arithmetic is unchanged (same kernel, same step count)."""
        params = [p for p in params]
        known = {id(p) for p in self.params}
        for p in params:
            if id(p) not in known:
                raise ValueError("Adam.overlap: parameter is not managed by this optimizer")
        for h in self._hooks:
            h.remove()
        self._hooks = []
        self._early = {id(p): i for i, p in enumerate(params)}
        self._early_stream = stream if stream is not None else torch.cuda.Stream(device=params[0].device)
        for p in params:
            self._hooks.append(p.register_post_accumulate_grad_hook(self._early_hook))

    @torch.no_grad()
    def begin_step(self) -> None:
        """Start of a step in overlap mode: forget which parameters were updated early and, in capturable mode,
        advance the device step counter once for all launches of this step."""
        self._early_done = set()
        self._ticked = False
        if self._early and self.capturable:
            if self._dev_step is None:
                dev = self.params[0].device
                t0 = max((st["step"] for st in self.state.values()), default=0)
                self._dev_step = torch.zeros(4, dtype=torch.int32, device=dev)
                self._dev_step[0] = t0
            lib().call("vc_adam_tick", ptr(self._dev_step), self.betas[0], self.betas[1], stream_ptr())
            self._ticked = True

    def _early_hook(self, p) -> None:
        if self._early_done is None or p.grad is None:   # begin_step() not called: step() handles p as usual
            return
        cur = torch.cuda.current_stream(p.device)
        side = self._early_stream
        side.wait_stream(cur)
        with torch.cuda.stream(side), torch.no_grad():
            self._update([p], slot=("early", self._early[id(p)]))
        self._early_done.add(id(p))
        self._early_pending = True

    # ----------------------------------------------------------------------------------------------- the update
    def _update(self, plist, slot) -> None:
        """One launch per distinct step count over the chunk table of `plist` on the current stream."""
        by_step: Dict[int, list] = {}
        for p in plist:
            g = p.grad
            if not g.is_contiguous():
                g = g.contiguous()
                p.grad = g
            if not p.data.is_contiguous():
                raise VaeclustError("vaeclust.optim.Adam: non-contiguous parameter")
            st = self._state(p)
            st["step"] += 1
            by_step.setdefault(st["step"], []).append((p, g, st))
        for t, items in by_step.items():
            rows = []
            for p, g, st in items:
                n = p.numel()
                base = np.arange(0, n, _CHUNK, dtype=np.int64)
                cnt = np.minimum(_CHUNK, n - base)
                tab = np.empty((base.size, 5), dtype=np.int64)
                tab[:, 0] = p.data_ptr() + 4 * base
                tab[:, 1] = g.data_ptr() + 4 * base
                tab[:, 2] = st["exp_avg"].data_ptr() + 4 * base
                tab[:, 3] = st["exp_avg_sq"].data_ptr() + 4 * base
                tab[:, 4] = cnt
                rows.append(tab)
            table = np.concatenate(rows, axis=0)
            key = table.tobytes()
            dev = items[0][0].device
            b1, b2 = self.betas
            c = self._tables.setdefault(slot, {"key": None, "dev": None, "pinned": None})
            if self.capturable:
                if len(by_step) != 1:
                    raise VaeclustError("vaeclust.optim.Adam(capturable=True): every stepped parameter must share one "
                                        "step count")
                if self._dev_step is None:
                    self._dev_step = torch.zeros(4, dtype=torch.int32, device=dev)
                    self._dev_step[0] = t - 1
                if c["key"] != key or c["dev"] is None or c["dev"].device != dev:
                    # pinned + async so the upload is legal inside a capture (a replay re-reads the same host buffer);
                    # both buffers are allocated by the first eager step and only rewritten afterwards
                    if c["pinned"] is None or tuple(c["pinned"].shape) != table.shape:
                        c["pinned"] = torch.empty(table.shape, dtype=torch.int64).pin_memory()
                        c["dev"] = torch.empty(table.shape, dtype=torch.int64, device=dev)
                    c["pinned"].numpy()[...] = table
                    c["dev"].copy_(c["pinned"], non_blocking=True)
                    c["key"] = key
                if self._ticked:
                    lib().call("vc_adam_multi_at", ptr(c["dev"]), int(table.shape[0]), self.lr, b1, b2, self.eps,
                               ptr(self._dev_step), stream_ptr())
                else:
                    lib().call("vc_adam_multi_graph", ptr(c["dev"]), int(table.shape[0]), self.lr, b1, b2, self.eps,
                               ptr(self._dev_step), stream_ptr())
                continue
            if c["key"] != key or c["dev"] is None or c["dev"].device != dev:
                c["dev"] = torch.from_numpy(table).to(dev, non_blocking=False)
                c["key"] = key
            lib().call("vc_adam_multi", ptr(c["dev"]), int(table.shape[0]), self.lr, b1, b2, self.eps,
                       1.0 - b1 ** t, 1.0 - b2 ** t, stream_ptr())

    @torch.no_grad()
    def step(self) -> None:
        self.lr = float(self.param_groups[0]["lr"])
        done = self._early_done or ()
        rest = [p for p in self.params if p.grad is not None and id(p) not in done]
        if rest:
            self._update(rest, slot=("rest",))
        if self._early_pending:
            torch.cuda.current_stream(self.params[0].device).wait_stream(self._early_stream)
            self._early_pending = False
        self._early_done = None
        self._ticked = False

    def state_dict(self) -> dict:
        if self.capturable and self._dev_step is not None:   # replays advance the device counter only
            t = int(self._dev_step[0])
            for st in self.state.values():
                st["step"] = t
        return {"state": {i: {k: (v.clone() if torch.is_tensor(v) else v) for k, v in self._state(p).items()}
                          for i, p in enumerate(self.params)},
                "param_groups": [{"lr": self.lr, "betas": self.betas, "eps": self.eps}]}
