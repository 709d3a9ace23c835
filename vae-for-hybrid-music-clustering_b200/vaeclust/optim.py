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
        self._table = None
        self._table_key = None
        # capturable (torch.optim.Adam(capturable=True)): step counter and bias corrections live on the device, so a
        # step recorded in a CUDA graph advances correctly on every replay (vc_adam_multi_graph)
        self.capturable = bool(capturable)
        self._dev_step = None
        self._table_pinned = None
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

    @torch.no_grad()
    def step(self) -> None:
        self.lr = float(self.param_groups[0]["lr"])
        by_step: Dict[int, list] = {}
        for p in self.params:
            if p.grad is None:
                continue
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
            if self.capturable:
                if len(by_step) != 1:
                    raise VaeclustError("vaeclust.optim.Adam(capturable=True): every stepped parameter must share one "
                                        "step count")
                if self._dev_step is None:
                    self._dev_step = torch.zeros(4, dtype=torch.int32, device=dev)
                    self._dev_step[0] = t - 1
                if self._table_key != key or self._table is None or self._table.device != dev:
                    # pinned + async so the upload is legal inside a capture (a replay re-reads the same host buffer);
                    # both buffers are allocated by the first eager step and only rewritten afterwards
                    if self._table_pinned is None or tuple(self._table_pinned.shape) != table.shape:
                        self._table_pinned = torch.empty(table.shape, dtype=torch.int64).pin_memory()
                        self._table = torch.empty(table.shape, dtype=torch.int64, device=dev)
                    self._table_pinned.numpy()[...] = table
                    self._table.copy_(self._table_pinned, non_blocking=True)
                    self._table_key = key
                lib().call("vc_adam_multi_graph", ptr(self._table), int(table.shape[0]), self.lr, b1, b2, self.eps,
                           ptr(self._dev_step), stream_ptr())
                continue
            if self._table_key != key or self._table is None or self._table.device != dev:
                self._table = torch.from_numpy(table).to(dev, non_blocking=False)
                self._table_key = key
            lib().call("vc_adam_multi", ptr(self._table), int(table.shape[0]), self.lr, b1, b2, self.eps,
                       1.0 - b1 ** t, 1.0 - b2 ** t, stream_ptr())

    def state_dict(self) -> dict:
        if self.capturable and self._dev_step is not None:   # replays advance the device counter only
            t = int(self._dev_step[0])
            for st in self.state.values():
                st["step"] = t
        return {"state": {i: {k: (v.clone() if torch.is_tensor(v) else v) for k, v in self._state(p).items()}
                          for i, p in enumerate(self.params)},
                "param_groups": [{"lr": self.lr, "betas": self.betas, "eps": self.eps}]}
