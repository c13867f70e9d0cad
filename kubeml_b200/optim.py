"""Flat-arena optimizers with the torch.optim surface the reference functions use
(`SGD(self.parameters(), lr=..., momentum=..., weight_decay=...)`,
ml/experiments/kubeml/function_lenet.py:79-81, function_resnet34.py:62; Adam in function_vgg11.py).

When the model lives in a `WeightArena` (every parameter is a slice of one fp32 buffer and every
`.grad` a slice of another) the whole step is one kernel launch (csrc/optim.cu) instead of
torch's per-tensor loop; otherwise the same math runs per parameter.  `reset_state()` is the
reference's per-sync optimizer reset (python/kubeml/kubeml/network.py:121-128).
"""
from __future__ import annotations

from typing import Iterable, List, Optional

import torch

from . import ops


def _flat_span(params: List[torch.nn.Parameter], attr: str):
    """(ptr, numel) if the tensors named by attr ('data' or 'grad') tile one contiguous buffer
    in order (gaps of < 64 padding elements allowed), else None."""
    lo = hi = None
    prev_end = None
    for p in params:
        t = p.data if attr == "data" else p.grad
        if t is None or not t.is_contiguous() or t.dtype != torch.float32:
            return None
        a = t.data_ptr()
        if prev_end is not None and not (0 <= a - prev_end < 64 * 4):
            return None
        lo = a if lo is None else lo
        prev_end = a + t.numel() * 4
        hi = prev_end
    return lo, (hi - lo) // 4


class _FlatOptimizer:
    def __init__(self, params: Iterable[torch.nn.Parameter], defaults: dict):
        self.params: List[torch.nn.Parameter] = [p for p in params if p.requires_grad]
        self.defaults = defaults
        self.param_groups = [dict(defaults, params=self.params)]
        self.state: dict = {}
        self._arena = None
        self._flat = None

    # -- arena attachment (done by DeviceSync.attach / models) --------------------------------
    def attach_arena(self, arena) -> None:
        self._arena = arena
        self._flat = None

    def _flat_views(self):
        """(w, g) flat fp32 tensors when parameters and grads are arena slices."""
        if self._arena is None:
            return None
        if self._flat is None:
            a = self._arena
            n = a.layout.n_param
            if n == 0:
                return None
            first = self.params[0]
            if first.data_ptr() != a.weights.data_ptr():
                return None
            self._flat = (a.weights[:n], a.grads[:n])
        # grads must still be the arena slices (zero_grad(set_to_none=True) by user code breaks it)
        p0 = self.params[0]
        if p0.grad is None or p0.grad.data_ptr() != self._arena.grads.data_ptr():
            return None
        return self._flat

    def zero_grad(self, set_to_none: bool = False) -> None:
        if self._arena is not None:
            a = self._arena
            # a native network's backward *writes* every gradient (and clears what it accumulates
            # into itself), so the 87 MB fill is only needed for autograd-accumulated models
            if not a.layout.want_bf16:
                ops.fill_(a.grads[: max(a.layout.n_param, 4)], 0.0)
            for name, p in self._named():
                p.grad = a.grad(name)
            return
        for p in self.params:
            if p.grad is not None:
                if set_to_none:
                    p.grad = None
                elif p.grad._is_view():
                    p.grad.zero_()
                else:
                    p.grad.detach_().zero_()

    def _named(self):
        if self._arena is None:
            return []
        if not hasattr(self, "_names"):
            by_ptr = {self._arena.tensor(n).data_ptr(): n
                      for n, e in self._arena.layout.entries.items() if e.is_param}
            self._names = [(by_ptr[p.data_ptr()], p) for p in self.params if p.data_ptr() in by_ptr]
        return self._names

    def reset_state(self) -> None:
        raise NotImplementedError

    def state_dict(self) -> dict:
        return {"defaults": self.defaults}


class SGD(_FlatOptimizer):
    def __init__(self, params, lr: float, momentum: float = 0.0, weight_decay: float = 0.0,
                 nesterov: bool = False):
        super().__init__(params, dict(lr=lr, momentum=momentum, weight_decay=weight_decay,
                                      nesterov=nesterov))
        self._bufs: Optional[List[torch.Tensor]] = None
        self._steps_in_round = 0
        self._defer_at = 0
        self._deferred = False

    def defer_step_at(self, nsteps: int) -> None:
        """The `nsteps`-th step() from now is not launched: the K-AVG merge kernel applies it in
        the same pass that averages the replicas (only meaningful on the flat arena path)."""
        self._steps_in_round = 0
        self._defer_at = int(nsteps)
        self._deferred = False

    def reset_state(self) -> None:
        if self._arena is not None and self.defaults["momentum"]:
            ops.fill_(self._arena.state1[: max(self._arena.layout.n_param, 4)], 0.0)
        self._bufs = None
        self.state = {}

    @torch.no_grad()
    def step(self) -> None:
        g = self.param_groups[0]
        lr, mom, wd, nes = g["lr"], g["momentum"], g["weight_decay"], g["nesterov"]
        flat = self._flat_views()
        if flat is not None:
            w, gr = flat
            self._steps_in_round += 1
            if self._defer_at and self._steps_in_round == self._defer_at and not nes and w.is_cuda:
                self._deferred = True
                self._defer_at = 0
                return
            mb = self._arena.state1[: w.numel()] if mom else None
            ops.sgd_step_(w, gr, mb, lr, wd, mom, nes, w_bf16=getattr(self._arena, "wbf16", None))
            return
        if mom and self._bufs is None:
            self._bufs = [torch.zeros_like(p) for p in self.params]
        for i, p in enumerate(self.params):
            if p.grad is None:
                continue
            dense = p.data.is_contiguous() and p.grad.is_contiguous()
            w, gr = (p.data.view(-1), p.grad.view(-1)) if dense else (p.data, p.grad)
            mb = (self._bufs[i].view(-1) if dense else self._bufs[i]) if mom else None
            ok = dense and w.is_cuda and w.numel() % 4 == 0 and w.data_ptr() % 16 == 0 and gr.data_ptr() % 16 == 0 \
                and (mb is None or mb.data_ptr() % 16 == 0)
            if ok or (dense and not w.is_cuda):
                ops.sgd_step_(w, gr, mb, lr, wd, mom, nes)
            else:   # odd-sized or strided (channels-last) tensor outside an arena: same math with torch ops
                d = gr.add(w, alpha=wd) if wd else gr
                if mom:
                    mb.mul_(mom).add_(d)
                    d = d.add(mb, alpha=mom) if nes else mb
                w.add_(d, alpha=-lr)


class Adam(_FlatOptimizer):
    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self._t = 0
        self._m: Optional[List[torch.Tensor]] = None
        self._v: Optional[List[torch.Tensor]] = None

    def attach_arena(self, arena) -> None:
        super().attach_arena(arena)
        arena.opt_step.zero_()          # a fresh optimizer counts from step 1

    def reset_state(self) -> None:
        self._t = 0
        if self._arena is not None:
            n = max(self._arena.layout.n_param, 4)
            ops.fill_(self._arena.state1[:n], 0.0)
            ops.fill_(self._arena.state2[:n], 0.0)
            self._arena.opt_step.zero_()
        self._m = self._v = None
        self.state = {}

    @torch.no_grad()
    def step(self) -> None:
        g = self.param_groups[0]
        self._t += 1
        flat = self._flat_views()
        if flat is not None:
            w, gr = flat
            n = w.numel()
            ops.adam_step_(w, gr, self._arena.state1[:n], self._arena.state2[:n], self._t, g["lr"],
                           g["betas"], g["eps"], g["weight_decay"], w_bf16=getattr(self._arena, "wbf16", None),
                           step_dev=self._arena.opt_step if w.is_cuda else None)
            return
        if self._m is None:
            self._m = [torch.zeros_like(p) for p in self.params]
            self._v = [torch.zeros_like(p) for p in self.params]
        for i, p in enumerate(self.params):
            if p.grad is None:
                continue
            dense = p.data.is_contiguous() and p.grad.is_contiguous()
            w, gr = (p.data.view(-1), p.grad.view(-1)) if dense else (p.data, p.grad)
            m, v = (self._m[i].view(-1), self._v[i].view(-1)) if dense else (self._m[i], self._v[i])
            aligned = dense and all(t.data_ptr() % 16 == 0 for t in (w, gr, m, v)) and w.numel() % 4 == 0
            if (dense and not w.is_cuda) or aligned:
                ops.adam_step_(w, gr, m, v, self._t, g["lr"], g["betas"], g["eps"], g["weight_decay"])
            else:
                d = gr.add(w, alpha=g["weight_decay"]) if g["weight_decay"] else gr
                m.mul_(g["betas"][0]).add_(d, alpha=1 - g["betas"][0])
                v.mul_(g["betas"][1]).addcmul_(d, d, value=1 - g["betas"][1])
                bc1, bc2 = 1 - g["betas"][0] ** self._t, 1 - g["betas"][1] ** self._t
                w.addcdiv_(m, v.sqrt() / bc2 ** 0.5 + g["eps"], value=-g["lr"] / bc1)
