"""Flat-arena optimizers with the torch.optim surface the reference functions use
(`SGD(self.parameters(), lr=..., momentum=..., weight_decay=...)`,
ml/experiments/kubeml/function_lenet.py:79-81, function_resnet34.py:62; Adam in function_vgg11.py).

When the model lives in a `WeightArena` (every parameter is a slice of one fp32 buffer and every
`.grad` a slice of another) the whole step is one kernel launch (csrc/optim.cu) instead of
torch's per-tensor loop; otherwise the same math runs per parameter.  `reset_state()` is the
reference's per-sync optimizer reset (python/kubeml/kubeml/network.py:121-128).
"""
from __future__ import annotations

import os
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
    def __init__(self, params: Iterable, defaults: dict):
        # torch.optim surface: an iterable of parameters, or a list of group dicts with per-group
        # overrides (e.g. a no-decay group for BatchNorm and biases)
        plist = list(params)
        self.defaults = defaults
        if plist and isinstance(plist[0], dict):
            self.param_groups = []
            for grp in plist:
                grp = dict(grp)
                grp["params"] = [p for p in grp["params"] if p.requires_grad]
                self.param_groups.append(dict(defaults, **grp))
        else:
            self.param_groups = [dict(defaults, params=[p for p in plist if p.requires_grad])]
        self.params: List[torch.nn.Parameter] = [p for g in self.param_groups for p in g["params"]]
        self.state: dict = {}
        self._arena = None
        self._flat = None
        self._step_calls = 0          # step() invocations, counted by KubeModel around each train()
        self._armed = False
        self._deferred = False
        self._deferred_lo = None          # a deferred step covers only [0, lo): the tail was stepped early (SGD)
        self._state_clean = False     # the merge kernel zeroed the optimizer state: reset_state() has nothing to do

    # -- arena attachment (done by DeviceSync.attach / models) --------------------------------
    def attach_arena(self, arena) -> None:
        self._arena = arena
        self._flat = None

    def _flat_views(self):
        """(w, g) flat fp32 tensors when parameters and grads are arena slices."""
        if self._arena is None or self._flat is False:
            return None
        if self._flat is None:
            a = self._arena
            n = a.layout.n_param
            # The one-launch update covers the arena's whole parameter prefix with ONE set of
            # hyper-parameters: only valid with a single group that holds every arena parameter
            # (a frozen parameter would still get weight decay, a second group would lose its lr).
            n_arena_params = sum(1 for e in a.layout.entries.values() if e.is_param)
            if n == 0 or len(self.param_groups) != 1 or len(self.params) != n_arena_params \
                    or not self.params or self.params[0].data_ptr() != a.weights.data_ptr():
                self._flat = False
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

    # ---- the round's last step, left to the K-AVG merge kernel (csrc/merge.cu phase 0) -------------------
    def arm_deferral(self) -> bool:
        """The next step() is not launched: the K-AVG merge kernel applies it in the same pass that
        averages the replicas.  KubeModel arms this only around the last train() call of a sync round, and
        only for functions it has seen call step() exactly once per batch; a second step() before the merge
        flushes the pending one first (see step())."""
        ok = (self._arena is not None and self._flat is not False and len(self.param_groups) == 1
              and not self.param_groups[0].get("nesterov", False) and bool(self.params) and self.params[0].is_cuda)
        self._armed = bool(ok)
        return self._armed

    def disarm(self) -> None:
        self._armed = False

    def flush_deferred(self) -> None:
        """Apply a pending deferred update with the ordinary kernel (the merge did not take it:
        this rank sat the round out, or user code stepped again)."""
        if not self._deferred:
            return
        self._deferred = False
        flat = self._flat_views()
        if flat is not None:
            lo, self._deferred_lo = self._deferred_lo, None
            if lo:                       # the tail [lo, n) was already stepped when the deferral was taken
                w, gr = flat
                self._flat_step(w[:lo], gr[:lo])
            else:
                self._flat_step(*flat)

    def _take_deferral(self) -> bool:
        """Inside step() on the flat path: True when this call is the one the merge kernel will apply."""
        if self._deferred:                 # a second step() before the merge: the pending one goes first
            self.flush_deferred()
        if self._armed:
            self._armed = False
            self._deferred = True
            self._deferred_lo = None
            return True
        return False

    def _flat_step(self, w, gr) -> None:
        raise NotImplementedError

    def _params_with_groups(self):
        for grp in self.param_groups:
            for p in grp["params"]:
                yield p, grp

    def _after_param_steps(self) -> None:
        """Per-parameter updates (several groups, frozen parameters) bypass the one-launch kernel that
        also refreshes the bf16 tensor-core copies: refresh them here."""
        a = self._arena
        if a is not None and getattr(a, "wbf16", None) is not None:
            a.refresh_bf16()

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

    def reset_state(self) -> None:
        if self._arena is not None and self.defaults["momentum"] and not self._state_clean:
            ops.fill_(self._arena.state1[: max(self._arena.layout.n_param, 4)], 0.0)
        self._bufs = None
        self.state = {}

    def _flat_step(self, w, gr) -> None:
        g = self.param_groups[0]
        lr, mom, wd, nes = g["lr"], g["momentum"], g["weight_decay"], g["nesterov"]
        mb = self._arena.state1[: w.numel()] if mom else None
        wb = getattr(self._arena, "wbf16", None)
        lo = self._early_tail_step(w, gr)
        if lo:
            ops.sgd_step_(w[:lo], gr[:lo], mb[:lo] if mb is not None else None, lr, wd, mom, nes,
                          w_bf16=wb[:lo] if wb is not None else None)
            torch.cuda.current_stream(w.device).wait_stream(self._arena._opt_stream)
            return
        ops.sgd_step_(w, gr, mb, lr, wd, mom, nes, w_bf16=wb)

    def _early_tail_step(self, w, gr) -> int:
        """The network told us that the gradients of the parameter tail [lo, n) were final early in its backward
        pass (ResNet: layer4 + fc, 62 % of the bytes).  Their update runs on its own stream behind exactly those
        two events: under graph capture it becomes a node that depends only on them, so it overlaps the rest of the
        backward pass instead of trailing it.  Returns lo (0: nothing was launched); the caller joins the stream."""
        early = getattr(self._arena, "early_ready", None)
        if early is None or not w.is_cuda or os.environ.get("KUBEML_SGD_OVERLAP", "1") == "0":
            return 0
        g = self.param_groups[0]
        lr, mom, wd, nes = g["lr"], g["momentum"], g["weight_decay"], g["nesterov"]
        mb = self._arena.state1[: w.numel()] if mom else None
        wb = getattr(self._arena, "wbf16", None)
        lo, ev_main, ev_side = early
        self._arena.early_ready = None
        if lo <= 0 or lo >= w.numel() or lo % 4:
            return 0
        st = getattr(self._arena, "_opt_stream", None)
        if st is None:
            st = self._arena._opt_stream = torch.cuda.Stream(device=w.device)
        st.wait_event(ev_main)
        st.wait_event(ev_side)
        with torch.cuda.stream(st):
            ops.sgd_step_(w[lo:], gr[lo:], mb[lo:] if mb is not None else None, lr, wd, mom, nes,
                          w_bf16=wb[lo:] if wb is not None else None)
        return int(lo)

    @torch.no_grad()
    def step(self) -> None:
        g = self.param_groups[0]
        lr, mom, wd, nes = g["lr"], g["momentum"], g["weight_decay"], g["nesterov"]
        self._step_calls += 1
        self._state_clean = False
        flat = self._flat_views()
        if flat is not None:
            w, gr = flat
            if w.is_cuda and not nes and self._take_deferral():
                # The merge kernel applies this step - but only to the layers whose gradients arrive last: the
                # tail that was final early is stepped now, under the rest of the backward pass, which leaves
                # the merge 38 % of the bytes to touch (KUBEML_PARTIAL_DEFER=0: the merge takes all of it).
                if os.environ.get("KUBEML_PARTIAL_DEFER", "1") != "0":
                    lo = self._early_tail_step(w, gr)
                    if lo:
                        torch.cuda.current_stream(w.device).wait_stream(self._arena._opt_stream)
                        self._deferred_lo = lo
                return
            self._flat_step(w, gr)
            return
        if self._bufs is None and any(grp["momentum"] for grp in self.param_groups):
            self._bufs = [torch.zeros_like(p) for p in self.params]
        for i, (p, grp) in enumerate(self._params_with_groups()):
            if p.grad is None:
                continue
            lr, mom, wd, nes = grp["lr"], grp["momentum"], grp["weight_decay"], grp["nesterov"]
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
        self._after_param_steps()


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
        if self._arena is not None and not self._state_clean:
            n = max(self._arena.layout.n_param, 4)
            ops.fill_(self._arena.state1[:n], 0.0)
            ops.fill_(self._arena.state2[:n], 0.0)
            self._arena.opt_step.zero_()
        self._m = self._v = None
        self.state = {}

    def _flat_step(self, w, gr) -> None:
        g = self.param_groups[0]
        n = w.numel()
        ops.adam_step_(w, gr, self._arena.state1[:n], self._arena.state2[:n], self._t, g["lr"],
                       g["betas"], g["eps"], g["weight_decay"], w_bf16=getattr(self._arena, "wbf16", None),
                       step_dev=self._arena.opt_step if w.is_cuda else None)

    @torch.no_grad()
    def step(self) -> None:
        g = self.param_groups[0]
        self._t += 1
        self._step_calls += 1
        self._state_clean = False
        flat = self._flat_views()
        if flat is not None:
            w, gr = flat
            if w.is_cuda and self._take_deferral():
                return
            self._flat_step(w, gr)
            return
        if self._m is None:
            self._m = [torch.zeros_like(p) for p in self.params]
            self._v = [torch.zeros_like(p) for p in self.params]
        for i, (p, g) in enumerate(self._params_with_groups()):
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
        self._after_param_steps()
