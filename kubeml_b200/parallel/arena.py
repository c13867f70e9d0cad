"""Device-resident weight store: one flat arena per worker, addressed with the reference's key
layout.

Reference: every `state_dict` entry is a RedisAI tensor, `"{jobId}:{name}"` for the reference
model and `"{jobId}:{name}/{funcId}"` for worker replicas (python/kubeml/kubeml/network.py:444-461,
ml/pkg/model/utils.go:140-158), float32 for weights/buffers and int64 for BatchNorm's
`num_batches_tracked` (ml/pkg/model/model.go:225-238).

Here the same names index slices of a single allocation:

    [ fp32 weights+buffers | int64 counters | signal pad | fp32 grads | fp32 optimizer state x2 ]

The allocation comes from the CUDA VMM API (csrc/arena.cpp) so peers can map it and a multicast
object can be bound to it.  `bind()` re-points an `nn.Module`'s parameters, buffers and `.grad`s at
the slices, so the module trains *in* the arena: publishing a replica and loading the merged
model (the reference's per-tensor `.cpu()` + TENSORSET / TENSORGET + `load_state_dict`) become
no-ops, and optimizers / the merge kernel run over flat buffers in one launch.
"""
from __future__ import annotations

import ctypes
from collections import OrderedDict
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import torch

from ..model.store import weight_key

_F32_ALIGN = 64         # elements: tensors start on 256-byte (fp32) / 128-byte (bf16 copy) boundaries
_REGION_ALIGN = 256     # bytes
SIGNAL_WORDS = 64       # uint32 words of the signal pad (2 phases x 8 ranks used)


@dataclass
class Entry:
    name: str
    kind: str                 # "f32" | "i64"
    shape: Tuple[int, ...]
    offset: int               # elements from the start of its region
    numel: int
    is_param: bool
    # Physical order of the elements when it differs from the logical (state_dict) shape: the
    # native convolutions keep filters as [Cout][KH][KW][Cin]; the parameter the user sees is the
    # permuted view with torch's [Cout][Cin][KH][KW] shape, so checkpoints keep the reference layout.
    phys_shape: Optional[Tuple[int, ...]] = None
    perm: Optional[Tuple[int, ...]] = None


def _round_up(x: int, a: int) -> int:
    return (x + a - 1) // a * a


class ArenaLayout:
    """Where each `state_dict` entry lives.  Parameters come first in the fp32 region so that
    gradients and optimizer state only need `n_param` elements."""

    def __init__(self, module: torch.nn.Module):
        sd = module.state_dict()
        params = dict(module.named_parameters())
        param_names = set(params)
        self.order: List[str] = list(sd.keys())          # state_dict order = reference layer list
        self.entries: "OrderedDict[str, Entry]" = OrderedDict()
        off = 0
        for name in [n for n in self.order if n in param_names]:
            t = sd[name]
            if not t.is_floating_point():
                raise TypeError(f"parameter {name} is not floating point")
            e = Entry(name, "f32", tuple(t.shape), off, t.numel(), True)
            if getattr(params[name], "_kml_layout", None) == "OHWI" and t.dim() == 4:
                O, I, KH, KW = t.shape
                e.phys_shape, e.perm = (O, KH, KW, I), (0, 3, 1, 2)
            self.entries[name] = e
            off += _round_up(max(t.numel(), 1), _F32_ALIGN)
        self.n_param = _round_up(off, 64)
        self.want_bf16 = bool(getattr(module, "native", False))
        off = self.n_param
        ioff = 0
        for name in [n for n in self.order if n not in param_names]:
            t = sd[name]
            if t.is_floating_point():
                self.entries[name] = Entry(name, "f32", tuple(t.shape), off, t.numel(), False)
                off += _round_up(max(t.numel(), 1), _F32_ALIGN)
            else:
                self.entries[name] = Entry(name, "i64", tuple(t.shape), ioff, t.numel(), False)
                ioff += max(t.numel(), 1)
        self.n_f32 = _round_up(off, 64)
        self.n_i64 = ioff
        # byte offsets of the regions
        self.off_f32 = 0
        self.off_i64 = _round_up(self.n_f32 * 4, _REGION_ALIGN)
        self.off_flags = _round_up(self.off_i64 + self.n_i64 * 8, _REGION_ALIGN)
        self.off_status = self.off_flags + SIGNAL_WORDS * 4
        self.off_grad = _round_up(self.off_status + 64, _REGION_ALIGN)
        self.off_state1 = _round_up(self.off_grad + self.n_param * 4, _REGION_ALIGN)
        self.off_state2 = _round_up(self.off_state1 + self.n_param * 4, _REGION_ALIGN)
        self.off_bf16 = _round_up(self.off_state2 + self.n_param * 4, _REGION_ALIGN)
        self.nbytes = _round_up(self.off_bf16 + (self.n_param * 2 if self.want_bf16 else 0), _REGION_ALIGN)

    def signature(self) -> Tuple:
        return tuple((e.name, e.kind, e.shape, e.phys_shape) for e in self.entries.values())

    def index(self, job_id: str, func_id: int = -1) -> List[dict]:
        """Key -> slice table in the reference's naming (for checkpoints / debugging)."""
        return [{"key": weight_key(job_id, e.name, func_id), "dtype": e.kind, "shape": list(e.shape),
                 "offset": e.offset} for e in (self.entries[n] for n in self.order)]


class _DeviceSpan:
    """Minimal `__cuda_array_interface__` carrier so torch can view foreign device memory."""

    def __init__(self, ptr: int, shape: Tuple[int, ...], typestr: str, owner):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False),
                                         "version": 3, "strides": None}
        self._owner = owner


class WeightArena:
    def __init__(self, layout: ArenaLayout, device: torch.device, exportable: bool = True):
        self.layout = layout
        self.device = torch.device(device)
        self.handle = None
        self._host = None
        if self.device.type == "cuda":
            from .. import _native
            lib = _native.lib()
            h = ctypes.c_void_p()
            _native.check(lib.kml_arena_create(self.device.index or 0, layout.nbytes,
                                               1 if exportable else 0, ctypes.byref(h)),
                          "kml_arena_create")
            self.handle = h
            self.base = int(lib.kml_arena_ptr(h))
            self.nbytes = int(lib.kml_arena_size(h))
        else:   # CPU stand-in with the same layout (tests, oracle)
            self._host = torch.zeros(layout.nbytes // 8 + 1, dtype=torch.int64)
            self.base = self._host.data_ptr()
            self.nbytes = layout.nbytes
        L = layout
        self.weights = self._view(L.off_f32, (L.n_f32,), torch.float32)
        self.counters = self._view(L.off_i64, (max(L.n_i64, 1),), torch.int64)
        self.flags = self._view(L.off_flags, (SIGNAL_WORDS,), torch.int32)
        self.status = self._view(L.off_status, (1,), torch.int32)
        self.opt_step = self._view(L.off_status + 16, (2,), torch.int32)     # device-side Adam step count
        self.grads = self._view(L.off_grad, (max(L.n_param, 4),), torch.float32)
        self.state1 = self._view(L.off_state1, (max(L.n_param, 4),), torch.float32)
        self.state2 = self._view(L.off_state2, (max(L.n_param, 4),), torch.float32)
        self.wbf16 = (self._view(L.off_bf16, (L.n_param,), torch.bfloat16) if L.want_bf16 else None)

    # ---- raw views -----------------------------------------------------------------------
    def _view(self, byte_off: int, shape: Tuple[int, ...], dtype: torch.dtype) -> torch.Tensor:
        """A tensor over arena bytes.  Every call returns an *independent* tensor object (own
        autograd version counter): slices of one base tensor would share a counter, and BatchNorm's
        in-place running-stat update would then invalidate every weight saved for backward."""
        n = 1
        for s in shape:
            n *= s
        if self._host is not None:
            buf = (ctypes.c_char * self.layout.nbytes).from_address(self.base)
            if n == 0:
                return torch.empty(shape, dtype=dtype)
            return torch.frombuffer(buf, dtype=dtype, count=n, offset=byte_off).view(shape)
        if n == 0:
            return torch.empty(shape, dtype=dtype, device=self.device)
        if dtype == torch.bfloat16:      # no bf16 in the array interface: view int16 bits
            span = _DeviceSpan(self.base + byte_off, tuple(shape), "<i2", self)
            return torch.as_tensor(span, device=self.device).view(torch.bfloat16)
        typestr = {torch.float32: "<f4", torch.int64: "<i8", torch.int32: "<i4"}[dtype]
        span = _DeviceSpan(self.base + byte_off, tuple(shape), typestr, self)
        return torch.as_tensor(span, device=self.device)

    def ptr(self, byte_off: int = 0) -> int:
        return self.base + byte_off

    def tensor(self, name: str) -> torch.Tensor:
        e = self.layout.entries[name]
        if e.kind == "f32":
            return self._logical(e, self.layout.off_f32 + e.offset * 4, torch.float32)
        return self._view(self.layout.off_i64 + e.offset * 8, e.shape, torch.int64)

    def _logical(self, e: Entry, byte_off: int, dtype: torch.dtype) -> torch.Tensor:
        if e.phys_shape is not None:
            return self._view(byte_off, e.phys_shape, dtype).permute(*e.perm)
        return self._view(byte_off, e.shape, dtype)

    def grad(self, name: str) -> torch.Tensor:
        e = self.layout.entries[name]
        assert e.is_param
        return self._logical(e, self.layout.off_grad + e.offset * 4, torch.float32)

    def physical(self, name: str, region: str = "weights") -> torch.Tensor:
        """The entry in its physical element order (what the native kernels read):
        region in {"weights", "grads", "bf16"}."""
        e = self.layout.entries[name]
        shape = e.phys_shape or e.shape
        if region == "weights":
            return self._view(self.layout.off_f32 + e.offset * 4, shape, torch.float32)
        if region == "grads":
            return self._view(self.layout.off_grad + e.offset * 4, shape, torch.float32)
        return self._view(self.layout.off_bf16 + e.offset * 2, shape, torch.bfloat16)

    # ---- module binding ------------------------------------------------------------------
    def bind(self, module: torch.nn.Module, copy_in: bool = True, bind_grads: bool = True) -> None:
        """Make `module` live in the arena.  With copy_in the module's current values are written
        into the arena first; otherwise the arena's content (e.g. a merged model) wins."""
        params = dict(module.named_parameters())
        with torch.no_grad():
            for name, e in self.layout.entries.items():
                view = self.tensor(name)
                if e.is_param:
                    p = params[name]
                    if copy_in:
                        view.copy_(p.detach().to(view.device, non_blocking=True))
                    p.data = view
                    if bind_grads:
                        g = self.grad(name)
                        p.grad = g
                else:
                    owner, _, leaf = name.rpartition(".")
                    mod = module.get_submodule(owner) if owner else module
                    if copy_in:
                        view.copy_(mod._buffers[leaf].detach().to(view.device, non_blocking=True))
                    mod._buffers[leaf] = view
        self.refresh_bf16()
        hook = getattr(module, "_on_arena_bound", None)
        if callable(hook):
            hook(self)

    def refresh_bf16(self) -> None:
        """Re-derive the bf16 tensor-core copy of the parameters from the fp32 masters (after a
        bind, a checkpoint load or a peer copy; the optimizer and the merge keep it current)."""
        if self.wbf16 is not None and self.layout.n_param:
            from .. import ops
            ops.cast_bf16_(self.wbf16, self.weights[: self.layout.n_param])

    def load_state_dict(self, state: Dict[str, torch.Tensor]) -> None:
        with torch.no_grad():
            for name in self.layout.entries:
                self.tensor(name).copy_(state[name].to(self.device))
        self.refresh_bf16()

    def state_dict(self) -> "OrderedDict[str, torch.Tensor]":
        return OrderedDict((n, self.tensor(n).detach().clone()) for n in self.layout.order)

    def close(self) -> None:
        if self.handle is not None:
            from .. import _native
            for name in ("weights", "counters", "flags", "status", "grads", "state1", "state2", "wbf16"):
                setattr(self, name, None)
            _native.lib().kml_arena_destroy(self.handle)
            self.handle = None
