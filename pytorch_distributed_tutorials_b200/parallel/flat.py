"""Flat, bucketed storage for parameters, gradients, momentum and the bf16 compute copies.

Every parameter of a wrapped module becomes a *view* into a per-bucket flat fp32 buffer; its ``.grad`` is a view
into the matching flat gradient buffer (``gradient_as_bucket_view`` semantics of torch DDP, reference
resnet/main.py:80 -> c10d::Reducer buckets).  The gradient buffers are what the P2P all-reduce kernels read on
peer GPUs, so on CUDA with world_size > 1 they are carved out of symmetric (peer-mapped) memory; everything
else is ordinary device memory.  Conv weights keep torch's logical OIHW shape but are stored OHWI
(``channels_last``), which is exactly the layout the tcgen05 kernels consume, so the bf16 copy is an
element-wise cast of the flat buffer and the fused SGD kernel never needs to know about shapes.
"""
import ctypes
from typing import Callable, List, Optional, Sequence

import torch

from ..ops import _lib

ALIGN = 64  # elements; keeps every slot 256 B (fp32) / 128 B (bf16) aligned for TMA and 16-byte vector access


def _round_up(n, a=ALIGN):
    return (n + a - 1) // a * a


def plan_buckets(sizes_bytes: Sequence[int], first_cap_bytes: int, cap_bytes: int) -> List[int]:
    """Bucket index per parameter, parameters given in gradient-ready order (native planner, csrc/reducer.cpp)."""
    n = len(sizes_bytes)
    if n == 0:
        return []
    arr = (ctypes.c_longlong * n)(*sizes_bytes)
    out = (ctypes.c_int * n)()
    _lib.lib().b2_plan_buckets(arr, n, ctypes.c_longlong(first_cap_bytes), ctypes.c_longlong(cap_bytes), out)
    return list(out)


def _view_like(flat, off, p):
    """View of flat[off:off+numel] with p's logical shape; 4-D tensors are stored channels_last (OHWI)."""
    n = p.numel()
    seg = flat[off:off + n]
    if p.dim() == 4 and flat.is_cuda:  # CPU (gloo plumbing) keeps torch's native OIHW order
        o, i, kh, kw = p.shape
        return seg.view(o, kh, kw, i).permute(0, 3, 1, 2)
    return seg.view(p.shape)


class Bucket:
    def __init__(self, index, params, device, grad_alloc):
        self.index = index
        self.params = list(params)
        self.offsets = []
        off = 0
        for p in self.params:
            self.offsets.append(off)
            off += _round_up(p.numel())
        self.numel = off
        self.used = sum(p.numel() for p in self.params)
        self.p32 = torch.zeros(off, dtype=torch.float32, device=device)
        self.g32 = grad_alloc(off) if grad_alloc is not None else torch.zeros(off, dtype=torch.float32,
                                                                                device=device)
        self.mom = None  # allocated lazily by the optimizer
        self.shadow = torch.zeros(off, dtype=torch.bfloat16, device=device) if device.type == "cuda" else None


class FlatParameterStore:
    """Owns the flat buffers of a list of parameters split into buckets (``bucket_of_param[i]`` per parameter)."""

    def __init__(self, params: Sequence[torch.nn.Parameter], bucket_of_param: Sequence[int], device,
                 grad_alloc: Optional[Callable[[int], torch.Tensor]] = None):
        self.params = list(params)
        self.device = torch.device(device)
        nb = (max(bucket_of_param) + 1) if len(bucket_of_param) else 0
        groups = [[] for _ in range(nb)]
        for p, b in zip(self.params, bucket_of_param):
            groups[b].append(p)
        self.buckets = [Bucket(i, g, self.device, grad_alloc) for i, g in enumerate(groups)]
        self.slot = {}  # id(param) -> (bucket, offset)
        self.index = {id(p): i for i, p in enumerate(self.params)}
        self.bucket_of_param = list(bucket_of_param)
        self.on_ready: Optional[Callable[[torch.nn.Parameter], None]] = None
        self.optimizer = None  # set by optim.SGD when it is asked to fuse the update into the reduction
        self._shadow_versions = None
        self.shadow_fresh = False
        self._numel = {id(p): p.numel() for p in self.params}
        with torch.no_grad():
            for b in self.buckets:
                for p, off in zip(b.params, b.offsets):
                    view = _view_like(b.p32, off, p)
                    view.copy_(p.data)
                    p.data = view
                    p._b2_store = self
                    self.slot[id(p)] = (b, off)
                    if p.grad is not None:
                        gv = _view_like(b.g32, off, p)
                        gv.copy_(p.grad)
                        p.grad = gv

    def rebuild(self, params, bucket_of_param, grad_alloc=None):
        """Re-bucket the same parameters in a new order (DDP's one-off rebuild in observed gradient-ready order):
        fresh flat buffers, values, live gradients and momentum carried over, every parameter re-pointed."""
        old_slot = self.slot
        old_mom = {pid: (b.mom[off:off + self._numel[pid]] if b.mom is not None else None)
                   for pid, (b, off) in old_slot.items()} if hasattr(self, "_numel") else {}
        self.params = list(params)
        nb = (max(bucket_of_param) + 1) if len(bucket_of_param) else 0
        groups = [[] for _ in range(nb)]
        for p, b in zip(self.params, bucket_of_param):
            groups[b].append(p)
        had_mom = any(b.mom is not None for b in self.buckets)
        self.buckets = [Bucket(i, g, self.device, grad_alloc) for i, g in enumerate(groups)]
        self.slot = {}
        self.index = {id(p): i for i, p in enumerate(self.params)}
        self.bucket_of_param = list(bucket_of_param)
        with torch.no_grad():
            for b in self.buckets:
                if had_mom:
                    b.mom = torch.zeros(b.numel, dtype=torch.float32, device=self.device)
                for p, off in zip(b.params, b.offsets):
                    view = _view_like(b.p32, off, p)
                    view.copy_(p.data)
                    p.data = view
                    self.slot[id(p)] = (b, off)
                    if p.grad is not None:
                        gv = _view_like(b.g32, off, p)
                        gv.copy_(p.grad)
                        p.grad = gv
                    om = old_mom.get(id(p))
                    if om is not None and b.mom is not None:
                        b.mom[off:off + p.numel()].copy_(om)
        self._numel = {id(p): p.numel() for p in self.params}
        self.shadow_fresh = False
        self._shadow_versions = None

    # ------------------------------------------------------------------ views
    def grad_view(self, p):
        b, off = self.slot[id(p)]
        return _view_like(b.g32, off, p)

    def shadow_view(self, p):
        b, off = self.slot[id(p)]
        return _view_like(b.shadow, off, p)

    def mom_view(self, p):
        b, off = self.slot[id(p)]
        return _view_like(b.mom, off, p)

    def ensure_momentum(self):
        for b in self.buckets:
            if b.mom is None:
                b.mom = torch.zeros(b.numel, dtype=torch.float32, device=self.device)

    def manages(self, p):
        return id(p) in self.slot

    # ------------------------------------------------------------------ bf16 compute copies
    def sync_shadow(self, force=False):
        """Refresh the bf16 copies if any parameter was modified outside the fused kernels (init, load_state_dict,
        a stock optimizer).  The fused SGD / all-reduce kernels refresh them in the same pass and set
        ``shadow_fresh``; torch's per-tensor version counters catch every other in-place write."""
        if self.device.type != "cuda":
            return
        versions = [p._version for p in self.params]
        if not force and self.shadow_fresh and versions == self._shadow_versions:
            return
        from ..ops import native
        for b in self.buckets:
            native.cast_f32_bf16(b.p32, b.shadow)
        self._shadow_versions = versions
        self.shadow_fresh = True

    def mark_params_updated_natively(self):
        """Called by kernels that rewrote p32 *and* shadow through raw pointers."""
        self._shadow_versions = [p._version for p in self.params]
        self.shadow_fresh = True

    # ------------------------------------------------------------------ gradient delivery (engine -> reducer)
    def grad_slot(self, p):
        """(view, accumulate) for writing p's gradient straight into its bucket slot. ``accumulate`` is True when
        a live gradient is already there (gradient accumulation / no_sync)."""
        gv = self.grad_view(p)
        if p.grad is None:
            return gv, False
        if p.grad.data_ptr() != gv.data_ptr():
            # a foreign gradient tensor is attached (e.g. set by user code): fold it in and take over
            gv.copy_(p.grad)
            p.grad = gv
        return gv, True

    def publish_grad(self, p, gv):
        if p.grad is None:
            p.grad = gv
        if self.on_ready is not None:
            self.on_ready(p)


def store_of(p):
    return getattr(p, "_b2_store", None)
