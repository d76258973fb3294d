"""SGD with momentum and weight decay, same call shape and state_dict as ``torch.optim.SGD``
(reference resnet/main.py:103 ``optim.SGD(ddp_model.parameters(), lr, momentum=0.9, weight_decay=1e-5)`` and
:120,:124 ``zero_grad()`` / ``step()``).

When the parameters live in a :class:`FlatParameterStore` on CUDA the whole update of a bucket is one kernel
(weight decay, momentum, parameter update and the bf16 compute-copy refresh in a single pass), and with
``fuse_with_backward=True`` that kernel is the epilogue of the bucket's gradient all-reduce: the update is
applied while backward is still running and ``step()`` only completes the handshake.  Update rule (torch's):
``g += wd*p; buf = g (first step) | mu*buf + g; p -= lr*buf``.
"""
import torch

from .ops import _lib, native
from .parallel.flat import store_of


class SGD(torch.optim.SGD):
    def __init__(self, params, lr=1e-3, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False,
                 fuse_with_backward=False):
        super().__init__(params, lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                         nesterov=nesterov)
        self.fuse_with_backward = bool(fuse_with_backward)
        self._steps = 0
        self._stores = []
        seen = set()
        for g in self.param_groups:
            for p in g["params"]:
                st = store_of(p)
                if st is not None and id(st) not in seen:
                    seen.add(id(st))
                    self._stores.append(st)
        if self.fuse_with_backward:
            for st in self._stores:
                st.optimizer = self
        self._fused_done = set()  # bucket ids already updated by the reduce kernel in this iteration
        self._hyper_dev = {}      # device -> (tensor {lr, momentum, weight_decay}, values it holds)

    # ------------------------------------------------------------------ fused path plumbing
    def _native_ok(self):
        if not self._stores or not _lib.use_native():
            return False
        if len(self.param_groups) != 1:
            return False
        g = self.param_groups[0]
        if g["dampening"] != 0 or g["nesterov"] or g.get("maximize", False):
            return False
        managed = all(store_of(p) is not None and p.is_cuda for p in g["params"])
        return managed

    def hyper(self):
        g = self.param_groups[0]
        return float(g["lr"]), float(g["momentum"]), float(g["weight_decay"])

    def hyper_dev(self, device):
        """Device-side copy of (lr, momentum, weight_decay). The native kernels read their hyper-parameters from
        this block, so a learning-rate schedule (``param_groups[0]["lr"] = ...``, any torch scheduler) changes
        three floats on the device and a captured CUDA graph of the step stays valid (utils/graph.py)."""
        vals = self.hyper()
        ent = self._hyper_dev.get(device)
        if ent is None:
            ent = [torch.tensor(vals, dtype=torch.float32, device=device), vals]
            self._hyper_dev[device] = ent
        elif ent[1] != vals:
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError("optimizer hyper-parameters changed inside a CUDA graph capture")
            ent[0].copy_(torch.tensor(vals, dtype=torch.float32))
            ent[1] = vals
        return ent[0]

    def has_momentum_state(self, st):
        """True once momentum buffers exist for this store's parameters (after the first step or a loaded
        state_dict); the kernels then compute ``mu*buf + g`` instead of initialising ``buf = g``."""
        if self.param_groups[0]["momentum"] == 0:
            return False
        return any(self.state[p].get("momentum_buffer") is not None for p in st.params)

    def prepare_fused(self, st):
        """Called by the DDP reducer before it launches fused reduce+update kernels for this iteration.
        Returns (lr, momentum, weight_decay, first_step, device hyper-parameter block)."""
        st.ensure_momentum()
        has = self.has_momentum_state(st)
        if has:
            self._bind_state(st)
        lr, mu, wd = self.hyper()
        return lr, mu, wd, (not has), self.hyper_dev(st.device)

    def _bind_state(self, st):
        """Expose the flat momentum buffer through torch's per-parameter state (state_dict parity)."""
        if self.param_groups[0]["momentum"] == 0:
            return
        for p in st.params:
            s = self.state[p]
            if "momentum_buffer" not in s or s["momentum_buffer"] is None or \
                    s["momentum_buffer"].data_ptr() != st.mom_view(p).data_ptr():
                old = s.get("momentum_buffer")
                mv = st.mom_view(p)
                if old is not None:
                    mv.copy_(old)
                s["momentum_buffer"] = mv

    def note_bucket_updated(self, bucket):
        self._fused_done.add(id(bucket))

    def zero_grad(self, set_to_none: bool = True):
        super().zero_grad(set_to_none=set_to_none)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if not self._native_ok():
            out = super().step()
            self._steps += 1
            return loss if loss is not None else out
        lr, mu, wd = self.hyper()
        for st in self._stores:
            st.ensure_momentum()
            has = self.has_momentum_state(st)
            if has:
                self._bind_state(st)  # adopt buffers that came from load_state_dict
            first = not has
            hyper = self.hyper_dev(st.device)
            for b in st.buckets:
                if id(b) in self._fused_done:
                    continue  # already applied by the all-reduce epilogue during backward
                if any(p.grad is None for p in b.params):
                    self._step_bucket_sparse(st, b, lr, mu, wd, first)
                    continue
                native.sgd_step(b.p32, b.g32, b.mom, b.shadow, lr, mu, wd, 1.0, first, hyper)
            st.mark_params_updated_natively()
            self._bind_state(st)
        self._fused_done.clear()
        self._steps += 1
        return loss

    def _step_bucket_sparse(self, st, b, lr, mu, wd, first):
        # Some parameter of the bucket got no gradient: update the others one by one (torch skips grad=None).
        hyper = self.hyper_dev(st.device)
        for p, off in zip(b.params, b.offsets):
            if p.grad is None:
                continue
            n = p.numel()
            native.sgd_step(b.p32[off:off + n], b.g32[off:off + n], b.mom[off:off + n], b.shadow[off:off + n], lr,
                            mu, wd, 1.0, first, hyper)
