"""Whole-step CUDA graph capture.

The reference's hot loop (reference resnet/main.py:118-124) launches every kernel from Python each step; for its
own workload (ResNet-18 on 32x32 images) a step is a couple of hundred tiny kernels and the GPU sits idle between
launches.  ``GraphedTrainStep`` captures ``zero_grad -> forward -> loss -> backward (with the bucketed P2P
all-reduces and their fused SGD epilogues on the communication stream) -> optimizer.step`` once per input shape
and replays it with a single launch.  Everything the step needs is device-side state (BN counters, barrier
epochs, momentum), so replays need no host bookkeeping.

    step = GraphedTrainStep(ddp_model, criterion, optimizer)
    for x, y in loader:
        loss = step(x, y)          # loss: device scalar, valid until the next call

Hyper-parameters: the native SGD kernels (stand-alone and as the all-reduce epilogue) read learning rate, momentum
and weight decay from a three-float device block owned by the optimizer (``optim.SGD.hyper_dev``), refreshed before
every replay, so per-step learning-rate schedules cost one tiny host->device copy and never a re-capture.  Only a
stock optimizer (values baked into its kernels' arguments) makes the step re-capture itself on a change.
"""
import warnings

import torch

from ..ops import _lib


class GraphedTrainStep:
    def __init__(self, model, criterion, optimizer, warmup=3, enabled=True):
        self.model, self.criterion, self.optimizer = model, criterion, optimizer
        self.warmup = int(warmup)
        self.enabled = bool(enabled) and torch.cuda.is_available()
        self._graphs = {}   # (shape, dtype, label shape) -> dict(graph, x, y, loss, hyper)
        self._eager_steps = 0
        self.captures = 0

    # ------------------------------------------------------------------ eager step (also the capture body)
    def _eager(self, x, y):
        self.optimizer.zero_grad(set_to_none=True)
        loss = self.criterion(self.model(x), y)
        loss.backward()
        self.optimizer.step()
        # detached: a caller that keeps the loss around must not keep this step's autograd graph (and with it the
        # per-parameter AccumulateGrad nodes, which remember the stream they were created on) alive into a capture
        return loss.detach()

    def _hyper(self):
        g = self.optimizer.param_groups[0]
        return (g["lr"], g.get("momentum", 0.0), g.get("weight_decay", 0.0))

    def _capture(self, x, y):
        ddp = self.model
        old_watchdog = getattr(ddp, "_watchdog", False)
        if hasattr(ddp, "_watchdog"):
            ddp._watchdog = False            # the watchdog's event cannot be recorded inside a capture
        sx, sy = torch.empty_like(x), torch.empty_like(y)
        sx.copy_(x)
        sy.copy_(y)
        # input pointer slot: the stem kernel of our ResNets reads the batch through a device-side pointer, so a
        # replay consumes the caller's tensor in place (e.g. the buffer the DevicePrefetcher's host->device copy
        # landed in) instead of a copy in a static buffer - 154 MB less traffic per step for ImageNet-shaped fp32
        root = getattr(ddp, "module", ddp)
        slot = None
        if hasattr(root, "_b2_input_slot") and x.is_contiguous() and x.dtype in (torch.float32, torch.bfloat16):
            from ..ops import native
            slot = torch.zeros(1, dtype=torch.int64, device=x.device)
            native.set_ptr_slot(slot, sx)
            root._b2_input_slot = slot
            root._b2_input_slot_used = False
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        n0 = _lib.launches()
        try:
            self.optimizer.zero_grad(set_to_none=True)
            # thread_local: DataLoader pin-memory threads keep calling cudaHostAlloc while we capture
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                loss = self.criterion(self.model(sx), sy)
                loss.backward()
                self.optimizer.step()
        finally:
            if slot is not None:
                root._b2_input_slot = None
                if not getattr(root, "_b2_input_slot_used", False):
                    slot = None              # this input shape went through a path that reads the tensor directly
            if hasattr(ddp, "_watchdog"):
                ddp._watchdog = old_watchdog
        self.captures += 1
        return {"graph": g, "x": sx, "y": sy, "loss": loss.detach(), "hyper": self._hyper(),
                "launches": _lib.launches() - n0, "slot": slot}

    def __call__(self, x, y):
        if not self.enabled or not x.is_cuda:
            return self._eager(x, y)
        if self._eager_steps < self.warmup:
            # first steps run eagerly: lazy initialisation (engine, momentum buffers, "first step" kernels) must
            # not be baked into the graph
            self._eager_steps += 1
            return self._eager(x, y)
        key = (tuple(x.shape), x.dtype, tuple(y.shape), bool(self.model.training))
        entry = self._graphs.get(key)
        dev_hyper = getattr(self.optimizer, "hyper_dev", None) is not None and self.optimizer._native_ok()
        if dev_hyper:
            self.optimizer.hyper_dev(x.device)   # refresh the device-side {lr, momentum, weight_decay} block
        elif entry is not None and entry["hyper"] != self._hyper():
            entry = None                       # stock optimizer: values are baked into the capture -> re-capture
        if entry is None:
            try:
                entry = self._capture(x, y)
            except Exception as e:  # noqa: BLE001
                import traceback
                warnings.warn("CUDA graph capture of the training step failed (%s: %s); running eagerly\n%s" %
                              (type(e).__name__, str(e).splitlines()[0] if str(e) else "", traceback.format_exc()))
                self.enabled = False
                torch.cuda.synchronize()
                return self._eager(x, y)
            self._graphs[key] = entry
            # the capture itself does not execute anything: fall through to the first replay
        slot = entry.get("slot")
        if slot is not None and x.is_contiguous() and x.dtype == entry["x"].dtype:
            from ..ops import native
            native.set_ptr_slot(slot, x)            # the replay reads this batch where it is
            self._keep = x                          # until the next call: the replay's first kernel reads it
        else:
            entry["x"].copy_(x, non_blocking=True)
            if slot is not None:
                from ..ops import native
                native.set_ptr_slot(slot, entry["x"])
        entry["y"].copy_(y, non_blocking=True)
        entry["graph"].replay()
        _lib.count_launch(entry["launches"])   # kernels of ours that one replay runs
        return entry["loss"]
