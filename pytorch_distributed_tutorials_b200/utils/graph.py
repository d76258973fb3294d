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

Hyper-parameters are baked into the captured kernels' arguments; the step re-captures itself when the learning
rate, momentum or weight decay of the optimizer changes (epoch-level schedules cost one capture per change).
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
        old_measure = getattr(ddp, "measure_exposed", False)
        old_watchdog = getattr(ddp, "_watchdog", False)
        if hasattr(ddp, "measure_exposed"):
            ddp.measure_exposed = False      # timing events cannot be recorded inside a capture
            ddp._watchdog = False
        sx, sy = torch.empty_like(x), torch.empty_like(y)
        sx.copy_(x)
        sy.copy_(y)
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
            if hasattr(ddp, "measure_exposed"):
                ddp.measure_exposed = old_measure
                ddp._watchdog = old_watchdog
        self.captures += 1
        return {"graph": g, "x": sx, "y": sy, "loss": loss.detach(), "hyper": self._hyper(),
                "launches": _lib.launches() - n0}

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
        if entry is not None and entry["hyper"] != self._hyper():
            entry = None                       # learning rate (or momentum / weight decay) changed: re-capture
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
        entry["x"].copy_(x, non_blocking=True)
        entry["y"].copy_(y, non_blocking=True)
        entry["graph"].replay()
        _lib.count_launch(entry["launches"])   # kernels of ours that one replay runs
        return entry["loss"]
