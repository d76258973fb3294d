"""Cross-entropy (mean reduction) with a fused native forward/backward.

Drop-in for ``nn.CrossEntropyLoss()`` as the reference uses it (reference resnet/main.py:102,122): same call
shape ``criterion(outputs, labels)``, fp32 scalar result.  On CUDA the log-softmax + NLL run as one kernel per
direction on bf16 logits; elsewhere it defers to ``F.cross_entropy``.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib, native


class _CEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels):
        l16 = logits.detach().to(torch.bfloat16).contiguous()
        labels = labels.contiguous()
        loss, lse = native.ce_fwd(l16, labels)
        ctx.save_for_backward(l16, labels, lse)
        ctx.in_dtype = logits.dtype
        return loss

    @staticmethod
    def backward(ctx, g):
        l16, labels, lse = ctx.saved_tensors
        g = g.detach().to(torch.float32).contiguous()
        d = native.ce_bwd(l16, labels, lse, g)
        return d.to(ctx.in_dtype), None


class CrossEntropyLoss(nn.Module):
    def forward(self, outputs, labels):
        if _lib.use_native(outputs) and outputs.dim() == 2 and labels.dtype == torch.int64:
            return _CEFn.apply(outputs, labels)
        return F.cross_entropy(outputs, labels)
