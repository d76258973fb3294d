"""Cross-entropy (mean reduction) with a fused native forward/backward.

Drop-in for ``nn.CrossEntropyLoss()`` as the reference uses it (reference resnet/main.py:102,122): same call
shape ``criterion(outputs, labels)``, fp32 scalar result.  On CUDA the log-softmax + NLL (+ the batch mean) run as
one kernel on the fp32 logits the classifier head produced - no cast in between - and the backward is one kernel that
emits bf16 ``dlogits`` together with the classifier's bias gradient.  When the logits come from our ResNet engine
they carry a private link (``_b2_head``): the backward then leaves ``dlogits`` where the engine's classifier-backward
GEMMs read them and hands autograd a zero-stride placeholder, so no fp32 copy of the gradient is ever materialised.
Any other criterion still works on the same logits (the engine then converts the fp32 gradient it receives).
Elsewhere (CPU, foreign dtypes) it defers to ``F.cross_entropy``.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib, native


class _CEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, link):
        lg = logits.detach()
        if lg.dtype != torch.float32 or lg.stride(1) != 1:
            lg = lg.float().contiguous()
        labels = labels.contiguous()
        loss, lse = native.ce_fwd_f32(lg, labels)
        ctx.save_for_backward(lg, labels, lse)
        ctx.in_dtype = logits.dtype
        ctx.link = link
        return loss

    @staticmethod
    def backward(ctx, g):
        lg, labels, lse = ctx.saved_tensors
        g = g.detach().to(torch.float32).contiguous()
        link = ctx.link
        if link is not None and link.accepts(lg):
            dl16 = link.dlogits_buffer(lg.shape)
            dbias, acc = link.bias_slot()
            native.ce_bwd_fused(lg, labels, lse, g, dl16, dbias, acc)
            link.ready = True
            return link.placeholder(lg.shape), None, None
        dl16 = torch.empty(lg.shape, dtype=torch.bfloat16, device=lg.device)
        native.ce_bwd_fused(lg, labels, lse, g, dl16)
        return dl16.to(ctx.in_dtype), None, None


class CrossEntropyLoss(nn.Module):
    def forward(self, outputs, labels):
        if _lib.use_native(outputs) and outputs.dim() == 2 and labels.dtype == torch.int64:
            link = getattr(outputs, "_b2_head", None) if outputs.requires_grad else None
            return _CEFn.apply(outputs, labels, link)
        return F.cross_entropy(outputs, labels)
