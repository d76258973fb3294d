"""Thin tensor-level wrappers over the C ABI (include/b200ddp.h) for everything except convolution
(see ops/conv.py).  All tensors are CUDA; activations are bf16 ``channels_last`` (NHWC in memory).

Each wrapper names the aten op(s) of the reference's dependency stack it stands in for; the reference
itself only reaches them through torchvision / nn / optim one-liners (reference resnet/main.py:76,102,103).
"""
import ctypes

import torch

from . import _lib
from ._lib import c_ll, c_f, c_p, ptr, stream_ptr, check, count_launch

_scratch = {}


def scratch(device, key, numel, dtype=torch.float32):
    """Persistent per-device scratch buffers (no per-step allocation; CUDA-graph friendly)."""
    k = (device, key)
    t = _scratch.get(k)
    if t is None or t.numel() < numel or t.dtype != dtype:
        t = torch.empty(numel, dtype=dtype, device=device)
        _scratch[k] = t
    return t


def bn_partial(device):
    return scratch(device, "bn_partial", 1184 * 2 * 2048)   # >= kMaxPartialBlocks * 2 * C_max


class XgpuCtx(ctypes.Structure):
    """Mirror of ``b2_xgpu_ctx``."""
    _fields_ = [("rank", ctypes.c_int), ("world", ctypes.c_int), ("data", ctypes.c_void_p * 8),
                ("flags", ctypes.c_void_p * 8), ("epoch", ctypes.c_void_p)]


def _xg(xg):
    return ctypes.byref(xg) if xg is not None else None


def _mc(t):
    """[N,C,H,W] channels_last -> (M, C)."""
    n, c, h, w = t.shape
    return n * h * w, c


# ------------------------------------------------------------------------------------------------ batch norm
def bn_fwd_train(y, res, gamma, beta, running_mean, running_var, nbt, momentum, eps, relu, mean, invstd,
                 scale_shift, xg=None, xg_slot=0, mask=None, partial_blocks=0, apply=True):
    """aten::native_batch_norm (training) + relu_ + add_ fused. Returns z. ``mask`` (uint8 [M*C/8], optional)
    receives one bit per element where the ReLU passed. ``apply=False``: statistics, running statistics and the
    scale/shift vector only (returns None) - the consumer applies them itself (``bn_relu_maxpool_fwd``)."""
    M, C = _mc(y)
    z = torch.empty_like(y) if apply else None
    if partial_blocks > 0:   # statistics partials already produced by the convolution's epilogue
        e = _lib.lib().b2_bn_fwd_train_partial(ptr(y), ptr(res), ptr(z), ptr(mask), c_ll(M), C, ptr(gamma), ptr(beta),
                                               ptr(running_mean), ptr(running_var), ptr(nbt), c_f(momentum),
                                               c_f(eps), int(relu), ptr(mean), ptr(invstd), ptr(scale_shift),
                                               ptr(bn_partial(y.device)), int(partial_blocks), _xg(xg), xg_slot,
                                               stream_ptr())
        check(e, "b2_bn_fwd_train_partial")
        count_launch(2 if apply else 1)
        return z
    e = _lib.lib().b2_bn_fwd_train(ptr(y), ptr(res), ptr(z), ptr(mask), c_ll(M), C, ptr(gamma), ptr(beta), ptr(running_mean),
                                   ptr(running_var), ptr(nbt), c_f(momentum), c_f(eps), int(relu), ptr(mean),
                                   ptr(invstd), ptr(scale_shift), ptr(bn_partial(y.device)), _xg(xg), xg_slot,
                                   stream_ptr())
    check(e, "b2_bn_fwd_train")
    count_launch(3 if apply else 2)
    return z


def bn_fwd_eval(y, res, gamma, beta, running_mean, running_var, eps, relu, scale_shift, apply=True):
    M, C = _mc(y)
    z = torch.empty_like(y) if apply else None
    e = _lib.lib().b2_bn_fwd_eval(ptr(y), ptr(res), ptr(z), c_ll(M), C, ptr(gamma), ptr(beta), ptr(running_mean),
                                  ptr(running_var), c_f(eps), int(relu), ptr(scale_shift), stream_ptr())
    check(e, "b2_bn_fwd_eval")
    count_launch(2 if apply else 1)
    return z


def bn_partial_b(device):
    """Second partial-sum scratch: holds the BatchNorm-backward sums a dgrad produced for the NEXT unit while other
    BatchNorm kernels (e.g. the downsample branch) use bn_partial in between."""
    return scratch(device, "bn_partial_b", 592 * 2 * 2048)


def bn_bwd_partial(dz, y, gamma, mean, invstd, dgamma, dbeta, accumulate, relu, coef, partial, partial_blocks,
                   xg=None, xg_slot=0, mask=None):
    """bn_bwd for a single incoming gradient whose reduce pass was done by the dgrad that produced it
    (native_batch_norm_backward inside loss.backward(), reference resnet/main.py:123)."""
    M, C = _mc(y)
    dy = torch.empty_like(y)
    e = _lib.lib().b2_bn_bwd_partial(ptr(dz), None, ptr(mask), ptr(y), ptr(dy), c_ll(M), C, ptr(gamma), ptr(mean),
                                     ptr(invstd), ptr(dgamma), ptr(dbeta), int(accumulate), int(relu), ptr(coef),
                                     ptr(partial), int(partial_blocks), _xg(xg), xg_slot, stream_ptr())
    check(e, "b2_bn_bwd_partial")
    count_launch(2)
    return dy


def bn_bwd(dz_a, dz_b, z, y, gamma, mean, invstd, dgamma, dbeta, accumulate, relu, coef, want_dres, xg=None,
           xg_slot=0, mask=None, dz_b_half_res=False):
    """aten::native_batch_norm_backward + threshold_backward + the gradient fan-in add, fused.
    Returns (dy, dres) where dres (optional) is the masked incoming gradient for the residual branch.
    ``dz_b_half_res``: dz_b is the compact [N, C, ceil(H/2), ceil(W/2)] input gradient of a 1x1 / stride 2
    convolution (non-zero at the even pixels only)."""
    M, C = _mc(y)
    dy = torch.empty_like(y)
    dres = torch.empty_like(y) if want_dres else None
    if dz_b_half_res:
        assert mask is not None and want_dres
        e = _lib.lib().b2_bn_bwd_q(ptr(dz_a), ptr(dz_b), int(y.shape[2]), int(y.shape[3]), ptr(mask), ptr(y), ptr(dy),
                                   ptr(dres), c_ll(M), C, ptr(gamma), ptr(mean), ptr(invstd), ptr(dgamma), ptr(dbeta),
                                   int(accumulate), int(relu), ptr(coef), ptr(bn_partial(y.device)), _xg(xg), xg_slot,
                                   stream_ptr())
        check(e, "b2_bn_bwd_q")
        count_launch(3)
        return dy, dres
    e = _lib.lib().b2_bn_bwd(ptr(dz_a), ptr(dz_b), ptr(z), ptr(mask), ptr(y), ptr(dy), ptr(dres), c_ll(M), C, ptr(gamma),
                             ptr(mean), ptr(invstd), ptr(dgamma), ptr(dbeta), int(accumulate), int(relu), ptr(coef),
                             ptr(bn_partial(y.device)), _xg(xg), xg_slot, stream_ptr())
    check(e, "b2_bn_bwd")
    count_launch(3)
    return dy, dres


# ------------------------------------------------------------------------------------------------ pooling
def maxpool_fwd(x, with_index=True):
    """aten::max_pool2d_with_indices 3x3/2/1. Returns (y, idx) where idx holds the winning tap (u8) per element."""
    n, c, h, w = x.shape
    ho, wo = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    y = torch.empty((n, c, ho, wo), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
    idx = torch.empty((n, ho, wo, c), dtype=torch.uint8, device=x.device) if with_index else None
    check(_lib.lib().b2_maxpool_fwd(ptr(x), ptr(y), ptr(idx), n, h, w, c, stream_ptr()), "b2_maxpool_fwd")
    count_launch()
    return y, idx


def bn_relu_maxpool_fwd(x, scale_shift, with_index=True):
    """maxpool(relu(x * scale + shift)) in one pass over the raw convolution output (the normalised full-resolution
    tensor and its ReLU mask never exist). ``scale_shift``: fp32 [2C] as left by ``bn_fwd_train(apply=False)``.
    ``maxpool_bwd`` of the returned index yields the gradient of the BatchNorm OUTPUT (ReLU backward included)."""
    n, c, h, w = x.shape
    ho, wo = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    y = torch.empty((n, c, ho, wo), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
    idx = torch.empty((n, ho, wo, c), dtype=torch.uint8, device=x.device) if with_index else None
    check(_lib.lib().b2_bn_relu_maxpool_fwd(ptr(x), ptr(scale_shift), ptr(scale_shift[c:]), ptr(y), ptr(idx), n, h, w,
                                            c, stream_ptr()), "b2_bn_relu_maxpool_fwd")
    count_launch()
    return y, idx


def maxpool_bwd(idx, x_shape, dy_a, dy_b=None):
    n, c, h, w = x_shape
    dx = torch.empty(x_shape, dtype=dy_a.dtype, device=dy_a.device, memory_format=torch.channels_last)
    check(_lib.lib().b2_maxpool_bwd(ptr(idx), ptr(dy_a), ptr(dy_b), ptr(dx), n, h, w, c, stream_ptr()),
          "b2_maxpool_bwd")
    count_launch()
    return dx


def avgpool_fwd(x):
    """aten::mean.dim over H,W -> [N, C] bf16."""
    n, c, h, w = x.shape
    y = torch.empty((n, c), dtype=x.dtype, device=x.device)
    check(_lib.lib().b2_avgpool_fwd(ptr(x), ptr(y), n, h * w, c, stream_ptr()), "b2_avgpool_fwd")
    count_launch()
    return y


def avgpool_bwd(dy, shape):
    n, c, h, w = shape
    dx = torch.empty(shape, dtype=dy.dtype, device=dy.device, memory_format=torch.channels_last)
    check(_lib.lib().b2_avgpool_bwd(ptr(dy), ptr(dx), n, h * w, c, stream_ptr()), "b2_avgpool_bwd")
    count_launch()
    return dx


# ------------------------------------------------------------------------------------------------ loss
def ce_fwd(logits, labels):
    """aten::_log_softmax + nll_loss_forward (mean). logits bf16 [rows, ncls]. Returns (loss fp32 scalar, lse)."""
    rows, ncls = logits.shape
    lse = torch.empty(rows, dtype=torch.float32, device=logits.device)
    loss_row = torch.empty(rows, dtype=torch.float32, device=logits.device)
    loss = torch.empty((), dtype=torch.float32, device=logits.device)
    check(_lib.lib().b2_ce_fwd(ptr(logits), ptr(labels), ptr(lse), ptr(loss_row), ptr(loss), rows, ncls,
                               logits.stride(0), stream_ptr()), "b2_ce_fwd")
    count_launch(2)
    return loss, lse


def ce_bwd(logits, labels, lse, gscale, dbias=None, accumulate_bias=False):
    rows, ncls = logits.shape
    dlogits = torch.empty_like(logits)
    check(_lib.lib().b2_ce_bwd(ptr(logits), ptr(labels), ptr(lse), ptr(gscale), ptr(dlogits), ptr(dbias),
                               int(accumulate_bias), rows, ncls, logits.stride(0), stream_ptr()), "b2_ce_bwd")
    count_launch(2 if dbias is not None else 1)
    return dlogits


# ------------------------------------------------------------------------------------------------ fused head
def head_supported(c, ncls):
    return bool(_lib.lib().b2_head_supported(int(c), int(ncls)))


def head_fwd(x, w16, bias):
    """aten::mean (global average pool) + addmm (fc) in one kernel (csrc/head.cu): x [N,C,H,W] channels_last bf16,
    w16 [ncls,C] bf16, bias fp32 -> (logits [N,ncls] fp32, pooled [N,C] bf16 kept for the backward pass)."""
    n, c, h, w = x.shape
    ncls = w16.shape[0]
    logits = torch.empty((n, ncls), dtype=torch.float32, device=x.device)
    pooled = torch.empty((n, c), dtype=torch.bfloat16, device=x.device)
    check(_lib.lib().b2_head_fwd(ptr(x), ptr(w16), ptr(bias), ptr(pooled), ptr(logits), n, h * w, c, ncls, ncls,
                                 stream_ptr()), "b2_head_fwd")
    count_launch()
    return logits, pooled


_zero_scratch = {}


def zero_scratch(device, key, numel, dtype):
    """Persistent scratch that starts zeroed (and that its kernels leave zeroed)."""
    k = (device, key)
    t = _zero_scratch.get(k)
    if t is None or t.numel() < numel:
        t = _zero_scratch[k] = torch.zeros(numel, dtype=dtype, device=device)
    return t


def ce_fwd_f32(logits, labels):
    """aten::_log_softmax + nll_loss_forward (mean) on fp32 logits, one launch. Returns (loss scalar, lse)."""
    rows, ncls = logits.shape
    lse = torch.empty(rows, dtype=torch.float32, device=logits.device)
    loss_row = torch.empty(rows, dtype=torch.float32, device=logits.device)
    loss = torch.empty((), dtype=torch.float32, device=logits.device)
    counter = zero_scratch(logits.device, "ce_counter", 1, torch.int32)
    check(_lib.lib().b2_ce_fwd_f32(ptr(logits), ptr(labels), ptr(lse), ptr(loss_row), ptr(loss), ptr(counter), rows,
                                   ncls, logits.stride(0), stream_ptr()), "b2_ce_fwd_f32")
    count_launch()
    return loss, lse


def ce_bwd_fused(logits, labels, lse, gscale, dlogits16, dbias=None, accumulate_bias=False):
    """_log_softmax_backward + nll_loss_backward: bf16 dlogits and (optionally) the classifier's bias gradient in
    one launch."""
    rows, ncls = logits.shape
    check(_lib.lib().b2_ce_bwd_fused(ptr(logits), ptr(labels), ptr(lse), ptr(gscale), ptr(dlogits16), ptr(dbias),
                                     int(accumulate_bias), rows, ncls, logits.stride(0), dlogits16.stride(0),
                                     stream_ptr()), "b2_ce_bwd_fused")
    count_launch()
    return dlogits16


def colsum_bf16(m, out, accumulate=False):
    rows, cols = m.shape
    check(_lib.lib().b2_colsum_bf16(ptr(m), ptr(out), rows, cols, m.stride(0), int(accumulate), stream_ptr()),
          "b2_colsum_bf16")
    count_launch()
    return out


# ------------------------------------------------------------------------------------------------ optimizer
def sgd_step(p, g, buf, shadow, lr, momentum, weight_decay, grad_scale, first_step, hyper=None):
    """torch.optim.SGD (momentum, weight decay, dampening 0) on flat fp32 buffers + bf16 shadow refresh.
    ``hyper``: optional device tensor {lr, momentum, weight_decay} the kernel reads instead of the scalars."""
    n = p.numel()
    check(_lib.lib().b2_sgd_step(ptr(p), ptr(g), ptr(buf), ptr(shadow), c_ll(n), c_f(lr), c_f(momentum),
                                 c_f(weight_decay), c_f(grad_scale), int(first_step), ptr(hyper), stream_ptr()),
          "b2_sgd_step")
    count_launch()


def set_ptr_slot(slot, tensor):
    """``slot[0] = tensor.data_ptr()`` in stream order (one-thread kernel; usable between CUDA-graph replays)."""
    check(_lib.lib().b2_set_ptr_slot(ptr(slot), ptr(tensor), stream_ptr()), "b2_set_ptr_slot")
    count_launch()


def cast_f32_bf16(src, dst):
    check(_lib.lib().b2_cast_f32_bf16(ptr(src), ptr(dst), c_ll(src.numel()), stream_ptr()), "b2_cast_f32_bf16")
    count_launch()


# ------------------------------------------------------------------------------------------------ stem / layout
def stem_im2col(x):
    """[N,3,H,W] fp32/bf16 NCHW -> im2col matrix [N*Ho*Wo, 192] bf16 for the 7x7/2 stem (K padded 147 -> 192)."""
    n, c, h, w = x.shape
    assert c == 3 and x.is_contiguous()
    ho, wo = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    a = torch.empty((n * ho * wo, 192), dtype=torch.bfloat16, device=x.device)
    check(_lib.lib().b2_stem_im2col(ptr(x), int(x.dtype == torch.bfloat16), ptr(a), n, h, w, stream_ptr()),
          "b2_stem_im2col")
    count_launch()
    return a, ho, wo


def stem_pack_w(w_fp32, out):
    check(_lib.lib().b2_stem_pack_w(ptr(w_fp32), ptr(out), w_fp32.shape[0], stream_ptr()), "b2_stem_pack_w")
    count_launch()
    return out


def stem_unpack_dw(dwp, dw, accumulate):
    check(_lib.lib().b2_stem_unpack_dw(ptr(dwp), ptr(dw), dw.shape[0], int(accumulate), stream_ptr()),
          "b2_stem_unpack_dw")
    count_launch()


def stem_supported(h, w):
    return bool(_lib.lib().b2_stem_supported(int(h), int(w)))


def stem_pack_input(x, slot=None):
    """[N,3,H,W] fp32/bf16 NCHW -> zero-padded NHWC4 bf16 image (flat) for the direct stem kernels
    (the 7x7/2 conv1 of the torchvision ResNet the reference builds at resnet/main.py:76).
    ``slot`` (device int64[1], optional): the kernel reads the batch from the address stored there instead of from
    ``x`` (which then only supplies shape and dtype); see ``set_ptr_slot``."""
    n, c, h, w = x.shape
    assert c == 3 and x.is_contiguous()
    lib = _lib.lib()
    lib.b2_stem_padded_elems.restype = ctypes.c_longlong
    xp = torch.empty(lib.b2_stem_padded_elems(n, h, w), dtype=torch.bfloat16, device=x.device)
    if slot is not None:
        check(lib.b2_stem_pack_input_indirect(ptr(slot), int(x.dtype == torch.bfloat16), ptr(xp), n, h, w,
                                              stream_ptr()), "b2_stem_pack_input_indirect")
    else:
        check(lib.b2_stem_pack_input(ptr(x), int(x.dtype == torch.bfloat16), ptr(xp), n, h, w, stream_ptr()),
              "b2_stem_pack_input")
    count_launch()
    return xp


def stem_pack_weight(w_ohwi_fp32, out):
    """fp32 OHWI [64,7,7,3] -> bf16 [64,224] (7 filter rows x 8 columns x 4 channels, zero padded)."""
    check(_lib.lib().b2_stem_pack_weight(ptr(w_ohwi_fp32), ptr(out), stream_ptr()), "b2_stem_pack_weight")
    count_launch()
    return out


def stem_fprop(xp, wp, n, h, w, partial=None):
    """7x7/2 stem convolution from the padded image (forward of reference resnet/main.py:121 for conv1).
    Returns (y [N,64,Ho,Wo] channels_last bf16, stats blocks)."""
    ho, wo = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    y = torch.empty((n, 64, ho, wo), dtype=torch.bfloat16, device=xp.device, memory_format=torch.channels_last)
    blocks = ctypes.c_int(0)
    check(_lib.lib().b2_stem_fprop(ptr(xp), ptr(wp), ptr(y), n, h, w, ptr(partial),
                                   ctypes.byref(blocks) if partial is not None else None, stream_ptr()),
          "b2_stem_fprop")
    count_launch()
    return y, blocks.value


def stem_wgrad(xp, dy, dw, n, h, w, accumulate):
    """dW (fp32 OHWI [64,7,7,3], written or accumulated) of the stem from the padded image and dy (NHWC bf16):
    conv1's share of loss.backward(), reference resnet/main.py:123."""
    lib = _lib.lib()
    lib.b2_stem_wgrad_workspace_floats.restype = ctypes.c_longlong
    ws = scratch(xp.device, "stem_wgrad_ws", lib.b2_stem_wgrad_workspace_floats())
    check(lib.b2_stem_wgrad(ptr(xp), ptr(dy), ptr(dw), ptr(ws), n, h, w, int(accumulate), stream_ptr()),
          "b2_stem_wgrad")
    count_launch(2)


def nchw_to_nhwc_bf16(x):
    n, c, h, w = x.shape
    assert x.is_contiguous()
    y = torch.empty((n, c, h, w), dtype=torch.bfloat16, device=x.device, memory_format=torch.channels_last)
    check(_lib.lib().b2_nchw_to_nhwc_bf16(ptr(x), int(x.dtype == torch.bfloat16), ptr(y), n, c, h * w,
                                          stream_ptr()), "b2_nchw_to_nhwc_bf16")
    count_launch()
    return y
