"""Convolution on the tcgen05 implicit-GEMM kernels (csrc/conv_igemm.cu) with a torch oracle path.

Stands in for ``aten::convolution`` / ``convolution_backward`` that the reference reaches through
torchvision's ResNet (reference resnet/main.py:76,121,123).  Activations are bf16 NHWC (torch
``channels_last``), weights bf16 OHWI (``channels_last`` of the OIHW parameter), weight gradients fp32 OHWI.
"""
import ctypes
import os

import torch
import torch.nn.functional as F

from . import _lib

_WS = {}
_WS_BYTES = 256 << 20
_WGRAD_RED = os.environ.get("B200DDP_WGRAD_RED", "1") != "0"


def workspace(device):
    """Per-device split-K scratch for wgrad (allocated once, reused by every layer)."""
    ws = _WS.get(device)
    if ws is None:
        ws = torch.empty(_WS_BYTES // 4, dtype=torch.float32, device=device)
        _WS[device] = ws
    return ws


def _geom(x, w, stride, pad):
    N, Cin, H, W = x.shape
    Cout, _, KH, KW = w.shape
    return N, H, W, Cin, Cout, KH, KW, stride, pad


def _is_cl(t):
    return t.is_contiguous(memory_format=torch.channels_last)


def native_ok(x, w):
    return (_lib.use_native(x) and x.dtype == torch.bfloat16 and x.shape[1] % 64 == 0 and w.shape[0] % 8 == 0)


def gemm_nt(a, w, bias=None):
    """a [M,K] bf16 @ w[N,K]^T (+bias) -> [M,N] bf16 on the fprop kernel (1x1 conv view). K % 64 == 0, N % 8 == 0."""
    M, K = a.shape
    N = w.shape[0]
    y = torch.empty((M, N), dtype=torch.bfloat16, device=a.device)
    e = _lib.lib().b2_conv_fprop(_lib.ptr(a), _lib.ptr(w), _lib.ptr(bias), _lib.ptr(y), M, 1, 1, K, N, 1, 1, 1, 0,
                                 _lib.stream_ptr())
    _lib.check(e, "b2_conv_fprop(gemm_nt)")
    _lib.count_launch()
    return y


def gemm_nn(dy, w):
    """dy [M,N] bf16 @ w[N,K] -> [M,K] bf16 on the dgrad kernel. K % 64 == 0, N % 8 == 0."""
    M, N = dy.shape
    K = w.shape[1]
    dx = torch.empty((M, K), dtype=torch.bfloat16, device=dy.device)
    e = _lib.lib().b2_conv_dgrad(_lib.ptr(dy), _lib.ptr(w), _lib.ptr(dx), M, 1, 1, K, N, 1, 1, 1, 0,
                                 _lib.stream_ptr())
    _lib.check(e, "b2_conv_dgrad(gemm_nn)")
    _lib.count_launch()
    return dx


def _wgrad_flags(accumulate):
    """(flags for b2_conv_wgrad, launches it makes). Split-K sums normally form in L2 (fp32 reductions straight into
    the gradient, order not fixed: the last bits can differ from run to run, as with cuDNN's default wgrad).
    ``torch.use_deterministic_algorithms(True)`` or ``B200DDP_WGRAD_RED=0`` selects fp32 slabs + a fixed-order reduce
    kernel instead."""
    fixed_order = torch.are_deterministic_algorithms_enabled() or not _WGRAD_RED
    return int(bool(accumulate)) | (2 if fixed_order else 0), (2 if fixed_order else 1)


def gemm_tn(dy, a, out, accumulate=False):
    """out[N,K] fp32 (+)= dy[M,N]^T @ a[M,K] on the wgrad kernel (split-K over M). K % 64 == 0, N % 8 == 0."""
    M, N = dy.shape
    K = a.shape[1]
    ws = workspace(a.device)
    flags, nlaunch = _wgrad_flags(accumulate)
    e = _lib.lib().b2_conv_wgrad(_lib.ptr(a), _lib.ptr(dy), _lib.ptr(out), _lib.ptr(ws), _lib.c_ll(ws.numel() * 4),
                                 M, 1, 1, K, N, 1, 1, 1, 0, flags, _lib.stream_ptr())
    _lib.check(e, "b2_conv_wgrad(gemm_tn)")
    _lib.count_launch(nlaunch)
    return out


def conv_fprop(x, w, stride, pad, bias=None):
    """x: [N,Cin,H,W] bf16 channels_last; w: [Cout,Cin,KH,KW] bf16 channels_last -> y bf16 channels_last.
    ``bias`` (fp32 [Cout], optional) is added in the epilogue (classifier)."""
    N, H, W, Cin, Cout, KH, KW, s, p = _geom(x, w, stride, pad)
    Ho, Wo = (H + 2 * p - KH) // s + 1, (W + 2 * p - KW) // s + 1
    assert _is_cl(x) and _is_cl(w), "conv_fprop expects channels_last operands"
    y = torch.empty((N, Cout, Ho, Wo), dtype=torch.bfloat16, device=x.device, memory_format=torch.channels_last)
    e = _lib.lib().b2_conv_fprop(_lib.ptr(x), _lib.ptr(w), _lib.ptr(bias), _lib.ptr(y), N, H, W, Cin, Cout, KH, KW, s, p,
                                 _lib.stream_ptr())
    _lib.check(e, "b2_conv_fprop")
    _lib.count_launch()
    return y


def conv_fprop_stats(x, w, stride, pad, partial):
    """conv_fprop whose epilogue also accumulates the BatchNorm statistics of y into ``partial``.
    Returns (y, blocks): ``blocks`` rows of [2][Cout] partial sums were written (0: not produced by this variant)."""
    N, H, W, Cin, Cout, KH, KW, s, p = _geom(x, w, stride, pad)
    Ho, Wo = (H + 2 * p - KH) // s + 1, (W + 2 * p - KW) // s + 1
    assert _is_cl(x) and _is_cl(w), "conv_fprop expects channels_last operands"
    y = torch.empty((N, Cout, Ho, Wo), dtype=torch.bfloat16, device=x.device, memory_format=torch.channels_last)
    blocks = ctypes.c_int(0)
    e = _lib.lib().b2_conv_fprop_stats(_lib.ptr(x), _lib.ptr(w), _lib.ptr(y), N, H, W, Cin, Cout, KH, KW, s, p,
                                       _lib.ptr(partial), ctypes.byref(blocks), _lib.stream_ptr())
    _lib.check(e, "b2_conv_fprop_stats")
    _lib.count_launch()
    return y, blocks.value


def gemm_nt_stats(a, w, partial):
    """gemm_nt (no bias) + per-column statistics partials of the bf16 result (stem convolution)."""
    M, K = a.shape
    N = w.shape[0]
    y = torch.empty((M, N), dtype=torch.bfloat16, device=a.device)
    blocks = ctypes.c_int(0)
    e = _lib.lib().b2_conv_fprop_stats(_lib.ptr(a), _lib.ptr(w), _lib.ptr(y), M, 1, 1, K, N, 1, 1, 1, 0,
                                       _lib.ptr(partial), ctypes.byref(blocks), _lib.stream_ptr())
    _lib.check(e, "b2_conv_fprop_stats(gemm_nt)")
    _lib.count_launch()
    return y, blocks.value


def conv_dgrad_fused(dy, w, x_shape, stride, pad, add, mask, bn_y, bn_mean, partial):
    """Residual fan-in fused into dgrad: dx = mask ? (dgrad + add) : 0, finished by the kernel's statistics warps on
    the output panel in shared memory, which also leave the BatchNorm-backward sums of dx (against bn_y / bn_mean) in
    ``partial`` (bn_y None: fan-in only, blocks = 0). Returns (dx, blocks), or (None, 0) when the active kernel
    variant cannot do it."""
    N, Cin, H, W = x_shape
    Cout, _, KH, KW = w.shape
    assert _is_cl(dy) and _is_cl(w) and _is_cl(add) and tuple(add.shape) == tuple(x_shape)
    assert bn_y is None or (_is_cl(bn_y) and tuple(bn_y.shape) == tuple(x_shape))
    dx = torch.empty((N, Cin, H, W), dtype=torch.bfloat16, device=dy.device, memory_format=torch.channels_last)
    blocks = ctypes.c_int(0)
    e = _lib.lib().b2_conv_dgrad_fused(_lib.ptr(dy), _lib.ptr(w), _lib.ptr(dx), N, H, W, Cin, Cout, KH, KW, stride,
                                       pad, _lib.ptr(add), _lib.ptr(mask), _lib.ptr(bn_y), _lib.ptr(bn_mean),
                                       _lib.ptr(partial), ctypes.byref(blocks), _lib.stream_ptr())
    if e == -1:
        return None, 0
    _lib.check(e, "b2_conv_dgrad_fused")
    _lib.count_launch()
    return dx, blocks.value


def conv_dgrad_bnstats(dy, w, x_shape, stride, pad, bn_y, bn_mask, bn_mean, partial):
    """conv_dgrad (cuDNN dgrad in the reference's loss.backward(), resnet/main.py:123) whose statistics warps also
    produce the reduce pass of the BatchNorm backward that consumes dx
    (dx = gradient of relu(bn(bn_y))). Returns (dx, blocks); blocks == 0: not produced, run the normal bn_bwd."""
    N, Cin, H, W = x_shape
    Cout, _, KH, KW = w.shape
    assert _is_cl(dy) and _is_cl(w) and _is_cl(bn_y) and tuple(bn_y.shape) == tuple(x_shape)
    dx = torch.empty((N, Cin, H, W), dtype=torch.bfloat16, device=dy.device, memory_format=torch.channels_last)
    blocks = ctypes.c_int(0)
    e = _lib.lib().b2_conv_dgrad_bnstats(_lib.ptr(dy), _lib.ptr(w), _lib.ptr(dx), N, H, W, Cin, Cout, KH, KW, stride,
                                         pad, _lib.ptr(bn_y), _lib.ptr(bn_mask), _lib.ptr(bn_mean), _lib.ptr(partial),
                                         ctypes.byref(blocks), _lib.stream_ptr())
    _lib.check(e, "b2_conv_dgrad_bnstats")
    _lib.count_launch()
    return dx, blocks.value


def conv_dgrad(dy, w, x_shape, stride, pad):
    N, Cin, H, W = x_shape
    Cout, _, KH, KW = w.shape
    assert _is_cl(dy) and _is_cl(w)
    dx = torch.empty((N, Cin, H, W), dtype=torch.bfloat16, device=dy.device, memory_format=torch.channels_last)
    e = _lib.lib().b2_conv_dgrad(_lib.ptr(dy), _lib.ptr(w), _lib.ptr(dx), N, H, W, Cin, Cout, KH, KW, stride, pad,
                                 _lib.stream_ptr())
    _lib.check(e, "b2_conv_dgrad")
    _lib.count_launch()
    return dx


def conv_wgrad(x, dy, w_shape, stride, pad, out=None, accumulate=False):
    """Returns (or fills ``out``, e.g. a gradient-bucket slot) the fp32 OHWI weight gradient."""
    N, Cin, H, W = x.shape
    Cout, _, KH, KW = w_shape
    assert _is_cl(x) and _is_cl(dy)
    if out is None:
        out = torch.empty(w_shape, dtype=torch.float32, device=x.device).contiguous(
            memory_format=torch.channels_last)
        accumulate = False
    ws = workspace(x.device)
    flags, nlaunch = _wgrad_flags(accumulate)
    e = _lib.lib().b2_conv_wgrad(_lib.ptr(x), _lib.ptr(dy), _lib.ptr(out), _lib.ptr(ws),
                                 _lib.c_ll(ws.numel() * 4), N, H, W, Cin, Cout, KH, KW, stride, pad,
                                 flags, _lib.stream_ptr())
    _lib.check(e, "b2_conv_wgrad")
    _lib.count_launch(nlaunch)
    return out


# ------------------------------------------------------------------------------------------ torch oracle
class _true_fp32:
    """cuDNN convolutions of fp32 tensors run as TF32 by default (10-bit mantissa - about what bf16 inputs carry);
    an oracle has to be real fp32, so TF32 is switched off around every reference call."""

    def __enter__(self):
        self.old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False

    def __exit__(self, *exc):
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = self.old


def ref_fprop(x, w, stride, pad):
    with _true_fp32():
        return F.conv2d(x, w, None, stride, pad)


def ref_dgrad(dy, w, x_shape, stride, pad):
    with _true_fp32():
        return torch.nn.grad.conv2d_input(x_shape, w, dy, stride, pad)


def ref_wgrad(x, dy, w_shape, stride, pad):
    with _true_fp32():
        return torch.nn.grad.conv2d_weight(x, w_shape, dy, stride, pad)
