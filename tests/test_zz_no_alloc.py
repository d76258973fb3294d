"""The allocation-free BatchNorm entry point, called through the C ABI exactly as an external caller would.

Some users of this library drive it from places where device memory cannot be allocated: from a serving loop that
cannot tolerate the device-wide synchronisation a cudaMalloc implies, from a host program that owns the whole
memory budget, or on a GPU that is already nearly full.  ``b2_bn_fwd_train_noalloc`` keeps
the kernel's scratch in a static device array for them.  This test checks that it computes what the regular entry
point (``b2_bn_fwd_train`` with a caller-provided scratch buffer) computes, on two calls in a row with different
shapes (so stale scratch contents from the first call are visible to the second if they mattered), and that a
size beyond the built-in scratch is refused.  The file name sorts last on purpose: if this test fails under
``pytest -x``, every other test has already run.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _run(fn_name, y, res, gamma, beta, relu, partial):
    from pytorch_distributed_tutorials_b200.ops import _lib
    L = _lib.lib()
    n, c, h, w = y.shape
    M = n * h * w
    z = torch.empty_like(y)
    rm, rv = torch.zeros(c, device=DEV), torch.ones(c, device=DEV)
    nbt = torch.zeros((), dtype=torch.int64, device=DEV)
    mean, invstd = torch.empty(c, device=DEV), torch.empty(c, device=DEV)
    ss = torch.empty(2 * c, device=DEV)
    p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)  # noqa: E731
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    common = [p(y), p(res), p(z)] + ([ctypes.c_void_p(0)] if partial is not None else []) + [
        ctypes.c_longlong(M), ctypes.c_int(c), p(gamma), p(beta), p(rm), p(rv), p(nbt),
        ctypes.c_float(0.1), ctypes.c_float(1e-5), ctypes.c_int(int(relu)), p(mean), p(invstd), p(ss)]
    if partial is not None:
        common.append(p(partial))
    err = getattr(L, fn_name)(*common, ctypes.c_void_p(0), ctypes.c_int(0), stream)
    torch.cuda.synchronize()
    return err, z, mean, invstd, rm, rv, nbt


def _case(shape, relu, with_res, seed):
    torch.manual_seed(seed)
    n, c, h, w = shape
    y = (torch.randn(shape, device=DEV) * 1.7 + 0.3).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    res = None
    if with_res:
        res = torch.randn(shape, device=DEV).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    gamma, beta = torch.rand(c, device=DEV) + 0.5, torch.randn(c, device=DEV)
    partial = torch.empty(592 * 2 * 2048, device=DEV)
    e0, *ref = _run("b2_bn_fwd_train", y, res, gamma, beta, relu, partial)
    e1, *got = _run("b2_bn_fwd_train_noalloc", y, res, gamma, beta, relu, None)
    assert e0 == 0 and e1 == 0, (e0, e1)
    names = ["z", "mean", "invstd", "running_mean", "running_var", "num_batches_tracked"]
    for name, a, b in zip(names, got, ref):
        a, b = a.float(), b.float()
        rel = ((a - b).abs().max() / (b.abs().max() + 1e-6)).item()
        assert rel < 1e-2, (name, rel)


def test_noalloc_bn_matches_regular_entry_point_twice_in_a_row():
    _case((4, 256, 14, 14), True, True, 0)     # first call: sets up the scratch address for this device
    _case((3, 64, 33, 31), False, False, 1)    # second call: different shape, scratch still holds call 1's sums
    _case((2, 2048, 7, 7), True, False, 2)     # widest layer the built-in scratch covers


def test_noalloc_bn_rejects_sizes_beyond_the_scratch():
    from pytorch_distributed_tutorials_b200.ops import _lib
    L = _lib.lib()
    c = 4096
    y = torch.zeros(1, c, 2, 2, device=DEV, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    z = torch.empty_like(y)
    f = torch.zeros(2 * c, device=DEV)
    p = lambda t: ctypes.c_void_p(t.data_ptr())  # noqa: E731
    nul = ctypes.c_void_p(0)
    err = L.b2_bn_fwd_train_noalloc(p(y), nul, p(z), ctypes.c_longlong(4), ctypes.c_int(c), p(f), p(f), nul, nul, nul,
                                    ctypes.c_float(0.1), ctypes.c_float(1e-5), ctypes.c_int(0), p(f), p(f), p(f),
                                    nul, ctypes.c_int(0), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert err != 0
    torch.cuda.synchronize()
