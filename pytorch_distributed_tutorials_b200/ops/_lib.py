"""ctypes binding of the native library (include/b200ddp.h).

The library is built in-tree by ``build.py`` (``__graft_entry__.build()``).  On a machine with a GPU the
native path is mandatory: a missing ``libb200ddp.so`` raises instead of silently falling back to eager
PyTorch.  On a CPU-only machine the library still loads (host-side runtime: bucket planner, sampler),
and the CUDA entry points are simply never called.
"""
import ctypes
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libb200ddp.so")

_lib = None
_lock = threading.Lock()

c_int, c_ll, c_f, c_p = ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_void_p


class NativeError(RuntimeError):
    pass


def _try_build():
    root = os.path.dirname(os.path.dirname(_HERE))
    script = os.path.join(root, "build.py")
    if not os.path.exists(script):
        return False
    import importlib.util

    spec = importlib.util.spec_from_file_location("_b200_build", script)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.build(quiet=True)
    return os.path.exists(LIB_PATH)


def lib():
    """Return the loaded ctypes library (loading / building it on first use)."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            ok = False
            if os.environ.get("B200DDP_NO_AUTOBUILD", "0") != "1":
                try:
                    ok = _try_build()
                except Exception as e:  # noqa: BLE001
                    raise NativeError("libb200ddp.so is missing and building it failed: %s" % e) from e
            if not ok:
                raise NativeError("libb200ddp.so not found at %s — run `python build.py`" % LIB_PATH)
        _lib = ctypes.CDLL(LIB_PATH)
        # tuning switches of the kernel library that have an environment spelling (A/B runs without code changes)
        for env, setter in (("B200DDP_BN_APPLY_BATCHED", "b2_bn_set_apply_batched"),
                            ("B200DDP_BN_REDUCE_VARIANT", "b2_bn_set_reduce_variant"),
                            ("B200DDP_CONV_HALO", "b2_conv_set_halo"),
                            ("B200DDP_WGRAD_RED", "b2_conv_set_wgrad_red"),
                            ("B200DDP_CONV_PAIRS_AUTO", "b2_conv_set_pairs_auto")):
            v = os.environ.get(env)
            if v is not None and hasattr(_lib, setter):
                getattr(_lib, setter)(int(v))
        return _lib


def available():
    try:
        lib()
        return True
    except (NativeError, OSError):
        return False


def use_native(t=None):
    """Native kernels are used for CUDA tensors unless B200DDP_FORCE_TORCH=1 (oracle / debugging switch)."""
    if os.environ.get("B200DDP_FORCE_TORCH", "0") == "1":
        return False
    if t is not None and not t.is_cuda:
        return False
    return torch.cuda.is_available()


def stream_ptr():
    return c_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    return c_p(t.data_ptr()) if t is not None else c_p(0)


def check(err, what):
    if err != 0:
        raise NativeError("%s failed with CUDA error %d" % (what, err))


_counter = {"launches": 0}


def count_launch(n=1):
    _counter["launches"] += n


def launches():
    return _counter["launches"]


def reset_launches():
    _counter["launches"] = 0
