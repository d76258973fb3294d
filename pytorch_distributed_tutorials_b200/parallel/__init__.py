"""Process-group set-up, the DDP-style wrapper, sampler and SyncBatchNorm."""
import datetime
import os

import torch
import torch.distributed as dist

from .ddp import DistributedDataParallel  # noqa: F401
from .sampler import DistributedSampler  # noqa: F401
from .syncbn import SyncBatchNorm, convert_sync_batchnorm  # noqa: F401


def init_process_group(backend=None, local_rank=None, timeout_s=1800):
    """One-liner equivalent of the reference's ``torch.distributed.init_process_group(backend="nccl")``
    (reference resnet/main.py:74), env:// rendezvous.

    ``backend``: ``"b200"`` / ``"nccl"`` / None on a GPU machine -> the torch process group is only the control
    plane (gloo for CPU objects, nccl registered for CUDA tensors so user code and the test oracle can still call
    ``dist.all_reduce``); gradient traffic goes through our P2P kernels.  ``"gloo"`` -> CPU plumbing config.
    Also pins this process to ``cuda:<local_rank>`` (the reference forgets ``set_device``).
    """
    if dist.is_initialized():
        return
    if local_rank is None:
        local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    use_cuda = torch.cuda.is_available() and backend not in ("gloo",)
    if use_cuda:
        torch.cuda.set_device(local_rank)
        be = "cpu:gloo,cuda:nccl"
    else:
        be = "gloo"
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("WORLD_SIZE", "1")
    kw = {}
    if use_cuda:
        kw["device_id"] = torch.device("cuda", local_rank)
    try:
        dist.init_process_group(backend=be, timeout=datetime.timedelta(seconds=timeout_s), **kw)
    except TypeError:
        dist.init_process_group(backend=be, timeout=datetime.timedelta(seconds=timeout_s))


def destroy_process_group():
    if dist.is_initialized():
        dist.destroy_process_group()
