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

    ``backend``:
      * ``"b200"`` / None on a GPU machine (the product path): the torch process group is **gloo only** and is
        nothing but the CPU control plane (shape checks, handle exchange, barriers); every byte of device traffic
        goes through our peer-memory kernels.  No NCCL communicator is ever created in the process.
      * ``"nccl"``: as above plus torch's NCCL backend registered for CUDA tensors, for user code (or a test
        oracle) that wants to call ``dist.all_reduce`` on device tensors itself; the wrapper still never uses it.
      * ``"gloo"``: the reference's CPU plumbing configuration (no GPU involved).
    Also pins this process to ``cuda:<local_rank>`` (the reference forgets ``set_device``) and to the CPUs of that
    GPU's NUMA node, so that pinned host batches are allocated next to the GPU they are copied to.
    """
    if dist.is_initialized():
        return
    if local_rank is None:
        local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    use_cuda = torch.cuda.is_available() and backend not in ("gloo",)
    be = "gloo"
    if use_cuda:
        torch.cuda.set_device(local_rank)
        if os.environ.get("B200DDP_NUMA_PIN", "1") == "1":
            from ..utils.affinity import pin_to_gpu_numa_node
            pin_to_gpu_numa_node(local_rank)
        if backend == "nccl":
            be = "cpu:gloo,cuda:nccl"
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("WORLD_SIZE", "1")
    kw = {}
    if use_cuda and backend == "nccl":
        kw["device_id"] = torch.device("cuda", local_rank)
    try:
        dist.init_process_group(backend=be, timeout=datetime.timedelta(seconds=timeout_s), **kw)
    except TypeError:
        dist.init_process_group(backend=be, timeout=datetime.timedelta(seconds=timeout_s))


def destroy_process_group():
    """Release every wrapper's peer-memory window (unmap, address-range free, handle release, multicast unbind -
    ``DistributedDataParallel.close``), then tear the control plane down."""
    from .ddp import close_all
    close_all()
    if dist.is_initialized():
        dist.destroy_process_group()
