"""Host -> device copy bandwidth per rank when all ranks of the node copy at once (what bounds the end-to-end step
when the per-step input is 154 MB): torchrun --nproc-per-node N tools/h2d_probe.py"""
import os
import time

import torch
import torch.distributed as dist

lr = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(lr)
if int(os.environ.get("WORLD_SIZE", "1")) > 1:
    dist.init_process_group("gloo")
x = torch.empty(256, 3, 224, 224).pin_memory()
d = torch.empty_like(x, device="cuda")
for _ in range(2):
    d.copy_(x, non_blocking=True)
torch.cuda.synchronize()
if dist.is_initialized():
    dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    d.copy_(x, non_blocking=True)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print("rank %d: %.2f ms per 154 MB copy = %.1f GB/s" % (lr, ms, x.numel() * 4 / ms / 1e6), flush=True)
if dist.is_initialized():
    dist.barrier()
    dist.destroy_process_group()
