"""Time the fan-in dgrad (dx = mask ? dgrad + skip : 0, + BN-backward sums) against the plain dgrad for a shape.
Usage: python tools/fanin_bench.py ci co k pad H N   (ci = channels of dx)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_tutorials_b200.ops import conv, native  # noqa: E402

ci, co, k, p, H, N = [int(v) for v in sys.argv[1:7]]
dev = "cuda"
cl = torch.channels_last
w = (torch.randn(co, ci, k, k, device=dev) / (ci * k * k) ** 0.5).to(torch.bfloat16).contiguous(memory_format=cl)
dy = torch.randn(N, co, H, H, device=dev).to(torch.bfloat16).contiguous(memory_format=cl)
add = torch.randn(N, ci, H, H, device=dev).to(torch.bfloat16).contiguous(memory_format=cl)
y = torch.randn(N, ci, H, H, device=dev).to(torch.bfloat16).contiguous(memory_format=cl)
mean = torch.zeros(ci, device=dev)
bits = torch.randint(0, 256, (N, H, H, ci // 8), device=dev, dtype=torch.uint8)
partial = native.bn_partial_b(torch.device(dev))


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


t_plain = timeit(lambda: conv.conv_dgrad(dy, w, (N, ci, H, H), 1, p))
t_fan = timeit(lambda: conv.conv_dgrad_fused(dy, w, (N, ci, H, H), 1, p, add, bits, y, mean, partial))
t_fan0 = timeit(lambda: conv.conv_dgrad_fused(dy, w, (N, ci, H, H), 1, p, add, bits, None, None, None))
b_fan0 = N * H * H * (co + 2 * ci) * 2 + N * H * H * ci // 8
px = N * H * H
b_plain = px * (co + ci) * 2
b_fan = px * (co + 3 * ci) * 2 + px * ci // 8
print("dx %d ch <- dy %d ch k=%d H=%d N=%d: plain %.3f ms (%.2f TB/s)  fan-in+sums %.3f ms (%.2f TB/s)  "
      "fan-in only %.3f ms (%.2f TB/s)" %
      (ci, co, k, H, N, t_plain, b_plain / t_plain / 1e9, t_fan, b_fan / t_fan / 1e9, t_fan0, b_fan0 / t_fan0 / 1e9))
