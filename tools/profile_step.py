"""Kernel-level time breakdown of one training step (torch.profiler / CUPTI; no nsys on the box).
Usage: python tools/profile_step.py [--impl ours|reference] [--model resnet50] [--batch 256] [--steps 3]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--image-size", type=int, default=224)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--rows", type=int, default=40)
    ap.add_argument("--wgrad-waves", type=int, default=0, help="experiment: split-K items per SM of the wgrad kernels")
    a = ap.parse_args()
    if a.wgrad_waves:
        from pytorch_distributed_tutorials_b200.ops import _lib
        _lib.lib().b2_conv_set_wgrad_waves(a.wgrad_waves)
    args = argparse.Namespace(model=a.model, batch=a.batch, image_size=a.image_size, sync_bn=False,
                              bucket_cap_mb=None, no_fuse_optimizer=False, ref_variant="r1", no_graph=True)
    torch.cuda.set_device(0)
    step, ddp = (bench.build_ours if a.impl == "ours" else bench.build_reference)(args, 0, 1)
    dev = torch.device("cuda", 0)
    x = torch.randn(a.batch, 3, a.image_size, a.image_size, device=dev)
    y = torch.randint(0, 1000, (a.batch,), device=dev)
    for _ in range(3):
        step(x, y)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(a.steps):
            step(x, y)
        torch.cuda.synchronize()
    ka = prof.key_averages()
    rows = [(k.key, k.device_time_total / a.steps / 1e3, k.count / a.steps) for k in ka if k.device_time_total > 0
            and k.device_type.name == "CUDA"]
    rows.sort(key=lambda r: -r[1])
    tot = sum(r[1] for r in rows)
    print("total device kernel time per step: %.3f ms over %d kernel kinds" % (tot, len(rows)))
    for name, ms, cnt in rows[: a.rows]:
        print("%8.3f ms %5.1f%% x%-6.1f %s" % (ms, 100 * ms / tot, cnt, name[:150]))


if __name__ == "__main__":
    main()
