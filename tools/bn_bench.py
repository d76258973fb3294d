"""Times the BatchNorm kernels on ResNet-50 activation shapes (batch 256) and reports achieved HBM bandwidth.
Usage: python tools/bn_bench.py [--shape C,H] [--iters 20]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from pytorch_distributed_tutorials_b200.ops import native  # noqa: E402


def timeit(fn, iters, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.zero_()  # evict L2 between iterations (buffer larger than the 126 MB L2)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default=None)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--batch", type=int, default=256)
    a = ap.parse_args()
    dev = "cuda:0"
    shapes = [(64, 112), (256, 56), (64, 56), (512, 28), (1024, 14), (2048, 7)]
    if a.shape:
        shapes = [tuple(int(v) for v in a.shape.split(","))]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for C, H in shapes:
        N = a.batch
        y = torch.randn(N, C, H, H, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        r = torch.randn_like(y)
        dza, dzb = torch.randn_like(y), torch.randn_like(y)
        gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
        rm, rv, nbt = torch.zeros(C, device=dev), torch.ones(C, device=dev), torch.tensor(0, device=dev)
        mean, invstd = torch.empty(C, device=dev), torch.empty(C, device=dev)
        ss, coef = torch.empty(2 * C, device=dev), torch.empty(3 * C, device=dev)
        mask = torch.empty(y.numel() // 8, dtype=torch.uint8, device=dev)
        dg, db = torch.empty(C, device=dev), torch.empty(C, device=dev)
        nel = y.numel()
        t_f = timeit(lambda: native.bn_fwd_train(y, r, gamma, beta, rm, rv, nbt, 0.1, 1e-5, True, mean, invstd, ss,
                                                 mask=mask), a.iters, flush)
        t_b = timeit(lambda: native.bn_bwd(dza, dzb, None, y, gamma, mean, invstd, dg, db, False, True, coef, True,
                                           mask=mask), a.iters, flush)
        bf = nel * (2 + 2 + 2 + 2 + 0.125)          # stats read y; apply read y,res, write z, mask
        bb = nel * (2 * 2 + 2 + 0.125) + nel * (2 * 2 + 2 + 0.125 + 4)  # reduce + apply(+dres)
        print("C=%4d H=%3d  fwd(stats+apply+res+relu) %.3f ms %.2f TB/s | bwd(reduce+apply, 2 inputs, dres) %.3f ms "
              "%.2f TB/s" % (C, H, t_f, bf / t_f / 1e9, t_b, bb / t_b / 1e9), flush=True)


if __name__ == "__main__":
    main()
