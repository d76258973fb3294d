"""GPU check of the tcgen05 implicit-GEMM conv kernels against an fp32 torch reference, shape by shape.
Each (kind, shape) prints one line; a CUDA fault in one kind does not hide the others (use --kind and run
each in its own process).  Usage: python tools/conv_check.py --kind fprop|dgrad|wgrad [--batch 8] [--bn 0]

--vs-cudnn [--out FILE.md]: head-to-head timing against cuDNN (what the reference's torchvision convolutions resolve
to, reference resnet/main.py:76 -> aten::convolution[_backward]) for every ResNet-50 shape x {fprop, dgrad, wgrad}
at --batch (256 = the benchmark's): both sides bf16 channels_last, cudnn.benchmark on, each candidate captured in a
CUDA graph of 20 back-to-back launches (no launch gaps), replayed 5 times, CUDA events; the kernel names cuDNN picked
come from torch.profiler.  Writes a markdown table.
"""
import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_tutorials_b200.ops import _lib, conv  # noqa: E402

# (Cin, Cout, k, stride, pad, H)  -- ResNet-50 @224 and ResNet-18 @32 shapes with Cin % 64 == 0
SHAPES = [
    (64, 64, 1, 1, 0, 56), (64, 64, 3, 1, 1, 56), (64, 256, 1, 1, 0, 56), (256, 64, 1, 1, 0, 56),
    (256, 128, 1, 1, 0, 56), (128, 128, 3, 2, 1, 56), (128, 512, 1, 1, 0, 28), (256, 512, 1, 2, 0, 56),
    (512, 128, 1, 1, 0, 28), (128, 128, 3, 1, 1, 28), (512, 256, 1, 1, 0, 28), (256, 256, 3, 2, 1, 28),
    (256, 1024, 1, 1, 0, 14), (512, 1024, 1, 2, 0, 28), (1024, 256, 1, 1, 0, 14), (256, 256, 3, 1, 1, 14),
    (1024, 512, 1, 1, 0, 14), (512, 512, 3, 2, 1, 14), (512, 2048, 1, 1, 0, 7), (1024, 2048, 1, 2, 0, 14),
    (2048, 512, 1, 1, 0, 7), (512, 512, 3, 1, 1, 7),
    # ResNet-18 @32 (CIFAR): tiny spatial extents
    (64, 64, 3, 1, 1, 8), (64, 128, 3, 2, 1, 8), (128, 128, 3, 1, 1, 4), (64, 128, 1, 2, 0, 8),
    (256, 256, 3, 1, 1, 2), (512, 512, 3, 1, 1, 1), (256, 512, 1, 2, 0, 2),
]


def _graph_time(fn, reps=20, replays=5):
    """ms per call of fn, timed as a captured graph of `reps` back-to-back launches."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(replays):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / reps)
    return best


def _kernel_names(fn):
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        fn()
        torch.cuda.synchronize()
    rows = sorted(((k.device_time_total, k.key) for k in prof.key_averages() if k.device_time_total > 0),
                  reverse=True)
    return [n for _, n in rows[:3]]


# how often each of the 22 non-stem shapes occurs in ResNet-50
R50_MULT = {(64, 64, 1, 1, 0, 56): 1, (64, 64, 3, 1, 1, 56): 3, (64, 256, 1, 1, 0, 56): 4, (256, 64, 1, 1, 0, 56): 2,
            (256, 128, 1, 1, 0, 56): 1, (128, 128, 3, 2, 1, 56): 1, (128, 512, 1, 1, 0, 28): 4,
            (256, 512, 1, 2, 0, 56): 1, (512, 128, 1, 1, 0, 28): 3, (128, 128, 3, 1, 1, 28): 3,
            (512, 256, 1, 1, 0, 28): 1, (256, 256, 3, 2, 1, 28): 1, (256, 1024, 1, 1, 0, 14): 6,
            (512, 1024, 1, 2, 0, 28): 1, (1024, 256, 1, 1, 0, 14): 5, (256, 256, 3, 1, 1, 14): 5,
            (1024, 512, 1, 1, 0, 14): 1, (512, 512, 3, 2, 1, 14): 1, (512, 2048, 1, 1, 0, 7): 3,
            (1024, 2048, 1, 2, 0, 14): 1, (2048, 512, 1, 1, 0, 7): 2, (512, 512, 3, 1, 1, 7): 2}


def vs_cudnn(a):
    """Per-shape head-to-head: ours vs cuDNN, fprop / dgrad / wgrad."""
    import torch.nn.functional as F
    dev = torch.device("cuda:0")
    torch.backends.cudnn.benchmark = True
    aten = torch.ops.aten
    shapes = list(SHAPES[:22])
    if a.shape:
        shapes = [tuple(int(v) for v in a.shape.split(','))]
    lines = ["| Cin->Cout k/s @H | op | ours ms | cuDNN ms | ours/cuDNN speed | ours TFLOP/s | cuDNN kernel |",
             "|---|---|---|---|---|---|---|"]
    wins = total = 0
    sum_ours = sum_cudnn = 0.0
    for (ci, co, k, s, p, H) in shapes:
        N = a.batch
        x = torch.randn(N, ci, H, H, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(co, ci, k, k, device=dev) / (ci * k * k) ** 0.5).to(torch.bfloat16).contiguous(
            memory_format=torch.channels_last)
        Ho = (H + 2 * p - k) // s + 1
        dy = torch.randn(N, co, Ho, Ho, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        fl = 2.0 * N * Ho * Ho * co * ci * k * k

        def bwd(mask):
            return aten.convolution_backward(dy, x, w, None, [s, s], [p, p], [1, 1], False, [0, 0], 1, mask)

        cands = {
            "fprop": (lambda: conv.conv_fprop(x, w, s, p), lambda: F.conv2d(x, w, None, s, p)),
            "dgrad": (lambda: conv.conv_dgrad(dy, w, x.shape, s, p), lambda: bwd([True, False, False])),
            "wgrad": (lambda: conv.conv_wgrad(x, dy, w.shape, s, p), lambda: bwd([False, True, False])),
        }
        for op in ("fprop", "dgrad", "wgrad"):
            ours, ref = cands[op]
            try:
                t_o = _graph_time(ours)
                t_c = _graph_time(ref)
                names = _kernel_names(ref)
            except Exception as e:  # noqa: BLE001
                lines.append("| %d->%d %d/%d @%d | %s | EXC %r | | | | |" % (ci, co, k, s, H, op, e))
                continue
            total += 1
            wins += t_o <= t_c
            m = R50_MULT.get((ci, co, k, s, p, H), 1)
            sum_ours += m * t_o
            sum_cudnn += m * t_c
            lines.append("| %d->%d %d/%d @%d | %s | %.4f | %.4f | %.2fx | %.0f | %s |" %
                         (ci, co, k, s, H, op, t_o, t_c, t_c / t_o, fl / t_o / 1e9, names[0][:70] if names else "?"))
            print(lines[-1], flush=True)
    head = ["# Convolution kernels vs cuDNN, per ResNet-50 shape (batch %d, bf16 NHWC, B200)" % a.batch, "",
            "Timing: CUDA graph of 20 back-to-back launches, best of 5 replays, CUDA events. `speed` > 1: ours faster.",
            "ours wins %d of %d rows; time weighted by how often each shape occurs in ResNet-50: ours %.3f ms, "
            "cuDNN %.3f ms per step (fprop+dgrad+wgrad of the 52 non-stem convolutions)." % (wins, total, sum_ours,
                                                                                          sum_cudnn), ""]
    text = "\n".join(head + lines) + "\n"
    if a.out:
        with open(a.out, "w") as fh:
            fh.write(text)
    print(text)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--vs-cudnn", action="store_true")
    ap.add_argument("--out", default=None)
    ap.add_argument("--kind", default="fprop")
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--bn", type=int, default=0)
    ap.add_argument("--limit", type=int, default=0)
    ap.add_argument("--time", action="store_true")
    ap.add_argument("--act-mode", type=int, default=-1)
    ap.add_argument("--impl", type=int, default=2)
    ap.add_argument("--debug-bits", type=int, default=0, help="timing experiments: 1 skip loads, 2 skip MMAs")
    ap.add_argument("--shape", default=None, help="ci,co,k,s,p,H: run only this shape")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    if a.vs_cudnn:
        return vs_cudnn(a)
    _lib.lib().b2_conv_set_tile_n(a.bn)
    _lib.lib().b2_conv_set_act_mode(a.act_mode)
    _lib.lib().b2_conv_set_impl(a.impl)
    _lib.lib().b2_conv_set_debug(a.debug_bits)
    shapes = SHAPES[: a.limit] if a.limit else SHAPES
    if a.shape:
        shapes = [tuple(int(v) for v in a.shape.split(','))]
    bad = 0
    for (ci, co, k, s, p, H) in shapes:
        N = a.batch
        x = torch.randn(N, ci, H, H, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(co, ci, k, k, device=dev) * (1.0 / (ci * k * k) ** 0.5)).to(torch.bfloat16).contiguous(
            memory_format=torch.channels_last)
        Ho = (H + 2 * p - k) // s + 1
        dy = torch.randn(N, co, Ho, Ho, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        xf, wf, dyf = x.float(), w.float(), dy.float()
        try:
            if a.kind == "fprop_stats":
                from pytorch_distributed_tutorials_b200.ops import native
                got = conv.conv_fprop_stats(x, w, s, p, native.bn_partial(dev))[0].float()
                ref = conv.ref_fprop(xf, wf, s, p)
            elif a.kind == "fprop":
                got = conv.conv_fprop(x, w, s, p).float()
                ref = conv.ref_fprop(xf, wf, s, p)
            elif a.kind == "dgrad":
                got = conv.conv_dgrad(dy, w, x.shape, s, p).float()
                ref = conv.ref_dgrad(dyf, wf, x.shape, s, p)
            else:
                got = conv.conv_wgrad(x, dy, w.shape, s, p).float()
                ref = conv.ref_wgrad(xf, dyf, w.shape, s, p)
            torch.cuda.synchronize()
            err = (got - ref).abs().max().item()
            scale = ref.abs().max().item() + 1e-6
            rel = err / scale
            cos = torch.nn.functional.cosine_similarity(got.flatten(), ref.flatten(), dim=0).item()
            ok = rel < 2e-2
            bad += (not ok)
            ms = ""
            if a.time and (ok or a.debug_bits):
                from pytorch_distributed_tutorials_b200.ops import native
                part = native.bn_partial(dev)
                fn = {"fprop": lambda: conv.conv_fprop(x, w, s, p),
                      "fprop_stats": lambda: conv.conv_fprop_stats(x, w, s, p, part),
                      "dgrad": lambda: conv.conv_dgrad(dy, w, x.shape, s, p),
                      "wgrad": lambda: conv.conv_wgrad(x, dy, w.shape, s, p)}[a.kind]
                for _ in range(3):
                    fn()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(10):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                t = e0.elapsed_time(e1) / 10
                fl = 2.0 * N * Ho * Ho * co * ci * k * k
                ms = " %.3f ms %.1f TF/s" % (t, fl / t / 1e9)
            print("%s %-5s ci=%4d co=%4d k=%d s=%d H=%3d N=%d  maxerr=%.4g rel=%.3g cos=%.5f%s" %
                  ("OK  " if ok else "FAIL", a.kind, ci, co, k, s, H, N, err, rel, cos, ms), flush=True)
        except Exception as e:  # noqa: BLE001
            print("EXC  %s ci=%d co=%d k=%d s=%d H=%d: %r" % (a.kind, ci, co, k, s, H, e), flush=True)
            bad += 1
            break
    print("%s: %d/%d failed" % (a.kind, bad, len(shapes)))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
