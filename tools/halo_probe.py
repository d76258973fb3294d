"""Probes for the weight-stationary 3x3 kernel (csrc/conv3x3.cu).
  (no argument)  correctness of fprop / dgrad against the general im2col kernel and an fp32 reference;
  timing         CUDA-graph timing of the product kernel and of its timing-probe variants;
  profile        per-role cycle counters of one launch (which warp role waits for which).
History: the shifted-window A descriptors work with the matrix-descriptor base-offset field left ZERO (the 128-byte
swizzle is a function of the address bits); writing the row phase there gave garbage. The first version was bound by
the single MMA-issuing thread (~95 cycles per MMA rebuilding descriptors; `profile` showed it), not by memory."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from pytorch_distributed_tutorials_b200.ops import _lib, conv  # noqa: E402


def rel(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-6)).item()


def main():
    dev = "cuda:0"
    L = _lib.lib()
    for (ci, co, H, N) in [(64, 64, 56, 4), (128, 128, 28, 4), (64, 128, 14, 3)]:
        torch.manual_seed(0)
        x = torch.randn(N, ci, H, H, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(co, ci, 3, 3, device=dev) / (ci * 9) ** 0.5).to(torch.bfloat16).contiguous(
            memory_format=torch.channels_last)
        dy = torch.randn(N, co, H, H, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        ref = conv.ref_fprop(x.float(), w.float(), 1, 1)
        refd = conv.ref_dgrad(dy.float(), w.float(), x.shape, 1, 1)
        for bits in (0,):
            L.b2_conv_set_halo(1)
            L.b2_conv_set_halo_debug(bits)
            y = conv.conv_fprop(x, w, 1, 1)
            dx = conv.conv_dgrad(dy, w, x.shape, 1, 1)
            torch.cuda.synchronize()
            print("ci=%d co=%d H=%d dbg=%d  fprop rel=%.4g  dgrad rel=%.4g" % (ci, co, H, bits, rel(y, ref),
                                                                               rel(dx, refd)), flush=True)
        L.b2_conv_set_halo(0)
        y = conv.conv_fprop(x, w, 1, 1)
        dx = conv.conv_dgrad(dy, w, x.shape, 1, 1)
        torch.cuda.synchronize()
        print("ci=%d co=%d H=%d general kernel  fprop rel=%.4g  dgrad rel=%.4g" % (ci, co, H, rel(y, ref),
                                                                                    rel(dx, refd)), flush=True)
        L.b2_conv_set_halo(1)
        L.b2_conv_set_halo_debug(0)


def timing():
    """Where the halo kernel's time goes: debug bits 1-3 are timing probes (wrong results)."""
    dev = "cuda:0"
    L = _lib.lib()
    for (ci, co, H, N) in [(64, 64, 56, 256), (64, 64, 56, 64), (64, 64, 56, 16), (128, 128, 28, 256),
                           (128, 128, 28, 32)]:
        x = torch.randn(N, ci, H, H, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(co, ci, 3, 3, device=dev) / (ci * 9) ** 0.5).to(torch.bfloat16).contiguous(
            memory_format=torch.channels_last)
        for bits, what in ((0, "product"), (32 + 64 + 128, "MMAs + accumulator read-out only")):
            L.b2_conv_set_halo_debug(bits)
            for _ in range(3):
                conv.conv_fprop(x, w, 1, 1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            g = torch.cuda.CUDAGraph()
            torch.cuda.synchronize()
            with torch.cuda.graph(g):
                for _ in range(20):
                    conv.conv_fprop(x, w, 1, 1)
            g.replay()
            torch.cuda.synchronize()
            e0.record()
            g.replay()
            e1.record()
            torch.cuda.synchronize()
            print("ci=%d co=%d H=%d N=%d  %-26s %.4f ms" % (ci, co, H, N, what, e0.elapsed_time(e1) / 20), flush=True)
        L.b2_conv_set_halo_debug(0)


def profile():
    """Per-role cycle counters of one launch (CTA 0 and the mean over CTAs)."""
    dev = "cuda:0"
    L = _lib.lib()
    ci, co, H, N = 64, 64, 56, 256
    x = torch.randn(N, ci, H, H, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(co, ci, 3, 3, device=dev) / (ci * 9) ** 0.5).to(torch.bfloat16).contiguous(
        memory_format=torch.channels_last)
    buf = torch.zeros(148 * 8, dtype=torch.int64, device=dev)
    names = ["producer wait slab-empty", "mma wait acc-empty", "mma wait slab-full", "mma issue+commit",
             "mma wait weights", "mma loop total", "epilogue wait acc-full", "epilogue loop total"]
    for bits in (0, 32 + 64 + 128):
        L.b2_conv_set_halo_debug(bits)
        conv.conv_fprop(x, w, 1, 1)
        torch.cuda.synchronize()
        L.b2_conv_set_halo_profile(ctypes.c_void_p(buf.data_ptr()))
        conv.conv_fprop(x, w, 1, 1)
        torch.cuda.synchronize()
        L.b2_conv_set_halo_profile(ctypes.c_void_p(0))
        v = buf.view(148, 8).double()
        print("dbg=%d (cycles; ~49 tiles per CTA)" % bits)
        for i, n in enumerate(names):
            print("  %-28s cta0 %9d   mean %9d" % (n, int(v[0, i]), int(v[:, i].mean())), flush=True)
    L.b2_conv_set_halo_debug(0)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "profile":
        import ctypes
        profile()
    elif len(sys.argv) > 1 and sys.argv[1] == "timing":
        timing()
    else:
        main()
