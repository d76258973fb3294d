"""Launches the round-2 kernels a few times each at the benchmark's shapes (batch 256) so that ncu can capture them:
    python tools/ncu_targets.py && ncu --set full --clock-control none --import-source on \\
        -k regex:"conv3x3_halo|head_fwd|bn_bwd_reduce_batched|ce_bwd_fused" -c 12 -o gpurun_out/prof_r2 \\
        python tools/ncu_targets.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from pytorch_distributed_tutorials_b200.ops import conv, native  # noqa: E402


def late_round2(dev):
    """Kernels changed at the end of round 2: class-strided dgrad (transposed epilogue), the CTA-pair kernel on a plain
    1x1 GEMM, bn1 + ReLU inside the stem's maxpool.
        ncu ... -k regex:"igemm2_kernel<1, 2|igemm3_kernel|maxpool_fwd" -c 9 python tools/ncu_targets.py late"""
    N, cl = 256, torch.channels_last
    dy = torch.randn(N, 128, 28, 28, device=dev).to(torch.bfloat16).contiguous(memory_format=cl)
    w = (torch.randn(128, 128, 3, 3, device=dev) / 34).to(torch.bfloat16).contiguous(memory_format=cl)
    x = torch.randn(N, 1024, 14, 14, device=dev).to(torch.bfloat16).contiguous(memory_format=cl)
    w1 = (torch.randn(512, 1024, 1, 1, device=dev) / 32).to(torch.bfloat16).contiguous(memory_format=cl)
    y = torch.randn(N, 64, 112, 112, device=dev).to(torch.bfloat16).contiguous(memory_format=cl)
    ss = torch.cat([torch.ones(64, device=dev), torch.zeros(64, device=dev)])
    for _ in range(2):
        conv.conv_dgrad(dy, w, (N, 128, 56, 56), 2, 1)      # 4 parity-class launches
        conv.conv_fprop(x, w1, 1, 0)                         # plain 1x1, K = 1024 -> CTA pairs
        pooled, pidx = native.bn_relu_maxpool_fwd(y, ss)
        native.maxpool_bwd(pidx, y.shape, pooled, pooled)
    torch.cuda.synchronize()
    print("ok")


def main():
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    if len(sys.argv) > 1 and sys.argv[1] == "late":
        return late_round2(dev)
    N = 256
    cl = torch.channels_last
    # weight-stationary 3x3, 64 -> 64 @ 56x56: fprop with BN statistics, dgrad with BN-backward sums
    x = torch.randn(N, 64, 56, 56, device=dev).to(torch.bfloat16).contiguous(memory_format=cl)
    w = (torch.randn(64, 64, 3, 3, device=dev) / 24).to(torch.bfloat16).contiguous(memory_format=cl)
    dy = torch.randn_like(x)
    mask = torch.randint(0, 255, (x.numel() // 8,), device=dev, dtype=torch.uint8)
    mean = torch.zeros(64, device=dev)
    for _ in range(3):
        conv.conv_fprop_stats(x, w, 1, 1, native.bn_partial(dev))
        conv.conv_dgrad_bnstats(dy, w, x.shape, 1, 1, x, mask, mean, native.bn_partial_b(dev))
    # fused head: avgpool + fc, 2048 -> 1000
    f = torch.randn(N, 2048, 7, 7, device=dev).to(torch.bfloat16).contiguous(memory_format=cl)
    wf = (torch.randn(1000, 2048, device=dev) / 45).to(torch.bfloat16)
    b = torch.zeros(1000, device=dev)
    labels = torch.randint(0, 1000, (N,), device=dev)
    for _ in range(3):
        logits, pooled = native.head_fwd(f, wf, b)
        loss, lse = native.ce_fwd_f32(logits, labels)
        native.ce_bwd_fused(logits, labels, lse, None, torch.empty(N, 1000, device=dev, dtype=torch.bfloat16),
                            torch.empty(1000, device=dev))
    # BatchNorm backward of a residual-block output, 256 channels @ 56x56 (reduce with two inputs + mask, then apply)
    y = torch.randn(N, 256, 56, 56, device=dev).to(torch.bfloat16).contiguous(memory_format=cl)
    dza, dzb = torch.randn_like(y), torch.randn_like(y)
    m2 = torch.randint(0, 255, (y.numel() // 8,), device=dev, dtype=torch.uint8)
    g, mu, istd = torch.ones(256, device=dev), torch.zeros(256, device=dev), torch.ones(256, device=dev)
    dg, db, coef = torch.empty(256, device=dev), torch.empty(256, device=dev), torch.empty(768, device=dev)
    for _ in range(3):
        native.bn_bwd(dza, dzb, None, y, g, mu, istd, dg, db, False, True, coef, True, mask=m2)
    torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    main()
