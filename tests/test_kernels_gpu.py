"""Per-kernel numerics on a B200: every native op against a plain PyTorch fp32 reference of the same op
(bf16 tolerances where the kernel's inputs/outputs are bf16)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _cl(t):
    return t.contiguous(memory_format=torch.channels_last)


def _bf(t):
    return _cl(t.to(torch.bfloat16))


def _rel(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-6)).item()


CONV_SHAPES = [  # ci, co, k, s, p, H, N
    (64, 64, 1, 1, 0, 56, 4), (64, 64, 3, 1, 1, 56, 4), (256, 128, 1, 1, 0, 56, 2), (128, 128, 3, 2, 1, 56, 2),
    (256, 512, 1, 2, 0, 56, 2), (512, 2048, 1, 1, 0, 7, 8), (512, 512, 3, 1, 1, 7, 8), (1024, 256, 1, 1, 0, 14, 4),
    (64, 128, 3, 2, 1, 8, 8), (512, 512, 3, 1, 1, 1, 8), (256, 512, 1, 2, 0, 2, 8), (128, 256, 3, 2, 1, 28, 3),
    # the benchmark's own batch (256 per GPU): persistent-kernel tile loops, split-K plans and statistics warps at
    # the sizes that are timed - 1x1 64->256 @56, 3x3 256 @14, strided 3x3 128 @56, strided 1x1 256->512 @56
    (64, 256, 1, 1, 0, 56, 256), (256, 256, 3, 1, 1, 14, 256), (128, 128, 3, 2, 1, 56, 256),
    (256, 512, 1, 2, 0, 56, 256),
    # weight-stationary halo kernel (csrc/conv3x3.cu): two k-blocks, rows that do not divide into tiles, two column
    # tiles in dgrad, a frame narrower than 128 rows
    (128, 128, 3, 1, 1, 28, 8), (64, 128, 3, 1, 1, 14, 5), (128, 64, 3, 1, 1, 9, 2), (128, 128, 3, 1, 1, 28, 256),
]


@pytest.mark.parametrize("ci,co,k,s,p,H,N", CONV_SHAPES)
def test_conv_fprop_dgrad_wgrad(ci, co, k, s, p, H, N):
    from pytorch_distributed_tutorials_b200.ops import conv
    torch.manual_seed(0)
    x = _bf(torch.randn(N, ci, H, H, device=DEV))
    w = _bf(torch.randn(co, ci, k, k, device=DEV) / (ci * k * k) ** 0.5)
    Ho = (H + 2 * p - k) // s + 1
    dy = _bf(torch.randn(N, co, Ho, Ho, device=DEV))
    xf, wf, dyf = x.float(), w.float(), dy.float()
    y = conv.conv_fprop(x, w, s, p)
    assert y.shape == (N, co, Ho, Ho) and y.is_contiguous(memory_format=torch.channels_last)
    assert _rel(y, conv.ref_fprop(xf, wf, s, p)) < 1e-2
    dx = conv.conv_dgrad(dy, w, x.shape, s, p)
    assert _rel(dx, conv.ref_dgrad(dyf, wf, x.shape, s, p)) < 1e-2
    dw = conv.conv_wgrad(x, dy, w.shape, s, p)
    ref_dw = conv.ref_wgrad(xf, dyf, w.shape, s, p)
    assert _rel(dw, ref_dw) < 2e-3
    # accumulate into an existing gradient
    dw2 = conv.conv_wgrad(x, dy, w.shape, s, p, out=dw.clone(), accumulate=True)
    assert _rel(dw2, 2 * ref_dw) < 2e-3


@pytest.mark.parametrize("rows", [37, 256])
def test_gemm_helpers_classifier_shapes(rows):
    """rows = 256 is the benchmark's batch: the classifier's input-gradient GEMM (K = 1000, not a multiple of 64: the
    last k-block is zero-filled by TMA) then runs on the CTA-pair kernel like every plain 1x1 GEMM with M >= 256."""
    from pytorch_distributed_tutorials_b200.ops import conv
    torch.manual_seed(1)
    a = torch.randn(rows, 2048, device=DEV).to(torch.bfloat16)
    w = (torch.randn(1000, 2048, device=DEV) / 45).to(torch.bfloat16)
    b = torch.randn(1000, device=DEV)
    y = conv.gemm_nt(a, w, b)
    assert _rel(y, a.float() @ w.float().t() + b) < 1e-2
    dy = torch.randn(rows, 1000, device=DEV).to(torch.bfloat16)
    assert _rel(conv.gemm_nn(dy, w), dy.float() @ w.float()) < 1e-2
    out = torch.zeros(1000, 2048, device=DEV)
    conv.gemm_tn(dy, a, out)
    assert _rel(out, dy.float().t() @ a.float()) < 2e-3


@pytest.mark.parametrize("C,H,N", [(64, 16, 4), (256, 14, 3), (2048, 7, 2), (128, 5, 1)])
@pytest.mark.parametrize("relu,res", [(True, False), (True, True), (False, False)])
def test_bn_forward_backward(C, H, N, relu, res):
    from pytorch_distributed_tutorials_b200.ops import native
    torch.manual_seed(2)
    y = _bf(torch.randn(N, C, H, H, device=DEV) * 2 + 0.5)
    r = _bf(torch.randn(N, C, H, H, device=DEV)) if res else None
    gamma = torch.rand(C, device=DEV) + 0.5
    beta = torch.randn(C, device=DEV)
    rm, rv = torch.randn(C, device=DEV), torch.rand(C, device=DEV) + 0.5
    nbt = torch.tensor(3, device=DEV)
    rm0, rv0 = rm.clone(), rv.clone()
    mean, invstd = torch.empty(C, device=DEV), torch.empty(C, device=DEV)
    ss, coef = torch.empty(2 * C, device=DEV), torch.empty(3 * C, device=DEV)
    mask = torch.zeros(y.numel() // 8, dtype=torch.uint8, device=DEV)
    z = native.bn_fwd_train(y, r, gamma, beta, rm, rv, nbt, 0.1, 1e-5, relu, mean, invstd, ss, mask=mask)
    # reference (fp32 math on the same bf16 inputs)
    yf = y.float().requires_grad_(True)
    g32, b32 = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    rf = r.float().requires_grad_(True) if res else None
    rm_r, rv_r = rm0.clone(), rv0.clone()
    o = F.batch_norm(yf, rm_r, rv_r, g32, b32, True, 0.1, 1e-5)
    if res:
        o = o + rf
    if relu:
        o = F.relu(o)
    assert _rel(z, o) < 1e-2
    assert torch.allclose(rm, rm_r, atol=1e-4, rtol=1e-4) and torch.allclose(rv, rv_r, atol=1e-4, rtol=1e-3)
    assert int(nbt.item()) == 4
    # backward with a two-way gradient fan-in
    dza, dzb = _bf(torch.randn_like(o)), _bf(torch.randn_like(o))
    dgamma, dbeta = torch.empty(C, device=DEV), torch.empty(C, device=DEV)
    dy, dres = native.bn_bwd(dza, dzb, z if relu else None, y, gamma, mean, invstd, dgamma, dbeta, False, relu, coef,
                             True)
    # reference gradient must use the same relu mask as the kernel (bf16 z): build it from z
    up = dza.float() + dzb.float()
    if relu:
        up = up * (z.float() > 0)
    ob = F.batch_norm(yf, None, None, g32, b32, True, 0.1, 1e-5)
    ob.backward(up)
    assert _rel(dy, yf.grad) < 2e-2
    assert _rel(dgamma, g32.grad) < 1e-2 and _rel(dbeta, b32.grad) < 1e-2
    assert _rel(dres, up) < 1e-2
    if relu:  # same result when the kernel is given the 1-bit mask instead of z
        dg2, db2 = torch.empty(C, device=DEV), torch.empty(C, device=DEV)
        dy2, dres2 = native.bn_bwd(dza, dzb, None, y, gamma, mean, invstd, dg2, db2, False, relu, coef, True, mask=mask)
        assert torch.equal(dy2, dy) and torch.equal(dres2, dres) and torch.equal(dg2, dgamma)


@pytest.mark.parametrize("C,H,W,N", [(256, 56, 56, 4), (512, 28, 28, 3), (64, 8, 6, 5), (1024, 14, 14, 2)])
def test_bn_backward_with_half_resolution_skip_gradient(C, H, W, N):
    """The second incoming gradient given compactly at the even pixels (input gradient of a 1x1 / stride 2
    downsample) must give exactly what the zero-filled full-resolution tensor gives."""
    from pytorch_distributed_tutorials_b200.ops import native
    torch.manual_seed(11)
    y = _bf(torch.randn(N, C, H, W, device=DEV))
    dza = _bf(torch.randn(N, C, H, W, device=DEV))
    q = _bf(torch.randn(N, C, H // 2, W // 2, device=DEV))
    full = torch.zeros(N, C, H, W, device=DEV)
    full[:, :, ::2, ::2] = q.float()
    full = _bf(full)
    mask = torch.randint(0, 256, (y.numel() // 8,), device=DEV, dtype=torch.uint8)
    gamma, mean, invstd = torch.rand(C, device=DEV) + 0.5, torch.randn(C, device=DEV) * 0.1, torch.rand(C, device=DEV) + 0.5
    outs = []
    for dzb, half in ((full, False), (q, True)):
        dg, db, coef = torch.empty(C, device=DEV), torch.empty(C, device=DEV), torch.empty(3 * C, device=DEV)
        dy, dres = native.bn_bwd(dza, dzb, None, y, gamma, mean, invstd, dg, db, False, True, coef, True, mask=mask,
                                 dz_b_half_res=half)
        torch.cuda.synchronize()
        outs.append((dy.clone(), dres.clone(), dg.clone(), db.clone()))
    for a_, b_ in zip(*outs):
        assert torch.equal(a_, b_)


def test_bn_eval():
    from pytorch_distributed_tutorials_b200.ops import native
    C = 128
    y = _bf(torch.randn(2, C, 9, 9, device=DEV))
    gamma, beta = torch.rand(C, device=DEV) + 0.5, torch.randn(C, device=DEV)
    rm, rv = torch.randn(C, device=DEV), torch.rand(C, device=DEV) + 0.5
    z = native.bn_fwd_eval(y, None, gamma, beta, rm, rv, 1e-5, True, torch.empty(2 * C, device=DEV))
    assert _rel(z, F.relu(F.batch_norm(y.float(), rm, rv, gamma, beta, False, 0.1, 1e-5))) < 1e-2


@pytest.mark.parametrize("H", [112, 16, 7])
def test_maxpool(H):
    from pytorch_distributed_tutorials_b200.ops import native
    torch.manual_seed(3)
    x = _bf(F.relu(torch.randn(3, 64, H, H, device=DEV)))  # relu -> many ties at 0, like the real stem
    y, idx = native.maxpool_fwd(x)
    xf = x.float().requires_grad_(True)
    yr = F.max_pool2d(xf, 3, 2, 1)
    assert torch.equal(y.float(), yr)
    da, db = _bf(torch.randn_like(yr)), _bf(torch.randn_like(yr))
    dx = native.maxpool_bwd(idx, x.shape, da, db)
    yr.backward(da.float() + db.float())
    assert _rel(dx, xf.grad) < 1e-2


@pytest.mark.parametrize("H", [56, 31])
def test_bn_relu_maxpool_fused_equals_apply_then_pool(H):
    """BatchNorm affine + ReLU applied inside the pooling kernel: same pooled values and winners as the apply kernel
    followed by maxpool, bit for bit; windows whose maximum is 0 carry tap code 9, and maxpool_bwd of that index is the
    gradient of the BatchNorm output (ReLU backward included)."""
    from pytorch_distributed_tutorials_b200.ops import native
    torch.manual_seed(5)
    C = 64
    y = _bf(torch.randn(3, C, H, H, device=DEV))
    gamma, beta = torch.rand(C, device=DEV) + 0.5, torch.randn(C, device=DEV) * 0.3
    rm, rv = torch.randn(C, device=DEV) * 0.2, torch.rand(C, device=DEV) + 0.5
    ss = torch.empty(2 * C, device=DEV)
    z = native.bn_fwd_eval(y, None, gamma, beta, rm, rv, 1e-5, True, ss)
    p_ref, idx_ref = native.maxpool_fwd(z)
    ss2 = torch.empty(2 * C, device=DEV)
    assert native.bn_fwd_eval(y, None, gamma, beta, rm, rv, 1e-5, True, ss2, apply=False) is None
    assert torch.equal(ss, ss2)
    p, idx = native.bn_relu_maxpool_fwd(y, ss2)
    assert torch.equal(p, p_ref)
    dead = (p_ref.permute(0, 2, 3, 1) == 0)
    assert torch.equal(idx[~dead], idx_ref[~dead]) and bool((idx[dead] == 9).all()) and dead.any()
    da, db = _bf(torch.randn_like(p.float())), _bf(torch.randn_like(p.float()))
    dz = native.maxpool_bwd(idx, y.shape, da, db)
    dz_ref = native.maxpool_bwd(idx_ref, y.shape, da, db).float() * (z > 0)
    assert torch.equal(dz.float(), dz_ref)
    # fp32 torch for the values (winners of near-ties legitimately differ between bf16 and fp32, so not the gradient)
    pr = F.max_pool2d(F.relu(F.batch_norm(y.float(), rm, rv, gamma, beta, False, 0.0, 1e-5)), 3, 2, 1)
    assert _rel(p, pr) < 1e-2


def test_avgpool():
    from pytorch_distributed_tutorials_b200.ops import native
    x = _bf(torch.randn(5, 2048, 7, 7, device=DEV))
    y = native.avgpool_fwd(x)
    assert _rel(y, x.float().mean((2, 3))) < 1e-2
    dy = torch.randn(5, 2048, device=DEV).to(torch.bfloat16)
    dx = native.avgpool_bwd(dy, x.shape)
    assert _rel(dx, (dy.float() / 49)[:, :, None, None].expand(5, 2048, 7, 7)) < 1e-2


@pytest.mark.parametrize("N,C,HW,ncls", [(256, 2048, 49, 1000), (37, 2048, 49, 1000), (16, 512, 1, 1000),
                                          (5, 512, 4, 10), (40, 1024, 16, 1000), (3, 64, 9, 24)])
def test_fused_head_avgpool_fc(N, C, HW, ncls):
    """avgpool -> fc in one kernel (cluster DSMEM pooled operand, tcgen05) against fp32 torch: logits fp32, pooled
    bf16. The oracle pools in fp32, rounds the mean to bf16 (as the kernel does, and as a bf16 avgpool would) and
    multiplies in fp32."""
    from pytorch_distributed_tutorials_b200.ops import native
    torch.manual_seed(N + C)
    side = int(HW ** 0.5)
    assert side * side == HW
    x = _bf(torch.randn(N, C, side, side, device=DEV) + 0.5)
    w = (torch.randn(ncls, C, device=DEV) / C ** 0.5).to(torch.bfloat16)
    b = torch.randn(ncls, device=DEV)
    assert native.head_supported(C, ncls)
    logits, pooled = native.head_fwd(x, w, b)
    torch.cuda.synchronize()
    pm = x.float().mean((2, 3))
    assert torch.isfinite(logits).all()
    assert _rel(pooled, pm) < 5e-3
    ref = pm.to(torch.bfloat16).float() @ w.float().t() + b
    assert _rel(logits, ref) < 2e-3, _rel(logits, ref)


def test_cross_entropy_backward_emits_bias_gradient():
    from pytorch_distributed_tutorials_b200.ops import native
    torch.manual_seed(9)
    for rows, ncls in ((256, 1000), (33, 1000), (7, 10)):
        logits = torch.randn(rows, ncls, device=DEV) * 3
        labels = torch.randint(0, ncls, (rows,), device=DEV)
        loss, lse = native.ce_fwd_f32(logits, labels)
        ref = logits.clone().requires_grad_(True)
        lr = F.cross_entropy(ref, labels)
        lr.backward()
        assert abs(loss.item() - lr.item()) < 1e-4 * max(1.0, abs(lr.item()))
        assert torch.allclose(lse, torch.logsumexp(logits, 1), atol=1e-4)
        g = torch.tensor(2.0, device=DEV)
        dl = torch.full((rows, ncls), float("nan"), device=DEV, dtype=torch.bfloat16)
        db = torch.full((ncls,), float("nan"), device=DEV)
        native.ce_bwd_fused(logits, labels, lse, g, dl, db, False)
        assert _rel(dl, 2.0 * ref.grad) < 1e-2
        assert torch.allclose(db, dl.float().sum(0), atol=1e-6, rtol=1e-5)
        native.ce_bwd_fused(logits, labels, lse, g, dl, db, True)
        assert torch.allclose(db, 2 * dl.float().sum(0), atol=1e-6, rtol=1e-5)
        loss2, _ = native.ce_fwd_f32(logits, labels)           # the completion counter reset itself
        assert loss2.item() == loss.item()


def test_cross_entropy():
    from pytorch_distributed_tutorials_b200.ops.loss import CrossEntropyLoss
    torch.manual_seed(4)
    logits = (torch.randn(33, 1000, device=DEV) * 3).to(torch.bfloat16).float().requires_grad_(True)
    labels = torch.randint(0, 1000, (33,), device=DEV)
    l = CrossEntropyLoss()(logits, labels)
    (l * 2.0).backward()
    ref = logits.detach().clone().requires_grad_(True)
    lr = F.cross_entropy(ref, labels)
    (lr * 2.0).backward()
    assert abs(l.item() - lr.item()) < 1e-3
    assert _rel(logits.grad, ref.grad) < 1e-2


def test_sgd_matches_torch():
    from pytorch_distributed_tutorials_b200.ops import native
    torch.manual_seed(5)
    n = 100003
    p = torch.randn(n, device=DEV)
    pr = p.clone().requires_grad_(True)
    buf = torch.zeros(n, device=DEV)
    sh = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    opt = torch.optim.SGD([pr], lr=0.01, momentum=0.9, weight_decay=1e-5)
    for it in range(3):
        g = torch.randn(n, device=DEV)
        native.sgd_step(p, g, buf, sh, 0.01, 0.9, 1e-5, 1.0, it == 0)
        pr.grad = g.clone()
        opt.step()
        assert torch.allclose(p, pr.detach(), atol=1e-6, rtol=1e-5)
        assert torch.allclose(buf, opt.state[pr]["momentum_buffer"], atol=1e-6, rtol=1e-5)
        assert torch.equal(sh, p.to(torch.bfloat16))


def test_stem_conv_via_im2col():
    from pytorch_distributed_tutorials_b200.ops import conv, native
    torch.manual_seed(6)
    x = torch.randn(3, 3, 64, 64, device=DEV)
    w = torch.randn(64, 3, 7, 7, device=DEV) * 0.1
    w_ohwi = w.permute(0, 2, 3, 1).contiguous()
    wp = torch.zeros(64, 192, device=DEV, dtype=torch.bfloat16)
    native.stem_pack_w(w_ohwi, wp)
    a, ho, wo = native.stem_im2col(x)
    y = conv.gemm_nt(a, wp).view(3, ho, wo, 64).permute(0, 3, 1, 2)
    ref = F.conv2d(x.to(torch.bfloat16).float(), w.to(torch.bfloat16).float(), None, 2, 3)
    assert _rel(y, ref) < 1e-2
    # weight gradient: dW = dy^T @ A, unpacked to OHWI
    dy = torch.randn(3 * ho * wo, 64, device=DEV).to(torch.bfloat16)
    dwp = torch.zeros(64, 192, device=DEV)
    conv.gemm_tn(dy, a, dwp)
    dw = torch.zeros(64, 7, 7, 3, device=DEV)
    native.stem_unpack_dw(dwp, dw, False)
    ref_dw = conv.ref_wgrad(x.to(torch.bfloat16).float(), dy.float().view(3, ho, wo, 64).permute(0, 3, 1, 2),
                            w.shape, 2, 3)
    assert _rel(dw.permute(0, 3, 1, 2), ref_dw) < 5e-3


@pytest.mark.parametrize("ci,co,k,s,p,H,N", [(64, 256, 1, 1, 0, 56, 4), (256, 256, 3, 1, 1, 14, 8),
                                               (512, 2048, 1, 1, 0, 7, 5), (128, 128, 3, 2, 1, 28, 3),
                                               (256, 64, 1, 1, 0, 56, 2), (64, 64, 3, 1, 1, 56, 16),
                                               (512, 2048, 1, 1, 0, 7, 64), (256, 512, 1, 2, 0, 56, 16),
                                               (128, 128, 3, 1, 1, 28, 5), (64, 128, 3, 1, 1, 14, 3),
                                               (128, 64, 3, 1, 1, 9, 2), (64, 64, 3, 1, 1, 56, 256)])
def test_conv_epilogue_bn_statistics(ci, co, k, s, p, H, N):
    """The persistent fprop kernel also accumulates per-channel sum / sum of squares of the bf16 output."""
    from pytorch_distributed_tutorials_b200.ops import conv, native
    torch.manual_seed(7)
    x = _bf(torch.randn(N, ci, H, H, device=DEV))
    w = _bf(torch.randn(co, ci, k, k, device=DEV) / (ci * k * k) ** 0.5)
    partial = native.bn_partial(torch.device(DEV))
    partial.fill_(float("nan"))                       # the kernel must zero what it uses
    y, blocks = conv.conv_fprop_stats(x, w, s, p, partial)
    assert blocks > 0
    torch.cuda.synchronize()
    part = partial[: blocks * 2 * co].view(blocks, 2, co).sum(0)
    yf = y.float()
    ref_s, ref_ss = yf.sum((0, 2, 3)), (yf * yf).sum((0, 2, 3))
    assert torch.allclose(part[0], ref_s, rtol=1e-3, atol=1e-2), (part[0] - ref_s).abs().max()
    assert torch.allclose(part[1], ref_ss, rtol=1e-3, atol=1e-2), (part[1] - ref_ss).abs().max()
    assert _rel(y, conv.ref_fprop(x.float(), w.float(), s, p)) < 1e-2


@pytest.mark.parametrize("N,H,W", [(4, 224, 224), (3, 32, 32), (2, 64, 96), (5, 57, 45)])
def test_direct_stem_conv(N, H, W):
    """7x7/2 stem straight from a padded NHWC4 image (overlapping-window TMA boxes): fprop, BN statistics, wgrad."""
    from pytorch_distributed_tutorials_b200.ops import native
    torch.manual_seed(11)
    assert native.stem_supported(H, W)
    x = torch.randn(N, 3, H, W, device=DEV)
    w = torch.randn(64, 3, 7, 7, device=DEV) / 147 ** 0.5
    w_ohwi = w.permute(0, 2, 3, 1).contiguous()
    wp = torch.empty(64, 224, dtype=torch.bfloat16, device=DEV)
    native.stem_pack_weight(w_ohwi, wp)
    xp = native.stem_pack_input(x)
    partial = native.bn_partial(torch.device(DEV))
    partial.fill_(float("nan"))
    y, blocks = native.stem_fprop(xp, wp, N, H, W, partial)
    torch.cuda.synchronize()
    xr, wr = x.to(torch.bfloat16).float(), w.to(torch.bfloat16).float()
    ref = torch.nn.functional.conv2d(xr, wr, stride=2, padding=3)
    assert y.shape == ref.shape
    assert _rel(y, ref) < 1e-2
    assert blocks > 0
    part = partial[: blocks * 2 * 64].view(blocks, 2, 64).sum(0)
    yf = y.float()
    assert torch.allclose(part[0], yf.sum((0, 2, 3)), rtol=1e-3, atol=5e-2)
    assert torch.allclose(part[1], (yf * yf).sum((0, 2, 3)), rtol=1e-3, atol=5e-2)
    # weight gradient
    dy = _bf(torch.randn_like(ref))
    dw = torch.full((64, 7, 7, 3), float("nan"), device=DEV)
    native.stem_wgrad(xp, dy, dw, N, H, W, False)
    ref_dw = torch.nn.grad.conv2d_weight(xr, w.shape, dy.float(), stride=2, padding=3)
    assert _rel(dw.permute(0, 3, 1, 2), ref_dw) < 5e-3
    native.stem_wgrad(xp, dy, dw, N, H, W, True)
    assert _rel(dw.permute(0, 3, 1, 2), 2 * ref_dw) < 5e-3


@pytest.mark.parametrize("ci,co,k,p,H,N", [(256, 64, 1, 0, 28, 4), (64, 64, 3, 1, 14, 3), (512, 128, 1, 0, 7, 5),
                                            (128, 64, 1, 0, 28, 3), (1024, 256, 1, 0, 14, 6)])
def test_conv_dgrad_fan_in_relu_mask_and_bn_backward_sums(ci, co, k, p, H, N):
    """dx = mask ? (dgrad + add) : 0 finished by the statistics warps (residual fan-in + ReLU backward), which also
    leave sum(dx), sum(dx * (y - mean)) per channel."""
    from pytorch_distributed_tutorials_b200.ops import conv, native
    torch.manual_seed(5)
    w = _bf(torch.randn(co, ci, k, k, device=DEV) / (ci * k * k) ** 0.5)
    dy = _bf(torch.randn(N, co, H, H, device=DEV))
    add = _bf(torch.randn(N, ci, H, H, device=DEV))
    y = _bf(torch.randn(N, ci, H, H, device=DEV))
    mean = torch.randn(ci, device=DEV) * 0.1
    keep = torch.rand(N, H, H, ci, device=DEV) > 0.4                     # NHWC, as the mask is laid out
    bits = (keep.view(N, H, H, ci // 8, 8).to(torch.uint8) << torch.arange(8, device=DEV, dtype=torch.uint8)).sum(
        -1).to(torch.uint8)
    partial = native.bn_partial_b(torch.device(DEV))
    for mask, keep_nchw in ((bits, keep.permute(0, 3, 1, 2)), (None, None)):
        partial.fill_(float("nan"))
        got, blocks = conv.conv_dgrad_fused(dy, w, (N, ci, H, H), 1, p, add, mask, y, mean, partial)
        torch.cuda.synchronize()
        assert got is not None and blocks > 0
        # the kernel rounds the dgrad to bf16 before the add (as the unfused path does)
        ref = conv.ref_dgrad(dy.float(), w.float(), (N, ci, H, H), 1, p).to(torch.bfloat16).float() + add.float()
        if keep_nchw is not None:
            ref = torch.where(keep_nchw, ref, torch.zeros_like(ref))
            assert (got.float()[~keep_nchw] == 0).all()
        assert _rel(got, ref) < 1.5e-2
        part = partial[: blocks * 2 * ci].view(blocks, 2, ci).sum(0)
        g = got.float()
        ref0, ref1 = g.sum((0, 2, 3)), (g * (y.float() - mean.view(1, -1, 1, 1))).sum((0, 2, 3))
        assert torch.allclose(part[0], ref0, rtol=1e-3, atol=5e-2), (part[0] - ref0).abs().max()
        assert torch.allclose(part[1], ref1, rtol=1e-3, atol=5e-2), (part[1] - ref1).abs().max()
        # fan-in only (no BatchNorm sums): same dx, bit for bit
        got0, blocks0 = conv.conv_dgrad_fused(dy, w, (N, ci, H, H), 1, p, add, mask, None, None, None)
        assert got0 is not None and blocks0 == 0 and torch.equal(got0, got)


@pytest.mark.parametrize("ci,co,k,p,H,N", [(64, 256, 1, 0, 28, 4), (128, 128, 3, 1, 14, 3), (256, 1024, 1, 0, 14, 5),
                                            (512, 512, 3, 1, 7, 6), (64, 64, 3, 1, 56, 2), (128, 64, 3, 1, 28, 3),
                                            (64, 128, 3, 1, 9, 2), (128, 128, 3, 1, 28, 64)])
def test_conv_dgrad_produces_bn_backward_sums(ci, co, k, p, H, N):
    """The dgrad kernel's statistics warps leave sum(g), sum(g*(y-mean)) per channel of dx (g = mask ? dx : 0)."""
    from pytorch_distributed_tutorials_b200.ops import conv, native
    torch.manual_seed(9)
    w = _bf(torch.randn(co, ci, k, k, device=DEV) / (ci * k * k) ** 0.5)
    dy = _bf(torch.randn(N, co, H, H, device=DEV))
    y = _bf(torch.randn(N, ci, H, H, device=DEV))
    mean = torch.randn(ci, device=DEV) * 0.1
    keep = torch.rand(N, H, H, ci, device=DEV) > 0.5
    bits = (keep.view(N, H, H, ci // 8, 8).to(torch.uint8) << torch.arange(8, device=DEV, dtype=torch.uint8)).sum(
        -1).to(torch.uint8)
    partial = native.bn_partial_b(torch.device(DEV))
    partial.fill_(float("nan"))
    dx, blocks = conv.conv_dgrad_bnstats(dy, w, (N, ci, H, H), 1, p, y, bits, mean, partial)
    torch.cuda.synchronize()
    assert blocks > 0
    assert _rel(dx, conv.ref_dgrad(dy.float(), w.float(), (N, ci, H, H), 1, p)) < 1e-2
    part = partial[: blocks * 2 * ci].view(blocks, 2, ci).sum(0)
    g = torch.where(keep.permute(0, 3, 1, 2), dx.float(), torch.zeros((), device=DEV))
    ref0 = g.sum((0, 2, 3))
    ref1 = (g * (y.float() - mean.view(1, -1, 1, 1))).sum((0, 2, 3))
    assert torch.allclose(part[0], ref0, rtol=1e-3, atol=5e-2), (part[0] - ref0).abs().max()
    assert torch.allclose(part[1], ref1, rtol=1e-3, atol=5e-2), (part[1] - ref1).abs().max()
