"""ResNet-18/34/50/101/152 with torchvision's parameter names, shapes and initialisation, executed by a
hand-scheduled sm_100a engine instead of autograd-over-aten.

The module tree is ordinary ``nn.Conv2d`` / ``nn.BatchNorm2d`` / ``nn.Linear`` containers, so ``state_dict()`` has
exactly torchvision's keys (reference resnet/main.py:76 uses ``torchvision.models.resnet18``; checkpoints of the
reference load here and vice versa, modulo the ``module.`` prefix the DDP wrapper adds, resnet/main.py:84,112).
On CUDA the forward pass is ONE autograd node (``_EngineFn``): its forward runs conv (tcgen05 implicit GEMM) ->
BN statistics -> BN-apply+ReLU(+residual) kernels and records a tape; its backward walks the tape with fused
BN-backward(+ReLU mask, + gradient fan-in) -> dgrad / wgrad kernels, writes parameter gradients straight into
their DDP bucket slots and tells the reducer as each one lands.  On CPU (gloo plumbing config) or with
``B200DDP_FORCE_TORCH=1`` the same modules run through stock PyTorch ops, which is also the numerical oracle.
"""
import os
from typing import List, Optional, Type, Union

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import _lib, conv as convops, native
from ..parallel.flat import store_of

_NVTX = os.environ.get("B200DDP_NVTX", "0") == "1"   # NVTX ranges around forward / backward for nsys / ncu


def conv3x3(inp, out, stride=1):
    return nn.Conv2d(inp, out, 3, stride, 1, bias=False)


def conv1x1(inp, out, stride=1):
    return nn.Conv2d(inp, out, 1, stride, 0, bias=False)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes, planes, stride=1, downsample=None, norm_layer=nn.BatchNorm2d):
        super().__init__()
        self.conv1 = conv3x3(inplanes, planes, stride)
        self.bn1 = norm_layer(planes)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = conv3x3(planes, planes)
        self.bn2 = norm_layer(planes)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        identity = x
        out = self.relu(self.bn1(self.conv1(x)))
        out = self.bn2(self.conv2(out))
        if self.downsample is not None:
            identity = self.downsample(x)
        return self.relu(out + identity)

    def units(self):
        return [(self.conv1, self.bn1), (self.conv2, self.bn2)]


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None, norm_layer=nn.BatchNorm2d):
        super().__init__()
        self.conv1 = conv1x1(inplanes, planes)
        self.bn1 = norm_layer(planes)
        self.conv2 = conv3x3(planes, planes, stride)  # stride on the 3x3 ("v1.5"), as torchvision
        self.bn2 = norm_layer(planes)
        self.conv3 = conv1x1(planes, planes * self.expansion)
        self.bn3 = norm_layer(planes * self.expansion)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        identity = x
        out = self.relu(self.bn1(self.conv1(x)))
        out = self.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        if self.downsample is not None:
            identity = self.downsample(x)
        return self.relu(out + identity)

    def units(self):
        return [(self.conv1, self.bn1), (self.conv2, self.bn2), (self.conv3, self.bn3)]


class ResNet(nn.Module):
    def __init__(self, block: Type[Union[BasicBlock, Bottleneck]], layers: List[int], num_classes: int = 1000,
                 norm_layer=nn.BatchNorm2d):
        super().__init__()
        self._norm_layer = norm_layer
        self.inplanes = 64
        self.conv1 = nn.Conv2d(3, 64, 7, 2, 3, bias=False)
        self.bn1 = norm_layer(64)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(3, 2, 1)
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], 2)
        self.layer3 = self._make_layer(block, 256, layers[2], 2)
        self.layer4 = self._make_layer(block, 512, layers[3], 2)
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Linear(512 * block.expansion, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, (nn.BatchNorm2d, nn.GroupNorm)):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)
        self._engine = None
        # device int64[1] holding the address of the input batch, or None: when set (utils/graph.py), the stem reads
        # the batch through it, so a captured step can be pointed at a fresh input buffer before every replay
        self._b2_input_slot = None

    def _make_layer(self, block, planes, blocks, stride=1):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(conv1x1(self.inplanes, planes * block.expansion, stride),
                                       self._norm_layer(planes * block.expansion))
        layers = [block(self.inplanes, planes, stride, downsample, self._norm_layer)]
        self.inplanes = planes * block.expansion
        for _ in range(1, blocks):
            layers.append(block(self.inplanes, planes, norm_layer=self._norm_layer))
        return nn.Sequential(*layers)

    # ------------------------------------------------------------------ stock path (CPU / oracle)
    def forward_reference(self, x):
        x = self.maxpool(self.relu(self.bn1(self.conv1(x))))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        return self.fc(torch.flatten(self.avgpool(x), 1))

    def _engine_applies(self):
        """The engine implements training-mode BatchNorm backward only. A forward that will be differentiated while
        (some) BatchNorm layers normalise with their running statistics - ``model.eval()`` with grad enabled
        (saliency maps, input gradients) or frozen-BN fine-tuning where single layers are in eval mode - goes through
        the stock autograd path instead, which differentiates those layers correctly."""
        if not torch.is_grad_enabled():
            return True
        bns = getattr(self, "_bn_modules", None)
        if bns is None:
            bns = self._bn_modules = [m for m in self.modules() if isinstance(m, nn.modules.batchnorm._BatchNorm)]
        if not self.training:
            return not any(p.requires_grad for p in self.parameters())
        return all(m.training for m in bns)

    def forward(self, x):
        if _lib.use_native(x) and next(self.parameters()).is_cuda and self._engine_applies():
            if self._engine is None:
                self._engine = ResNetEngine(self)
                self._engine.xg = getattr(self, "_b2_xgpu_ctx", None)
                self._engine.xg_slot = getattr(self, "_b2_xgpu_slot", 0)
            return self._engine(x)
        return self.forward_reference(x)

    def grad_ready_order(self):
        """Parameters in the order the engine's backward produces their gradients (last layer first)."""
        order = [self.fc.bias, self.fc.weight]
        for layer in (self.layer4, self.layer3, self.layer2, self.layer1):
            for blk in reversed(list(layer)):
                us = blk.units()
                last_conv, last_bn = us[-1]
                order += [last_bn.weight, last_bn.bias, last_conv.weight]
                for c, b in reversed(us[1:-1]):
                    order += [b.weight, b.bias, c.weight]
                if blk.downsample is not None:
                    order += [blk.downsample[1].weight, blk.downsample[1].bias, blk.downsample[0].weight]
                c, b = us[0]     # the first conv's dgrad absorbs the skip-path gradient, so it runs last
                order += [b.weight, b.bias, c.weight]
        order += [self.bn1.weight, self.bn1.bias, self.conv1.weight]
        return order


def _resnet(block, layers, **kw):
    return ResNet(block, layers, **kw)


def resnet18(pretrained=False, **kw):
    return _resnet(BasicBlock, [2, 2, 2, 2], **kw)


def resnet34(pretrained=False, **kw):
    return _resnet(BasicBlock, [3, 4, 6, 3], **kw)


def resnet50(pretrained=False, **kw):
    return _resnet(Bottleneck, [3, 4, 6, 3], **kw)


def resnet101(pretrained=False, **kw):
    return _resnet(Bottleneck, [3, 4, 23, 3], **kw)


def resnet152(pretrained=False, **kw):
    return _resnet(Bottleneck, [3, 8, 36, 3], **kw)


MODELS = {"resnet18": resnet18, "resnet34": resnet34, "resnet50": resnet50, "resnet101": resnet101,
          "resnet152": resnet152}


# =====================================================================================================
# Native engine
# =====================================================================================================
class _BnState:
    """Persistent per-BatchNorm device buffers (saved statistics and kernel coefficient scratch)."""

    def __init__(self, bn, device):
        c = bn.num_features
        self.mean = torch.empty(c, dtype=torch.float32, device=device)
        self.invstd = torch.empty(c, dtype=torch.float32, device=device)
        self.scale_shift = torch.empty(2 * c, dtype=torch.float32, device=device)
        self.coef = torch.empty(3 * c, dtype=torch.float32, device=device)


class _Rec:
    __slots__ = ("x", "y", "mask", "conv", "bn", "relu", "stem_a")

    def __init__(self, x, y, mask, conv, bn, relu, stem_a=None):
        self.x, self.y, self.mask, self.conv, self.bn, self.relu, self.stem_a = x, y, mask, conv, bn, relu, stem_a


class _HeadLink:
    """Private channel between the engine and ``ops.loss.CrossEntropyLoss``. The model returns fp32 logits and the
    criterion is a separate call (reference resnet/main.py:121-122), so by autograd's contract the engine would
    receive an fp32 ``dlogits`` tensor and have to cast it to bf16 for its GEMMs. With this link (attached to the
    logits tensor as ``_b2_head``) our loss kernel writes bf16 ``dlogits`` and the classifier's bias gradient
    straight where the engine's backward wants them and returns a zero-stride placeholder to autograd."""

    def __init__(self, engine):
        self.engine = engine
        self.ready = False
        self.logits_ptr = None
        self.dl = None
        self.slot = None
        self._placeholder = None

    def arm(self, logits):
        self.ready, self.slot, self.dl = False, None, None
        self.logits_ptr = logits.data_ptr()

    def accepts(self, lg):
        return lg.data_ptr() == self.logits_ptr and lg.dtype == torch.float32

    def dlogits_buffer(self, shape):
        n = shape[0] * shape[1]
        self.dl = native.scratch(self.engine.device, "head_dlogits", n, torch.bfloat16)[:n].view(shape)
        return self.dl

    def bias_slot(self):
        self.slot = self.engine._slot(self.engine.model.fc.bias)
        return self.slot[0], self.slot[1]

    def placeholder(self, shape):
        if self._placeholder is None:
            self._placeholder = torch.zeros((), dtype=torch.float32, device=self.engine.device)
        return self._placeholder.expand(shape)


class ResNetEngine:
    """Runs a :class:`ResNet` on the native kernels. One instance per model; holds only scratch state."""

    def __init__(self, model: ResNet):
        self.model = model
        self.device = next(model.parameters()).device
        self.bn_state = {}
        self.local_shadow = {}     # bf16 copies for parameters that are not in a FlatParameterStore
        self.local_versions = {}
        self.stem_wp = torch.zeros(64, 192, dtype=torch.bfloat16, device=self.device)
        self.stem_dwp = torch.zeros(64, 192, dtype=torch.float32, device=self.device)
        self.stem_wp_direct = torch.zeros(64, 224, dtype=torch.bfloat16, device=self.device)
        # direct stem (no im2col matrix); B200DDP_STEM=im2col selects the materialised-im2col GEMM path instead
        self.stem_direct = os.environ.get("B200DDP_STEM", "direct") != "im2col"
        # B200DDP_FUSE_FANIN=1: the first conv's dgrad of every block also finishes the residual fan-in (skip gradient
        # add + ReLU backward of the block input, done by the kernel's statistics warps on the output panel in shared
        # memory) and produces the previous block's BatchNorm-backward sums. Correct (tests) but off by default: the
        # panel hand-offs (epilogue -> statistics warps -> store warp, two panel buffers) leave the kernel at
        # ~3 TB/s, 0.44 ms on the 256-channel 56x56 layer against 0.09 + 0.27 ms for dgrad + the two-input reduce pass.
        self.fuse_fanin = os.environ.get("B200DDP_FUSE_FANIN", "0") == "1"
        # the dgrad that produces a BatchNorm's output gradient also produces that BatchNorm's backward sums
        self.fuse_bn_bwd = os.environ.get("B200DDP_FUSE_BN_BWD", "1") != "0"
        # strided 1x1 downsample: half-resolution input gradient, read compactly by the BatchNorm backward it feeds
        self.compact_ds_grad = os.environ.get("B200DDP_COMPACT_DS_GRAD", "1") != "0"
        # bn1 + ReLU applied inside the maxpool kernel (the normalised 112x112 tensor and its mask never exist)
        self.fuse_stem_pool = os.environ.get("B200DDP_FUSE_STEM_POOL", "1") != "0"
        self._bn_blocks = 0
        self.params = list(model.parameters())
        self.xg = None             # peer-memory context for SyncBatchNorm (handed over by the DDP wrapper)
        self.xg_slot = 0           # float offset of the statistics exchange area inside the symmetric window
        # avgpool + fc as one kernel (csrc/head.cu); B200DDP_HEAD=split selects avgpool kernel + GEMM instead
        self.fused_head = os.environ.get("B200DDP_HEAD", "fused") != "split"
        self.head_link = _HeadLink(self)

    # ------------------------------------------------------------------ helpers
    def _bn(self, bn):
        st = self.bn_state.get(id(bn))
        if st is None:
            st = _BnState(bn, self.device)
            self.bn_state[id(bn)] = st
        return st

    def _w16(self, p):
        st = store_of(p)
        if st is not None and st.device.type == "cuda":
            return st.shadow_view(p)
        v = self.local_versions.get(id(p))
        if v != p._version or id(p) not in self.local_shadow:
            self.local_shadow[id(p)] = p.detach().to(torch.bfloat16).contiguous(
                memory_format=torch.channels_last if p.dim() == 4 else torch.contiguous_format)
            self.local_versions[id(p)] = p._version
        return self.local_shadow[id(p)]

    def _sync_shadows(self):
        seen = set()
        for p in self.params:
            st = store_of(p)
            if st is not None and id(st) not in seen:
                seen.add(id(st))
                st.sync_shadow()

    def _sync_bn_ctx(self, bn):
        return self.xg if getattr(bn, "_b2_sync", False) and self.xg is not None else None

    XG_SLOT_FLOATS = 2 * 2048 + 256     # one exchange area: (sum, sumsq | sum_dy, sum_dy_xmu) + per-CTA counts

    def _next_xg_slot(self, bn):
        """Exchange area of the next cross-GPU BatchNorm call: two areas used alternately, so the kernels need one
        device barrier per call, not two (csrc/bn.cu). A pure call counter: identical on every rank, and a captured
        step contains an even number of calls (forward + backward of every BatchNorm), so replays keep alternating."""
        if self._sync_bn_ctx(bn) is None:
            return self.xg_slot
        self._xg_calls = getattr(self, "_xg_calls", 0) + 1
        return self.xg_slot + (self._xg_calls & 1) * self.XG_SLOT_FLOATS

    # ------------------------------------------------------------------ forward units
    def _conv(self, x, conv):
        return convops.conv_fprop(x, self._w16(conv.weight), conv.stride[0], conv.padding[0])

    def _unit(self, x, conv, bn, relu, res, training, tape):
        blocks = 0
        if training or bn.running_mean is None:
            # training-mode BN: the conv epilogue accumulates the statistics, no separate pass over y
            y, blocks = convops.conv_fprop_stats(x, self._w16(conv.weight), conv.stride[0], conv.padding[0],
                                                 native.bn_partial(self.device))
        else:
            y = self._conv(x, conv)
        mask = self._new_mask(y) if (tape is not None and relu) else None
        z = self._bn_apply(y, bn, relu, res, training, mask, blocks)
        if tape is not None:
            tape.append(_Rec(x, y, mask, conv, bn, relu))
        return z

    @staticmethod
    def _new_mask(y):
        """ReLU bit mask (1 bit / element): what backward needs instead of re-reading the bf16 output."""
        return torch.empty(y.numel() // 8, dtype=torch.uint8, device=y.device)

    def _bn_apply(self, y, bn, relu, res, training, mask=None, partial_blocks=0, apply=True):
        """``apply=False``: statistics / running statistics / scale+shift only (returns None)."""
        st = self._bn(bn)
        use_batch_stats = training or bn.running_mean is None
        if use_batch_stats:
            track = bn.track_running_stats and bn.running_mean is not None and training
            mom = bn.momentum if bn.momentum is not None else 0.1
            return native.bn_fwd_train(y, res, bn.weight, bn.bias, bn.running_mean if track else None,
                                       bn.running_var if track else None,
                                       bn.num_batches_tracked if track else None, mom, bn.eps, relu, st.mean,
                                       st.invstd, st.scale_shift, self._sync_bn_ctx(bn), self._next_xg_slot(bn), mask,
                                       partial_blocks, apply=apply)
        return native.bn_fwd_eval(y, res, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.eps, relu,
                                  st.scale_shift, apply=apply)

    def _stem(self, x, training, tape):
        m = self.model
        xin = x if x.dtype in (torch.float32, torch.bfloat16) else x.float()
        xin = xin.contiguous()
        w_ohwi = m.conv1.weight.detach().permute(0, 2, 3, 1)   # a free view when the weight is stored OHWI
        if not w_ohwi.is_contiguous():
            w_ohwi = w_ohwi.contiguous()
        n, _, h, w = x.shape
        want_stats = training or m.bn1.running_mean is None
        blocks = 0
        if self.stem_direct and native.stem_supported(h, w):
            # tensor-core convolution straight from a padded NHWC4 copy of the batch (csrc/stem.cu)
            xp = native.stem_pack_input(xin, slot=m._b2_input_slot)
            if m._b2_input_slot is not None:
                m._b2_input_slot_used = True
            native.stem_pack_weight(w_ohwi, self.stem_wp_direct)
            y, blocks = native.stem_fprop(xp, self.stem_wp_direct, n, h, w,
                                          native.bn_partial(self.device) if want_stats else None)
            a = ("direct", xp, (n, h, w))
        else:
            a, ho, wo = native.stem_im2col(xin)
            native.stem_pack_w(w_ohwi, self.stem_wp)            # fp32 OHWI [64,147] -> bf16 [64,192]
            if want_stats:
                y2d, blocks = convops.gemm_nt_stats(a, self.stem_wp, native.bn_partial(self.device))
            else:
                y2d = convops.gemm_nt(a, self.stem_wp)
            y = y2d.view(n, ho, wo, 64).permute(0, 3, 1, 2)
        if self.fuse_stem_pool:
            # coefficients only; the pooling kernel normalises on load and folds the ReLU backward into its index
            self._bn_apply(y, m.bn1, True, None, training, None, blocks, apply=False)
            if tape is not None:
                tape.append(_Rec(None, y, None, m.conv1, m.bn1, True, stem_a=a))
            return native.bn_relu_maxpool_fwd(y, self._bn(m.bn1).scale_shift, with_index=tape is not None)
        mask = self._new_mask(y) if tape is not None else None
        z = self._bn_apply(y, m.bn1, True, None, training, mask, blocks)
        if tape is not None:
            tape.append(_Rec(None, y, mask, m.conv1, m.bn1, True, stem_a=a))
        return native.maxpool_fwd(z, with_index=tape is not None)

    def _run_forward(self, x, training, tape):
        m = self.model
        self._sync_shadows()
        p, pool_idx = self._stem(x, training, tape)
        block_tapes = []
        cur = p
        for layer in (m.layer1, m.layer2, m.layer3, m.layer4):
            for blk in layer:
                bt = [] if tape is not None else None
                us = blk.units()
                h = cur
                for conv, bn in us[:-1]:
                    h = self._unit(h, conv, bn, True, None, training, bt)
                if blk.downsample is not None:
                    idn = self._unit(cur, blk.downsample[0], blk.downsample[1], False, None, training, bt)
                else:
                    idn = cur
                conv, bn = us[-1]
                cur = self._unit(h, conv, bn, True, idn, training, bt)
                if tape is not None:
                    block_tapes.append((blk, bt))
        w_fc = self._w16(m.fc.weight)
        if self.fused_head and native.head_supported(cur.shape[1], w_fc.shape[0]):
            logits, pooled = native.head_fwd(cur, w_fc, m.fc.bias.detach())     # fp32 logits, no cast afterwards
        else:
            pooled = native.avgpool_fwd(cur)
            logits = convops.gemm_nt(pooled, w_fc, m.fc.bias.detach()).float()
        if tape is not None:
            tape.append(("head", tape[0].y.shape, pool_idx, block_tapes, cur.shape, pooled))
        return logits

    # ------------------------------------------------------------------ backward units
    def _deliver(self, p, writer):
        """Write p's gradient (writer(out, accumulate)) into its bucket slot, or return a fresh tensor."""
        st = store_of(p)
        if st is not None and p.requires_grad:
            gv, acc = st.grad_slot(p)
            writer(gv, acc)
            st.publish_grad(p, gv)
            return None
        out = torch.empty(p.shape, dtype=torch.float32, device=self.device)
        if p.dim() == 4:
            out = out.contiguous(memory_format=torch.channels_last)
        writer(out, False)
        return out

    def _unit_bwd(self, rec, dz_a, dz_b, need_dx, want_dres, grads, pre_masked=False, fan=None,
                  given_blocks=0, next_rec=None, dz_b_half_res=False, compact_dx=False):
        """Backward of one conv+BN(+ReLU) unit. ``pre_masked``: dz_a already is the complete, ReLU-masked gradient
        of this unit's output (produced by a fan-in dgrad), so it doubles as the residual-branch gradient.
        ``fan`` = (skip gradient, record of the previous block's last unit): fuse the residual fan-in, the ReLU
        backward of this unit's input and the previous block's BatchNorm-backward sums into this unit's dgrad; the
        second return value then is the number of partial rows produced (0: not fused, plain dgrad returned).
        ``given_blocks`` > 0: the dgrad that produced dz_a already left this BatchNorm's backward sums in
        ``native.bn_partial_b`` (the reduce pass is skipped). ``next_rec``: the unit whose BatchNorm consumes this
        unit's dx - its backward sums are produced by this unit's dgrad, the block count is left in
        ``self._bn_blocks``."""
        conv, bn = rec.conv, rec.bn
        st = self._bn(bn)
        # BN backward needs dgamma/dbeta destinations up front (the finalize kernel writes them).
        dg_slot = self._slot(bn.weight)
        db_slot = self._slot(bn.bias)
        if given_blocks > 0:
            dy = native.bn_bwd_partial(dz_a, rec.y, bn.weight, st.mean, st.invstd, dg_slot[0], db_slot[0],
                                       dg_slot[1] and db_slot[1], rec.relu and not pre_masked, st.coef,
                                       native.bn_partial_b(self.device), given_blocks,
                                       self._sync_bn_ctx(bn), self._next_xg_slot(bn), None if pre_masked else rec.mask)
            dres = dz_a if pre_masked else None
        elif pre_masked:
            dy, _ = native.bn_bwd(dz_a, None, None, rec.y, bn.weight, st.mean, st.invstd,
                                  dg_slot[0], db_slot[0], dg_slot[1] and db_slot[1], False, st.coef, False,
                                  self._sync_bn_ctx(bn), self._next_xg_slot(bn), None)
            dres = dz_a
        else:
            dy, dres = native.bn_bwd(dz_a, dz_b, None, rec.y, bn.weight, st.mean, st.invstd,
                                     dg_slot[0], db_slot[0], dg_slot[1] and db_slot[1], rec.relu, st.coef, want_dres,
                                     self._sync_bn_ctx(bn), self._next_xg_slot(bn), rec.mask,
                                     dz_b_half_res=dz_b_half_res)
        self._publish(bn.weight, dg_slot, grads)
        self._publish(bn.bias, db_slot, grads)
        w = conv.weight
        if isinstance(rec.stem_a, tuple):
            _, xp, (n, h, wd) = rec.stem_a
            g = self._deliver(w, lambda out, acc: native.stem_wgrad(xp, dy, out, n, h, wd, acc))
            grads[id(w)] = g
            return None, dres
        if rec.stem_a is not None:
            dy2d = dy.permute(0, 2, 3, 1).reshape(-1, 64)
            convops.gemm_tn(dy2d, rec.stem_a, self.stem_dwp)
            g = self._deliver(w, lambda out, acc: native.stem_unpack_dw(self.stem_dwp, out, acc))
            grads[id(w)] = g
            return None, dres
        s, pd = conv.stride[0], conv.padding[0]
        # dgrad first: delivering the weight gradient may complete a bucket, and the fused all-reduce + SGD kernel
        # that this triggers rewrites the bucket's weights — every reader of the old weights must already be queued
        dx, fan_blocks = None, 0
        if need_dx and fan is not None:
            skip, prev = fan
            dx, fan_blocks = convops.conv_dgrad_fused(dy, self._w16(w), rec.x.shape, s, pd, skip, prev.mask, prev.y,
                                                      self._bn(prev.bn).mean, native.bn_partial_b(self.device))
        self._bn_blocks = 0
        if need_dx and dx is None and next_rec is not None and self.fuse_bn_bwd and next_rec.relu:
            nst = self._bn(next_rec.bn)
            dx, self._bn_blocks = convops.conv_dgrad_bnstats(dy, self._w16(w), rec.x.shape, s, pd, next_rec.y,
                                                             next_rec.mask, nst.mean,
                                                             native.bn_partial_b(self.device))
        if need_dx and dx is None and compact_dx:
            # 1x1 / stride 2 (downsample): the input gradient is non-zero at the even pixels only - compute just those,
            # as a plain GEMM on the half-resolution grid; the consumer (the previous block's BatchNorm backward)
            # reads the compact tensor (native.bn_bwd dz_b_half_res) instead of a zero-filled full-size one
            n_, c_ = rec.x.shape[0], rec.x.shape[1]
            dx = convops.conv_dgrad(dy, self._w16(w), (n_, c_, dy.shape[2], dy.shape[3]), 1, 0)
        if need_dx and dx is None:
            dx = convops.conv_dgrad(dy, self._w16(w), rec.x.shape, s, pd)
        g = self._deliver(w, lambda out, acc: convops.conv_wgrad(rec.x, dy, w.shape, s, pd, out=out, accumulate=acc))
        grads[id(w)] = g
        if fan is not None:
            return dx, fan_blocks
        return dx, dres

    def _slot(self, p):
        st = store_of(p)
        if st is not None and p.requires_grad:
            gv, acc = st.grad_slot(p)
            return (gv, acc, st)
        return (torch.empty(p.shape, dtype=torch.float32, device=self.device), False, None)

    def _publish(self, p, slot, grads):
        gv, _, st = slot
        if st is not None:
            st.publish_grad(p, gv)
            grads[id(p)] = None
        else:
            grads[id(p)] = gv

    def _run_backward(self, tape, dlogits):
        m = self.model
        grads = {}
        head = tape[-1]
        _, stem_shape, pool_idx, block_tapes, feat_shape, pooled = head
        link = self.head_link
        if link.ready:
            # our criterion already left bf16 dlogits and the bias gradient in place (dlogits is a placeholder)
            dl, bias_slot = link.dl, link.slot
            link.ready = False
        else:
            dl = dlogits if dlogits.dtype == torch.bfloat16 else dlogits.to(torch.bfloat16)
            dl, bias_slot = dl.contiguous(), None
        # classifier: bias grad (column sum), weight grad (dl^T @ pooled), input grad (dl @ W)
        dpooled = convops.gemm_nn(dl, self._w16(m.fc.weight))   # before the fc gradients are delivered (see _unit_bwd)
        if bias_slot is not None:
            self._publish(m.fc.bias, bias_slot, grads)
        else:
            grads[id(m.fc.bias)] = self._deliver(m.fc.bias, lambda out, acc: native.colsum_bf16(dl, out, acc))
        grads[id(m.fc.weight)] = self._deliver(m.fc.weight, lambda out, acc: convops.gemm_tn(dl, pooled, out, acc))
        da, db = native.avgpool_bwd(dpooled, feat_shape), None
        rev = list(reversed(block_tapes))
        pre_blocks = 0       # > 0: da is the finished fan-in gradient of the block output, sums are in bn_partial_b
        db_half = False      # db is a compact half-resolution downsample gradient (see _unit_bwd compact_dx)
        for i, (blk, bt) in enumerate(rev):
            n_main = len(blk.units())
            recs_main = bt[:n_main - 1]
            rec_ds = bt[n_main - 1] if blk.downsample is not None else None
            rec_last = bt[-1]
            # each dgrad of the main path also produces the BatchNorm-backward sums of the unit it feeds
            chain = list(recs_main)            # first .. last-but-one unit of the main path
            d, dres = self._unit_bwd(rec_last, da, db, True, True, grads, pre_masked=pre_blocks > 0,
                                     given_blocks=pre_blocks, next_rec=chain[-1], dz_b_half_res=db_half)
            db_half = False
            blocks_first = self._bn_blocks
            for j in range(len(chain) - 1, 0, -1):
                d, _ = self._unit_bwd(chain[j], d, None, True, False, grads, given_blocks=self._bn_blocks,
                                      next_rec=chain[j - 1])
                blocks_first = self._bn_blocks
            # gradient of the block input = dgrad(first conv) + skip-path gradient, masked by the ReLU that produced
            # the input (the previous block's output ReLU): finished inside the first conv's dgrad kernel, which
            # also produces the BatchNorm-backward sums of the previous block's last unit
            skip = dres
            if rec_ds is not None:
                # a strided 1x1 downsample hands its input gradient on in compact form when the consumer is another
                # block's BatchNorm backward with a ReLU mask (always, except in front of the stem's max-pool)
                compact = (self.compact_ds_grad and not self.fuse_fanin and rec_ds.conv.kernel_size == (1, 1)
                           and rec_ds.conv.stride == (2, 2) and rec_ds.conv.padding == (0, 0) and i + 1 < len(rev)
                           and rev[i + 1][1][-1].relu and rev[i + 1][1][-1].mask is not None
                           and rec_ds.x.shape[2] % 2 == 0 and rec_ds.x.shape[3] % 2 == 0)
                skip, _ = self._unit_bwd(rec_ds, dres, None, True, False, grads, compact_dx=compact)
                db_half = compact
            prev_last = rev[i + 1][1][-1] if i + 1 < len(rev) else None
            pre_blocks = 0
            if self.fuse_fanin and prev_last is not None and prev_last.relu and prev_last.mask is not None:
                d, pre_blocks = self._unit_bwd(recs_main[0], d, None, True, False, grads, fan=(skip, prev_last),
                                               given_blocks=blocks_first)
            else:
                d, _ = self._unit_bwd(recs_main[0], d, None, True, False, grads, given_blocks=blocks_first)
            da, db = (d, None) if pre_blocks > 0 else (d, skip)
        dz = native.maxpool_bwd(pool_idx, stem_shape, da, db)
        # fused stem pooling: dead windows scatter nothing, dz already is the gradient of bn1's output
        self._unit_bwd(tape[0], dz, None, False, False, grads, pre_masked=self.fuse_stem_pool)
        return grads

    # ------------------------------------------------------------------ entry
    def __call__(self, x):
        m = self.model
        training = m.training
        if _NVTX:
            torch.cuda.nvtx.range_push("b200ddp.resnet.forward")
            try:
                return self._call(x, training)
            finally:
                torch.cuda.nvtx.range_pop()
        return self._call(x, training)

    def _call(self, x, training):
        m = self.model
        if torch.is_grad_enabled() and any(p.requires_grad for p in self.params):
            out = _EngineFn.apply(self, x, training, *self.params)
            self.head_link.arm(out)
            out._b2_head = self.head_link
            return out
        return self._run_forward(x, training, None)


class _EngineFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, engine, x, training, *params):
        tape = []
        logits = engine._run_forward(x, training, tape)
        ctx.engine = engine
        ctx.tape = tape
        ctx.nparams = len(params)
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        engine = ctx.engine
        tape, ctx.tape = ctx.tape, None
        if tape is None:
            raise RuntimeError("the native engine's activations were already consumed by a previous backward()")
        if _NVTX:
            torch.cuda.nvtx.range_push("b200ddp.resnet.backward")
        try:
            grads = engine._run_backward(tape, dlogits)
        finally:
            if _NVTX:
                torch.cuda.nvtx.range_pop()
        out = []
        for p in engine.params:
            g = grads.get(id(p))
            out.append(g if (g is not None and p.requires_grad) else None)
        return (None, None, None) + tuple(out)
