"""Config 1 of BASELINE.json: ResNet-18 on CIFAR-shaped data, DDP on CPU/gloo, world_size 2.
Our wrapper + native reducer must reproduce torch DDP's loss curve and weights, its collective sequence rules,
checkpoint format and the "eval issues no collectives" design rule."""
import os

import torch
import torch.nn as nn

from _dist_utils import run_ranks


def _train_compare(rank, world, steps, bs):
    import torch.distributed as dist
    from torch.nn.parallel import DistributedDataParallel as TorchDDP
    from pytorch_distributed_tutorials_b200 import optim
    from pytorch_distributed_tutorials_b200.models.resnet import resnet18
    from pytorch_distributed_tutorials_b200.parallel import DistributedDataParallel

    torch.manual_seed(0)
    torch.set_num_threads(2)
    ours = resnet18()
    ref = resnet18()
    ref.load_state_dict(ours.state_dict())
    if rank == 1:  # perturb rank 1: the ctor broadcast from rank 0 must undo it
        with torch.no_grad():
            ours.fc.bias.add_(1.0)
            ours.bn1.running_mean.add_(3.0)
    d_ours = DistributedDataParallel(ours)
    d_ref = TorchDDP(ref)
    o_ours = optim.SGD(d_ours.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-5)
    o_ref = torch.optim.SGD(d_ref.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-5)
    crit = nn.CrossEntropyLoss()
    g = torch.Generator().manual_seed(100 + rank)
    losses = []
    for it in range(steps):
        n = bs if it != steps - 1 else bs // 2 + 3  # ragged last batch
        x = torch.randn(n, 3, 32, 32, generator=g)
        y = torch.randint(0, 10, (n,), generator=g)
        o_ours.zero_grad()
        l1 = crit(d_ours(x), y)
        l1.backward()
        o_ours.step()
        o_ref.zero_grad()
        l2 = crit(d_ref(x), y)
        l2.backward()
        o_ref.step()
        losses.append((l1.item(), l2.item()))
    sd1, sd2 = d_ours.state_dict(), d_ref.state_dict()
    assert list(sd1.keys()) == list(sd2.keys())
    maxdiff = max((sd1[k].float() - sd2[k].float()).abs().max().item() for k in sd1)
    # every rank must hold identical weights
    flat = torch.cat([p.detach().float().reshape(-1) for p in d_ours.parameters()])  # (BN buffers are rank-local)
    other = flat.clone()
    dist.broadcast(other, 0)
    drift = (flat - other).abs().max().item()
    log = d_ours.collective_log()
    # eval forward: zero collectives
    n_before = len(d_ours.collective_log())
    d_ours.eval()
    with torch.no_grad():
        d_ours(torch.randn(2, 3, 32, 32))
    n_after = len(d_ours.collective_log())
    return {"losses": losses, "maxdiff": maxdiff, "drift": drift, "log": log, "eval_colls": n_after - n_before,
            "buckets": d_ours.bucket_sizes(), "keys": list(sd1.keys())[:3], "nkeys": len(sd1)}


def test_matches_torch_ddp_on_gloo():
    steps = 3
    out = run_ranks(_train_compare, world=2, backend="gloo", args=(steps, 8), timeout=600)
    for r in out:
        for l1, l2 in r["losses"]:
            assert abs(l1 - l2) < 1e-3 * max(1.0, abs(l2)), r["losses"]
        assert r["maxdiff"] < 1e-4, r["maxdiff"]
        assert r["drift"] == 0.0
        assert r["eval_colls"] == 0
        assert r["nkeys"] == 122 and r["keys"][0] == "module.conv1.weight"
        # 3 buckets for ResNet-18 at the default cap (1 MiB first bucket = fc, then 25 MiB cap)
        assert len(r["buckets"]) == 3 and r["buckets"][0] == 513000, r["buckets"]
        assert sum(r["buckets"]) == 11689512
        # collective sequence: ctor broadcast, then per step one buffer broadcast + 3 bucket all-reduces in order
        log = r["log"]
        assert log[0][0] == "broadcast"
        per_step = log[1:1 + 4]
        assert [k for k, _ in per_step] == ["broadcast", "allreduce", "allreduce", "allreduce"], log[:8]
        assert [i for _, i in per_step[1:]] == [0, 1, 2]


def _no_sync_and_ckpt(rank, world, tmp):
    import torch.distributed as dist
    from pytorch_distributed_tutorials_b200 import optim
    from pytorch_distributed_tutorials_b200.models.resnet import resnet18
    from pytorch_distributed_tutorials_b200.parallel import DistributedDataParallel
    from pytorch_distributed_tutorials_b200.utils.checkpoint import load_checkpoint, save_checkpoint

    torch.manual_seed(1)
    torch.set_num_threads(2)
    m = DistributedDataParallel(resnet18())
    opt = optim.SGD(m.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-5)
    crit = nn.CrossEntropyLoss()
    g = torch.Generator().manual_seed(5 + rank)
    x, y = torch.randn(4, 3, 32, 32, generator=g), torch.randint(0, 10, (4,), generator=g)
    opt.zero_grad()
    n0 = len(m.collective_log())
    with m.no_sync():
        crit(m(x), y).backward()
    n1 = len([c for c in m.collective_log()[n0:] if c[0] == "allreduce"])
    crit(m(x), y).backward()
    n2 = len([c for c in m.collective_log()[n0:] if c[0] == "allreduce"])
    opt.step()
    path = os.path.join(tmp, "ck", "resnet_distributed.pth")
    if rank == 0:
        save_checkpoint(m, path)
    dist.barrier()
    sd = torch.load(path, weights_only=True)
    m2 = DistributedDataParallel(resnet18())
    load_checkpoint(m2, path)
    sd2 = m2.state_dict()  # (the file holds rank 0's BN buffers; every rank must end up with exactly those)
    same = list(sd2.keys()) == list(sd.keys()) and all(torch.equal(sd[k], sd2[k]) for k in sd)
    import torchvision
    tv = torchvision.models.resnet18()
    tv.load_state_dict({k[len("module."):]: v for k, v in sd.items()})
    dts = sorted({str(v.dtype) for v in sd.values()})
    return {"nosync_colls": n1, "sync_colls": n2, "same": same, "dtypes": dts, "nkeys": len(sd)}


def test_no_sync_and_checkpoint_roundtrip(tmp_path):
    out = run_ranks(_no_sync_and_ckpt, world=2, backend="gloo", args=(str(tmp_path),), timeout=600)
    for r in out:
        assert r["nosync_colls"] == 0
        assert r["sync_colls"] == 3
        assert r["same"]
        assert r["dtypes"] == ["torch.float32", "torch.int64"]
        assert r["nkeys"] == 122


class _OutOfOrder(nn.Module):
    """Registration order (head, tail, mid) differs from the order of use (tail -> mid -> head): reversed
    registration order is the wrong guess for the gradient-ready order."""

    def __init__(self):
        super().__init__()
        self.head = nn.Linear(48, 10)
        self.tail = nn.Linear(20, 64)
        self.mid = nn.Linear(64, 48)

    def forward(self, x):
        return self.head(torch.tanh(self.mid(torch.relu(self.tail(x)))))


def _rebuild(rank, world):
    import torch.distributed as dist
    from torch.nn.parallel import DistributedDataParallel as TorchDDP
    from pytorch_distributed_tutorials_b200 import optim
    from pytorch_distributed_tutorials_b200.parallel import DistributedDataParallel
    torch.manual_seed(0)
    net = _OutOfOrder()
    ref = _OutOfOrder()
    ref.load_state_dict(net.state_dict())
    ours = DistributedDataParallel(net, bucket_cap_mb=0.004)        # ~4 KB buckets: several of them
    theirs = TorchDDP(ref, bucket_cap_mb=0.004)
    o1 = optim.SGD(ours.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    o2 = torch.optim.SGD(theirs.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    names = {id(p): n for n, p in net.named_parameters()}
    before = [names[id(p)] for p in ours.ordered_params]
    g = torch.Generator().manual_seed(5 + rank)
    worst = 0.0
    for _ in range(4):
        x = torch.randn(16, 20, generator=g)
        y = torch.randint(0, 10, (16,), generator=g)
        for m, o in ((ours, o1), (theirs, o2)):
            o.zero_grad()
            nn.functional.cross_entropy(m(x), y).backward()
            o.step()
        worst = max(worst, max((a - b).abs().max().item() for a, b in zip(net.parameters(), ref.parameters())))
    after = [names[id(p)] for p in ours.ordered_params]
    flat = torch.cat([p.detach().reshape(-1) for p in net.parameters()])
    other = flat.clone()
    dist.broadcast(other, 0)
    sd = o1.state_dict()
    return {"before": before, "after": after, "rebuilt": ours.rebuilt_buckets, "maxdiff": worst,
            "ident": bool(torch.equal(flat, other)), "nbuckets": len(ours.store.buckets),
            "order_bcasts": len([c for c in ours.collective_log() if c[0] == "broadcast_order"]),
            "momentum_ok": all("momentum_buffer" in s for s in sd["state"].values())}


def test_buckets_are_rebuilt_in_observed_gradient_ready_order():
    """K5: a generic module starts from reversed registration order; after the first backward the buckets follow
    the order the gradients really arrived in (rank 0's, broadcast), and training still tracks torch DDP."""
    for r in run_ranks(_rebuild, world=2, backend="gloo", timeout=300):
        assert r["before"] == ["mid.bias", "mid.weight", "tail.bias", "tail.weight", "head.bias", "head.weight"], r
        assert r["rebuilt"] and r["order_bcasts"] == 1
        # gradients arrive last layer first: head, then mid, then tail (bias / weight order is autograd's)
        assert [n.split(".")[0] for n in r["after"]] == ["head", "head", "mid", "mid", "tail", "tail"], r
        assert r["ident"] and r["momentum_ok"] and r["nbuckets"] > 1
        assert r["maxdiff"] < 1e-5, r
