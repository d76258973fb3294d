"""Tier-1 unit tests that need no GPU and no ranks: bucket planner vs torch's own, sampler vs torch's, the CLI
contract of the reference (with its documented defects fixed), state_dict parity with torchvision, optimizer
fallback parity, the reducer state machine, and the training entry point end to end on synthetic data."""
import ctypes
import os
import subprocess
import sys

import pytest
import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torch_buckets(sizes_bytes, first, cap):
    import torch.distributed as dist
    tensors = [torch.empty(s // 4) for s in sizes_bytes]
    limits = [first, cap] if first > 0 else [cap]
    idx, _ = dist._compute_bucket_assignment_by_size(tensors, limits)
    return idx


def _autograd_ready_order(m):
    seen = []
    hooks = [p.register_post_accumulate_grad_hook(lambda q: seen.append(q)) for p in m.parameters()]
    m(torch.randn(2, 3, 32, 32)).sum().backward()
    for h in hooks:
        h.remove()
    return seen


@pytest.mark.parametrize("name,ngold", [("resnet18", [513000, 7213056, 3963456]),
                                        ("resnet50", [2049000, 7875584, 6563840, 6637568, 2431040])])
def test_bucket_plan_matches_torch_and_golden_tables(name, ngold):
    """steady-state DDP buckets (1 MiB first bucket, 25 MiB cap). The golden tables were measured with
    torch DDP, i.e. in autograd's gradient-ready order of the stock path; our planner must reproduce them on that
    order, and agree with torch's own planner on the engine's static order too."""
    from pytorch_distributed_tutorials_b200.models.resnet import MODELS
    from pytorch_distributed_tutorials_b200.parallel.flat import plan_buckets
    m = MODELS[name]()
    for order, gold in ((_autograd_ready_order(m), ngold), (m.grad_ready_order(), None)):
        sizes = [p.numel() * 4 for p in order]
        ours = plan_buckets(sizes, 1 << 20, 25 << 20)
        ref = _torch_buckets(sizes, 1 << 20, 25 << 20)
        groups = [[i for i, b in enumerate(ours) if b == k] for k in range(max(ours) + 1)]
        assert groups == ref
        elems = [sum(order[i].numel() for i in g) for g in groups]
        if gold is not None:
            assert elems == gold
        assert len(elems) == len(ngold) and elems[0] == ngold[0] and sum(elems) == sum(ngold)


def test_bucket_plan_explicit_cap_has_no_small_first_bucket():
    from pytorch_distributed_tutorials_b200.models.resnet import resnet152
    from pytorch_distributed_tutorials_b200.parallel.flat import plan_buckets
    m = resnet152()
    sizes = [p.numel() * 4 for p in m.grad_ready_order()]
    ours = plan_buckets(sizes, -1, 100 << 20)
    elems = [sum(s // 4 for s, b in zip(sizes, ours) if b == k) for k in range(max(ours) + 1)]
    assert len(elems) == 3 and sum(elems) == 60192808
    assert elems[0] >= 26214400 and elems[0] - 26214400 < 2400000   # overshoots the cap by at most one tensor
    ref = _torch_buckets(sizes, -1, 100 << 20)
    assert [len(g) for g in ref] == [ours.count(k) for k in range(3)]


def test_engine_ready_order_is_what_autograd_produces():
    """The static order the DDP wrapper plans with must be (up to ties inside a unit) the order in which autograd
    delivers gradients for the stock path, so that buckets fill front to back."""
    from pytorch_distributed_tutorials_b200.models.resnet import resnet18
    m = resnet18()
    seen = []
    for p in m.parameters():
        p.register_post_accumulate_grad_hook(lambda q: seen.append(id(q)))
    m(torch.randn(2, 3, 32, 32)).sum().backward()
    static = [id(p) for p in m.grad_ready_order()]
    assert set(seen) == set(static)
    pos = {pid: i for i, pid in enumerate(seen)}
    # bucket-level agreement: walking our order, autograd positions never jump back by more than one block
    worst = max(max(0, max(pos[q] for q in static[:i + 1]) - i) for i in range(len(static)))
    assert worst <= 12, worst


@pytest.mark.parametrize("n,world", [(203, 2), (50000, 8), (10, 4), (7, 8)])
def test_sampler_matches_torch(n, world):
    from torch.utils.data.distributed import DistributedSampler as TorchSampler
    from pytorch_distributed_tutorials_b200.parallel.sampler import DistributedSampler
    ds = list(range(n))
    for rank in range(world):
        for epoch in (0, 3):
            a = DistributedSampler(ds, num_replicas=world, rank=rank)
            b = TorchSampler(ds, num_replicas=world, rank=rank)
            a.set_epoch(epoch)
            b.set_epoch(epoch)
            assert list(a) == list(b) and len(a) == len(b)
    a = DistributedSampler(ds, num_replicas=world, rank=0, shuffle=False, drop_last=True)
    b = TorchSampler(ds, num_replicas=world, rank=0, shuffle=False, drop_last=True)
    assert list(a) == list(b)


def test_cli_contract_of_the_reference_with_defects_fixed():
    from pytorch_distributed_tutorials_b200.train import build_parser
    p = build_parser()
    d = p.parse_args([])
    assert (d.num_epochs, d.batch_size, d.learning_rate, d.seed) == (10000, 256, 0.01, 0)
    assert (d.model_dir, d.model_filename, d.resume, d.local_rank) == ("saved_models", "resnet_distributed.pth",
                                                                       False, None)
    assert p.parse_args(["--learning_rate", "0.1"]).learning_rate == 0.1            # D4: was type=int
    assert p.parse_args(["--local_rank", "3"]).local_rank == 3
    assert p.parse_args(["--local-rank=5"]).local_rank == 5                         # D5: torch>=2 launch spelling
    assert p.parse_args(["--batch-size", "64", "--resume"]).resume is True


def test_state_dict_parity_with_torchvision():
    import torchvision
    from pytorch_distributed_tutorials_b200.models.resnet import MODELS
    for name in ("resnet18", "resnet50", "resnet152"):
        ours, tv = MODELS[name](), getattr(torchvision.models, name)()
        a, b = ours.state_dict(), tv.state_dict()
        assert list(a.keys()) == list(b.keys())
        assert all(a[k].shape == b[k].shape and a[k].dtype == b[k].dtype for k in a)
        tv.load_state_dict(a)
    assert sum(p.numel() for p in MODELS["resnet50"]().parameters()) == 25557032


def test_reducer_state_machine_launches_in_order():
    from pytorch_distributed_tutorials_b200.ops import _lib
    L = _lib.lib()
    L.b2_reducer_create.restype = ctypes.c_void_p
    bucket_of = (ctypes.c_int * 6)(0, 0, 1, 1, 2, 2)
    r = ctypes.c_void_p(L.b2_reducer_create(6, bucket_of, 3))
    out = (ctypes.c_int * 3)()
    L.b2_reducer_reset(r)
    assert L.b2_reducer_mark_ready(r, 4, out) == 0          # bucket 2 filling first launches nothing
    assert L.b2_reducer_mark_ready(r, 5, out) == 0
    assert L.b2_reducer_mark_ready(r, 0, out) == 0
    assert L.b2_reducer_mark_ready(r, 2, out) == 0
    assert L.b2_reducer_mark_ready(r, 1, out) == 1 and out[0] == 0
    assert L.b2_reducer_mark_ready(r, 1, out) == -1         # marked twice
    assert L.b2_reducer_mark_ready(r, 3, out) == 2 and list(out[:2]) == [1, 2]   # unblocks 1 and the waiting 2
    assert L.b2_reducer_unlaunched(r) == 0
    order = (ctypes.c_int * 6)()
    assert L.b2_reducer_ready_order(r, order) == 6 and list(order) == [4, 5, 0, 2, 1, 3]
    L.b2_reducer_destroy(r)


def test_sgd_flat_store_cpu_falls_back_to_torch_math():
    from pytorch_distributed_tutorials_b200 import optim
    from pytorch_distributed_tutorials_b200.parallel.flat import FlatParameterStore
    torch.manual_seed(0)
    a, b = nn.Linear(7, 5), nn.Linear(7, 5)
    b.load_state_dict(a.state_dict())
    params = list(a.parameters())
    FlatParameterStore(params, [0, 0], "cpu")
    o1 = optim.SGD(a.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-3)
    o2 = torch.optim.SGD(b.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-3)
    x = torch.randn(3, 7)
    for _ in range(3):
        for m, o in ((a, o1), (b, o2)):
            o.zero_grad()
            m(x).pow(2).sum().backward()
            o.step()
    assert all(torch.allclose(p, q) for p, q in zip(a.parameters(), b.parameters()))


def test_train_entry_point_end_to_end_gloo(tmp_path):
    """BASELINE config 1 through the real entry point: torchrun, 2 ranks, gloo, synthetic CIFAR-shaped data,
    evaluation + checkpoint at epoch 0, then --resume from that checkpoint."""
    env = dict(os.environ, OMP_NUM_THREADS="2", B200DDP_NO_AUTOBUILD="1")
    base = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
            "127.0.0.1", "--master-port", "29871", os.path.join(ROOT, "resnet", "main.py"), "--backend", "gloo",
            "--synthetic", "--dataset-size", "48", "--batch-size", "8", "--num-workers", "0", "--num_epochs", "1",
            "--model_dir", str(tmp_path / "saved_models"), "--max-steps", "2"]
    r = subprocess.run(base, env=env, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "Local Rank: 0, Epoch: 0, Training ..." in r.stdout and "Local Rank: 1, Epoch: 0, Training ..." in r.stdout
    assert "Epoch: 0, Accuracy:" in r.stdout
    ck = tmp_path / "saved_models" / "resnet_distributed.pth"
    sd = torch.load(ck, weights_only=True)
    assert len(sd) == 122 and all(k.startswith("module.") for k in sd)
    r2 = subprocess.run(base + ["--resume"], env=env, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r2.returncode == 0, r2.stdout[-2000:] + r2.stderr[-4000:]


def test_tail_bucket_split_keeps_order_and_sizes():
    """The exposed-tail cut only re-labels the parameters at the END of the last bucket: order, totals and every
    other bucket are untouched, and what is cut off is at most the requested size."""
    from pytorch_distributed_tutorials_b200.models.resnet import resnet50
    from pytorch_distributed_tutorials_b200.parallel.ddp import split_tail_bucket
    from pytorch_distributed_tutorials_b200.parallel.flat import plan_buckets
    m = resnet50()
    sizes = [p.numel() * 4 for p in m.grad_ready_order()]
    plan = plan_buckets(sizes, 1 << 20, 25 << 20)
    cut = split_tail_bucket(plan, sizes, 1 << 20)
    cut2 = split_tail_bucket(cut, sizes, 1 << 16)
    for a, b in ((plan, cut), (cut, cut2)):
        assert len(a) == len(b) and max(b) == max(a) + 1
        assert all(x <= y for x, y in zip(b, b[1:]))                      # still contiguous runs in ready order
        last = max(a)
        assert [x for x in b if x < last] == [x for x in a if x < last]   # earlier buckets untouched
    tail_bytes = sum(s for s, k in zip(sizes, cut2) if k == max(cut2))
    mid_bytes = sum(s for s, k in zip(sizes, cut2) if k == max(cut2) - 1)
    assert 0 < tail_bytes <= 1 << 16 and 0 < mid_bytes <= 1 << 20
    assert split_tail_bucket(plan, sizes, 0) == plan                       # switched off
    assert split_tail_bucket([0, 0], [8, 8], 4) == [0, 0]                  # nothing small enough to cut off


def test_device_prefetcher_ring_preserves_order_and_values_on_cpu():
    """The prefetcher's ring / look-ahead bookkeeping (device side is plain tensors on CPU): every batch comes out
    once, in order, including a ragged last batch that does not fit the ring's buffers."""
    from pytorch_distributed_tutorials_b200.utils.data import DevicePrefetcher
    batches = [(torch.full((4, 3), float(i)), torch.full((4,), i)) for i in range(7)] + \
              [(torch.full((2, 3), 7.0), torch.full((2,), 7))]
    for depth in (1, 2, 3):
        out = [(x.clone(), y.clone()) for x, y in DevicePrefetcher(batches, "cpu", depth=depth)]
        assert len(out) == 8
        for i, (x, y) in enumerate(out):
            assert x.shape[0] == (4 if i < 7 else 2) and bool((x == float(i)).all()) and bool((y == i).all())


def test_wgrad_flags_follow_the_determinism_switch():
    """b2_conv_wgrad's flag word: bit 0 = accumulate, bit 1 = "sum the split-K partials in a fixed order". The second
    bit follows torch.use_deterministic_algorithms (and B200DDP_WGRAD_RED=0); the launch count the engine reports
    drops from 2 (GEMM + reduce kernel) to 1 when the sums form in L2."""
    import torch
    from pytorch_distributed_tutorials_b200.ops import conv
    was = torch.are_deterministic_algorithms_enabled()
    try:
        torch.use_deterministic_algorithms(False)
        if conv._WGRAD_RED:
            assert conv._wgrad_flags(False) == (0, 1) and conv._wgrad_flags(True) == (1, 1)
        torch.use_deterministic_algorithms(True, warn_only=True)
        assert conv._wgrad_flags(False) == (2, 2) and conv._wgrad_flags(True) == (3, 2)
    finally:
        torch.use_deterministic_algorithms(was)

