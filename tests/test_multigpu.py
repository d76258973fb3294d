"""Multi-GPU checks (>= 2 B200s on one node): P2P all-reduce / broadcast kernels against torch.distributed (NCCL),
the fused reduce+SGD epilogue against unfused torch ops, the DDP wrapper against torch DDP, and SyncBatchNorm with
in-kernel cross-GPU statistics against torch.nn.SyncBatchNorm."""
import pytest
import torch

from _dist_utils import run_ranks

pytestmark = pytest.mark.gpu


def _need(n):
    if torch.cuda.device_count() < n:
        pytest.skip("needs %d GPUs" % n)


def _collectives(rank, world, skew_us=0):
    import torch.distributed as dist
    from pytorch_distributed_tutorials_b200.parallel.comm import P2PComm
    from pytorch_distributed_tutorials_b200.parallel.flat import Bucket
    dev = torch.device("cuda", rank)
    comm = P2PComm(dev, None, 600 << 20)
    # skew injection: a pseudo-random subset of the CTAs sleeps before / between the phases of every collective
    # kernel, so a kernel that consumes what a differently-numbered CTA of a peer produced reads stale data
    comm.set_debug_skew_us(skew_us)
    res = {"transport": comm.window.transport, "multicast": bool(comm.mc_base)}
    sizes = [128, 4096, 129 * 1024, 513000 + 24, 7213056, 25 * 1024 * 1024 // 4]
    algos = ["oneshot", "twoshot"] + (["multimem"] if comm.mc_base else [])

    class FakeBucket:
        pass

    for n in sizes:
        n = (n + 63) // 64 * 64
        b = FakeBucket()
        b.index, b.numel = 0, n
        b.g32 = comm.grad_alloc(n)
        for algo in algos:
            if algo == "oneshot" and n > 524288:
                continue
            g = torch.Generator(device=dev).manual_seed(1000 * rank + n % 977)
            local = torch.randn(n, device=dev, generator=g)
            b.g32.copy_(local)
            ref = local.clone()
            dist.all_reduce(ref)
            ref /= world
            torch.cuda.synchronize()
            dist.barrier()
            comm.allreduce_bucket(b, sgd=None, algo=algo)
            comm.finalize()
            torch.cuda.synchronize()
            err = (b.g32 - ref).abs().max().item()
            res["%s_%d" % (algo, n)] = err
            # every rank must hold bit-identical results
            mine = b.g32.clone()
            other = mine.clone()
            dist.broadcast(other, 0)
            res["%s_%d_ident" % (algo, n)] = bool(torch.equal(mine, other))
            dist.barrier()
        # fused SGD epilogue vs unfused torch math (two steps: first-step buffer init, then momentum), with every
        # algorithm that can carry it - multimem is the default for world > 2 (B2_AR_AUTO)
        sgd_algos = ["twoshot" if n > 4096 else "oneshot"] + (["multimem"] if comm.mc_base and n * 4 >= 32768 else [])
        for wire in ("fp32", "bf16"):
            comm.wire = wire
            for algo in sgd_algos:
                p = torch.randn(n, device=dev, generator=torch.Generator(device=dev).manual_seed(7))
                b.p32, b.mom = p.clone(), torch.zeros(n, device=dev)
                b.shadow = torch.empty(n, device=dev, dtype=torch.bfloat16)
                pr, mr = p.clone(), None
                for step in range(2):
                    g = torch.Generator(device=dev).manual_seed(31 * rank + step)
                    local = torch.randn(n, device=dev, generator=g)
                    b.g32.copy_(local)
                    if wire == "fp32":
                        avg = local.clone()
                        dist.all_reduce(avg)
                        avg /= world
                    else:
                        parts = [torch.empty_like(local) for _ in range(world)]
                        dist.all_gather(parts, local)
                        acc = torch.zeros_like(local)
                        for q in parts:
                            acc += q.to(torch.bfloat16).float()
                        avg = (acc / world).to(torch.bfloat16).float()
                    gg = avg + 1e-5 * pr
                    mr = gg.clone() if mr is None else 0.9 * mr + gg
                    pr = pr - 0.01 * mr
                    torch.cuda.synchronize()
                    dist.barrier()
                    comm.allreduce_bucket(b, sgd=(0.01, 0.9, 1e-5, step == 0), algo=algo)
                    comm.finalize()
                    torch.cuda.synchronize()
                tag = "sgd%s_%s_%d" % ("16" if wire == "bf16" else "", algo, n)
                res[tag] = (b.p32 - pr).abs().max().item()
                res[tag + "_shadow"] = bool(torch.equal(b.shadow, b.p32.to(torch.bfloat16)))
                res[tag + "_grad"] = (b.g32 - avg).abs().max().item()
                mine = b.p32.clone()
                other = mine.clone()
                dist.broadcast(other, 0)
                res[tag + "_ident"] = bool(torch.equal(mine, other))
        comm.wire = "fp32"
    # bf16 wire format: fp32 accumulation of bf16-rounded inputs in rank order, rounded to bf16 once
    comm16 = comm
    comm16.wire = "bf16"
    for n in [4096, 513024, 7213056]:
        b = FakeBucket()
        b.index, b.numel = 0, n
        b.g32 = comm.grad_alloc(n)
        for algo in algos:
            if algo == "oneshot" and n > 524288:
                continue
            g = torch.Generator(device=dev).manual_seed(77 * rank + n % 13)
            local = torch.randn(n, device=dev, generator=g)
            b.g32.copy_(local)
            parts = [torch.empty_like(local) for _ in range(world)]
            dist.all_gather(parts, local)
            acc = torch.zeros_like(local)
            for q in parts:
                acc += q.to(torch.bfloat16).float()
            ref = (acc / world).to(torch.bfloat16).float()
            torch.cuda.synchronize()
            dist.barrier()
            comm16.allreduce_bucket(b, sgd=None, algo=algo)
            comm16.finalize()
            torch.cuda.synchronize()
            res["bf16wire_%s_%d" % (algo, n)] = ((b.g32 - ref).abs().max() / ref.abs().max()).item()
            dist.barrier()
    comm.wire = "fp32"
    # broadcast of arbitrary tensors through the staging window
    t = [torch.full((1000, 3), float(rank + 1), device=dev), torch.arange(10, device=dev) + 100 * rank]
    comm.broadcast_(t, src=0)
    torch.cuda.synchronize()
    res["bcast"] = bool((t[0] == 1.0).all().item() and (t[1] == torch.arange(10, device=dev)).all().item())
    return res


@pytest.mark.parametrize("world,skew_us", [(2, 0), (2, 300), (4, 300), (8, 300)])
def test_p2p_collectives_match_nccl(world, skew_us):
    """skew_us > 0: same checks with CTAs of one rank pushed far apart in progress (csrc/xgpu.cuh debug_skew) - the
    cross-CTA race of the multimem reduce+SGD kernel found in review passes the skew-free run by lock-step luck."""
    _need(world)
    out = run_ranks(_collectives, world=world, backend="nccl", args=(skew_us,), timeout=900)
    for r in out:
        for k, v in r.items():
            if k.endswith("_ident") or k.endswith("_shadow") or k == "bcast":
                assert v is True, (k, r)
            elif k.startswith(("oneshot", "twoshot", "multimem")):
                assert v < 1e-5, (k, v)
            elif k.startswith("sgd"):
                # fp32 wire: exact up to fp32 rounding; bf16 wire (values pre-scaled): the averaged gradient may
                # differ by one bf16 ulp where the in-switch reduction order differs from the reference order
                if k.startswith("sgd_"):
                    assert v < 1e-5, (k, v)
                elif k.endswith("_grad"):
                    assert v < 5e-2, (k, v)          # one bf16 ulp of an O(1) gradient
                else:
                    assert v < 2e-3, (k, v)          # lr * (1 + mu + 1) * one ulp
            elif k.startswith("bf16wire_"):
                assert v < 1e-2, (k, v)      # at most one bf16 ulp (multimem's in-switch order may differ)


def _ddp_train(rank, world, sync_bn):
    import copy
    import torch.distributed as dist
    import torch.nn as nn
    from torch.nn.parallel import DistributedDataParallel as TorchDDP
    from pytorch_distributed_tutorials_b200 import optim
    from pytorch_distributed_tutorials_b200.models.resnet import resnet18
    from pytorch_distributed_tutorials_b200.ops.loss import CrossEntropyLoss
    from pytorch_distributed_tutorials_b200.parallel import DistributedDataParallel, convert_sync_batchnorm
    dev = torch.device("cuda", rank)
    torch.manual_seed(0)
    base = resnet18()
    ref = copy.deepcopy(base)
    if sync_bn:
        base = convert_sync_batchnorm(base)
        ref = nn.SyncBatchNorm.convert_sync_batchnorm(ref)
    base, ref = base.to(dev), ref.to(dev)
    if rank == 1:
        with torch.no_grad():
            base.fc.bias.add_(1.0)   # ctor broadcast from rank 0 must undo this
    ours = DistributedDataParallel(base, device_ids=[rank])
    theirs = TorchDDP(ref, device_ids=[rank])
    # small learning rate: five steps on one batch must not memorise it, or the two bf16 trajectories (same maths,
    # different rounding) drift apart chaotically and the comparison tests luck instead of kernels
    o1 = optim.SGD(ours.parameters(), lr=0.005, momentum=0.9, weight_decay=1e-5, fuse_with_backward=True)
    o2 = torch.optim.SGD(theirs.parameters(), lr=0.005, momentum=0.9, weight_decay=1e-5)
    crit = CrossEntropyLoss()
    g = torch.Generator(device=dev).manual_seed(10 + rank)
    x = torch.randn(32, 3, 32, 32, device=dev, generator=g)
    y = torch.randint(0, 10, (32,), device=dev, generator=g)
    l1s, l2s = [], []
    for _ in range(5):
        o1.zero_grad()
        l1 = crit(ours(x), y)
        l1.backward()
        o1.step()
        o2.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = theirs(x)
        l2 = nn.functional.cross_entropy(out.float(), y)
        l2.backward()
        o2.step()
        l1s.append(l1.item())
        l2s.append(l2.item())
    torch.cuda.synchronize()
    flat = torch.cat([p.detach().reshape(-1) for p in ours.parameters()])
    other = flat.clone()
    dist.broadcast(other, 0)
    log = ours.collective_log()
    return {"ours": l1s, "theirs": l2s, "ident": bool(torch.equal(flat, other)), "backend": ours.backend,
            "n_allreduce": len([c for c in log if c[0] == "allreduce"]), "exposed": ours.exposed_comm_ms(),
            "buckets": ours.bucket_sizes()}


@pytest.mark.parametrize("world,sync_bn", [(2, False), (2, True), (4, False), (8, False), (8, True)])
def test_ddp_p2p_trains_like_torch_ddp(world, sync_bn):
    _need(world)
    out = run_ranks(_ddp_train, world=world, backend="nccl", args=(sync_bn,), timeout=600)
    for r in out:
        assert r["backend"] == "p2p"
        assert r["ident"], "replicas drifted apart"
        # torch's bucketing rule in the engine's gradient-ready order, the last bucket cut twice more so that only
        # the stem's gradients (<= 64 KiB) wait for the very end of backward
        bk = r["buckets"]
        assert bk[0] == 513000 and sum(bk) == 11689512 and len(bk) == 5 and bk[-1] * 4 <= 64 << 10 \
            and bk[-2] * 4 <= 1 << 20, bk
        assert r["n_allreduce"] == 5 * len(bk)
        assert r["exposed"] is not None and r["exposed"] >= 0.0
        assert r["ours"][-1] < r["ours"][0]
        # against torch DDP under bf16 autocast, step by step: 5 % on the first two steps, 10 % afterwards (two bf16
        # implementations of the same maths drift apart as training proceeds: measured 0.6 % / 0.8 % / 3.2 % / 5.5 % /
        # 8 % over five steps with SyncBatchNorm; the old bound was 35 % on the last step only)
        for i, (a_, b_) in enumerate(zip(r["ours"], r["theirs"])):
            assert abs(a_ - b_) < max((0.05 if i < 2 else 0.10) * b_, 0.03), r


def _syncbn_stats(rank, world):
    import torch.distributed as dist
    import torch.nn as nn
    from pytorch_distributed_tutorials_b200.ops import native
    from pytorch_distributed_tutorials_b200.parallel.comm import P2PComm
    dev = torch.device("cuda", rank)
    comm = P2PComm(dev, None, 64 << 20)
    xg = comm.window.ctx(P2PComm.PAD_COMPUTE)
    area = 2 * 2048 + 256
    slot0 = comm.window.alloc(2 * 4 * area) // 4     # two exchange areas, alternated call by call (one barrier each)
    calls = [0]

    def next_slot():
        calls[0] += 1
        return slot0 + (calls[0] & 1) * area
    out = {}
    for C, H, N in [(64, 9, 3 + rank), (512, 5, 2), (2048, 3, 4)]:
        torch.manual_seed(100 + rank)
        y = (torch.randn(N, C, H, H, device=dev) * (1 + rank) + rank).to(torch.bfloat16).contiguous(
            memory_format=torch.channels_last)
        gamma, beta = torch.rand(C, device=dev) + 0.5, torch.randn(C, device=dev)
        dist.broadcast(gamma, 0)
        dist.broadcast(beta, 0)
        rm, rv, nbt = torch.zeros(C, device=dev), torch.ones(C, device=dev), torch.tensor(0, device=dev)
        mean, invstd = torch.empty(C, device=dev), torch.empty(C, device=dev)
        ss, coef = torch.empty(2 * C, device=dev), torch.empty(3 * C, device=dev)
        z = native.bn_fwd_train(y, None, gamma, beta, rm, rv, nbt, 0.1, 1e-5, True, mean, invstd, ss, xg, next_slot())
        ref = nn.SyncBatchNorm(C).to(dev)
        with torch.no_grad():
            ref.weight.copy_(gamma)
            ref.bias.copy_(beta)
        yf = y.float().requires_grad_(True)
        zr = torch.relu(ref(yf))
        out["fwd_%d" % C] = ((z.float() - zr).abs().max() / zr.abs().max()).item()
        out["rm_%d" % C] = (rm - ref.running_mean).abs().max().item()
        out["rv_%d" % C] = ((rv - ref.running_var).abs().max() / ref.running_var.abs().max()).item()
        dz = torch.randn_like(zr).to(torch.bfloat16)
        dg, db = torch.empty(C, device=dev), torch.empty(C, device=dev)
        dy, _ = native.bn_bwd(dz, None, z, y, gamma, mean, invstd, dg, db, False, True, coef, False, xg, next_slot())
        zr.backward(dz.float() * (z.float() > 0) / 1.0)
        out["bwd_%d" % C] = ((dy.float() - yf.grad).abs().max() / yf.grad.abs().max()).item()
        out["dg_%d" % C] = ((dg - ref.weight.grad).abs().max() / ref.weight.grad.abs().max()).item()
        torch.cuda.synchronize()
        dist.barrier()
    return out


def test_syncbn_in_kernel_stats_match_torch():
    _need(2)
    outs = run_ranks(_syncbn_stats, world=2, backend="nccl", timeout=600)
    for r in outs:
        for k, v in r.items():
            assert v < 3e-2, (k, v, r)


def _graph_ddp(rank, world):
    import copy
    import torch.distributed as dist
    from pytorch_distributed_tutorials_b200 import optim
    from pytorch_distributed_tutorials_b200.models.resnet import resnet18
    from pytorch_distributed_tutorials_b200.ops.loss import CrossEntropyLoss
    from pytorch_distributed_tutorials_b200.parallel import DistributedDataParallel
    from pytorch_distributed_tutorials_b200.utils.graph import GraphedTrainStep
    torch.use_deterministic_algorithms(True, warn_only=True)   # graph == eager bit for bit needs fixed-order wgrad sums
    dev = torch.device("cuda", rank)
    torch.manual_seed(0)
    base = resnet18().to(dev)
    twin = copy.deepcopy(base)
    g = torch.Generator(device=dev).manual_seed(10 + rank)
    xs = [torch.randn(16, 3, 32, 32, device=dev, generator=g) for _ in range(2)]
    ys = [torch.randint(0, 10, (16,), device=dev, generator=g) for _ in range(2)]
    out = []
    for net, graph in ((base, True), (twin, False)):
        ddp = DistributedDataParallel(net, device_ids=[rank])
        opt = optim.SGD(ddp.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-5, fuse_with_backward=True)
        step = GraphedTrainStep(ddp, CrossEntropyLoss(), opt, warmup=2, enabled=graph)
        losses = [float(step(xs[i % 2], ys[i % 2]).item()) for i in range(7)]
        torch.cuda.synchronize()
        dist.barrier()
        out.append((losses, torch.cat([p.detach().reshape(-1) for p in ddp.parameters()]).clone(), step.captures,
                    ddp.exposed_comm_ms(last=3)))
    (lg, pg, caps, exposed), (le, pe, _, _) = out
    other = pg.clone()
    dist.broadcast(other, 0)
    return {"captures": caps, "same_as_eager": bool(torch.equal(pg, pe)), "ident": bool(torch.equal(pg, other)),
            "lg": lg, "le": le, "exposed": exposed}


@pytest.mark.parametrize("world", [2, 4, 8])
def test_graph_captured_step_with_p2p_allreduce_matches_eager(world):
    _need(world)
    outs = run_ranks(_graph_ddp, world=world, backend="b200", timeout=600)   # gloo control plane: no NCCL at all
    for r in outs:
        assert r["captures"] == 1
        assert r["ident"], "replicas drifted apart under graph replay"
        assert r["same_as_eager"], (r["lg"], r["le"])
        # exposed communication is measured by device timestamps inside the replayed graph (not only in eager steps)
        assert r["exposed"] is not None and 0.0 <= r["exposed"] < 50.0, r["exposed"]


def _fault(rank, world):
    """Rank 1 dies mid-training; rank 0's next collective waits for it, the device-side spin bound trips and the
    process must surface an error instead of hanging."""
    import os
    import time
    from pytorch_distributed_tutorials_b200 import optim
    from pytorch_distributed_tutorials_b200.models.resnet import resnet18
    from pytorch_distributed_tutorials_b200.ops.loss import CrossEntropyLoss
    from pytorch_distributed_tutorials_b200.parallel import DistributedDataParallel
    dev = torch.device("cuda", rank)
    torch.manual_seed(0)
    ddp = DistributedDataParallel(resnet18().to(dev), device_ids=[rank])
    ddp.comm.set_spin_timeout_ms(2000)
    opt = optim.SGD(ddp.parameters(), lr=0.01, momentum=0.9)
    crit = CrossEntropyLoss()
    x = torch.randn(8, 3, 32, 32, device=dev)
    y = torch.randint(0, 10, (8,), device=dev)
    t0 = time.time()
    try:
        for it in range(4):
            if rank == 1 and it == 2:
                torch.cuda.synchronize()
                os._exit(0)          # simulated crash: no clean-up, no goodbye
            opt.zero_grad()
            crit(ddp(x), y).backward()
            opt.step()
            torch.cuda.synchronize()
        return {"error": None, "seconds": time.time() - t0}
    except Exception as e:  # noqa: BLE001
        return {"error": type(e).__name__ + ": " + str(e)[:200], "seconds": time.time() - t0}


def test_peer_failure_is_detected_not_hung():
    _need(2)
    import torch.multiprocessing as mp
    from _dist_utils import _entry, free_port
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=_entry, args=(r, 2, port, _fault, "b200", (), q)) for r in range(2)]
    for p in procs:
        p.start()
    rank, status, out = q.get(timeout=120)       # only rank 0 reports; rank 1 has exited
    for p in procs:
        p.join(timeout=30)
        if p.is_alive():
            p.terminate()
    assert rank == 0
    # either our wrapper returned the CUDA error, or the harness caught it: both mean "detected, not hung"
    msg = out["error"] if status == "ok" else out
    assert msg is not None and ("CUDA" in msg or "launch failure" in msg or "trap" in msg.lower() or "error" in msg.lower()), out
    if status == "ok":
        assert out["seconds"] < 60


def _generic_module(rank, world):
    """An arbitrary nn.Module (no native engine): autograd hooks feed the same buckets / P2P kernels."""
    import copy
    import torch.distributed as dist
    import torch.nn as nn
    from torch.nn.parallel import DistributedDataParallel as TorchDDP
    from pytorch_distributed_tutorials_b200 import optim
    from pytorch_distributed_tutorials_b200.parallel import DistributedDataParallel
    dev = torch.device("cuda", rank)
    torch.manual_seed(0)
    net = nn.Sequential(nn.Linear(64, 256), nn.ReLU(), nn.Linear(256, 256), nn.Tanh(), nn.Linear(256, 10)).to(dev)
    ref = copy.deepcopy(net)
    ours = DistributedDataParallel(net, device_ids=[rank], bucket_cap_mb=0.1)
    theirs = TorchDDP(ref, device_ids=[rank])
    o1 = optim.SGD(ours.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4, fuse_with_backward=True)
    o2 = torch.optim.SGD(theirs.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    g = torch.Generator(device=dev).manual_seed(3 + rank)
    worst = 0.0
    for _ in range(5):
        x = torch.randn(32, 64, device=dev, generator=g)
        y = torch.randint(0, 10, (32,), device=dev, generator=g)
        for m, o in ((ours, o1), (theirs, o2)):
            o.zero_grad()
            nn.functional.cross_entropy(m(x), y).backward()
            o.step()
        torch.cuda.synchronize()
        worst = max(worst, max((a - b).abs().max().item() for a, b in zip(ours.parameters(), theirs.parameters())))
    flat = torch.cat([p.detach().reshape(-1) for p in ours.parameters()])
    other = flat.clone()
    dist.broadcast(other, 0)
    return {"maxdiff": worst, "ident": bool(torch.equal(flat, other)), "buckets": len(ours.store.buckets)}


def test_generic_module_through_p2p_backend_matches_torch_ddp():
    _need(2)
    for r in run_ranks(_generic_module, world=2, backend="nccl", timeout=600):
        assert r["ident"] and r["buckets"] > 1
        assert r["maxdiff"] < 1e-4, r


def _teardown(rank, world):
    """Construct / destroy the wrapper many times: the symmetric windows (VMM allocations mapped on every peer,
    multicast binding) must really be released - device memory held outside torch's caching allocator stays flat."""
    import gc
    import torch.distributed as dist
    from pytorch_distributed_tutorials_b200.models.resnet import resnet18
    from pytorch_distributed_tutorials_b200.parallel import DistributedDataParallel
    from pytorch_distributed_tutorials_b200.parallel.ddp import _LIVE
    dev = torch.device("cuda", rank)
    x = torch.randn(4, 3, 32, 32, device=dev)
    outside, errors, live, win_mib = [], [], [], 0
    for it in range(12):
        net = resnet18().to(dev)
        ddp = DistributedDataParallel(net, device_ids=[rank])
        ddp(x).sum().backward()
        torch.cuda.synchronize()
        dist.barrier()
        res = ddp.comm.window._res
        win_mib = res["nbytes"] >> 20
        if it % 2 == 0:
            ddp.close()                       # explicit
            rm = net.bn1.running_mean.clone() # the module survives its wrapper: buffers were moved out of the window
            assert torch.isfinite(rm).all()
        del ddp, net                          # odd iterations: garbage collection alone must release the window
        gc.collect()
        torch.cuda.synchronize()
        dist.barrier()
        errors += res.get("errors", [])
        live.append((len(_LIVE), bool(res.get("released"))))
        torch.cuda.empty_cache()
        free, total = torch.cuda.mem_get_info(dev)
        outside.append(((total - free) - torch.cuda.memory_reserved(dev)) >> 20)
    return {"outside_mib": outside, "errors": errors, "live": live, "window_mib": win_mib}


def test_symmetric_windows_are_released():
    _need(2)
    for r in run_ranks(_teardown, world=2, backend="b200", timeout=600):
        u = r["outside_mib"]
        assert not r["errors"], r
        assert all(rel for _, rel in r["live"]), r
        # memory outside torch's allocator (CUDA context + windows) must not grow by a window per iteration
        assert max(u[2:]) - min(u[2:]) < max(32, r["window_mib"] // 2), r


def _slow_rank0(rank, world):
    """Rank 0 is busy for 25 s before a step (the reference evaluates the whole test set and writes the checkpoint
    on rank 0 alone, resnet/main.py:109-115) while rank 1 already sits in the step's first collective. With the
    default device-side spin bound (30 min, as NCCL's process-group timeout) nothing traps."""
    import os
    import time
    os.environ.pop("B200DDP_SPIN_TIMEOUT_MS", None)
    from pytorch_distributed_tutorials_b200 import optim
    from pytorch_distributed_tutorials_b200.models.resnet import resnet18
    from pytorch_distributed_tutorials_b200.ops.loss import CrossEntropyLoss
    from pytorch_distributed_tutorials_b200.parallel import DistributedDataParallel
    dev = torch.device("cuda", rank)
    torch.manual_seed(0)
    ddp = DistributedDataParallel(resnet18().to(dev), device_ids=[rank])
    opt = optim.SGD(ddp.parameters(), lr=0.01, momentum=0.9, fuse_with_backward=True)
    crit = CrossEntropyLoss()
    x = torch.randn(8, 3, 32, 32, device=dev)
    y = torch.randint(0, 10, (8,), device=dev)
    t0 = time.time()
    for it in range(3):
        if rank == 0 and it == 1:
            time.sleep(25)
        opt.zero_grad()
        loss = crit(ddp(x), y)
        loss.backward()
        opt.step()
        torch.cuda.synchronize()
    return {"seconds": time.time() - t0, "loss": float(loss.item())}


def test_slow_rank0_does_not_trip_the_spin_bound():
    _need(2)
    for r in run_ranks(_slow_rank0, world=2, backend="b200", timeout=300):
        assert r["seconds"] > 24 and r["loss"] == r["loss"], r
