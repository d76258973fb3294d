#!/usr/bin/env python
"""Headline benchmark (BASELINE.json): ResNet-50 training images/sec, ImageNet-shaped synthetic data
(3x224x224, 1000 classes), batch 256 per GPU, bf16 compute, SGD-momentum, DDP over N GPUs of one node.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # stock torch DDP + torchvision + cuDNN/NCCL, same metric

Rank 0 prints ONE JSON line. `value` is the whole-job images/s with the K timed steps bracketed by
barrier + cuda.synchronize and timed by CUDA events (max over ranks), inputs resident on the device.
`e2e` is the same metric through the public API (DevicePrefetcher -> DDP -> loss -> backward -> step) with the
per-step pinned-host -> device input copy and a device -> host read of the loss inside the timed region.
Per-step inputs (154 MB fp32 per GPU) are larger than the 126 MB L2, so no explicit L2 flush is needed.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--batch", type=int, default=256, help="per-GPU batch")
    ap.add_argument("--image-size", type=int, default=224)
    ap.add_argument("--ref-variant", default="r1", choices=["r0", "r1"],
                    help="r0 = as written (fp32, deterministic cuDNN); r1 = tuned stock (bf16 autocast, "
                         "channels_last, cudnn.benchmark, bf16 gradient compression, fused SGD)")
    ap.add_argument("--sync-bn", action="store_true")
    ap.add_argument("--bucket-cap-mb", type=float, default=None)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-bf16-input-e2e", action="store_true", help="skip the additional bf16-host-input e2e line")
    ap.add_argument("--no-fuse-optimizer", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel from Python instead of replaying "
                    "the captured whole-step CUDA graph")
    ap.add_argument("--no-comm-probe", action="store_true", help="skip the backward with/without-sync probe")
    ap.add_argument("--kernel-breakdown", default=None, metavar="FILE",
                    help="after the timed region, run 3 more steps under torch.profiler (CUPTI) and write rank 0's "
                         "per-kernel device-time table to FILE (not part of any reported number)")
    ap.add_argument("--reference-strict", action="store_true",
                    help="report the reference as unavailable instead of running the minimally-fixed harness")
    return ap.parse_args()


def dist_setup():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    return rank, local_rank, world


_CPU_PLANE = False   # ours arm: the process group is gloo only (no NCCL communicator exists in the process)


def barrier_sync(world):
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()          # gloo (CPU) for our arm, NCCL for the reference arm
    torch.cuda.synchronize()


def max_over_ranks(v, world):
    if world == 1:
        return v
    t = torch.tensor([v], dtype=torch.float64, device="cpu" if _CPU_PLANE else "cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_objects(obj, world):
    if world == 1:
        return [obj]
    out = [None] * world
    dist.all_gather_object(out, obj)
    return out


def timed_region(step_fn, steps, world):
    """Exactly `steps` calls of step_fn between barrier+sync, CUDA-event timed, max over ranks (ms)."""
    barrier_sync(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step_fn(i)
    e1.record()
    e1.synchronize()
    barrier_sync(world)
    return max_over_ranks(e0.elapsed_time(e1), world)


def make_host_batches(n, batch, size, seed):
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n):
        x = torch.randn(batch, 3, size, size, generator=g).pin_memory()
        y = torch.randint(0, 1000, (batch,), generator=g).pin_memory()
        out.append((x, y))
    return out


class HostLoader:
    """Cycles a few pinned host batches: stands in for a DataLoader with pin_memory=True."""

    def __init__(self, batches, steps):
        self.batches, self.steps = batches, steps

    def __len__(self):
        return self.steps

    def __iter__(self):
        for i in range(self.steps):
            yield self.batches[i % len(self.batches)]


# ------------------------------------------------------------------------------------------------ our arm
def build_ours(args, local_rank, world):
    from pytorch_distributed_tutorials_b200 import optim
    from pytorch_distributed_tutorials_b200.models.resnet import MODELS
    from pytorch_distributed_tutorials_b200.ops.loss import CrossEntropyLoss
    from pytorch_distributed_tutorials_b200.parallel import (DistributedDataParallel, convert_sync_batchnorm,
                                                             init_process_group)
    from pytorch_distributed_tutorials_b200.utils.seed import set_random_seeds
    global _CPU_PLANE
    set_random_seeds(0)
    if world > 1:
        init_process_group(backend="b200", local_rank=local_rank)   # gloo control plane only; pins CPU/NUMA
        _CPU_PLANE = True
    else:
        from pytorch_distributed_tutorials_b200.utils.affinity import pin_to_gpu_numa_node
        pin_to_gpu_numa_node(local_rank)
    dev = torch.device("cuda", local_rank)
    model = MODELS[args.model]()
    if args.sync_bn:
        model = convert_sync_batchnorm(model)
    model = model.to(dev)
    ddp = DistributedDataParallel(model, device_ids=[local_rank], output_device=local_rank,
                                  bucket_cap_mb=args.bucket_cap_mb)
    if world > 1:
        # the product path or nothing: a silent fallback to torch collectives would still print a number
        assert ddp.backend == "p2p", "expected the peer-memory backend, got %r" % ddp.backend
        ddp.comm.set_spin_timeout_ms(120000)   # a benchmark has no rank-0-only phases: fail fast on a lost peer
    opt = optim.SGD(ddp.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-5,
                    fuse_with_backward=not args.no_fuse_optimizer)
    crit = CrossEntropyLoss()
    ddp.train()
    from pytorch_distributed_tutorials_b200.utils.graph import GraphedTrainStep
    # public API for a whole training step: zero_grad -> forward -> loss -> backward (+ bucketed all-reduce with
    # the fused SGD epilogue) -> step, captured into one CUDA graph after three eager steps
    step = GraphedTrainStep(ddp, crit, opt, warmup=3, enabled=not args.no_graph)
    ddp._bench_step = step
    ddp._bench_parts = (crit, opt, lambda x: x)
    return step, ddp


def _numa_note():
    from pytorch_distributed_tutorials_b200.utils import affinity
    last = affinity.LAST
    if last["node"] is None:
        return "unpinned (platform does not report the GPU's NUMA node)"
    if last["cpus"] is None:
        return "node %d, affinity left alone" % last["node"]
    return "node %d, %d cpus" % (last["node"], last["cpus"])


def replica_checksum(ddp):
    """Bit-level checksum of this rank's flat fp32 parameters (wrapping int64 sum of the raw words)."""
    tot = 0
    for b in ddp.store.buckets:
        tot += int(b.p32.view(torch.int32).sum(dtype=torch.int64).item())
    return tot & 0xFFFFFFFFFFFFFFFF


def backward_comm_overhead(ddp, parts, x, y, world, autocast, n=3):
    """How much longer ``loss.backward()`` takes WITH gradient synchronisation than without (``no_sync()``), by
    CUDA events, eager launches, mean over ``n`` pairs, max over ranks. Same probe for both arms: it contains the
    exposed tail of the last reduction AND whatever the reductions cost backward's own kernels (SM / bandwidth
    contention) - the cross-check SURVEY 7.3 #9 asks for beside the tail measured with device timestamps."""
    if world == 1:
        return 0.0
    import contextlib
    crit, opt, prep = parts
    times = {True: [], False: []}
    for i in range(2 * n + 2):
        sync = (i % 2 == 0)
        opt.zero_grad()
        with (contextlib.nullcontext() if sync else ddp.no_sync()):
            with (torch.autocast("cuda", dtype=torch.bfloat16) if autocast else contextlib.nullcontext()):
                loss = crit(ddp(prep(x)), y)
            barrier_sync(world)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            loss.backward()
            e1.record()
            e1.synchronize()
        opt.step()
        if i >= 2:
            times[sync].append(e0.elapsed_time(e1))
    d = sum(times[True]) / len(times[True]) - sum(times[False]) / len(times[False])
    return max_over_ranks(max(0.0, d), world)


# ------------------------------------------------------------------------------------------------ reference arm
def load_reference_module():
    """The reference file cannot be imported as shipped (SyntaxError at resnet/main.py:111, KeyError at :58 ...).
    Apply the minimal documented fix set to the verbatim text in memory and exec it, so that the
    helpers that can be reused (set_random_seeds) are the reference's own."""
    path = os.path.join(ROOT, "baseline", "_ref", "resnet", "main.py")
    if not os.path.exists(path):
        alt = "/root/reference/resnet/main.py"
        path = alt if os.path.exists(alt) else None
    if path is None:
        return None
    src = open(path).read()
    src = src.replace("evaluate(model=ddp_model, device=device, test_loader)",
                      "evaluate(model=ddp_model, device=device, test_loader=test_loader)")
    src = src.replace('default=defaults["model_filename"]', 'default=defaults["filename"]')
    src = src.replace('"Epoch: {}, Accuracy: {}".format(epoch, accuracy)', '"Epoch: {}, Accuracy: {}".format(epoch, acc)')
    mod = {"__name__": "reference_main"}
    exec(compile(src, path, "exec"), mod)
    return mod


def build_reference(args, local_rank, world):
    """Stock path the reference resolves to: torchvision ResNet + torch DDP (NCCL) + cuDNN/cuBLAS + optim.SGD,
    in the order of the reference's main() (seed -> init_process_group -> model.to -> DDP -> loss/optimizer)."""
    import torch.nn as nn
    import torch.optim as optim
    import torchvision
    ref = load_reference_module()
    if ref is not None:
        ref["set_random_seeds"](0)          # reference resnet/main.py:16-21 (cudnn.deterministic=True, benchmark=False)
    if args.ref_variant == "r1":
        torch.backends.cudnn.deterministic = False
        torch.backends.cudnn.benchmark = True
    # the reference always initialises the process group and wraps in DDP, also on one GPU (resnet/main.py:74,80)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29531")
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("WORLD_SIZE", "1")
    dist.init_process_group(backend="nccl")
    dev = torch.device("cuda", local_rank)
    model = getattr(torchvision.models, args.model)(weights=None)
    if args.sync_bn:
        model = nn.SyncBatchNorm.convert_sync_batchnorm(model)
    model = model.to(dev)
    r1 = args.ref_variant == "r1"
    if r1:
        model = model.to(memory_format=torch.channels_last)
    kw = {}
    if args.bucket_cap_mb is not None:
        kw["bucket_cap_mb"] = args.bucket_cap_mb
    ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local_rank], output_device=local_rank,
                                                    gradient_as_bucket_view=r1, **kw)
    if r1 and world > 1:
        from torch.distributed.algorithms.ddp_comm_hooks import default_hooks
        ddp.register_comm_hook(None, default_hooks.bf16_compress_hook)
    crit = nn.CrossEntropyLoss()
    opt = optim.SGD(ddp.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-5, **({"fused": True} if r1 else {}))
    ddp.train()
    ddp._bench_parts = (crit, opt, (lambda x: x.contiguous(memory_format=torch.channels_last)) if r1 else
                        (lambda x: x))

    def step(x, y):
        opt.zero_grad()
        if r1:
            x = x.contiguous(memory_format=torch.channels_last)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = ddp(x)
                loss = crit(out, y)
        else:
            loss = crit(ddp(x), y)
        loss.backward()
        opt.step()
        return loss

    return step, ddp


# ------------------------------------------------------------------------------------------------ main
def main():
    args = parse()
    if not torch.cuda.is_available():
        print(json.dumps({"impl": args.impl, "unavailable": "no CUDA device on this machine"}))
        return 0
    if args.impl == "reference" and args.reference_strict:
        print(json.dumps({"impl": "reference", "unavailable":
                          "reference has no setup.py/pyproject (pip install fails) and resnet/main.py does not "
                          "compile (SyntaxError at line 111)"}))
        return 0
    rank, local_rank, world = dist_setup()
    if world != args.gpus and rank == 0:
        print("warning: --gpus %d but WORLD_SIZE=%d (launch with torchrun for N>1)" % (args.gpus, world),
              file=sys.stderr)
    dev = torch.device("cuda", local_rank)
    build = build_ours if args.impl == "ours" else build_reference
    step, ddp = build(args, local_rank, world)

    from pytorch_distributed_tutorials_b200.ops import _lib
    from pytorch_distributed_tutorials_b200.utils.data import DevicePrefetcher
    from pytorch_distributed_tutorials_b200.utils.timing import ClockSampler

    # ---- device-resident inputs (two distinct batches, alternated)
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    xs = [torch.randn(args.batch, 3, args.image_size, args.image_size, device=dev, generator=g) for _ in range(2)]
    ys = [torch.randint(0, 1000, (args.batch,), device=dev, generator=g) for _ in range(2)]

    def dev_step(i):
        return step(xs[i & 1], ys[i & 1])

    with ClockSampler(local_rank, 100) as clocks:   # started before warm-up: nvidia-smi needs a moment to come up
        for i in range(args.warmup):
            dev_step(i)
        # the whole-step CUDA graph is captured on the first call after three eager steps; if --warmup was too
        # short for that, keep warming up (untimed, and reported) so the capture never lands in the timed region
        extra_warmup = 0
        gstep = getattr(ddp, "_bench_step", None)
        while gstep is not None and gstep.enabled and gstep.captures == 0 and extra_warmup < 6:
            dev_step(args.warmup + extra_warmup)
            extra_warmup += 1
        barrier_sync(world)
        if hasattr(ddp, "reset_exposed_comm"):
            ddp.reset_exposed_comm()     # exposed communication is reported for the timed (graph-replayed) steps only
        _lib.reset_launches()
        t_begin = time.monotonic()
        ms_total = timed_region(dev_step, args.steps, world)
        t_end = time.monotonic()
        launches = _lib.launches()
        exposed = ddp.exposed_comm_ms(last=args.steps) if (args.impl == "ours" and hasattr(ddp, "exposed_comm_ms")) \
            else None
        anatomy = ddp.step_anatomy(last=args.steps) if (args.impl == "ours" and hasattr(ddp, "step_anatomy")) else None
        if ms_total < 500.0:
            # very short timed regions (ResNet-18 on 32x32 images: ~1 ms per step): keep the same load on for about
            # 0.8 s more so that the clock sampler (200 ms period) sees it. The number of extra steps is derived from
            # the max-over-ranks time, i.e. it is the SAME on every rank - every step is a collective.
            n_extra = int(800.0 / max(ms_total / args.steps, 0.05))
            for i in range(n_extra):
                dev_step(i)
            torch.cuda.synchronize()
            t_end = time.monotonic()
    clk = clocks.summary(t_begin, t_end)
    ms_step = ms_total / args.steps
    global_batch = args.batch * world
    value = global_batch / (ms_step / 1e3)

    # ---- end to end: pinned host batches -> prefetcher (H2D on a side stream) -> step -> loss read back
    e2e = None
    if not args.no_e2e:
        host = make_host_batches(3, args.batch, args.image_size, 99 + rank)
        loss_host = torch.zeros(args.steps + args.warmup, dtype=torch.float32).pin_memory()

        def run(nsteps, base, batches):
            it = DevicePrefetcher(HostLoader(batches, nsteps), dev)
            for k, (x, y) in enumerate(it):
                loss = step(x, y)
                loss_host[base + k].copy_(loss.detach().float(), non_blocking=True)   # device -> host, every step
            torch.cuda.current_stream().synchronize()

        def timed_e2e(batches):
            run(max(3, args.warmup), 0, batches)
            barrier_sync(world)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            e0.record()
            run(args.steps, 0, batches)
            e1.record()
            e1.synchronize()
            wall_ms = (time.perf_counter() - t0) * 1e3
            barrier_sync(world)
            return max_over_ranks(max(e0.elapsed_time(e1), wall_ms), world) / args.steps

        ms_e2e = timed_e2e(host)
        h2d = host[0][0].numel() * 4 + host[0][1].numel() * 8
        e2e = {"value": global_batch / (ms_e2e / 1e3), "unit": "images/s", "ms_per_step": ms_e2e,
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
               "final_loss": float(loss_host[args.steps - 1])}
        if args.impl == "ours" and hasattr(ddp, "step_anatomy"):
            e2e["step_anatomy_ms"] = ddp.step_anatomy(last=args.steps)   # where the end-to-end steps' time went
        if args.impl == "ours" and not args.no_bf16_input_e2e:
            # Second, separately labelled end-to-end figure: the same step fed with bf16 host batches (what a
            # DataLoader whose last transform is `.to(torch.bfloat16)` yields; the stem kernel converts to bf16
            # anyway). Half the host->device bytes: shows how much of the fp32 figure above is the input copy.
            # The headline e2e stays the fp32 one, the dtype the reference arm is fed with.
            host16 = [(x.to(torch.bfloat16).pin_memory(), y) for x, y in host]
            ms16 = timed_e2e(host16)
            e2e["bf16_input"] = {"value": global_batch / (ms16 / 1e3), "ms_per_step": ms16,
                                 "h2d_bytes_per_step": host16[0][0].numel() * 2 + host16[0][1].numel() * 8,
                                 "note": "same step, host batches already bf16; not the headline"}

    if args.kernel_breakdown:
        from torch.profiler import ProfilerActivity, profile
        barrier_sync(world)
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for i in range(3):
                dev_step(i)
            torch.cuda.synchronize()
        barrier_sync(world)
        if rank == 0:
            rows = [(k.key, k.device_time_total / 3e3, k.count / 3) for k in prof.key_averages()
                    if k.device_time_total > 0]
            rows.sort(key=lambda r: -r[1])
            with open(args.kernel_breakdown, "w") as fh:
                fh.write("# %s n_gpus=%d model=%s batch=%d: device time per step by kernel (torch.profiler/CUPTI; "
                         "streams overlap, so the sum exceeds the step time)\n" % (args.impl, world, args.model,
                                                                                  args.batch))
                fh.write("total %.3f ms over %d kernel kinds; step %.3f ms\n" % (sum(r[1] for r in rows), len(rows),
                                                                                ms_step))
                for name, ms, cnt in rows[:60]:
                    fh.write("%8.3f ms x%-6.1f %s\n" % (ms, cnt, name[:160]))
    if os.environ.get("B200DDP_TRACE", "0") == "1" and hasattr(ddp, "bucket_trace") and rank == 0:
        print("bucket trace (us): " + json.dumps(ddp.bucket_trace()), file=sys.stderr, flush=True)

    # ---- self-check: replicas bit-identical after all the timed steps (vouches for the collectives), loss finite
    replicas_identical = None
    if args.impl == "ours":
        sums = gather_objects(replica_checksum(ddp), world)
        replicas_identical = all(c == sums[0] for c in sums)
    # ---- communication overhead of backward, same probe on both arms (runs last: it desynchronises nothing that
    #      is measured afterwards)
    overhead = None
    if world > 1 and not args.no_comm_probe and args.impl == "reference":
        # (our arm's figure is measured inside the replayed graph; an eager with/without-sync probe would time Python
        #  launches and, without sync, an optimizer that is no longer fused into backward - not the product path)
        overhead = backward_comm_overhead(ddp, ddp._bench_parts, xs[0], ys[0], world,
                                          autocast=(args.impl == "reference" and args.ref_variant == "r1"))
        if args.impl == "reference":
            exposed = overhead     # torch DDP offers no per-step tail timestamp; the differential is its figure

    if rank == 0:
        out = {
            "metric": "resnet50_train_images_per_sec" if args.model == "resnet50" else args.model + "_train_images_per_sec",
            "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup + extra_warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if (args.impl == "ours" or args.ref_variant == "r1") else "fp32(tf32 conv)",
            "data": "synthetic (random-init weights, random 3x%dx%d images, 1000 classes)" % (args.image_size,
                                                                                              args.image_size),
            "impl": args.impl,
            "config": {"model": args.model, "global_batch": global_batch, "per_gpu_batch": args.batch,
                       "image_size": args.image_size, "seq_len": None,
                       "parallelism": "dp%d" % world, "optimizer": "sgd(momentum=0.9, wd=1e-5)",
                       "sync_bn": bool(args.sync_bn),
                       "l2": "per-step inputs (%.0f MB/GPU) exceed the 126 MB L2; no explicit flush" %
                             (args.batch * 3 * args.image_size ** 2 * 4 / 1e6),
                       "numa_pin": _numa_note() if args.impl == "ours" else "none (the reference does not pin)",
                       "backend": getattr(ddp, "backend", "torch-ddp/nccl" if world > 1 else "none"),
                       "cuda_graph": bool(getattr(getattr(ddp, "_bench_step", None), "captures", 0))},
            "clocks": clk, "e2e": e2e, "gpu_launches": launches if args.impl == "ours" else 0,
            "exposed_comm_ms_per_step": exposed,
            "exposed_comm_how": ("device %globaltimer stamps inside the replayed graph: end of last reduction minus "
                                 "end of backward compute" if args.impl == "ours" else
                                 "backward(sync) - backward(no_sync), CUDA events"),
            "comm_overhead_bwd_ms": overhead,
            "step_anatomy_ms": anatomy,
            "replicas_identical": replicas_identical,
        }
        if args.impl == "reference":
            out["reference_variant"] = args.ref_variant
            out["reference_note"] = ("resnet/main.py cannot be installed (no setup.py/pyproject) nor compiled as "
                                     "shipped (SyntaxError :111); this arm runs the stock path it resolves to "
                                     "(torchvision + torch DDP/NCCL + cuDNN + optim.SGD) in the reference's order, "
                                     "with the reference's own set_random_seeds when the file is present")
        print(json.dumps(out), flush=True)
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
