"""All-reduce sweep: our P2P kernels (one-shot / two-shot / multimem) vs NCCL (torch.distributed.all_reduce) and
torch's symmetric-memory all-reduces, on the DDP bucket sizes of ResNet-18/50/152 plus a 1 KiB..256 MiB sweep.
Launch: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/comm_bench.py
Reports device-timed latency (max over ranks) and bus bandwidth 2(W-1)/W * S / t.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from pytorch_distributed_tutorials_b200.parallel import init_process_group  # noqa: E402
from pytorch_distributed_tutorials_b200.parallel.comm import P2PComm  # noqa: E402


class _B:
    pass


def timed(fn, iters, world):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    e1.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64)     # CPU tensor: gloo control plane
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    init_process_group(backend="b200", local_rank=lr)   # gloo control plane; the product path creates no NCCL
    rank, world = dist.get_rank(), dist.get_world_size()
    dev = torch.device("cuda", lr)
    # the baselines get their own NCCL group: dist.all_reduce over it, and torch's symmetric-memory all-reduces
    # (one_shot_all_reduce / two_shot_all_reduce_ / multimem_all_reduce_, "the kernels to beat")
    nccl = dist.new_group(backend="nccl", device_id=dev)
    symm, symm_err = None, None
    try:
        import torch.distributed._symmetric_memory as symm
        try:
            symm.enable_symm_mem_for_group(nccl.group_name)
        except Exception:  # noqa: BLE001  (deprecated / implicit in newer torch)
            pass
    except Exception as e:  # noqa: BLE001
        symm, symm_err = None, repr(e)
    comm = P2PComm(dev, None, 1200 << 20)
    if rank == 0:
        print(json.dumps({"world": world, "transport": comm.window.transport, "multicast": bool(comm.mc_base),
                          "vmm_error": getattr(comm.window, "_vmm_error", None)}), flush=True)
    named = [("syncbn_516B", 132), ("r18_b0", 513024), ("r50_b0", 2049024), ("r18_b2", 3963456),
             ("r18_b1", 7213056), ("r50_b1", 7875584), ("r152_100MB", 26215424), ("256MiB", 64 << 20)]
    sweep = [(("%dKiB" % (1 << k)), (1 << k) * 256) for k in range(0, 19, 3)]
    rows = []
    only = os.environ.get("COMM_BENCH_ONLY")        # comma-separated row names: a quick re-measurement of a few sizes
    todo = [(nm, n) for nm, n in sweep + named if not only or nm in only.split(",")]
    for name, n in todo:
        n = (n + 63) // 64 * 64
        b = _B()
        b.index, b.numel = 0, n
        b.g32 = comm.grad_alloc(n)
        b.g32.normal_()
        ref = torch.randn(n, device=dev)
        iters = 50 if n < (1 << 22) else 15
        row = {"name": name, "bytes": n * 4}
        row["nccl_us"] = 1e3 * timed(lambda: dist.all_reduce(ref, group=nccl), iters, world)
        if symm is not None:
            try:
                st = symm.empty(n, dtype=torch.float32, device=dev)
                symm.rendezvous(st, nccl.group_name)
                st.normal_()
                gname = nccl.group_name
                if n * 4 <= (8 << 20):
                    row["torch_symm_one_shot_us"] = 1e3 * timed(
                        lambda: torch.ops.symm_mem.one_shot_all_reduce(st, "sum", gname), iters, world)
                row["torch_symm_two_shot_us"] = 1e3 * timed(
                    lambda: torch.ops.symm_mem.two_shot_all_reduce_(st, "sum", gname), iters, world)
                try:
                    row["torch_symm_multimem_us"] = 1e3 * timed(
                        lambda: torch.ops.symm_mem.multimem_all_reduce_(st, "sum", gname), iters, world)
                except Exception as e:  # noqa: BLE001
                    row["torch_symm_multimem_err"] = repr(e)[:120]
                del st
            except Exception as e:  # noqa: BLE001
                row["torch_symm_err"] = repr(e)[:160]
                symm = None
        elif symm_err:
            row["torch_symm_err"] = symm_err[:160]
        for algo in ["oneshot", "twoshot"] + (["multimem"] if comm.mc_base else []):
            if algo == "oneshot" and n > 524288:
                continue
            def fn(a=algo):
                comm.allreduce_bucket(b, sgd=None, algo=a, on_stream=torch.cuda.current_stream())
            row[algo + "_us"] = 1e3 * timed(fn, iters, world)
        # fused reduce + SGD update (the product path) vs NCCL all-reduce followed by torch's fused SGD
        b.p32, b.mom = torch.randn(n, device=dev), torch.zeros(n, device=dev)
        b.shadow = torch.empty(n, device=dev, dtype=torch.bfloat16)
        def fused():
            comm.allreduce_bucket(b, sgd=(0.01, 0.9, 1e-5, False), algo="auto", on_stream=torch.cuda.current_stream())
        row["fused_auto_sgd_us"] = 1e3 * timed(fused, iters, world)
        p = torch.nn.Parameter(torch.randn(n, device=dev))
        p.grad = torch.randn(n, device=dev)
        opt = torch.optim.SGD([p], lr=0.01, momentum=0.9, weight_decay=1e-5, fused=True)
        def unfused():
            p.grad.div_(world)
            dist.all_reduce(p.grad, group=nccl)
            opt.step()
        row["nccl_plus_fused_sgd_us"] = 1e3 * timed(unfused, iters, world)
        k = 2.0 * (world - 1) / world * n * 4
        for key in list(row):
            if key.endswith("_us") and "sgd" not in key:
                row[key.replace("_us", "_busGBs")] = k / (row[key] * 1e-6) / 1e9
        rows.append(row)
        if rank == 0:
            print(json.dumps({k: (round(v, 2) if isinstance(v, float) else v) for k, v in row.items()}), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
