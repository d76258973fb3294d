"""Helpers to run a function on N ranks (spawned processes, gloo or nccl control plane, 127.0.0.1 rendezvous)."""
import os
import socket
import sys
import traceback

import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _entry(rank, world, port, fn, backend, args, q):
    os.environ.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank),
                       "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank)})
    # a protocol bug must fail a test within a minute, not sit in a spin loop for the 30-minute production default
    os.environ.setdefault("B200DDP_SPIN_TIMEOUT_MS", "60000")
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    try:
        from pytorch_distributed_tutorials_b200.parallel import destroy_process_group, init_process_group
        init_process_group(backend=backend, local_rank=rank)
        out = fn(rank, world, *args)
        q.put((rank, "ok", out))
        destroy_process_group()
    except Exception:  # noqa: BLE001
        q.put((rank, "err", traceback.format_exc()))


def run_ranks(fn, world=2, backend="gloo", args=(), timeout=300):
    """Runs fn(rank, world, *args) on `world` processes; returns the list of per-rank return values."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=_entry, args=(r, world, port, fn, backend, args, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = {}
    try:
        for _ in range(world):
            rank, status, out = q.get(timeout=timeout)
            if status != "ok":
                raise AssertionError("rank %d failed:\n%s" % (rank, out))
            results[rank] = out
    finally:
        for p in procs:
            p.join(timeout=20)
            if p.is_alive():
                p.terminate()
    return [results[r] for r in range(world)]
