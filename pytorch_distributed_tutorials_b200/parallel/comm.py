"""Collective back ends behind the DDP wrapper.

``TorchComm`` drives any ``torch.distributed`` process group (gloo on CPU: the reference's plumbing config;
also the oracle the P2P kernels are tested against).  ``P2PComm`` is the product path on B200: buckets live in a
symmetric window and are reduced by the kernels of csrc/allreduce.cu on a high-priority stream, optionally with
the SGD update as their epilogue.  ``LocalComm`` is world_size == 1.
(reference resnet/main.py:74 ``init_process_group(backend="nccl")``, :80 DDP's bucket all-reduce / broadcasts.)
"""
import ctypes
import os

import torch
import torch.distributed as dist

from ..ops import _lib
from ..ops._lib import c_f, c_ll, c_p, check, ptr

ALGOS = {"auto": 0, "oneshot": 1, "twoshot": 2, "multimem": 3}


class _Base:
    name = "base"

    def __init__(self, group=None):
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.launch_log = []   # (kind, detail) per collective, for tests and the exposed-comm report

    def grad_alloc(self, total_numel):
        return None

    def begin_backward(self):
        pass

    def finalize(self):
        pass


class LocalComm(_Base):
    """world_size == 1: nothing to reduce; the fused path is just the optimizer kernel."""
    name = "local"

    def allreduce_bucket(self, bucket, sgd=None):
        self.launch_log.append(("allreduce", bucket.index))
        if sgd is not None:
            from ..ops import native
            lr, mu, wd, first, hyper = (tuple(sgd) + (None,))[:5]
            native.sgd_step(bucket.p32, bucket.g32, bucket.mom, bucket.shadow, lr, mu, wd, 1.0, first, hyper)

    def broadcast_(self, tensors, src=0):
        self.launch_log.append(("broadcast", len(tensors)))


class TorchComm(_Base):
    """Any torch.distributed backend. Mean all-reduce = pre-divide then SUM, as DDP's default hook does."""
    name = "torch"

    def __init__(self, group=None):
        super().__init__(group)
        self._works = []

    def allreduce_bucket(self, bucket, sgd=None):
        self.launch_log.append(("allreduce", bucket.index))
        bucket.g32.div_(self.world)
        self._works.append(dist.all_reduce(bucket.g32, group=self.group, async_op=True))
        if sgd is not None:
            raise RuntimeError("fused optimizer epilogue needs the P2P or local backend")

    def broadcast_(self, tensors, src=0):
        """Coalesced broadcast from ``src`` (``_broadcast_coalesced`` of torch DDP)."""
        self.launch_log.append(("broadcast", len(tensors)))
        by_dtype = {}
        for t in tensors:
            by_dtype.setdefault(t.dtype, []).append(t)
        for ts in by_dtype.values():
            flat = torch.cat([t.reshape(-1) for t in ts])
            dist.broadcast(flat, src, group=self.group)
            off = 0
            for t in ts:
                n = t.numel()
                t.copy_(flat[off:off + n].view_as(t))
                off += n

    def finalize(self):
        for w in self._works:
            w.wait()
        self._works = []


class P2PComm(_Base):
    """Hand-written kernels over NVLink peer memory (csrc/allreduce.cu, csrc/symm.cu). Never calls NCCL."""
    name = "p2p"
    PAD_COMM = 0      # signal pad used by kernels on the communication stream
    PAD_COMPUTE = 1   # signal pad used by kernels on the compute stream (SyncBN statistics, buffer broadcast)

    def __init__(self, device, group=None, window_bytes=None, algo=None, nblocks=None, wire=None):
        super().__init__(group)
        self.wire = (wire or os.environ.get("B200DDP_WIRE", "fp32")).lower()   # "fp32" | "bf16"
        if self.wire not in ("fp32", "bf16"):
            raise ValueError("gradient wire format must be fp32 or bf16")
        from .symm import SymmetricWindow
        self.device = torch.device(device)
        self.window_bytes = int(window_bytes)
        self.window = SymmetricWindow(self.window_bytes, self.device, group)
        self.algo = ALGOS[algo or os.environ.get("B200DDP_ALGO", "auto")]
        self.nblocks = int(nblocks or os.environ.get("B200DDP_COMM_BLOCKS", "64"))
        lo, hi = torch.cuda.Stream.priority_range()
        self.stream = torch.cuda.Stream(device=self.device, priority=hi)
        self._grad_region = None
        self.write_grad = os.environ.get("B200DDP_WRITE_GRAD", "1") == "1"
        self.use_mc_broadcast = os.environ.get("B200DDP_MC_BROADCAST", "1") == "1"
        self._events = []
        t = os.environ.get("B200DDP_SPIN_TIMEOUT_MS")
        if t:
            self.set_spin_timeout_ms(int(t))

    def set_debug_skew_us(self, us):
        """Protocol debugging: delay a pseudo-random subset of the CTAs of every collective kernel by ``us``
        microseconds before / between the phases (0 = off). See csrc/xgpu.cuh ``debug_skew``."""
        check(_lib.lib().b2_set_debug_skew_ns_allreduce(ctypes.c_ulonglong(int(us) * 1000)),
              "b2_set_debug_skew_ns_allreduce")

    def set_spin_timeout_ms(self, ms):
        """How long a kernel may wait for a peer before it traps (failure detection; default 1800 s, the
        process-group timeout of the reference stack)."""
        ns = ctypes.c_ulonglong(int(ms) * 1000000)
        check(_lib.lib().b2_set_spin_timeout_ns_allreduce(ns), "b2_set_spin_timeout_ns_allreduce")
        check(_lib.lib().b2_set_spin_timeout_ns_bn(ns), "b2_set_spin_timeout_ns_bn")

    # gradient buckets are carved out of the window so that peers can read them
    def grad_alloc(self, numel):
        off = self.window.alloc(numel * 4)
        return self.window.tensor(off, numel, torch.float32)

    def alloc_tensor(self, numel, dtype):
        esz = torch.empty((), dtype=dtype).element_size()
        off = self.window.alloc(numel * esz)
        return self.window.tensor(off, numel, dtype)

    @property
    def mc_base(self):
        return self.window.mc_ptr

    def begin_backward(self):
        self._first_launch = True

    def allreduce_bucket(self, bucket, sgd=None, algo=None, on_stream=None):
        """Launch the fused reduce(+update) kernel for one bucket. The communication stream first waits for the
        work already queued on the current (compute) stream, i.e. for the kernels that produced the gradients."""
        self.launch_log.append(("allreduce", bucket.index))
        cur = torch.cuda.current_stream(self.device)
        st = on_stream or self.stream
        if st is not cur:
            ev = torch.cuda.Event()
            ev.record(cur)
            st.wait_event(ev)
        a = self.algo if algo is None else ALGOS[algo]
        if a == ALGOS["multimem"] and not self.mc_base:
            a = ALGOS["twoshot"]
        if self.wire == "bf16":
            return self._allreduce_bucket_bf16(bucket, sgd, a, st)
        off = self.window.offset_of(bucket.g32)
        args = self._sgd_args(bucket, sgd)
        e = _lib.lib().b2_allreduce(ctypes.byref(self.window.ctx(self.PAD_COMM)), c_p(self.mc_base or 0), c_ll(off),
                                    c_ll(bucket.numel), a, c_f(1.0 / self.world), *args, int(self.write_grad),
                                    self.nblocks, c_p(st.cuda_stream))
        check(e, "b2_allreduce")
        _lib.count_launch()

    @staticmethod
    def _sgd_args(bucket, sgd):
        """(p32, momentum, bf16 shadow, lr, mu, wd, first step, device hyper-parameter block) of the fused update."""
        if sgd is None:
            return (c_p(0), c_p(0), c_p(0), c_f(0.0), c_f(0.0), c_f(0.0), 0, c_p(0))
        lr, mu, wd, first, hyper = (tuple(sgd) + (None,))[:5]
        return (ptr(bucket.p32), ptr(bucket.mom), ptr(bucket.shadow), c_f(lr), c_f(mu), c_f(wd), int(first),
                ptr(hyper))

    def _allreduce_bucket_bf16(self, bucket, sgd, a, st):
        """bf16 wire: the kernel casts the fp32 bucket into a bf16 staging area of the window and reduces that."""
        wire = getattr(bucket, "wire16", None)
        if wire is None:
            wire = bucket.wire16 = self.alloc_tensor(bucket.numel, torch.bfloat16)
        args = self._sgd_args(bucket, sgd)
        g_out = ptr(bucket.g32) if (self.write_grad or sgd is None) else c_p(0)
        e = _lib.lib().b2_allreduce_bf16(ctypes.byref(self.window.ctx(self.PAD_COMM)), c_p(self.mc_base or 0),
                                         c_ll(self.window.offset_of(wire)), c_ll(bucket.numel), a,
                                         c_f(1.0 / self.world), ptr(bucket.g32), g_out, *args, self.nblocks,
                                         c_p(st.cuda_stream))
        check(e, "b2_allreduce_bf16")
        _lib.count_launch()

    def broadcast_window_(self, offset, nbytes, src=0, stream=None, pad=None):
        """In-place broadcast of a window region: every rank pulls [offset, offset+nbytes) from ``src``."""
        st = stream or torch.cuda.current_stream(self.device)
        pad = self.PAD_COMPUTE if pad is None else pad
        nb = (nbytes + 15) // 16 * 16
        blocks = min(self.nblocks, 16 if nb < (1 << 20) else 64)
        if self.mc_base and self.use_mc_broadcast and self.world > 2:
            # one multicast store per 16 bytes on the source instead of W-1 peer pulls
            e = _lib.lib().b2_broadcast_mc(ctypes.byref(self.window.ctx(pad)), c_p(self.mc_base), c_ll(offset),
                                           c_ll(nb), src, blocks, c_p(st.cuda_stream))
            check(e, "b2_broadcast_mc")
        else:
            e = _lib.lib().b2_broadcast(ctypes.byref(self.window.ctx(pad)), c_ll(offset), c_ll(nb), src, c_p(0),
                                        blocks, c_p(st.cuda_stream))
            check(e, "b2_broadcast")
        _lib.count_launch()

    def broadcast_(self, tensors, src=0):
        """Broadcast arbitrary device tensors from ``src`` through a staging area of the window."""
        self.launch_log.append(("broadcast", len(tensors)))
        total = sum((t.numel() * t.element_size() + 255) // 256 * 256 for t in tensors)
        if total == 0:
            return
        if getattr(self, "_stage", None) is None or self._stage.numel() < total:
            self._stage = self.alloc_tensor(max(total, 1 << 20), torch.uint8)
        off = 0
        views = []
        for t in tensors:
            nb = t.numel() * t.element_size()
            v = self._stage[off:off + nb]
            views.append((t, v))
            if self.rank == src:
                v.copy_(t.detach().contiguous().view(-1).view(torch.uint8))
            off += (nb + 255) // 256 * 256
        self.broadcast_window_(self.window.offset_of(self._stage), total, src)
        if self.rank != src:
            for t, v in views:
                t.detach().copy_(v.view(t.dtype).view(t.shape) if t.is_contiguous() else
                                 v.view(t.dtype).view(t.shape))

    def barrier(self, stream=None, pad=None):
        st = stream or torch.cuda.current_stream(self.device)
        pad = self.PAD_COMPUTE if pad is None else pad
        check(_lib.lib().b2_xgpu_barrier(ctypes.byref(self.window.ctx(pad)), 255, c_p(st.cuda_stream)),
              "b2_xgpu_barrier")

    def finalize(self):
        """Later work on the compute stream (next forward, optimizer.step) waits for the reductions."""
        torch.cuda.current_stream(self.device).wait_stream(self.stream)
