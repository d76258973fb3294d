"""DistributedDataParallel-style wrapper with a native reducer and pluggable collectives.

Same surface as ``torch.nn.parallel.DistributedDataParallel`` as the reference uses it (reference
resnet/main.py:80 ``DistributedDataParallel(model, device_ids=[local_rank], output_device=local_rank)``, then
``.parameters()``, ``.state_dict()`` / ``.load_state_dict()`` with ``module.``-prefixed keys at :84,:112,
``.train()`` / ``.eval()`` at :24,:117 and ``__call__`` at :31,:121), plus ``no_sync()``.

What happens underneath is different:
  * construction: parameter shapes are verified across ranks, parameters are moved into flat per-bucket buffers
    (gradient buckets in symmetric memory on CUDA), parameters and buffers are broadcast from rank 0;
  * buckets are planned once, in the order gradients become ready (known statically for our ResNets, reverse
    registration order otherwise), 1 MiB first bucket / 25 MiB cap by default like torch;
  * during ``loss.backward()`` each gradient lands directly in its bucket slot; the native reducer
    (csrc/reducer.cpp) launches, strictly in bucket order, one fused all-reduce(+SGD) kernel per full bucket on a
    high-priority stream, overlapped with the rest of backward;
  * a no-grad / eval forward issues **zero** collectives (the reference's rank-0-only evaluation at
    resnet/main.py:110-111 therefore cannot desynchronise the ranks).
"""
import contextlib
import ctypes
import os
import weakref

import torch
import torch.distributed as dist
import torch.nn as nn

from ..ops import _lib
from .comm import LocalComm, P2PComm, TorchComm
from .flat import FlatParameterStore, plan_buckets

MiB = 1 << 20


class _BucketDesc(ctypes.Structure):
    """Mirror of ``b2_bucket_desc`` (include/b200ddp.h): everything the native reducer needs to launch a bucket."""
    from ..ops.native import XgpuCtx as _Ctx
    _fields_ = [("ctx", _Ctx), ("mc_base", ctypes.c_void_p), ("offset_bytes", ctypes.c_longlong),
                ("n", ctypes.c_longlong), ("algo", ctypes.c_int), ("scale", ctypes.c_float),
                ("p32", ctypes.c_void_p), ("mom", ctypes.c_void_p), ("shadow", ctypes.c_void_p),
                ("write_grad", ctypes.c_int), ("nblocks", ctypes.c_int), ("wire_bf16", ctypes.c_int),
                ("wire_offset_bytes", ctypes.c_longlong), ("g32", ctypes.c_void_p)]


_LIVE = weakref.WeakSet()   # wrappers whose symmetric windows destroy_process_group() must release


def close_all():
    """Release the peer-memory windows of every live wrapper (called by ``destroy_process_group``)."""
    for d in list(_LIVE):
        d.close()


def split_tail_bucket(bucket_of_param, sizes_bytes, tail_bytes):
    """Give the parameters whose gradients arrive LAST their own small bucket.

    The last bucket of the plan (stem + first stage for a ResNet: 9.3 MiB for ResNet-50) is the only
    one whose reduction has no backward compute left to hide behind. Cutting it so that only the final
    ``tail_bytes`` of gradients wait for the very end of backward turns the exposed tail into one small,
    latency-bound kernel; the rest of the old last bucket is reduced while the first stage's backward still runs.
    Pure function of the plan (every rank computes the same split)."""
    if tail_bytes <= 0 or not bucket_of_param:
        return list(bucket_of_param)
    last = bucket_of_param[-1]
    idx = [i for i, b in enumerate(bucket_of_param) if b == last]
    total = sum(sizes_bytes[i] for i in idx)
    if len(idx) < 2 or total <= 2 * tail_bytes:
        return list(bucket_of_param)
    acc, cut = 0, None
    for i in reversed(idx):
        acc += sizes_bytes[i]
        if acc > tail_bytes:
            break
        cut = i
    if cut is None or cut == idx[0]:
        return list(bucket_of_param)
    out = list(bucket_of_param)
    for i in idx:
        if i >= cut:
            out[i] = last + 1
    return out


def _ready_order(module, params):
    """(parameters in the order their gradients become ready, whether that order is known statically)."""
    if hasattr(module, "grad_ready_order"):
        order = [p for p in module.grad_ready_order() if p.requires_grad]
        ids = {id(p) for p in order}
        if len(order) == len(params) and ids == {id(p) for p in params}:
            return order, True
    return list(reversed(params)), False


class DistributedDataParallel(nn.Module):
    def __init__(self, module, device_ids=None, output_device=None, dim=0, broadcast_buffers=True,
                 process_group=None, bucket_cap_mb=None, find_unused_parameters=False,
                 gradient_as_bucket_view=True, static_graph=False, backend=None, comm_algo=None,
                 init_sync=True, gradient_wire=None, tail_bucket_mb=None):
        super().__init__()
        if find_unused_parameters:
            raise NotImplementedError("find_unused_parameters is not supported (ResNets use every parameter)")
        self.module = module
        self.process_group = process_group
        self.broadcast_buffers = broadcast_buffers
        self.require_backward_grad_sync = True
        self.gradient_as_bucket_view = True
        params = [p for p in module.parameters() if p.requires_grad]
        if not params:
            raise RuntimeError("DistributedDataParallel needs at least one parameter that requires grad")
        self.device = params[0].device
        if device_ids is not None and len(device_ids) == 1 and self.device.type == "cuda":
            want = torch.device("cuda", device_ids[0]) if isinstance(device_ids[0], int) else torch.device(
                device_ids[0])
            if want.index is not None and want != self.device:
                raise ValueError("module is on %s but device_ids=%s" % (self.device, device_ids))
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(process_group) if dist.is_initialized() else 0

        # ---- bucket plan (gradient-ready order; torch's 1 MiB first bucket only with the default cap)
        ordered, self._static_order = _ready_order(module, params)
        if bucket_cap_mb is None:
            first_cap, cap = 1 * MiB, 25 * MiB
        else:
            first_cap, cap = -1, int(bucket_cap_mb * MiB)
        sizes = [p.numel() * p.element_size() for p in ordered]
        self.bucket_of_param = plan_buckets(sizes, first_cap, cap)
        self.bucket_cap_bytes = cap
        self._first_cap_bytes = first_cap
        self._tail_bucket_mb = tail_bucket_mb

        # ---- K1: every rank must hold the same parameter shapes
        if self.world > 1:
            sig = [tuple(p.shape) for p in ordered]
            sigs = [None] * self.world
            dist.all_gather_object(sigs, sig, group=process_group)
            for r, s in enumerate(sigs):
                if s != sig:
                    raise RuntimeError("parameter shapes differ between rank %d and rank %d" % (self.rank, r))

        # ---- collective back end
        backend = backend or os.environ.get("B200DDP_BACKEND")
        if backend is None:
            if self.world == 1:
                backend = "local"
            elif self.device.type == "cuda" and _lib.use_native():
                backend = "p2p"
            else:
                backend = "torch"
        buffers = [b for b in module.buffers()]
        # exposed-tail cut (P2P backend only; the CPU / torch back ends keep torch's exact bucket tables)
        tail = self._tail_bucket_mb
        if tail is None:
            tail = os.environ.get("B200DDP_TAIL_BUCKET_MB", "1,0.0625" if backend == "p2p" else "0")
        tails = [float(t) for t in str(tail).split(",")] if not isinstance(tail, (int, float)) else [float(tail)]
        self._tails = tails
        if self.world > 1:
            # successive cuts (1 MiB, then 64 KiB by default): what is left exposed after backward's last kernel is
            # one latency-bound one-shot reduction of the stem's gradients; the 1 MiB before it hides behind the
            # stem's own backward kernels
            for t in tails:
                if t > 0:
                    self.bucket_of_param = split_tail_bucket(self.bucket_of_param, sizes, int(t * MiB))
        if backend == "p2p":
            total = sum((p.numel() + 63) // 64 * 64 for p in ordered) * 4
            buf_bytes = sum((b.numel() * b.element_size() + 255) // 256 * 256 for b in buffers)
            extra = 16 * MiB + 2 * buf_bytes + 2 * MiB  # staging, buffer region, SyncBN slots
            wire = (gradient_wire or os.environ.get("B200DDP_WIRE", "fp32")).lower()
            if wire == "bf16":
                extra += total // 2 + MiB          # bf16 staging copy of every bucket
            if not self._static_order:
                extra += total + MiB               # room for the one-off bucket rebuild in observed ready order
            self.comm = P2PComm(self.device, process_group, total + extra, algo=comm_algo, wire=wire)
        elif backend == "local":
            self.comm = LocalComm(process_group)
        elif backend == "torch":
            self.comm = TorchComm(process_group)
        else:
            raise ValueError("unknown backend %r" % backend)
        self.backend = backend

        # ---- flat storage; gradients are bucket views from now on
        self.store = FlatParameterStore(ordered, self.bucket_of_param, self.device,
                                        grad_alloc=self.comm.grad_alloc if backend == "p2p" else None)
        # weak back-references only: the store, the parameters and their autograd hooks must not keep the wrapper
        # alive, so that dropping the last reference to it releases its peer-memory window (see __del__ / close)
        wself = weakref.ref(self)

        def _ready(p, _w=wself):
            d = _w()
            if d is not None:
                d._on_param_ready(p)

        def _hook(p, _w=wself):
            d = _w()
            if d is not None:
                d._autograd_hook(p)

        self._weak_ready, self._weak_hook = _ready, _hook
        self.store.on_ready = _ready
        self._param_index = {id(p): i for i, p in enumerate(ordered)}
        self.ordered_params = ordered

        # ---- module buffers (BN running stats): one flat region; symmetric on the P2P backend so that the
        #      per-iteration sync (K3) is a single in-place pull from rank 0
        self._buffers_flat = None
        self._setup_buffers(buffers)

        # ---- K2: initial parameter + buffer broadcast from rank 0
        if init_sync and self.world > 1:
            self._sync_module_states()

        L = _lib.lib()
        self._reducer = None
        self._make_reducer()
        # K5: a module without a statically known gradient-ready order starts from reversed registration order; the
        # order observed in the first synchronised backward (rank 0's, broadcast over the control plane) then
        # replaces it once, before the second training forward - as c10d's Reducer::rebuild_buckets does
        self._rebuild_pending = (not self._static_order) and self.world > 1 and \
            os.environ.get("B200DDP_REBUILD_BUCKETS", "1") == "1"
        self._observed_order = None
        self.rebuilt_buckets = False
        self._callback_queued = False
        self._fused_args = None
        self._in_backward_iteration = False
        self._ready_ids = set()
        self.debug = os.environ.get("B200DDP_DEBUG", "0") == "1"
        self._iter_log_start = 0
        self.measure_exposed = os.environ.get("B200DDP_MEASURE_EXPOSED", "1") == "1"
        # device-side timestamps (csrc/misc.cu timestamp_kernel): [count, stamp ring] on the compute stream at the
        # end of backward and on the communication stream after the last reduction. Marker kernels are part of a
        # captured graph like any other launch, so exposed communication is measured in steady-state replays too.
        self._ts_cap = 256
        self._ts_compute = self._ts_comm = self._ts_start = None
        if self.measure_exposed and self.device.type == "cuda" and backend in ("p2p", "local") and _lib.use_native():
            self._ts_compute = torch.zeros(self._ts_cap + 1, dtype=torch.int64, device=self.device)
            self._ts_comm = torch.zeros(self._ts_cap + 1, dtype=torch.int64, device=self.device)
            self._ts_start = torch.zeros(self._ts_cap + 1, dtype=torch.int64, device=self.device)
        self.trace_buckets = os.environ.get("B200DDP_TRACE", "0") == "1" and backend == "p2p"
        # native launch path (csrc/reducer.cpp): marking a gradient ready launches the fused reductions from C++
        self._native_launch = backend == "p2p" and not self.trace_buckets and \
            os.environ.get("B200DDP_NATIVE_LAUNCH", "1") == "1"
        self._native_key = None
        self._bucket_rings = {}
        self._buffers_in_sync = bool(init_sync) or self.world == 1   # constructor broadcast done; False after load_state_dict
        self._post_forward_sync = backend == "p2p" and os.environ.get("B200DDP_BUFFER_SYNC", "post") == "post"
        self._buffer_sync_pending = False
        self._hooks = [p.register_post_accumulate_grad_hook(self._weak_hook) for p in ordered]
        self.iterations = 0
        self.xgpu_compute_ctx = self.comm.window.ctx(P2PComm.PAD_COMPUTE) if backend == "p2p" else None
        self._install_syncbn_ctx()
        _LIVE.add(self)
        self._closed = False
        self._watchdog_evt = None
        wd = int(os.environ.get("B200DDP_WATCHDOG_SEC", "0"))
        if wd > 0 and self.device.type == "cuda":
            L.b2_watchdog_start(wd, self.device.index or 0, 1)
            self._watchdog = True
        else:
            self._watchdog = False

    # ------------------------------------------------------------------ reducer / bucket rebuild
    def _make_reducer(self):
        L = _lib.lib()
        if self._reducer is not None:
            self._finalizer.detach()
            L.b2_reducer_destroy(self._reducer)
        n = len(self.ordered_params)
        arr = (ctypes.c_int * n)(*self.bucket_of_param)
        L.b2_reducer_create.restype = ctypes.c_void_p
        self._reducer = ctypes.c_void_p(L.b2_reducer_create(n, arr, len(self.store.buckets)))
        self._launch_buf = (ctypes.c_int * max(1, len(self.store.buckets)))()
        self._finalizer = weakref.finalize(self, L.b2_reducer_destroy, self._reducer)

    def _register_native_launch(self):
        """(Re-)register every bucket's launch descriptor with the native reducer when a buffer address changed
        (momentum buffers appear with the first fused step; a bucket rebuild replaces everything)."""
        comm = self.comm
        key = tuple((b.p32.data_ptr(), b.mom.data_ptr() if b.mom is not None else 0, b.g32.data_ptr())
                    for b in self.store.buckets)
        if key == self._native_key:
            return
        from .comm import ALGOS
        algo = comm.algo
        if algo == ALGOS["multimem"] and not comm.mc_base:
            algo = ALGOS["twoshot"]
        L = _lib.lib()
        for b in self.store.buckets:
            d = _BucketDesc()
            ctypes.memmove(ctypes.byref(d.ctx), ctypes.byref(comm.window.ctx(P2PComm.PAD_COMM)), ctypes.sizeof(d.ctx))
            d.mc_base = comm.mc_base or 0
            d.offset_bytes = comm.window.offset_of(b.g32)
            d.n, d.algo, d.scale = b.numel, algo, 1.0 / self.world
            d.p32, d.mom = b.p32.data_ptr(), (b.mom.data_ptr() if b.mom is not None else 0)
            d.shadow = b.shadow.data_ptr() if b.shadow is not None else 0
            d.write_grad, d.nblocks = int(comm.write_grad), comm.nblocks
            d.wire_bf16 = int(comm.wire == "bf16")
            if d.wire_bf16:
                if getattr(b, "wire16", None) is None:
                    b.wire16 = comm.alloc_tensor(b.numel, torch.bfloat16)
                d.wire_offset_bytes = comm.window.offset_of(b.wire16)
            d.g32 = b.g32.data_ptr()
            e = L.b2_reducer_set_bucket(self._reducer, b.index, ctypes.byref(d))
            if e:
                raise _lib.NativeError("b2_reducer_set_bucket failed with CUDA error %d" % e)
        self._native_key = key

    def _record_ready_order(self):
        """After the first synchronised backward: the order in which the gradients actually arrived."""
        n = len(self.ordered_params)
        out = (ctypes.c_int * n)()
        got = _lib.lib().b2_reducer_ready_order(self._reducer, out)
        if got == n:
            self._observed_order = list(out)

    def _rebuild_buckets(self):
        """K5 (``Reducer::rebuild_buckets`` + ``sync_bucket_indices`` of the reference stack): re-plan the buckets
        in rank 0's observed gradient-ready order so that reductions of a generic module overlap its backward."""
        self._rebuild_pending = False
        order = [self._observed_order]
        dist.broadcast_object_list(order, src=0, group=self.process_group)      # every rank adopts rank 0's order
        order = order[0]
        self.comm.launch_log.append(("broadcast_order", len(order)))
        if order is None or order == list(range(len(order))):
            return
        params = [self.ordered_params[i] for i in order]
        sizes = [p.numel() * p.element_size() for p in params]
        plan = plan_buckets(sizes, self._first_cap_bytes, self.bucket_cap_bytes)
        for t in self._tails:
            if t > 0:
                plan = split_tail_bucket(plan, sizes, int(t * MiB))
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        self.store.rebuild(params, plan, grad_alloc=self.comm.grad_alloc if self.backend == "p2p" else None)
        self.ordered_params = params
        self.bucket_of_param = plan
        self._param_index = {id(p): i for i, p in enumerate(params)}
        self._make_reducer()
        self._native_key = None
        self.rebuilt_buckets = True

    # ------------------------------------------------------------------ buffers
    def _setup_buffers(self, buffers):
        self._buffer_list = buffers
        if not buffers or self.backend != "p2p":
            return
        total = sum((b.numel() * b.element_size() + 255) // 256 * 256 for b in buffers)
        region = self.comm.alloc_tensor(total, torch.uint8)
        off = 0
        owners = {}
        for name, mod in self.module.named_modules():
            for bname, b in mod._buffers.items():
                if b is not None:
                    owners[id(b)] = (mod, bname)
        with torch.no_grad():
            for b in buffers:
                nb = b.numel() * b.element_size()
                v = region[off:off + nb].view(b.dtype).view(b.shape)
                v.copy_(b)
                mod, bname = owners[id(b)]
                mod._buffers[bname] = v
                off += (nb + 255) // 256 * 256
        self._buffers_flat = region
        self._buffer_list = [b for b in self.module.buffers()]

    def _sync_buffers(self, stream=None):
        """K3: broadcast BN buffers from rank 0 (before a training forward, or - P2P backend - right after it on the
        communication stream, see forward())."""
        if self.world == 1 or not self._buffer_list:
            return
        if self._buffers_flat is not None:
            self.comm.launch_log.append(("broadcast", len(self._buffer_list)))
            pad = None
            if stream is not None:
                ev = torch.cuda.Event()
                ev.record(torch.cuda.current_stream(self.device))
                stream.wait_event(ev)
                pad = P2PComm.PAD_COMM
            self.comm.broadcast_window_(self.comm.window.offset_of(self._buffers_flat), self._buffers_flat.numel(), 0,
                                        stream=stream, pad=pad)
        else:
            with torch.no_grad():
                self.comm.broadcast_(self._buffer_list, 0)

    def _sync_module_states(self):
        with torch.no_grad():
            if self.backend == "p2p":
                # stage parameters through the (symmetric) gradient buckets: copy, pull from rank 0, copy back
                self.comm.launch_log.append(("broadcast", len(self.ordered_params) + len(self._buffer_list)))
                for b in self.store.buckets:
                    if self.rank == 0:
                        b.g32.copy_(b.p32)
                    self.comm.broadcast_window_(self.comm.window.offset_of(b.g32), b.numel * 4, 0)
                    if self.rank != 0:
                        b.p32.copy_(b.g32)
                    b.g32.zero_()
                if self._buffers_flat is not None:
                    self.comm.broadcast_window_(self.comm.window.offset_of(self._buffers_flat),
                                                self._buffers_flat.numel(), 0)
                self.comm.barrier()
                torch.cuda.synchronize(self.device)
            else:
                self.comm.broadcast_([b.p32 for b in self.store.buckets] + list(self._buffer_list), 0)
        self.store.shadow_fresh = False

    def _install_syncbn_ctx(self):
        eng = getattr(self.module, "_engine", None)
        self._syncbn_ctx = self.xgpu_compute_ctx
        if hasattr(self.module, "_engine"):
            # the engine is created lazily; hand it the context through the model
            self.module._b2_xgpu_ctx = self.xgpu_compute_ctx
            slot = 0
            if self.backend == "p2p":
                # exchange area for per-GPU BN statistics: (sum, sumsq | sum_dy, sum_dy_xmu) + per-CTA counts
                slot = self.comm.window.alloc(2 * 4 * (2 * 2048 + 256)) // 4   # two areas, used alternately
            self.module._b2_xgpu_slot = slot
            if eng is not None:
                eng.xg = self.xgpu_compute_ctx
                eng.xg_slot = slot

    # ------------------------------------------------------------------ forward
    def _to_device(self, obj):
        """torch DDP moves the inputs to ``device_ids[0]`` (``_to_kwargs``); same here for (nested) tensors."""
        if isinstance(obj, torch.Tensor):
            return obj.to(self.device, non_blocking=True) if obj.device != self.device else obj
        if isinstance(obj, (list, tuple)):
            return type(obj)(self._to_device(o) for o in obj)
        if isinstance(obj, dict):
            return {k: self._to_device(v) for k, v in obj.items()}
        return obj

    def forward(self, *inputs, **kwargs):
        if self.device.type == "cuda":
            inputs, kwargs = self._to_device(inputs), self._to_device(kwargs)
        grad_mode = torch.is_grad_enabled()
        training_fwd = grad_mode and self.module.training
        if training_fwd:
            if self._ts_start is not None and self.require_backward_grad_sync:
                _lib.lib().b2_timestamp(ctypes.c_void_p(self._ts_start.data_ptr()), self._ts_cap, ctypes.c_void_p(
                    torch.cuda.current_stream(self.device).cuda_stream))
                _lib.count_launch()
            if self.require_backward_grad_sync:
                if self._rebuild_pending and self._observed_order is not None:
                    self._rebuild_buckets()
                _lib.lib().b2_reducer_reset(self._reducer)
                self._callback_queued = False
                self._fused_args = None
                self._in_backward_iteration = True
                self._ready_ids = set()
                self.comm.begin_backward()
                self._trim_log()
            if self._buffer_sync_pending:
                # a buffer broadcast of the previous forward may still be in flight on the communication stream
                # (only when no backward ran in between: backward's finalize already joined the streams)
                torch.cuda.current_stream(self.device).wait_stream(self.comm.stream)
                self._buffer_sync_pending = False
            if self.broadcast_buffers and not (self._post_forward_sync and self._buffers_in_sync):
                self._sync_buffers()
        out = self.module(*inputs, **kwargs)
        if training_fwd and self.broadcast_buffers and self._post_forward_sync and self.world > 1 \
                and self._buffers_flat is not None:
            # K3, moved off the critical path: the running statistics this forward just updated are broadcast from
            # rank 0 on the communication stream while backward computes, instead of before the next forward on the
            # compute stream (two device barriers in front of the stem). Every training forward therefore still
            # starts from rank 0's buffers, as with torch DDP's pre-forward broadcast (reference resnet/main.py:121).
            self._sync_buffers(stream=self.comm.stream)
            self._buffers_in_sync = True
            self._buffer_sync_pending = True
        return out

    def _trim_log(self):
        """The collective log only ever needs the current iteration (debug order check, tests)."""
        log = self.comm.launch_log
        if len(log) > 4096:
            del log[:len(log) - 256]
            self._iter_log_start = min(self._iter_log_start, len(log))

    # ------------------------------------------------------------------ reducer hooks
    def _autograd_hook(self, p):
        """Generic path: autograd accumulated p.grad. Make it a bucket view (once), then mark ready."""
        gv = self.store.grad_view(p)
        if p.grad is not None and p.grad.data_ptr() != gv.data_ptr():
            gv.copy_(p.grad)
            p.grad = gv
        self._on_param_ready(p, from_autograd=True)

    def _on_param_ready(self, p, from_autograd=False):
        if not self.require_backward_grad_sync or not self._in_backward_iteration:
            return
        idx = self._param_index.get(id(p))
        if idx is None:
            return
        if from_autograd and id(p) in self._ready_ids:
            return  # the native engine already delivered this gradient; autograd's hook still fires afterwards
        self._ready_ids.add(id(p))
        if not self._callback_queued:
            self._callback_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(self._finalize_backward)
        if self._native_launch:
            return self._mark_ready_native(idx)
        n = _lib.lib().b2_reducer_mark_ready(self._reducer, idx, self._launch_buf)
        if n == -1:
            raise RuntimeError("a parameter was marked ready twice in one iteration (second backward() through "
                               "the same forward? use no_sync() for gradient accumulation)")
        for i in range(n):
            self._launch_bucket(self.store.buckets[self._launch_buf[i]])

    def _fused_sgd_args(self):
        opt = self.store.optimizer
        if opt is not None and opt.fuse_with_backward and self.backend in ("p2p", "local") and opt._native_ok():
            if self._fused_args is None:
                self._fused_args = opt.prepare_fused(self.store)
            return self._fused_args
        return None

    def _mark_ready_native(self, idx):
        """One C call: mark ready + launch every bucket that became launchable (events, stream wait, kernel)."""
        sgd = self._fused_sgd_args()
        if self._native_key is None or sgd is not None:
            self._register_native_launch()
        lr, mu, wd, first, hyper = (tuple(sgd) + (None,))[:5] if sgd is not None else (0.0, 0.0, 0.0, 0, None)
        cur = torch.cuda.current_stream(self.device)
        n = _lib.lib().b2_reducer_mark_ready_launch(
            self._reducer, idx, int(sgd is not None), ctypes.c_float(lr), ctypes.c_float(mu), ctypes.c_float(wd),
            int(first), ctypes.c_void_p(hyper.data_ptr() if hyper is not None else 0),
            ctypes.c_void_p(cur.cuda_stream), ctypes.c_void_p(self.comm.stream.cuda_stream), self._launch_buf)
        if n == -1:
            raise RuntimeError("a parameter was marked ready twice in one iteration (second backward() through "
                               "the same forward? use no_sync() for gradient accumulation)")
        if n < -1:
            raise _lib.NativeError("native bucket launch failed (%d)" % n)
        opt = self.store.optimizer
        for i in range(n):
            b = self.store.buckets[self._launch_buf[i]]
            self.comm.launch_log.append(("allreduce", b.index))
            _lib.count_launch()
            if sgd is not None:
                opt.note_bucket_updated(b)

    def _launch_bucket(self, bucket):
        opt = self.store.optimizer
        sgd = None
        if opt is not None and opt.fuse_with_backward and self.backend in ("p2p", "local") and opt._native_ok():
            if self._fused_args is None:
                self._fused_args = opt.prepare_fused(self.store)
            sgd = self._fused_args
        if self.world == 1 and sgd is None:
            self.comm.launch_log.append(("allreduce", bucket.index))
            return
        if self.trace_buckets:
            self._stamp((bucket.index, "ready"), torch.cuda.current_stream(self.device))
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            self.comm.stream.wait_event(ev)
            self._stamp((bucket.index, "launch"), self.comm.stream)
        self.comm.allreduce_bucket(bucket, sgd=sgd)
        if self.trace_buckets:
            self._stamp((bucket.index, "done"), self.comm.stream)
        if sgd is not None:
            opt.note_bucket_updated(bucket)

    def _finalize_backward(self):
        self._callback_queued = False
        self._in_backward_iteration = False
        left = _lib.lib().b2_reducer_unlaunched(self._reducer)
        if self._rebuild_pending and self._observed_order is None and not left:
            self._record_ready_order()
        if self._ts_compute is not None:
            # exposed communication = how long the reductions outlast backward's own kernels (device clock)
            L = _lib.lib()
            L.b2_timestamp(ctypes.c_void_p(self._ts_compute.data_ptr()), self._ts_cap,
                           ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream))
            comm_stream = getattr(self.comm, "stream", None) or torch.cuda.current_stream(self.device)
            L.b2_timestamp(ctypes.c_void_p(self._ts_comm.data_ptr()), self._ts_cap,
                           ctypes.c_void_p(comm_stream.cuda_stream))
            _lib.count_launch(2)
        if self._native_launch and self._native_key is not None:
            e = _lib.lib().b2_reducer_join(self._reducer, ctypes.c_void_p(torch.cuda.current_stream(
                self.device).cuda_stream), ctypes.c_void_p(self.comm.stream.cuda_stream))
            if e:
                raise _lib.NativeError("b2_reducer_join failed with CUDA error %d" % e)
        else:
            self.comm.finalize()
        self._buffer_sync_pending = False
        if self._fused_args is not None:
            self.store.mark_params_updated_natively()
        if self._watchdog:
            self._watchdog_evt = torch.cuda.Event()
            self._watchdog_evt.record()
            _lib.lib().b2_watchdog_watch(ctypes.c_void_p(self._watchdog_evt.cuda_event))
        self.iterations += 1
        if self.debug and self.world > 1 and not torch.cuda.is_current_stream_capturing():
            self._check_collective_order()
        if left:
            raise RuntimeError("%d gradient bucket(s) never became ready: some parameters took no part in the "
                               "loss (find_unused_parameters is not supported)" % left)

    def _check_collective_order(self):
        """Debug mode (B200DDP_DEBUG=1): the epoch-numbered device barriers require every rank to issue the same
        kernels in the same order; compare this iteration's collective sequence across ranks over the CPU control
        plane and fail loudly on the first divergence instead of hanging in a spin loop."""
        mine = [tuple(c) for c in self.comm.launch_log[self._iter_log_start:]]
        self._iter_log_start = len(self.comm.launch_log)
        seqs = [None] * self.world
        dist.all_gather_object(seqs, mine, group=self.process_group)
        for r, s in enumerate(seqs):
            if s != seqs[0]:
                raise RuntimeError("collective order diverged at iteration %d: rank 0 issued %s, rank %d issued %s"
                                   % (self.iterations, seqs[0][:8], r, s[:8]))

    @contextlib.contextmanager
    def no_sync(self):
        """Accumulate gradients locally; the next backward outside the context reduces the sum."""
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    def exposed_comm_ms(self, last=None):
        """Mean time per step by which the gradient reductions outlasted backward compute, from device-side
        timestamps (valid for eager steps and for CUDA-graph replays alike). ``last``: only the most recent N steps
        (at most 256 are kept)."""
        if self.world == 1:
            return 0.0
        if self._ts_compute is None:
            return None
        torch.cuda.synchronize(self.device)
        tc, tm = self._ts_compute.cpu(), self._ts_comm.cpu()
        n = int(min(tc[0].item(), tm[0].item()))
        if n == 0:
            return None
        k = min(n, self._ts_cap, last or self._ts_cap)
        vals = []
        for i in range(n - k, n):
            a, b = int(tc[1 + i % self._ts_cap]), int(tm[1 + i % self._ts_cap])
            vals.append(max(0.0, (b - a) / 1e6))
        return sum(vals) / len(vals)

    def reset_exposed_comm(self):
        if self._ts_compute is not None:
            self._ts_compute.zero_()
            self._ts_comm.zero_()
            self._ts_start.zero_()

    def step_anatomy(self, last=None):
        """Where a training step's time goes, from the device-side stamps of the most recent steps (ms, means):
        ``period`` = start of a step's forward to the start of the next one; ``fwd_bwd`` = start of forward to the end
        of backward's last kernel; ``tail`` = how long the last reduction outlasts that (exposed communication);
        ``gap`` = end of the step's device work to the start of the next step (launch / replay hand-over)."""
        if self._ts_start is None:
            return None
        torch.cuda.synchronize(self.device)
        ts, tc, tm = self._ts_start.cpu(), self._ts_compute.cpu(), self._ts_comm.cpu()
        n = int(min(ts[0].item(), tc[0].item(), tm[0].item()))
        k = min(n, self._ts_cap - 1, last or self._ts_cap)
        if k < 2:
            return None
        cap = self._ts_cap
        per, busy, tail, gap = [], [], [], []
        for i in range(n - k, n - 1):
            s0, s1 = int(ts[1 + i % cap]), int(ts[1 + (i + 1) % cap])
            c, m = int(tc[1 + i % cap]), int(tm[1 + i % cap])
            per.append((s1 - s0) / 1e6)
            busy.append((c - s0) / 1e6)
            tail.append(max(0, m - c) / 1e6)
            gap.append((s1 - max(c, m)) / 1e6)
        f = lambda v: round(sum(v) / len(v), 4)   # noqa: E731
        return {"period": f(per), "fwd_bwd": f(busy), "tail": f(tail), "gap": f(gap), "steps": len(per)}

    # ------------------------------------------------------------------ per-bucket trace (B200DDP_TRACE=1)
    def _stamp(self, key, stream):
        ring = self._bucket_rings.get(key)
        if ring is None:
            ring = self._bucket_rings[key] = torch.zeros(self._ts_cap + 1, dtype=torch.int64, device=self.device)
        _lib.lib().b2_timestamp(ctypes.c_void_p(ring.data_ptr()), self._ts_cap, ctypes.c_void_p(stream.cuda_stream))
        _lib.count_launch()

    def bucket_trace(self):
        """Per-bucket ``ready`` (compute stream: its last gradient was written), ``launch`` (communication stream
        reached the reduction) and ``done`` times of the most recent step in microseconds relative to the first
        bucket's ready time, plus the end of backward compute. Needs ``B200DDP_TRACE=1`` (three marker kernels per
        bucket, also inside graph replays). The flight-recorder view of ProcessGroupNCCL, per bucket."""
        if not self._bucket_rings:
            return None
        torch.cuda.synchronize(self.device)

        def last(ring):
            r = ring.cpu()
            n = int(r[0])
            return int(r[1 + (n - 1) % self._ts_cap]) if n else None

        rows, t0 = [], None
        for b in self.store.buckets:
            ent = {k: last(self._bucket_rings[(b.index, k)]) for k in ("ready", "launch", "done")
                   if (b.index, k) in self._bucket_rings}
            if ent.get("ready") is not None and t0 is None:
                t0 = ent["ready"]
            rows.append((b.index, b.numel * 4, ent))
        out = [{"bucket": i, "bytes": nb, **{k: (v - t0) / 1e3 for k, v in ent.items() if v is not None}}
               for i, nb, ent in rows]
        end = last(self._ts_compute) if self._ts_compute is not None else None
        return {"buckets": out, "backward_end_us": (end - t0) / 1e3 if end and t0 else None}

    # ------------------------------------------------------------------ teardown
    def close(self):
        """Release the symmetric window of this wrapper (P2P backend). The module stays usable: its buffers are
        moved back to ordinary device memory and gradients that live in the window are dropped first."""
        if self._closed:
            return
        self._closed = True
        if self.backend != "p2p":
            return
        torch.cuda.synchronize(self.device)
        with torch.no_grad():
            if self._buffers_flat is not None:
                for mod in self.module.modules():
                    for bname, b in list(mod._buffers.items()):
                        if b is not None and b.is_cuda:
                            mod._buffers[bname] = b.clone()
            for p in self.ordered_params:
                p.grad = None
        for b in self.store.buckets:
            b.g32 = torch.zeros(0, dtype=torch.float32, device=self.device)
            if hasattr(b, "wire16"):
                b.wire16 = None
        self._buffers_flat = None
        self._buffer_list = [b for b in self.module.buffers()]
        self.comm._stage = None
        if hasattr(self.module, "_b2_xgpu_ctx"):
            self.module._b2_xgpu_ctx = None
            eng = getattr(self.module, "_engine", None)
            if eng is not None:
                eng.xg = None
        self.xgpu_compute_ctx = None
        self.store.on_ready = None
        for h in self._hooks:
            h.remove()
        self.comm.window.close()

    def __del__(self):
        try:
            if not getattr(self, "_closed", True):
                self.close()
        except Exception:  # noqa: BLE001  (interpreter shutdown)
            pass

    def load_state_dict(self, state_dict, strict=True, **kw):
        out = super().load_state_dict(state_dict, strict=strict, **kw)
        self._buffers_in_sync = False    # ranks may have loaded different files: the next forward syncs first
        return out

    # ------------------------------------------------------------------ introspection
    def bucket_sizes(self):
        return [b.used for b in self.store.buckets]

    def collective_log(self):
        return list(self.comm.launch_log)
