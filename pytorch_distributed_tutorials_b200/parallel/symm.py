"""Symmetric (peer-mapped) device memory for one node: every rank allocates a window of the same size, exports
it, and maps all peers' windows, so kernels can load/store any rank's buffers over NVLink (csrc/symm.cu).

This is the framework's stand-in for ProcessGroupNCCL's communicator set-up (reference resnet/main.py:74): the
torch process group is used only as the control plane to exchange 64-byte IPC handles (or, for the VMM +
multicast transport, UNIX-socket names over which file descriptors are passed).
"""
import ctypes
import os
import weakref

import torch
import torch.distributed as dist

from ..ops import _lib
from ..ops.native import XgpuCtx

CHANNELS = 256            # B2_XGPU_CHANNELS
PAD_BYTES = CHANNELS * 8 * 4   # one signal pad: 256 channels x 8 ranks x u32
N_PADS = 8
HEADER_BYTES = 64 << 10    # pads live at the start of every window
MAX_WORLD = 8


class _RawCuda:
    """Minimal __cuda_array_interface__ carrier so torch can alias raw device memory."""

    def __init__(self, ptr, nbytes, owner):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}
        self._owner = owner


def _check(err, what):
    if err != 0:
        raise RuntimeError("%s failed (error %d)" % (what, err))


def _release(res):
    """Undo everything a window's set-up created (``res`` is the plain dict of raw handles and addresses, so this
    can run from a ``weakref.finalize`` after the window object itself is gone). Local only - no collectives: a
    peer that still maps this rank's memory keeps the physical pages alive through its own imported handle."""
    if res.get("released"):
        return
    res["released"] = True
    try:
        L = _lib.lib()
        errs = res.setdefault("errors", [])

        def call(what, rc):
            if rc != 0:
                errs.append((what, rc))

        if torch.cuda.is_available():
            torch.cuda.synchronize(res["device"])
        nbytes = ctypes.c_longlong(res["nbytes"])
        if res["transport"] == "vmm":
            if res.get("mc_handle") is not None:
                if res.get("mc_bound"):
                    call("mc_unbind", L.b2_mc_unbind(ctypes.c_ulonglong(res["mc_handle"]), res["device"], nbytes))
                if res.get("mc_ptr"):
                    call("mc_unmap", L.b2_vmm_unmap(ctypes.c_void_p(res["mc_ptr"]), nbytes))
                call("mc_release", L.b2_vmm_release(ctypes.c_ulonglong(res["mc_handle"])))
            for ptr_, h in res.get("peers", []):
                call("peer_unmap", L.b2_vmm_unmap(ctypes.c_void_p(ptr_), nbytes))
                call("peer_release", L.b2_vmm_release(ctypes.c_ulonglong(h)))
            if res.get("local_ptr"):
                call("local_unmap", L.b2_vmm_unmap(ctypes.c_void_p(res["local_ptr"]), nbytes))
            if res.get("local_handle") is not None:
                call("local_release", L.b2_vmm_release(ctypes.c_ulonglong(res["local_handle"])))
        elif res["transport"] == "ipc":
            for ptr_, _ in res.get("peers", []):
                call("ipc_close", L.b2_ipc_close(ctypes.c_void_p(ptr_)))
            if res.get("local_ptr"):
                call("ipc_free", L.b2_ipc_free(ctypes.c_void_p(res["local_ptr"])))
    except Exception:  # noqa: BLE001  (interpreter shutdown: the driver reclaims everything anyway)
        pass


class SymmetricWindow:
    def __init__(self, nbytes, device, group=None, transport=None):
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        if self.world > MAX_WORLD:
            raise ValueError("the P2P backend covers one NVSwitch domain (<= %d GPUs)" % MAX_WORLD)
        self.device = torch.device(device)
        self.transport = transport or os.environ.get("B200DDP_SYMM", "auto")
        self.mc_ptr = None
        self._cursor = HEADER_BYTES
        L = _lib.lib()
        torch.cuda.set_device(self.device)
        dev_index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        # raw resources, released by close() or, when the last tensor aliasing the window is gone, by the finalizer
        self._res = {"transport": None, "device": dev_index, "nbytes": 0, "peers": [], "local_ptr": None,
                     "local_handle": None, "mc_handle": None, "mc_ptr": None, "mc_bound": False}
        nbytes = (int(nbytes) + HEADER_BYTES + (2 << 20) - 1) // (2 << 20) * (2 << 20)
        self.nbytes = nbytes
        self.ptrs = [0] * self.world
        ok = False
        if self.transport in ("auto", "vmm"):
            try:
                ok = self._setup_vmm(L)
            except Exception as e:  # noqa: BLE001
                if self.transport == "vmm":
                    raise
                ok = False
                self._vmm_error = repr(e)
            # every rank must agree on the transport
            flags = [None] * self.world
            dist.all_gather_object(flags, bool(ok), group=group)
            if not all(flags):
                if ok:
                    raise RuntimeError("VMM transport came up on this rank but not on all ranks")
                ok = False
        if not ok:
            self._setup_ipc(L)
            self.transport = "ipc"
        else:
            self.transport = "vmm"
        self.epochs = torch.zeros(N_PADS * CHANNELS, dtype=torch.int32, device=self.device)
        self._ctx_cache = {}
        self._res.update(transport=self.transport, nbytes=self.nbytes, local_ptr=self.local_ptr, mc_ptr=self.mc_ptr)
        self._finalizer = weakref.finalize(self, _release, self._res)
        dist.barrier(group=group)

    @property
    def closed(self):
        return bool(self._res.get("released"))

    def close(self):
        """Release the window: multicast unbind, unmap + address-range free of every mapping (own, peers',
        multicast), release of every handle (``cuMemRelease``), or ``cudaIpcCloseMemHandle`` / ``cudaFree`` for
        the IPC transport. Every tensor that aliases the window is invalid afterwards (``DistributedDataParallel
        .close()`` moves the module's buffers and gradients out first). Local operation; idempotent. Without an
        explicit call the same happens when the last tensor aliasing the window is garbage-collected."""
        self._finalizer()
        self.mc_ptr = None
        self._ctx_cache = {}

    # ------------------------------------------------------------------ transports
    def _setup_ipc(self, L):
        p = ctypes.c_void_p()
        h = (ctypes.c_ubyte * 64)()
        _check(L.b2_ipc_alloc(ctypes.c_longlong(self.nbytes), ctypes.byref(p), h), "b2_ipc_alloc")
        self.local_ptr = p.value
        self._res.update(transport="ipc", nbytes=self.nbytes, local_ptr=self.local_ptr)
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(h), group=self.group)
        for r in range(self.world):
            if r == self.rank:
                self.ptrs[r] = self.local_ptr
            else:
                q = ctypes.c_void_p()
                hb = (ctypes.c_ubyte * 64).from_buffer_copy(handles[r])
                _check(L.b2_ipc_open(hb, ctypes.byref(q)), "b2_ipc_open(rank %d)" % r)
                self.ptrs[r] = q.value
                self._res["peers"].append((q.value, 0))

    def _setup_vmm(self, L):
        L.b2_vmm_round.restype = ctypes.c_longlong
        dev = self.device.index if self.device.index is not None else torch.cuda.current_device()
        rounded = L.b2_vmm_round(ctypes.c_longlong(self.nbytes), dev, self.world)
        if rounded <= 0:
            return False
        self.nbytes = int(rounded)
        p = ctypes.c_void_p()
        mh = ctypes.c_ulonglong()
        fd = ctypes.c_int()
        _check(L.b2_vmm_alloc(ctypes.c_longlong(self.nbytes), dev, ctypes.byref(p), ctypes.byref(mh),
                              ctypes.byref(fd)), "b2_vmm_alloc")
        self.local_ptr = p.value
        self._res.update(transport="vmm", nbytes=self.nbytes, local_ptr=self.local_ptr, local_handle=mh.value)
        want_mc = bool(L.b2_vmm_multicast_supported(dev)) and os.environ.get("B200DDP_NVLS", "1") == "1"
        fds = [fd.value]
        mc_handle = ctypes.c_ulonglong()
        if want_mc and self.rank == 0:
            mfd = ctypes.c_int()
            if L.b2_mc_create(ctypes.c_longlong(self.nbytes), self.world, ctypes.byref(mc_handle),
                              ctypes.byref(mfd)) == 0:
                fds.append(mfd.value)
            else:
                want_mc = False
        # serve my descriptors to every peer from a helper thread while I fetch theirs
        name = "b2symm-%s-%d-%d" % (os.environ.get("MASTER_PORT", "0"), os.getpid(), id(self) & 0xFFFF)
        sock = L.b2_fd_listen(name.encode())
        if sock < 0:
            raise RuntimeError("cannot open the descriptor-passing socket")
        info = [None] * self.world
        dist.all_gather_object(info, (name, len(fds), want_mc), group=self.group)
        want_mc = all(i[2] for i in info) and info[0][1] == 2
        import threading
        arr = (ctypes.c_int * len(fds))(*fds)
        n_serve = (self.world - 1) * (2 if (want_mc and self.rank == 0) else 1)
        srv = threading.Thread(target=lambda: L.b2_fd_serve(sock, arr, len(fds), n_serve), daemon=True)
        srv.start()
        for r in range(self.world):
            if r == self.rank:
                self.ptrs[r] = self.local_ptr
                continue
            pfd = L.b2_fd_fetch(info[r][0].encode(), 0, 30000)
            if pfd < 0:
                raise RuntimeError("fetching the window descriptor of rank %d failed (%d)" % (r, pfd))
            q = ctypes.c_void_p()
            ph = ctypes.c_ulonglong()
            _check(L.b2_vmm_import_h(pfd, ctypes.c_longlong(self.nbytes), dev, ctypes.byref(q), ctypes.byref(ph)),
                   "b2_vmm_import_h")
            L.b2_fd_close(pfd)
            self.ptrs[r] = q.value
            self._res["peers"].append((q.value, ph.value))
        if want_mc:
            if self.rank != 0:
                mfd = L.b2_fd_fetch(info[0][0].encode(), 1, 30000)
                if mfd < 0:
                    raise RuntimeError("fetching the multicast descriptor failed (%d)" % mfd)
                _check(L.b2_mc_import(mfd, ctypes.byref(mc_handle)), "b2_mc_import")
                L.b2_fd_close(mfd)
            self._res["mc_handle"] = mc_handle.value
            _check(L.b2_mc_add_device(mc_handle, dev), "b2_mc_add_device")
            dist.barrier(group=self.group)  # every device added before anyone binds
            mp = ctypes.c_void_p()
            _check(L.b2_mc_bind_and_map(mc_handle, mh, ctypes.c_longlong(self.nbytes), dev, ctypes.byref(mp)),
                   "b2_mc_bind_and_map")
            self.mc_ptr = mp.value
            self._res.update(mc_ptr=mp.value, mc_bound=True)
        elif self.rank == 0 and len(fds) == 2:
            self._res["mc_handle"] = mc_handle.value   # created but not used by every rank: still to be released
        srv.join(timeout=30)
        L.b2_fd_close(sock)
        for f in fds:
            L.b2_fd_close(f)
        return True

    # ------------------------------------------------------------------ sub-allocation (identical on all ranks)
    def alloc(self, nbytes, align=256):
        off = (self._cursor + align - 1) // align * align
        if off + nbytes > self.nbytes:
            raise MemoryError("symmetric window exhausted (%d + %d > %d)" % (off, nbytes, self.nbytes))
        self._cursor = off + nbytes
        return off

    def tensor(self, offset, numel, dtype):
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        raw = _RawCuda(self.local_ptr + offset, nbytes, self)
        return torch.as_tensor(raw, device=self.device).view(dtype)

    def peer_tensor(self, rank, offset, numel, dtype):
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        raw = _RawCuda(self.ptrs[rank] + offset, nbytes, self)
        return torch.as_tensor(raw, device=self.device).view(dtype)

    def offset_of(self, t):
        return t.data_ptr() - self.local_ptr

    def ctx(self, pad=0):
        """``b2_xgpu_ctx`` for kernels; ``pad`` selects an independent signal pad (one per concurrent stream)."""
        c = self._ctx_cache.get(pad)
        if c is None:
            c = XgpuCtx()
            c.rank, c.world = self.rank, self.world
            for r in range(self.world):
                c.data[r] = self.ptrs[r]
                c.flags[r] = self.ptrs[r] + pad * PAD_BYTES
            c.epoch = self.epochs.data_ptr() + pad * CHANNELS * 4
            self._ctx_cache[pad] = c
        return c
