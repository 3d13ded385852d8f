"""Symmetric device buffers: the same allocation on every rank, mapped into every process.

Two back-ends, tried in this order:

``vmm``   the native ``hpnn_symm`` object (csrc/core/symm.cu): CUDA virtual-memory-management
          allocations exported as POSIX file descriptors, which this module passes between the
          processes (``SCM_RIGHTS`` over unix sockets; the socket paths travel through
          ``torch.distributed``), plus -- where the fabric has NVLS -- one *multicast* mapping of the
          whole set: ``mc_ptr`` is an address through which one ``multimem.st`` / ``multimem.red``
          reaches every rank's copy and ``multimem.ld_reduce`` returns the sum over the ranks;
``ipc``   ``cudaMalloc`` + ``cudaIpc*`` handles (no multicast), the round-1 path, kept as fall-back.

``ptrs[q]`` is rank q's buffer as seen from this process: kernels load from / store to any of them
and the traffic goes over NVLink.  The reference's counterpart is a gather-to-GPU-0 + broadcast by
``cudaMemcpyAsync`` (src/cuda_ann.cu:610-666) and ``MPI_Allgather`` (src/ann.c:925)."""
from __future__ import annotations

import ctypes as C
import os
import socket
import threading
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from .. import _native as N
from ..models.net import _DevBuf

_seq = 0


def _bind_symm(lib):
    if getattr(lib, "_symm_bound", False):
        return
    vp, i = C.c_void_p, C.c_int
    lib.hpnn_symm_multicast_supported.restype, lib.hpnn_symm_multicast_supported.argtypes = i, [i]
    lib.hpnn_symm_create_shared.restype, lib.hpnn_symm_create_shared.argtypes = vp, [i, i, i, C.c_size_t, i]
    lib.hpnn_symm_create_local.restype, lib.hpnn_symm_create_local.argtypes = vp, [C.POINTER(i), i, C.c_size_t, i]
    lib.hpnn_symm_export_fd.restype, lib.hpnn_symm_export_fd.argtypes = i, [vp]
    lib.hpnn_symm_import_peer.restype, lib.hpnn_symm_import_peer.argtypes = i, [vp, i, i]
    lib.hpnn_symm_mc_create.restype, lib.hpnn_symm_mc_create.argtypes = i, [vp]
    lib.hpnn_symm_mc_import.restype, lib.hpnn_symm_mc_import.argtypes = i, [vp, i]
    lib.hpnn_symm_mc_add_device.restype, lib.hpnn_symm_mc_add_device.argtypes = i, [vp]
    lib.hpnn_symm_mc_bind.restype, lib.hpnn_symm_mc_bind.argtypes = i, [vp]
    lib.hpnn_symm_mc_drop.restype, lib.hpnn_symm_mc_drop.argtypes = None, [vp]
    lib.hpnn_symm_ptr.restype, lib.hpnn_symm_ptr.argtypes = vp, [vp, i]
    lib.hpnn_symm_mc_ptr.restype, lib.hpnn_symm_mc_ptr.argtypes = vp, [vp]
    lib.hpnn_symm_size.restype, lib.hpnn_symm_size.argtypes = C.c_size_t, [vp]
    lib.hpnn_symm_destroy.restype, lib.hpnn_symm_destroy.argtypes = None, [vp]
    lib._symm_bound = True


def _all_ok(flag: bool, group) -> bool:
    """Logical AND of a per-rank flag over the group (every rank gets the same answer)."""
    box = [None] * dist.get_world_size(group)
    dist.all_gather_object(box, bool(flag), group=group)
    return all(box)


def _pass_fds(my_fd: Optional[int], senders: List[int], rank: int, world: int, group) -> Dict[int, int]:
    """Every rank in ``senders`` hands ``my_fd`` to every other rank; returns {sender: received fd}.
    File descriptors cannot be pickled: they cross process boundaries as SCM_RIGHTS ancillary data on
    a unix socket; only the socket paths go through torch.distributed."""
    global _seq
    _seq += 1
    path = "/tmp/hpnn_symm_%d_%d_%d.sock" % (os.getuid(), os.getpid(), _seq)
    try:
        os.unlink(path)
    except OSError:
        pass
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    srv.bind(path)
    srv.listen(world)
    paths = [None] * world
    dist.all_gather_object(paths, path, group=group)
    expect = len([s for s in senders if s != rank])
    got: Dict[int, int] = {}

    def serve():
        for _ in range(expect):
            conn, _a = srv.accept()
            msg, fds, _f, _ad = socket.recv_fds(conn, 32, 1)
            got[int(msg.decode())] = fds[0]
            conn.close()

    th = threading.Thread(target=serve, daemon=True)
    th.start()
    if rank in senders:
        for q in range(world):
            if q == rank:
                continue
            c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            c.connect(paths[q])
            socket.send_fds(c, [str(rank).encode()], [my_fd])
            c.close()
    th.join(timeout=120)
    srv.close()
    try:
        os.unlink(path)
    except OSError:
        pass
    dist.barrier(group=group)
    return got


class SymmetricBuffer:
    """``n_doubles`` fp64 values on every rank of ``group``.  ``backend``: "auto" | "vmm" | "ipc"
    (env ``HPNN_SYMM`` overrides "auto"); ``multicast`` asks for the NVLS mapping (vmm only)."""

    def __init__(self, n_doubles: int, group=None, device=None, backend: str = "auto", multicast: bool = True):
        self.lib = N.load()
        _bind_symm(self.lib)
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        torch.cuda.set_device(self.device)
        self.n = n_doubles
        self.ptrs: List[int] = [0] * self.world
        self.mc_ptr: int = 0
        self._symm = None
        self._ptr = None
        self._opened: List[int] = []
        if backend == "auto":
            backend = os.environ.get("HPNN_SYMM", "auto")
        self.backend = None
        if backend in ("auto", "vmm") and self.world > 1:
            if self._try_vmm(multicast and os.environ.get("HPNN_MULTICAST", "1") != "0"):
                self.backend = "vmm"
            elif backend == "vmm":
                raise RuntimeError("symmetric memory through the VMM API is unavailable: " + N.last_error())
        if self.backend is None:
            self._ipc()
            self.backend = "ipc"
        self.local = torch.as_tensor(_DevBuf(self.ptrs[self.rank], n_doubles), device=self.device)

    # ------------------------------------------------------------------ back-ends
    def _try_vmm(self, want_mc: bool) -> bool:
        lib, g = self.lib, self.group
        dev = self.device.index if self.device.index is not None else torch.cuda.current_device()
        s = lib.hpnn_symm_create_shared(dev, self.rank, self.world, self.n * 8, int(want_mc))
        if not _all_ok(bool(s), g):
            if s:
                lib.hpnn_symm_destroy(s)
            return False
        fd = lib.hpnn_symm_export_fd(s)
        ok = _all_ok(fd >= 0, g)
        got: Dict[int, int] = {}
        if ok:
            got = _pass_fds(fd, list(range(self.world)), self.rank, self.world, g)
            ok = len(got) == self.world - 1
            for q, pfd in got.items():
                ok = ok and lib.hpnn_symm_import_peer(s, q, pfd) == 0
        for pfd in got.values():
            os.close(pfd)
        if fd >= 0:
            os.close(fd)
        if not _all_ok(ok, g):
            lib.hpnn_symm_destroy(s)
            return False
        self._symm = s
        for q in range(self.world):
            self.ptrs[q] = lib.hpnn_symm_ptr(s, q)
        # ---- multicast object: rank 0 creates, everybody adds its device, then binds
        if want_mc:
            mcfd = lib.hpnn_symm_mc_create(s) if self.rank == 0 else -1
            flag = [mcfd >= 0]
            dist.broadcast_object_list(flag, src=dist.get_global_rank(g, 0) if g is not None else 0, group=g)
            if flag[0]:
                got = _pass_fds(mcfd if self.rank == 0 else None, [0], self.rank, self.world, g)
                ok = True
                if self.rank != 0:
                    ok = 0 in got and lib.hpnn_symm_mc_import(s, got[0]) == 0
                    for pfd in got.values():
                        os.close(pfd)
                else:
                    os.close(mcfd)
                ok = ok and lib.hpnn_symm_mc_add_device(s) == 0
                ok = _all_ok(ok, g)  # (also the barrier: every device is added before anyone binds)
                if ok:
                    ok = lib.hpnn_symm_mc_bind(s) == 0
                    ok = _all_ok(ok, g)
                if ok:
                    self.mc_ptr = lib.hpnn_symm_mc_ptr(s) or 0
                else:
                    lib.hpnn_symm_mc_drop(s)
        torch.cuda.synchronize(self.device)
        dist.barrier(group=g)
        return True

    def _ipc(self):
        self._ptr = self.lib.hpnn_dev_alloc(self.n * 8)
        if not self._ptr:
            raise MemoryError(N.last_error())
        self.ptrs[self.rank] = self._ptr
        if self.world > 1:
            h = C.create_string_buffer(64)
            N.check(self.lib.hpnn_ipc_export(self._ptr, h), "hpnn_ipc_export")
            everyone = [None] * self.world
            dist.all_gather_object(everyone, (self.rank, bytes(h.raw)), group=self.group)
            for r, raw in everyone:
                if r != self.rank:
                    p = self.lib.hpnn_ipc_open(raw)
                    if not p:
                        raise RuntimeError("hpnn_ipc_open: " + N.last_error())
                    self.ptrs[r] = p
                    self._opened.append(p)
            dist.barrier(group=self.group)

    def close(self):
        if self._ptr is None and self._symm is None:
            return
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier(group=self.group)
        self.local = None
        if self._symm is not None:
            self.lib.hpnn_symm_destroy(self._symm)
            self._symm = None
        else:
            for p in self._opened:
                self.lib.hpnn_ipc_close(p)
            self.lib.hpnn_dev_free(self._ptr)
            self._ptr = None
        self.mc_ptr = 0
