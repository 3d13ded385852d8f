"""Symmetric device buffers: the same allocation on every rank, mapped into every process.

``cudaMalloc`` (through the native core, so the memory is IPC-exportable) + ``cudaIpc*`` handle
exchange over ``torch.distributed``.  ``ptrs[q]`` is rank q's buffer as seen from this process:
kernels can load from / store to any of them and the traffic goes over NVLink."""
from __future__ import annotations

import ctypes as C
from typing import List

import torch
import torch.distributed as dist

from .. import _native as N
from ..models.net import _DevBuf


class SymmetricBuffer:
    def __init__(self, n_doubles: int, group=None, device=None):
        self.lib = N.load()
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        torch.cuda.set_device(self.device)
        self.n = n_doubles
        self._ptr = self.lib.hpnn_dev_alloc(n_doubles * 8)
        if not self._ptr:
            raise MemoryError(N.last_error())
        self.local = torch.as_tensor(_DevBuf(self._ptr, n_doubles), device=self.device)
        self.ptrs: List[int] = [0] * self.world
        self.ptrs[self.rank] = self._ptr
        self._opened: List[int] = []
        if self.world > 1:
            h = C.create_string_buffer(64)
            N.check(self.lib.hpnn_ipc_export(self._ptr, h), "hpnn_ipc_export")
            everyone = [None] * self.world
            dist.all_gather_object(everyone, (self.rank, bytes(h.raw)), group=group)
            for r, raw in everyone:
                if r != self.rank:
                    p = self.lib.hpnn_ipc_open(raw)
                    if not p:
                        raise RuntimeError("hpnn_ipc_open: " + N.last_error())
                    self.ptrs[r] = p
                    self._opened.append(p)
            dist.barrier(group=group)

    def close(self):
        if self._ptr is None:
            return
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier(group=self.group)
        for p in self._opened:
            self.lib.hpnn_ipc_close(p)
        self.local = None
        self.lib.hpnn_dev_free(self._ptr)
        self._ptr = None
