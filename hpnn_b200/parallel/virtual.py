"""Virtual ranks: a P-rank row-split job on ONE GPU, all ranks inside one cooperative launch.

Every rank is a real :class:`HpnnNet` (own weight slice, own arena, own sync words) whose
``arena[q]`` pointers reference the other ranks' arenas on the same device.  The native entry
point ``hpnn_dev_train_sample_vranks`` gives each rank ``SMs / P`` CTAs of one cooperative grid, so
the persistent kernels of all ranks are co-resident and can wait for one another -- peer stores,
the cross-rank barrier and the tagged hand-over run exactly as they do across NVLink, only with
local addresses.  This is how the multi-rank protocols are verified on single-GPU test boxes; it
is the successor of the reference's DEBUG build that fakes three GPUs with three driver contexts
on device 0 (include/libhpnn/common.h:511-572, src/libhpnn.c:277-279).  It is not a performance
path: every rank has a fraction of the SMs.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import numpy as np
import torch

from .. import _native as N
from ..kernel_io import Kernel
from ..models.net import HpnnNet
from ..trainer import SampleResult


class VirtualRanksEngine:
    name = "virtual-ranks"

    def __init__(self, kernel: Kernel, nn_type: int, train: int, world: int = 2, device=None, lr: Optional[float] = None,
                 alpha: float = 0.2, delta: float = -1.0, min_iter: int = -1, max_iter: int = -1,
                 stable_softmax: bool = False, **_):
        if not 1 <= world <= N.MAX_WORLD:
            raise ValueError("world must be 1..%d" % N.MAX_WORLD)
        self.lib = N.load()
        self.world, self.type, self.train_kind = world, nn_type, train
        self.nets: List[HpnnNet] = [HpnnNet(kernel, nn_type, device=device, momentum=(train == N.TRAIN_BPM),
                                            virtual_rank=(r, world)) for r in range(world)]
        for a in self.nets:  # every rank sees every arena (same device: plain pointers)
            for q, b in enumerate(self.nets):
                a.net.arena[q] = b._arena_ptr
        self.opts = self.nets[0].make_opts(train, lr, alpha, delta, min_iter, max_iter, stable_softmax)
        self.stable = stable_softmax
        self.device = self.nets[0].device
        self._res = torch.zeros((world, C.sizeof(N.TrainResult) // 8), dtype=torch.float64, device=self.device)

    def _launch(self, mode: int, opts) -> None:
        arr = (N.DevNet * self.world)(*[n.net for n in self.nets])
        res = (C.c_void_p * self.world)(*[self._res[r].data_ptr() for r in range(self.world)])
        stream = torch.cuda.current_stream(self.device).cuda_stream
        N.check(self.lib.hpnn_dev_train_sample_vranks(arr, self.world, C.byref(opts), mode, res, stream),
                "hpnn_dev_train_sample_vranks")

    def _load(self, x, t) -> None:
        xt = torch.as_tensor(np.asarray(x), dtype=torch.float64).to(self.device)
        for n in self.nets:  # every rank loads the same sample
            n.input_buf.copy_(xt)
            if t is None:
                n.target_buf.zero_()
            else:
                n.target_buf.copy_(torch.as_tensor(np.asarray(t), dtype=torch.float64).to(self.device))

    def train(self, x, t) -> SampleResult:
        self._load(x, t)
        self._launch(0, self.opts)
        torch.cuda.synchronize(self.device)
        raw = self._res.cpu().numpy()
        results = [N.TrainResult.from_buffer_copy(raw[r].tobytes()) for r in range(self.world)]
        if any(r.status != 0 for r in results):
            self.last_status = [r.status for r in results]
            for n in self.nets:
                n.reset_sync()
            raise RuntimeError("hpnn_b200: sample abandoned on ranks %s (spin time-out)" %
                               [i for i, r in enumerate(results) if r.status != 0])
        first = results[0]
        for r in results[1:]:  # every rank takes the exit decision redundantly: they must agree
            if (r.n_iter, r.success, r.first_ok) != (first.n_iter, first.success, first.first_ok):
                raise RuntimeError("virtual ranks disagree on the loop exit")
        return SampleResult.from_native(first)

    def run(self, x) -> np.ndarray:
        self._load(x, None)
        o = self.nets[0].make_opts(N.TRAIN_BP, stable_softmax=self.stable)
        self._launch(1, o)
        torch.cuda.synchronize(self.device)
        return self.nets[0].output_buf.cpu().numpy()

    def kernel(self) -> Kernel:
        rows = [n.local_kernel_rows() for n in self.nets]
        ws = [np.concatenate([rows[g][l] for g in range(self.world)], axis=0) for l in range(self.nets[0].n_layers)]
        return Kernel(list(self.nets[0].dims), ws, self.nets[0].name)

    def close(self):
        for n in self.nets:
            n.close()
