"""Sample-parallel (data-parallel) gradient merge -- an EXTENSION, not reference semantics.

BASELINE.json lists an "MPI sample-parallel gradient merge"; the reference itself never splits
samples (every MPI rank trains the same sample).  Here every rank holds a full replica,
takes ONE BP step's worth of update ``lr * delta (x) h`` for its own sample from the shared weights,
and the per-rank updates are merged with kernel K-M: each rank reads all peers' update buffers over
NVLink (peer-mapped memory) and applies ``W += sum_r G_r`` with the ranks summed in index order, so
all replicas stay bit-identical without a broadcast.  ``merge="nccl"`` does the same with
``all_reduce`` (the baseline).  This changes training semantics (a synchronous mini-batch of P
samples per step instead of per-sample convergence) and is never used for parity."""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist

from ..kernel_io import Kernel
from ..ops import reference as R
from .symm import SymmetricBuffer


class DataParallelTrainer:
    def __init__(self, kernel: Kernel, nn_type: int, lr: float = 0.01, group=None, device=None, merge: str = "fused"):
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.type, self.lr, self.merge = nn_type, lr, merge
        self.dims, self.name_ = list(kernel.dims), kernel.name
        sizes = [w.size for w in kernel.weights]
        self.offsets = np.concatenate([[0], np.cumsum([(s + 1) & ~1 for s in sizes])]).astype(int)
        total = int(self.offsets[-1])
        # one flat parameter vector (views per layer) and one symmetric update buffer
        self.flat = torch.zeros(total, dtype=torch.float64, device=self.dev)
        self.W = []
        for l, w in enumerate(kernel.weights):
            v = self.flat[self.offsets[l]: self.offsets[l] + w.size].view(w.shape)
            v.copy_(torch.from_numpy(w))
            self.W.append(v)
        self.grads = SymmetricBuffer(total, group=group, device=self.dev)
        self.G = [self.grads.local[self.offsets[l]: self.offsets[l] + w.size].view(w.shape)
                  for l, w in enumerate(kernel.weights)]

    def step(self, x, t) -> float:
        """One merged BP step; ``x``/``t`` are THIS rank's sample.  Returns the local error before it."""
        from ..ops import cuda as ops

        x = torch.as_tensor(np.asarray(x), dtype=torch.float64, device=self.dev)
        t = torch.as_tensor(np.asarray(t), dtype=torch.float64, device=self.dev)
        acts = R.forward(self.W, x, self.type)
        E = float(R.error(acts[-1], t, self.type))
        d = R.deltas(self.W, acts, t, self.type)
        for l in range(len(self.W)):
            torch.outer(self.lr * d[l], acts[l], out=self.G[l])
        torch.cuda.synchronize(self.dev)
        if self.world > 1:
            dist.barrier(group=self.group)  # every rank's update buffer is complete
        if self.merge == "fused":
            ops.grad_merge_(self.flat, self.grads.ptrs, 1.0)  # peer loads over NVLink inside the kernel
        else:
            g = self.grads.local.clone()
            if self.world > 1:
                dist.all_reduce(g, group=self.group)
            self.flat += g
        torch.cuda.synchronize(self.dev)
        if self.world > 1:
            dist.barrier(group=self.group)  # nobody overwrites its buffer while a peer still reads it
        return E

    def kernel(self) -> Kernel:
        return Kernel(list(self.dims), [w.cpu().numpy().copy() for w in self.W], self.name_)

    def close(self):
        self.grads.close()
