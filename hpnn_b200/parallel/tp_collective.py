"""Row-split (tensor-parallel) training with ``torch.distributed`` collectives between layers.

This is the *internal baseline* of the fused path: the same
partition as :class:`hpnn_b200.models.net.HpnnNet` -- rank ``g`` owns rows
``[g*N/P, (g+1)*N/P)`` of every layer -- but every all-gather of activations and every
reduction of ``W^T delta`` partials is a separate NCCL (GPUs) or gloo (CPU) call, i.e. what the
reference does with ``MPI_Allgather`` (src/ann.c:925,1326) minus its per-iteration broadcast of
the updated weight rows (src/ann.c:1638), which a true partition does not need.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np
import torch
import torch.distributed as dist

from ..kernel_io import Kernel
from ..models.net import split_rows
from ..ops import reference as R
from ..trainer import SampleResult


class CollectiveTPEngine:
    name = "tp-collective"

    def __init__(self, kernel: Kernel, nn_type: int, train: int, group=None, device="cpu", lr: Optional[float] = None,
                 alpha: float = 0.2, delta: float = -1.0, min_iter: int = -1, max_iter: int = -1,
                 stable_softmax: bool = False, cuda_defaults: bool = True, **_):
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.dev = torch.device(device)
        self.type, self.train_kind = nn_type, train
        self.dims, self.name_ = list(kernel.dims), kernel.name
        self.splits = [[split_rows(n, g, self.world) for g in range(self.world)] for n in self.dims[1:]]
        self.W = []
        for l, w in enumerate(kernel.weights):
            r0, nr = self.splits[l][self.rank]
            self.W.append(torch.from_numpy(np.ascontiguousarray(w[r0:r0 + nr])).to(self.dev))
        self.mom = [torch.zeros_like(w) for w in self.W] if train == R.TRAIN_BPM else None
        self.lr = lr if lr and lr > 0 else R.default_lr(nn_type, train, cuda_defaults)
        self.alpha = alpha
        self.delta = delta if delta > 0 else 1e-6
        self.min_iter = min_iter if min_iter >= 0 else (15 if train == R.TRAIN_BPM else 31)
        self.max_iter = max_iter if max_iter >= 0 else 102399
        self.stable = stable_softmax
        self.n_collectives = 0

    # -- collectives ------------------------------------------------------------------
    def _all_gather_rows(self, local: torch.Tensor, l: int) -> torch.Tensor:
        """All-gather of the owned slice of a layer output (uneven slices are padded)."""
        mx = max(nr for _, nr in self.splits[l])
        buf = torch.zeros(mx, dtype=torch.float64, device=self.dev)
        buf[: local.numel()] = local
        outs = [torch.empty_like(buf) for _ in range(self.world)]
        dist.all_gather(outs, buf, group=self.group)
        self.n_collectives += 1
        return torch.cat([outs[g][: self.splits[l][g][1]] for g in range(self.world)])

    def _reduce_to_owner(self, part: torch.Tensor, l: int) -> torch.Tensor:
        """Sum the full-width partials over ranks and keep the slice this rank owns of layer ``l``."""
        dist.all_reduce(part, op=dist.ReduceOp.SUM, group=self.group)
        self.n_collectives += 1
        r0, nr = self.splits[l][self.rank]
        return part[r0:r0 + nr]

    # -- one sample ---------------------------------------------------------------------
    def _forward(self, x: torch.Tensor) -> List[torch.Tensor]:
        acts = [x]
        L = len(self.W)
        for l, W in enumerate(self.W):
            z = W @ acts[-1]
            raw = (l == L - 1 and self.type == R.TYPE_SNN)
            full = self._all_gather_rows(z if raw else R.act(z), l)
            acts.append(R.softmax_ref(full, self.stable) if raw else full)
        return acts

    def train(self, x, t) -> SampleResult:
        x = torch.as_tensor(np.asarray(x), dtype=torch.float64, device=self.dev)
        t = torch.as_tensor(np.asarray(t), dtype=torch.float64, device=self.dev)
        L = len(self.W)
        if self.mom is not None:
            for m in self.mom:
                m.zero_()
        acts = self._forward(x)
        E = float(R.error(acts[-1], t, self.type))
        init, first_ok, it, ok, dE = E, False, 0, False, 0.0
        p_trg = R.target_index(t)
        while True:
            it += 1
            r0, nr = self.splits[L - 1][self.rank]
            d = [None] * L
            d[L - 1] = R.output_delta(acts[L], t, self.type)[r0:r0 + nr]
            for l in range(L - 1, 0, -1):
                part = self.W[l].T @ d[l]                       # old weights, partial over owned rows
                own = self._reduce_to_owner(part, l - 1)
                p0, pn = self.splits[l - 1][self.rank]
                d[l - 1] = own * R.dact(acts[l][p0:p0 + pn])
            R.apply_update(self.W, self.mom, d, acts, self.lr, self.alpha)
            acts = self._forward(x)
            En = float(R.error(acts[-1], t, self.type))
            dE, E = E - En, En
            ok = R.argmax_first(acts[-1]) == p_trg
            if it == 1:
                first_ok = ok
            if it > self.max_iter:
                break
            ok = ok and it > self.min_iter
            if not (dE > self.delta or not ok):
                break
        return SampleResult(init, dE, E, it, first_ok, ok)

    def run(self, x) -> np.ndarray:
        x = torch.as_tensor(np.asarray(x), dtype=torch.float64, device=self.dev)
        return self._forward(x)[-1].cpu().numpy()

    def kernel(self) -> Kernel:
        rows = [w.cpu().numpy() for w in self.W]
        everyone = [None] * self.world
        dist.all_gather_object(everyone, rows, group=self.group)
        ws = [np.concatenate([everyone[g][l] for g in range(self.world)], axis=0) for l in range(len(self.W))]
        return Kernel(list(self.dims), ws, self.name_)

    def close(self):
        pass
