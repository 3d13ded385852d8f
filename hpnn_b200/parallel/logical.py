"""Multi-launch trainer over P *logical* ranks on one GPU.

Two jobs in one class:
  * the stand-alone phase kernels K-A..K-E driven by a host loop (one launch per phase, one
    D2H of the output statistics per iteration -- structurally what the reference does,
    src/cuda_ann.cu:2098-2895, minus its per-iteration allocations), used as the fall-back for
    layers too wide for the persistent kernel's shared memory;
  * the successor of the reference's DEBUG "fake multi-GPU" build (include/libhpnn/common.h:
    511-572, three driver contexts on device 0): the P-rank row partition, the per-rank partial
    ``W^T delta`` buffers and the ordered finalisation are exercised with P logical ranks on a
    single device, so the partition logic is testable without P GPUs.  Ranks run one after the
    other (kernels that wait for each other must never share a GPU).
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np
import torch

from .. import _native as N
from ..kernel_io import Kernel
from ..models.net import split_rows
from ..ops import cuda as ops
from ..ops import reference as R
from ..trainer import SampleResult


class LogicalRanksEngine:
    name = "logical-ranks"

    def __init__(self, kernel: Kernel, nn_type: int, train: int, world: int = 1, device="cuda", lr: Optional[float] = None,
                 alpha: float = 0.2, delta: float = -1.0, min_iter: int = -1, max_iter: int = -1,
                 stable_softmax: bool = False, **_):
        self.dev = torch.device(device)
        self.type, self.train_kind, self.world = nn_type, train, world
        self.dims, self.name_ = list(kernel.dims), kernel.name
        L = len(self.dims) - 1
        self.splits = [[split_rows(self.dims[l + 1], g, world) for g in range(world)] for l in range(L)]
        self.W: List[List[torch.Tensor]] = []
        self.M: List[List[Optional[torch.Tensor]]] = []
        for g in range(world):
            ws, ms = [], []
            for l in range(L):
                r0, nr = self.splits[l][g]
                w = ops.padded_weights(torch.from_numpy(np.ascontiguousarray(kernel.weights[l][r0:r0 + nr])).to(self.dev))
                ws.append(w)
                ms.append(torch.zeros_like(w) if train == N.TRAIN_BPM else None)
            self.W.append(ws)
            self.M.append(ms)
        self.lr = lr if lr and lr > 0 else 0.01
        self.alpha = alpha
        self.delta = delta if delta > 0 else 1e-6
        self.min_iter = min_iter if min_iter >= 0 else (15 if train == N.TRAIN_BPM else 31)
        self.max_iter = max_iter if max_iter >= 0 else 102399
        self.stable = stable_softmax
        self.upd = 2 if train == N.TRAIN_BPM else 1

    # ---- phases ----------------------------------------------------------------------
    def _forward_from(self, acts: List[torch.Tensor], first: int) -> None:
        L = len(self.dims) - 1
        for l in range(first, L):
            kind = 1 if (l == L - 1 and self.type == N.TYPE_SNN) else 0
            xp = ops.padded_vector(acts[l])
            parts = [ops.forward_layer(self.W[g][l], self.dims[l], xp, kind) for g in range(self.world)
                     if self.splits[l][g][1] > 0]
            acts[l + 1] = torch.cat(parts)  # "all-gather": owned slices in rank order

    def train(self, x, t) -> SampleResult:
        L = len(self.dims) - 1
        x = torch.as_tensor(np.asarray(x), dtype=torch.float64, device=self.dev)
        t = torch.as_tensor(np.asarray(t), dtype=torch.float64, device=self.dev)
        for ms in self.M:
            for m in ms:
                if m is not None:
                    m.zero_()
        acts: List[torch.Tensor] = [x] + [None] * L
        self._forward_from(acts, 0)
        o, d_out, E, amax, trg = ops.output_stats(self.type, acts[L], t, self.stable)
        init, it, first_ok, ok, dE = E, 0, False, False, 0.0
        while True:
            it += 1
            delta = d_out
            # top-down: partial W^T delta per logical rank (old weights) + update, then the ordered sum
            for l in range(L - 1, 0, -1):
                hp = ops.padded_vector(acts[l])
                parts = []
                for g in range(self.world):
                    r0, nr = self.splits[l][g]
                    if nr == 0:
                        parts.append(torch.zeros(self.dims[l], dtype=torch.float64, device=self.dev))
                        continue
                    parts.append(ops.backward_update(self.W[g][l], self.dims[l], delta[r0:r0 + nr].contiguous(), hp,
                                                     self.upd, self.lr, self.alpha, self.M[g][l]))
                delta_prev = ops.delta_finalize(torch.stack(parts), acts[l])
                delta = delta_prev
            # layer 0: update fused with its forward
            xp = ops.padded_vector(acts[0])
            outs = []
            for g in range(self.world):
                r0, nr = self.splits[0][g]
                if nr == 0:
                    continue
                outs.append(ops.update_forward(self.W[g][0], self.dims[0], delta[r0:r0 + nr].contiguous(), xp, self.upd,
                                               self.lr, self.alpha, self.M[g][0],
                                               1 if (L == 1 and self.type == N.TYPE_SNN) else 0))
            acts[1] = torch.cat(outs)
            self._forward_from(acts, 1)
            o, d_out, En, amax, trg = ops.output_stats(self.type, acts[L], t, self.stable)
            dE, E = E - En, En
            ok = amax == trg
            if it == 1:
                first_ok = ok
            if it > self.max_iter:
                break
            ok = ok and it > self.min_iter
            if not (dE > self.delta or not ok):
                break
        self._last_out = o
        return SampleResult(init, dE, E, it, first_ok, ok)

    def run(self, x) -> np.ndarray:
        L = len(self.dims) - 1
        x = torch.as_tensor(np.asarray(x), dtype=torch.float64, device=self.dev)
        acts = [x] + [None] * L
        self._forward_from(acts, 0)
        t = torch.zeros(self.dims[-1], dtype=torch.float64, device=self.dev)
        o = ops.output_stats(self.type, acts[L], t, self.stable)[0]
        return o.cpu().numpy()

    def kernel(self) -> Kernel:
        L = len(self.dims) - 1
        ws = []
        for l in range(L):
            rows = [self.W[g][l][:, : self.dims[l]].cpu().numpy() for g in range(self.world)]
            ws.append(np.concatenate(rows, axis=0))
        return Kernel(list(self.dims), ws, self.name_)

    def close(self):
        pass
