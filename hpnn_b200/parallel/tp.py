"""Batched inference on a row-split (tensor-parallel) network.

``fused``  every layer is ONE GEMM kernel per rank whose epilogue stores its slice of the result
           into all ranks' activation matrices over NVLink peer memory (GEMM fused with the
           all-gather of its output), followed by a one-CTA cross-rank barrier kernel;
``nccl``   same GEMM with a local epilogue, then ``dist.all_gather`` (the baseline);
``batch``  no tensor parallelism at all: every rank needs a FULL replica and takes a slice of the
           batch (the right choice for inference when the replica fits: zero communication).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from .. import _native as N
from ..models.net import HpnnNet
from ..ops import cuda as ops
from .symm import SymmetricBuffer


class TPBatchRunner:
    def __init__(self, net: HpnnNet, max_batch: int = 4096):
        if net.world < 2:
            raise ValueError("TPBatchRunner needs a row-split net (world > 1)")
        self.net, self.maxB = net, max_batch
        self.lib = N.load()
        self.wmax = max((d + 7) & ~7 for d in net.dims[1:])
        # two ping-pong activation matrices + the flag words, identical layout on every rank
        self.act_doubles = max_batch * self.wmax
        self.buf = SymmetricBuffer(2 * self.act_doubles + 16, group=net.group, device=net.device)
        self.epoch = 0

    def _ptrs(self, offset_doubles: int):
        arr = (C.c_void_p * self.net.world)(*[p + 8 * offset_doubles for p in self.buf.ptrs])
        return arr

    def forward(self, X: torch.Tensor, mode: str = "fused", stable_softmax: bool = False) -> torch.Tensor:
        net = self.net
        B = X.shape[0]
        assert B <= self.maxB
        NL = net.n_layers
        H = X.to(net.device, torch.float64)
        if H.stride(1) != 1 or H.stride(0) % 2 or H.data_ptr() % 16:
            Hp = torch.zeros((B, (H.shape[1] + 7) & ~7), dtype=torch.float64, device=net.device)
            Hp[:, : H.shape[1]].copy_(H)
            H = Hp
        stream = torch.cuda.current_stream(net.device).cuda_stream
        flags = self._ptrs(2 * self.act_doubles)
        for l in range(NL):
            kind = 1 if (l == NL - 1 and net.type == N.TYPE_SNN) else 0
            rows, r0 = net.net.rows[l], net.net.row0[l]
            n_full = net.dims[l + 1]
            ldy = (n_full + 7) & ~7
            off = (l & 1) * self.act_doubles
            Yfull = self.buf.local[off: off + B * ldy].view(B, ldy)
            if mode == "fused":
                # NVLS: one store reaches every rank.  At 2 ranks two plain stores (one local, one peer) are
                # faster than a trip through the switch (8.3 M vs 7.1 M samples/s at batch 4096), so the
                # multicast epilogue is used from 4 ranks on (HPNN_TP_BATCH_MC=0/1 overrides)
                use_mc = os.environ.get("HPNN_TP_BATCH_MC", "1" if net.world >= 4 else "0") == "1"
                mc = (self.buf.mc_ptr + 8 * off) if (self.buf.mc_ptr and use_mc) else None
                N.check(self.lib.hpnn_dev_forward_batch_f64_tp_mc(
                    net.W_full[l].data_ptr() if rows else None, net.net.ld[l], rows, net.dims[l], H.data_ptr(), H.stride(0), B,
                    self._ptrs(off), mc, net.world, r0, ldy, kind, stream), "hpnn_dev_forward_batch_f64_tp_mc")
                self.epoch += 1
                N.check(self.lib.hpnn_dev_world_barrier(flags, net.rank, net.world, self.epoch, stream),
                        "hpnn_dev_world_barrier")
            else:
                Yl = ops.forward_batch(net.W_full[l], net.dims[l], H, kind) if rows else \
                    torch.empty((B, 0), dtype=torch.float64, device=net.device)
                mx = max((g + 1) * n_full // net.world - g * n_full // net.world for g in range(net.world))
                pad = torch.zeros((B, mx), dtype=torch.float64, device=net.device)
                pad[:, :rows] = Yl
                outs = [torch.empty_like(pad) for _ in range(net.world)]
                dist.all_gather(outs, pad, group=net.group)
                for g in range(net.world):
                    g0 = g * n_full // net.world
                    gr = (g + 1) * n_full // net.world - g0
                    Yfull[:, g0:g0 + gr] = outs[g][:, :gr]
            H = Yfull[:, :n_full]
        out = H.clone()
        if net.type == N.TYPE_SNN:
            ops.softmax_rows_(out, stable_softmax)
        # Nobody may start the next forward before everyone has read its result: a later "fused" call
        # stores straight into the peers' buffer 0.  This holds after an "nccl" pass as well (its
        # last copies into the shared buffers run after the collective has returned on the peers;
        # seen as a 3e-3 error in one of five 4-GPU runs before the barrier covered both modes).
        self.epoch += 1
        N.check(self.lib.hpnn_dev_world_barrier(flags, net.rank, net.world, self.epoch, stream), "hpnn_dev_world_barrier")
        return out

    def close(self):
        self.buf.close()


def forward_batch_parallel(net_full: HpnnNet, X: torch.Tensor, group=None, gather: bool = True) -> torch.Tensor:
    """Batch-parallel inference: ``net_full`` is a single-rank (full replica) net on every rank."""
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    B = X.shape[0]
    b0, b1 = rank * B // world, (rank + 1) * B // world
    Y = net_full.forward_batch(X[b0:b1])
    if not gather or world == 1:
        return Y
    mx = max((g + 1) * B // world - g * B // world for g in range(world))
    pad = torch.zeros((mx, Y.shape[1]), dtype=torch.float64, device=Y.device)
    pad[: Y.shape[0]] = Y
    outs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(outs, pad, group=group)
    return torch.cat([outs[g][: (g + 1) * B // world - g * B // world] for g in range(world)])
