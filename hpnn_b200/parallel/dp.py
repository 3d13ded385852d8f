"""Sample-parallel (data-parallel) gradient merge -- an EXTENSION, not reference semantics.

BASELINE.json lists an "MPI sample-parallel gradient merge"; the reference itself never splits
samples (every MPI rank trains the same sample; its serial loop over samples is
src/libhpnn.c:1221-1288).  Here every rank holds a full replica and takes ONE BP step's worth of
update ``lr * delta (x) h`` for its own sample from the shared weights; the per-rank updates are
merged so that all replicas stay bit-identical without a broadcast.  This changes training semantics
(a synchronous mini-batch of P samples per step instead of per-sample convergence) and is never used
for parity.

``merge`` selects how the updates meet:

``factored``  (default) the device path: forward, output statistics and deltas by this repository's
              kernels (K-A, K-B, K-C), then the FACTORS ``lr*delta_l`` and ``h_(l-1)`` -- O(N+M) per
              layer instead of O(N*M) -- are published into every rank's factor buffer from inside a
              kernel (one ``multimem.st`` per element through the NVLS multicast mapping, peer stores
              otherwise; ``fence.sys`` + a step flag per rank), and one kernel per layer waits for the
              flags ON THE DEVICE and adds ``sum_r u_r (x) v_r`` to its replica in rank order
              (csrc/kernels/dp_merge.cu).  No host barrier, no matrix crosses NVLink.
``nvls``      matrix form, NVLS all-reduce: rank r reduces slice r of all gradient matrices with
              ``multimem.ld_reduce`` and writes the new weights into every replica with ``multimem.st``.
``fused``     matrix form, round-1 kernel: every rank loads all peers' gradient matrices (NVLink peer loads).
``nccl``      matrix form, ``all_reduce`` (the baseline).
The matrix forms bracket the merge with host barriers; they exist for comparison
(``bench/dp_merge.py``)."""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

import numpy as np
import torch
import torch.distributed as dist

from .. import _native as N
from ..kernel_io import Kernel
from ..ops import cuda as ops
from .symm import SymmetricBuffer


def _pad8(v: int) -> int:
    return (v + 7) & ~7


class DataParallelTrainer:
    def __init__(self, kernel: Kernel, nn_type: int, lr: float = 0.01, group=None, device=None, merge: str = "factored"):
        self.lib = N.load()
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        torch.cuda.set_device(self.dev)
        self.type, self.lr, self.merge = nn_type, lr, merge
        self.dims, self.name_ = list(kernel.dims), kernel.name
        L = len(self.dims) - 1
        self.ld = [_pad8(self.dims[l]) for l in range(L)]
        sizes = [self.dims[l + 1] * self.ld[l] for l in range(L)]
        self.offsets = np.concatenate([[0], np.cumsum(sizes)]).astype(int)
        total = int(self.offsets[-1])
        # one flat parameter vector of padded layers [N, ld]; symmetric (multicast-mapped) for "nvls"
        self._wbuf = SymmetricBuffer(total, group=group, device=self.dev) if merge == "nvls" else None
        self.flat = self._wbuf.local if self._wbuf is not None else torch.zeros(total, dtype=torch.float64, device=self.dev)
        self.flat.zero_()
        self.Wp: List[torch.Tensor] = []
        for l, w in enumerate(kernel.weights):
            v = self.flat[self.offsets[l]: self.offsets[l + 1]].view(self.dims[l + 1], self.ld[l])
            v[:, : self.dims[l]].copy_(torch.from_numpy(np.ascontiguousarray(w)))
            self.Wp.append(v)
        self.step_no = 0
        self._status = torch.zeros(1, dtype=torch.int32, device=self.dev)
        self.timeout_ms = int(os.environ.get("HPNN_SPIN_TIMEOUT_MS", "10000"))
        if merge == "factored":
            # factor buffer: 2 slots (step parity) x world ranks x S doubles, then 2 x world step flags
            self.u_off, self.v_off, off = [], [], 0
            for l in range(L):
                self.u_off.append(off)
                off += _pad8(self.dims[l + 1])
                self.v_off.append(off)
                off += self.ld[l]
            self.S = off
            self.flag_off = 2 * self.world * self.S
            self.fbuf = SymmetricBuffer(self.flag_off + 2 * N.MAX_WORLD, group=group, device=self.dev)
            self.mine = torch.zeros(self.S, dtype=torch.float64, device=self.dev)
            self.grads = None
        else:
            self.grads = SymmetricBuffer(total, group=group, device=self.dev)
            self.G = [self.grads.local[self.offsets[l]: self.offsets[l + 1]].view(self.dims[l + 1], self.ld[l])
                      for l in range(L)]

    @property
    def W(self) -> List[torch.Tensor]:
        """Unpadded views of the layers."""
        return [w[:, : self.dims[l]] for l, w in enumerate(self.Wp)]

    # ------------------------------------------------------------------ per-rank step by our kernels
    def _stream(self) -> int:
        return torch.cuda.current_stream(self.dev).cuda_stream

    def _local_factors(self, x: torch.Tensor, t: torch.Tensor):
        """Forward, output statistics and deltas of THIS rank's sample (K-A, K-B, K-C + finalize);
        returns (acts, deltas, stats) -- all device tensors, nothing is read back."""
        lib, L, s = self.lib, len(self.Wp), self._stream()
        if not hasattr(self, "_acts"):
            self._acts = [torch.zeros(self.ld[l], dtype=torch.float64, device=self.dev) for l in range(L)]
            self._acts.append(torch.zeros(_pad8(self.dims[L]), dtype=torch.float64, device=self.dev))
            self._deltas = [torch.zeros(_pad8(self.dims[l + 1]), dtype=torch.float64, device=self.dev) for l in range(L)]
            self._part = [torch.zeros(self.ld[l], dtype=torch.float64, device=self.dev) for l in range(L)]
            self._o = torch.zeros(_pad8(self.dims[L]), dtype=torch.float64, device=self.dev)
            self._stats = torch.zeros(3, dtype=torch.float64, device=self.dev)
        a, d = self._acts, self._deltas
        a[0][: self.dims[0]].copy_(x, non_blocking=True)
        for l in range(L):
            kind = 1 if (l == L - 1 and self.type == N.TYPE_SNN) else 0
            N.check(lib.hpnn_dev_forward_layer(self.Wp[l].data_ptr(), self.ld[l], self.dims[l + 1], self.dims[l],
                                               a[l].data_ptr(), a[l + 1].data_ptr(), kind, s), "hpnn_dev_forward_layer")
        N.check(lib.hpnn_dev_output_stats(self.type, 0, a[L].data_ptr(), t.data_ptr(), self.dims[L], self._o.data_ptr(),
                                          d[L - 1].data_ptr(), self._stats.data_ptr(), s), "hpnn_dev_output_stats")
        for l in range(L - 1, 0, -1):
            N.check(lib.hpnn_dev_backward_update(self.Wp[l].data_ptr(), None, self.ld[l], self.dims[l + 1], self.dims[l],
                                                 d[l].data_ptr(), None, self._part[l].data_ptr(), 0, 0.0, 0.0, s),
                    "hpnn_dev_backward_update")
            N.check(lib.hpnn_dev_delta_finalize(self._part[l].data_ptr(), 1, self.ld[l], a[l].data_ptr(), self.dims[l],
                                                d[l - 1].data_ptr(), s), "hpnn_dev_delta_finalize")
        return a, d, self._stats

    def step(self, x, t, want_error: bool = True):
        """One merged BP step; ``x``/``t`` are THIS rank's sample.  Returns the local error before the
        step (``want_error=False``: returns None and never synchronises with the host)."""
        x = torch.as_tensor(np.asarray(x), dtype=torch.float64).to(self.dev, non_blocking=True)
        t = torch.as_tensor(np.asarray(t), dtype=torch.float64).to(self.dev, non_blocking=True)
        acts, deltas, stats = self._local_factors(x, t)
        L = len(self.Wp)
        self.step_no += 1
        if self.merge == "factored":
            self._merge_factored(acts, deltas)
        else:
            for l in range(L):
                torch.outer(self.lr * deltas[l][: self.dims[l + 1]], acts[l], out=self.G[l])
            self._merge_matrices()
        return float(stats[0]) if want_error else None

    def _merge_factored(self, acts, deltas):
        lib, L, s, W = self.lib, len(self.Wp), self._stream(), self.world
        slot = self.step_no & 1
        for l in range(L):
            self.mine[self.u_off[l]: self.u_off[l] + self.dims[l + 1]] = self.lr * deltas[l][: self.dims[l + 1]]
            self.mine[self.v_off[l]: self.v_off[l] + self.ld[l]] = acts[l]
        base = (slot * W + self.rank) * self.S
        dst = (C.c_void_p * W)(*[p + 8 * base for p in self.fbuf.ptrs])
        foff = self.flag_off + slot * N.MAX_WORLD + self.rank
        flg = (C.c_void_p * W)(*[p + 8 * foff for p in self.fbuf.ptrs])
        mc = self.fbuf.mc_ptr
        N.check(lib.hpnn_dev_dp_publish(self.mine.data_ptr(), self.S, dst, (mc + 8 * base) if mc else None, flg,
                                        (mc + 8 * foff) if mc else None, W, self.step_no, s), "hpnn_dev_dp_publish")
        local = self.fbuf.ptrs[self.rank]
        flags = local + 8 * (self.flag_off + slot * N.MAX_WORLD)
        for l in range(L):
            u = local + 8 * (slot * W * self.S + self.u_off[l])
            v = local + 8 * (slot * W * self.S + self.v_off[l])
            N.check(lib.hpnn_dev_dp_apply(self.Wp[l].data_ptr(), self.ld[l], self.dims[l + 1], self.dims[l], u, self.S, v,
                                          self.S, W, flags, self.step_no, self.timeout_ms, self._status.data_ptr(), s),
                    "hpnn_dev_dp_apply")

    def check(self):
        """Raises if a merge kernel gave up waiting for a rank (synchronises)."""
        if int(self._status.item()) != 0:
            raise RuntimeError("hpnn_b200: data-parallel merge abandoned, a rank did not publish its update in time")

    def _merge_matrices(self):
        torch.cuda.synchronize(self.dev)
        if self.world > 1:
            dist.barrier(group=self.group)  # every rank's update buffer is complete
        if self.merge == "fused":
            ops.grad_merge_(self.flat, self.grads.ptrs, 1.0)  # peer loads over NVLink inside the kernel
        elif self.merge == "nvls":
            if not (self.grads.mc_ptr and self._wbuf.mc_ptr):
                raise RuntimeError("merge='nvls' needs NVLS multicast mappings (SymmetricBuffer backend vmm)")
            N.check(self.lib.hpnn_dev_grad_merge_nvls(self.flat.data_ptr(), self._wbuf.mc_ptr, self.grads.mc_ptr,
                                                      self.flat.numel(), self.rank, self.world, 1.0, self._stream()),
                    "hpnn_dev_grad_merge_nvls")
        else:
            g = self.grads.local.clone()
            if self.world > 1:
                dist.all_reduce(g, group=self.group)
            self.flat += g
        torch.cuda.synchronize(self.dev)
        if self.world > 1:
            dist.barrier(group=self.group)  # nobody overwrites its buffer while a peer still reads it

    def kernel(self) -> Kernel:
        return Kernel(list(self.dims), [w.cpu().numpy().copy() for w in self.W], self.name_)

    def close(self):
        torch.cuda.synchronize(self.dev)
        if getattr(self, "fbuf", None) is not None:
            self.fbuf.close()
        if self.grads is not None:
            self.grads.close()
        if self._wbuf is not None:
            self.flat = None
            self._wbuf.close()
