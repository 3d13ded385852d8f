"""``HpnnNet``: a feed-forward ANN / SNN resident on one B200, or row-split over the
GPUs of one NVLink domain (one process per GPU).

Design, on purpose different from the reference:
  * weights are *truly partitioned*: rank ``g`` holds rows ``[g*N/P, (g+1)*N/P)`` of every
    layer in its own HBM (the reference keeps everything on GPU 0 in P2P mode and a full
    replica per GPU in EXP mode, src/cuda_ann.cu:197-254);
  * every rank owns a small symmetric "arena" (activations, partial deltas, barrier
    flags) that all peers map through CUDA IPC; the persistent kernel all-gathers
    activations and reduce-scatters deltas by storing straight into the peers' arenas;
  * the whole per-sample convergence loop is one cooperative kernel launch.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np
import torch

from .. import _native as N
from ..kernel_io import Kernel


class _DevBuf:
    """Expose raw device memory to torch through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, n_doubles: int):
        self.__cuda_array_interface__ = {"shape": (n_doubles,), "typestr": "<f8", "data": (ptr, False),
                                         "version": 2, "strides": None}


def pad8(v: int) -> int:
    return (v + 7) & ~7


def split_rows(n: int, rank: int, world: int):
    """Balanced row split: rank owns ``[rank*n//world, (rank+1)*n//world)``."""
    r0 = rank * n // world
    return r0, (rank + 1) * n // world - r0


class HpnnNet(torch.nn.Module):
    """``torch.nn.Module`` whose parameters are this rank's padded weight rows (``requires_grad=False``:
    training is the library's own BP / BPM, not autograd) and whose buffers are the momenta, so
    ``parameters()``, ``state_dict()`` / ``load_state_dict()`` and ``forward(X)`` work as usual.  The
    kernels hold raw pointers into these tensors: ``.to()`` / ``.double()`` that keep device and dtype
    are fine (pointers are refreshed), anything else is refused.  (``net.type`` is the network type
    ANN / SNN as in the reference's ``nn_def``; it shadows ``Module.type()``.)"""

    def __init__(self, kernel: Kernel, nn_type: int, device: Optional[torch.device] = None,
                 momentum: bool = True, group=None, virtual_rank: Optional[tuple] = None):
        super().__init__()
        if not torch.cuda.is_available():
            raise RuntimeError("HpnnNet needs a CUDA device (use the 'cpu' or 'torch' engine otherwise)")
        self.lib = N.load()
        self.type = int(nn_type)
        self.dims = list(kernel.dims)
        self.name = kernel.name
        self.group = group
        self.virtual = virtual_rank is not None
        if self.virtual:
            # (rank, world) of a row-split job whose ranks all live on THIS device
            # (parallel/virtual.py links the arenas and launches every rank in one cooperative grid)
            self.rank, self.world = int(virtual_rank[0]), int(virtual_rank[1])
        elif group is not None:
            import torch.distributed as dist
            self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        else:
            self.rank, self.world = 0, 1
        if self.world > N.MAX_WORLD:
            raise ValueError(f"at most {N.MAX_WORLD} ranks per NVLink domain")
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        torch.cuda.set_device(self.device)
        NL = len(self.dims) - 1
        if NL < 2 or NL > N.MAX_LAYERS:
            raise ValueError("need 1..%d hidden layers" % (N.MAX_LAYERS - 1))
        net = N.DevNet()
        net.n_layers = NL
        for i, d in enumerate(self.dims):
            net.dims[i] = d
        net.rank, net.world = self.rank, self.world
        n_doubles = self.lib.hpnn_dev_arena_layout(C.byref(net))
        if n_doubles <= 0:
            raise ValueError("invalid network layout")
        self.net = net
        # ---- weights (+ momentum): private, padded leading dimension, owned rows only
        self._W = torch.nn.ParameterList()
        self._has_momentum = bool(momentum)
        for l in range(NL):
            Wl = torch.zeros((net.rows[l], net.ld[l]), dtype=torch.float64, device=self.device)
            r0 = net.row0[l]
            src = torch.from_numpy(np.ascontiguousarray(kernel.weights[l][r0:r0 + net.rows[l]]))
            Wl[:, : self.dims[l]].copy_(src)
            self._W.append(torch.nn.Parameter(Wl, requires_grad=False))
            net.W[l] = Wl.data_ptr() if Wl.numel() else None
            if momentum:
                Ml = torch.zeros_like(Wl)
                self.register_buffer(f"momentum_{l}", Ml)
                net.dW[l] = Ml.data_ptr() if Ml.numel() else None
            else:
                net.dW[l] = None
        # ---- symmetric arena + private sync words.  Several ranks: the arena is a SymmetricBuffer
        # (VMM allocation mapped into every process, NVLS multicast address when the fabric has one;
        # cudaMalloc + CUDA IPC as fall-back); one rank / virtual ranks: plain device memory.
        self._symm = None
        self._sync_ptr = self.lib.hpnn_dev_alloc(256 * 4)
        if not self._sync_ptr:
            raise MemoryError(N.last_error())
        for q in range(N.MAX_WORLD):
            net.arena[q] = None
        net.arena_mc = None
        if self.world > 1 and not self.virtual:
            from ..parallel.symm import SymmetricBuffer

            self._symm = SymmetricBuffer(n_doubles, group=group, device=self.device)
            self._arena_ptr = self._symm.ptrs[self.rank]
            for q in range(self.world):
                net.arena[q] = self._symm.ptrs[q]
            net.arena_mc = self._symm.mc_ptr or None
            self.arena = self._symm.local
        else:
            self._arena_ptr = self.lib.hpnn_dev_alloc(n_doubles * 8)
            if not self._arena_ptr:
                raise MemoryError(N.last_error())
            self.arena = torch.as_tensor(_DevBuf(self._arena_ptr, n_doubles), device=self.device)
            net.arena[self.rank] = self._arena_ptr
        net.local_sync = self._sync_ptr
        self._peer_ptrs: List[int] = []
        # ---- per-sample staging: pinned result ring + device result slots
        self._n_slots = 1024
        self._res_dev = torch.zeros(self._n_slots * C.sizeof(N.TrainResult) // 8, dtype=torch.float64,
                                    device=self.device)
        self._res_host = torch.zeros_like(self._res_dev, device="cpu").pin_memory()
        self._closed = False

    # ------------------------------------------------------------------ set-up
    def close(self):
        if getattr(self, "_closed", True):
            return
        self._closed = True
        torch.cuda.synchronize(self.device)
        self.arena = None
        if self._symm is not None:
            self._symm.close()  # (synchronises with the peers: nobody unmaps while another still runs)
            self._symm = None
        else:
            if self.world > 1 and not self.virtual:
                import torch.distributed as dist
                dist.barrier(group=self.group)
            self.lib.hpnn_dev_free(self._arena_ptr)
        self.lib.hpnn_dev_free(self._sync_ptr)

    def __del__(self):
        try:
            if self.world == 1 or self.virtual:
                self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ nn.Module plumbing
    @property
    def W_full(self) -> List[torch.Tensor]:
        """Owned rows of every layer in their padded storage ``[rows_l, ld_l]``."""
        return [p.data for p in self._W]

    @property
    def dW_full(self) -> List[Optional[torch.Tensor]]:
        return [getattr(self, f"momentum_{l}") if self._has_momentum else None for l in range(self.n_layers)]

    def _apply(self, fn, *args, **kwargs):
        out = super()._apply(fn, *args, **kwargs)
        for l, W in enumerate(self.W_full):  # the kernels address these tensors by pointer
            if W.device != self.device or W.dtype != torch.float64:
                raise RuntimeError("HpnnNet lives on %s in float64; build a new net for another device / dtype" % self.device)
            self.net.W[l] = W.data_ptr() if W.numel() else None
            M = self.dW_full[l]
            self.net.dW[l] = M.data_ptr() if (M is not None and M.numel()) else None
        self._w_bf16 = None
        return out

    def forward(self, X: torch.Tensor) -> torch.Tensor:
        """``[B, n_in] -> [B, n_out]`` (batched FP64 path) or ``[n_in] -> [n_out]`` (persistent kernel)."""
        if X.dim() == 2:
            return self.forward_batch(X)
        return self.run(X).clone()

    # ------------------------------------------------------------------ views
    @property
    def n_layers(self) -> int:
        return len(self.dims) - 1

    @property
    def weights(self) -> List[torch.Tensor]:
        """Owned rows of every layer, ``[rows_l, dims[l]]`` views into the padded storage."""
        return [W[:, : self.dims[l]] for l, W in enumerate(self.W_full)]

    def _slice(self, off: int, n: int) -> torch.Tensor:
        return self.arena[off: off + n]

    @property
    def input_buf(self) -> torch.Tensor:
        return self._slice(self.net.off_act[0], self.dims[0])

    @property
    def target_buf(self) -> torch.Tensor:
        return self._slice(self.net.off_target, self.dims[-1])

    @property
    def output_buf(self) -> torch.Tensor:
        return self._slice(self.net.off_out, self.dims[-1])

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    # ------------------------------------------------------------------ per-sample ops
    def make_opts(self, train: int, lr: Optional[float] = None, alpha: float = 0.2, delta: float = -1.0,
                  min_iter: int = -1, max_iter: int = -1, stable_softmax: bool = False) -> N.TrainOpts:
        o = N.TrainOpts()
        self.lib.hpnn_train_opts_default(C.byref(o), self.type, int(train), 1)
        if lr is not None and lr > 0:
            o.lr = lr
        o.alpha = alpha
        if delta > 0:
            o.delta = delta
        if min_iter >= 0:
            o.min_iter = min_iter
        if max_iter >= 0:
            o.max_iter = max_iter
        o.stable_softmax = int(stable_softmax)
        return o

    def load_sample(self, x: torch.Tensor, t: Optional[torch.Tensor] = None) -> None:
        """Stream-ordered copy of one sample into the arena (H2D when ``x`` is pinned host memory)."""
        self.input_buf.copy_(x, non_blocking=True)
        if t is not None:
            self.target_buf.copy_(t, non_blocking=True)

    def train_sample_async(self, opts: N.TrainOpts, slot: int = 0) -> None:
        """Enqueue the persistent trainer for the sample currently in the arena."""
        self._w_bf16 = None  # the weights are about to change
        res_ptr = self._res_dev.data_ptr() + (slot % self._n_slots) * C.sizeof(N.TrainResult)
        N.check(self.lib.hpnn_dev_train_sample(C.byref(self.net), C.byref(opts), 0, res_ptr, self._stream()),
                "hpnn_dev_train_sample")

    def fetch_results(self, n: int = 1) -> List[N.TrainResult]:
        """D2H of the first ``n`` result slots (synchronises the stream)."""
        n = min(n, self._n_slots)
        nd = n * C.sizeof(N.TrainResult) // 8
        self._res_host[:nd].copy_(self._res_dev[:nd], non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        buf = self._res_host.numpy()
        arr = list((N.TrainResult * n).from_buffer_copy(buf[:nd].tobytes()))
        if any(r.status != 0 for r in arr):
            # a wait inside the persistent kernel ran out of time (a rank that never launched, a dead
            # peer): every rank left its kernel; weights and sync words of this net are undefined now
            self.reset_sync()
            raise RuntimeError("hpnn_b200: sample abandoned, a wait for another CTA or GPU exceeded the spin "
                               "time-out (hpnn_dev_set_spin_timeout_ms / HPNN_SPIN_TIMEOUT_MS)")
        return arr

    def reset_sync(self) -> None:
        """Barrier counters, flag words, abort word and launch id back to the state of a fresh net
        (after an abandoned sample; every rank of the job must do this before the next launch).
        The whole arena is cleared: the launch id restarts from zero, so no {value, tag} pair of an
        earlier launch may survive (its tag could match a future one)."""
        torch.cuda.synchronize(self.device)
        self.lib.hpnn_dev_memset(self._sync_ptr, 0, 256 * 4, self._stream())
        self.arena.zero_()
        torch.cuda.synchronize(self.device)

    def train_sample(self, x: torch.Tensor, t: torch.Tensor, opts: N.TrainOpts) -> N.TrainResult:
        self.load_sample(x, t)
        self.train_sample_async(opts, 0)
        return self.fetch_results(1)[0]

    def run(self, x: torch.Tensor, t: Optional[torch.Tensor] = None, stable_softmax: bool = False) -> torch.Tensor:
        """Forward one sample (``nn_run_kernel`` inner step); returns the output vector (device)."""
        if t is None:
            self.target_buf.zero_()
        self.load_sample(x, t)
        o = self.make_opts(N.TRAIN_BP, stable_softmax=stable_softmax)
        N.check(self.lib.hpnn_dev_train_sample(C.byref(self.net), C.byref(o), 1, self._res_dev.data_ptr(),
                                               self._stream()), "hpnn_dev_train_sample(run)")
        return self.output_buf

    # ------------------------------------------------------------------ batched inference
    def forward_batch(self, X: torch.Tensor, stable_softmax: bool = False, precision: str = "fp64") -> torch.Tensor:
        """``[B, n_in] -> [B, n_out]`` (extension).  Single rank only; a row-split
        job gathers each layer with one collective per layer in
        :func:`hpnn_b200.parallel.tp.forward_batch_tp`."""
        from ..ops import cuda as ops

        if self.world != 1:
            raise RuntimeError("forward_batch on a row-split net: use parallel.tp.forward_batch_tp")
        H = X.to(self.device, torch.float64).contiguous()
        NL = self.n_layers
        if precision == "bf16":
            # tensor-core chain: bf16 weights are cast once and cached until the weights change; hidden
            # layers hand bf16 activations from GEMM to GEMM, only the last layer produces FP64
            if getattr(self, "_w_bf16", None) is None:
                self._w_bf16 = [ops.cast_bf16(self.W_full[l], self.dims[l]) for l in range(NL)]
            Hb = ops.cast_bf16(H, self.dims[0])
            for l in range(NL - 1):
                Hb = ops.gemm_bf16_chain(self._w_bf16[l], Hb, self.dims[l], 0)
            kind = 1 if self.type == N.TYPE_SNN else 0
            H = ops.gemm_bf16(self._w_bf16[NL - 1], Hb, self.dims[NL - 1], None, kind)
            if self.type == N.TYPE_SNN:
                ops.softmax_rows_(H, stable_softmax)
            return H
        for l in range(NL):
            last = l == NL - 1
            kind = 1 if (last and self.type == N.TYPE_SNN) else 0
            H = ops.forward_batch(self.W_full[l], self.dims[l], H, kind, precision=precision)
        if self.type == N.TYPE_SNN:
            ops.softmax_rows_(H, stable_softmax)
        return H

    # ------------------------------------------------------------------ export
    def local_kernel_rows(self) -> List[np.ndarray]:
        torch.cuda.synchronize(self.device)
        return [w.detach().cpu().numpy().copy() for w in self.weights]

    def to_kernel(self) -> Kernel:
        """Gather the row slices of all ranks into a host ``Kernel`` (every rank gets it)."""
        rows = self.local_kernel_rows()
        if self.world == 1:
            return Kernel(list(self.dims), rows, self.name)
        import torch.distributed as dist

        everyone = [None] * self.world
        dist.all_gather_object(everyone, rows, group=self.group)
        ws = [np.concatenate([everyone[g][l] for g in range(self.world)], axis=0) for l in range(self.n_layers)]
        return Kernel(list(self.dims), ws, self.name)

    def load_weights(self, kernel: Kernel) -> None:
        self._w_bf16 = None
        for l, W in enumerate(self.W_full):
            r0 = self.net.row0[l]
            src = torch.from_numpy(np.ascontiguousarray(kernel.weights[l][r0:r0 + self.net.rows[l]]))
            W[:, : self.dims[l]].copy_(src)
