"""Torch-tensor front end of the stand-alone sm_100a kernels (K-A..K-E, K-G, K-M).

Every function launches exactly one hand-written kernel on the current CUDA stream;
there is no PyTorch fallback here: a missing native library on a GPU box raises."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import torch

from .. import _native as N


def _s() -> int:
    return torch.cuda.current_stream().cuda_stream


def _chk(t: torch.Tensor, name: str) -> torch.Tensor:
    if not (t.is_cuda and t.dtype == torch.float64):
        raise TypeError(f"{name} must be a CUDA float64 tensor")
    return t


def padded_weights(W: torch.Tensor) -> torch.Tensor:
    """Copy ``[N, M]`` into zero-padded ``[N, pad8(M)]`` storage (64-byte aligned rows)."""
    n, m = W.shape
    ld = (m + 7) & ~7
    out = torch.zeros((n, ld), dtype=torch.float64, device=W.device)
    out[:, :m].copy_(W)
    return out


def padded_vector(x: torch.Tensor) -> torch.Tensor:
    m = x.numel()
    out = torch.zeros(((m + 7) & ~7,), dtype=torch.float64, device=x.device)
    out[:m].copy_(x)
    return out


def forward_layer(Wp: torch.Tensor, cols: int, xp: torch.Tensor, act_kind: int = 0) -> torch.Tensor:
    """K-A: ``y = f(W x)``; ``Wp`` padded ``[N, ld]``, ``xp`` padded ``[ld]``."""
    lib = N.load()
    _chk(Wp, "Wp"), _chk(xp, "xp")
    rows, ld = Wp.shape
    assert xp.numel() >= ld and Wp.stride(0) == ld
    y = torch.empty(rows, dtype=torch.float64, device=Wp.device)
    N.check(lib.hpnn_dev_forward_layer(Wp.data_ptr(), ld, rows, cols, xp.data_ptr(), y.data_ptr(), act_kind, _s()),
            "hpnn_dev_forward_layer")
    return y


def output_stats(nn_type: int, z: torch.Tensor, t: torch.Tensor, stable: bool = False):
    """K-B: returns ``(o, delta_out, E, argmax, target_index)``."""
    lib = N.load()
    _chk(z, "z"), _chk(t, "t")
    n = z.numel()
    o = torch.empty_like(z)
    d = torch.empty_like(z)
    st = torch.empty(3, dtype=torch.float64, device=z.device)
    N.check(lib.hpnn_dev_output_stats(nn_type, int(stable), z.data_ptr(), t.data_ptr(), n, o.data_ptr(), d.data_ptr(),
                                      st.data_ptr(), _s()), "hpnn_dev_output_stats")
    s = st.tolist()
    return o, d, s[0], int(s[1]), int(s[2])


def backward_update(Wp: torch.Tensor, cols: int, delta: torch.Tensor, h: Optional[torch.Tensor], update: int = 0,
                    lr: float = 0.0, alpha: float = 0.0, dWp: Optional[torch.Tensor] = None) -> torch.Tensor:
    """K-C fused with K-D/K-E: ``part = W^T delta`` from the old weights, then (update 1/2)
    the BP/BPM update of ``Wp`` (and ``dWp``) in the same pass.  ``h`` is the padded input
    vector of the layer."""
    lib = N.load()
    _chk(Wp, "Wp"), _chk(delta, "delta")
    rows, ld = Wp.shape
    part = torch.zeros(ld, dtype=torch.float64, device=Wp.device)
    N.check(lib.hpnn_dev_backward_update(Wp.data_ptr(), dWp.data_ptr() if dWp is not None else None, ld, rows, cols,
                                         delta.data_ptr(), h.data_ptr() if h is not None else None, part.data_ptr(),
                                         update, lr, alpha, _s()), "hpnn_dev_backward_update")
    return part[:cols]


def delta_finalize(parts: torch.Tensor, h: torch.Tensor) -> torch.Tensor:
    """``delta[r] = dact(h[r]) * sum_g parts[g, r]`` (fixed rank order)."""
    lib = N.load()
    parts = parts.reshape(-1, parts.shape[-1]).contiguous()
    g, n = parts.shape
    d = torch.empty(n, dtype=torch.float64, device=parts.device)
    N.check(lib.hpnn_dev_delta_finalize(parts.data_ptr(), g, n, h.data_ptr(), n, d.data_ptr(), _s()),
            "hpnn_dev_delta_finalize")
    return d


def update_forward(Wp: torch.Tensor, cols: int, delta: torch.Tensor, xp: torch.Tensor, update: int, lr: float,
                   alpha: float = 0.0, dWp: Optional[torch.Tensor] = None, act_kind: int = 0) -> torch.Tensor:
    """K-D / K-E: row-wise update fused with the forward of the same layer."""
    lib = N.load()
    rows, ld = Wp.shape
    y = torch.empty(rows, dtype=torch.float64, device=Wp.device)
    N.check(lib.hpnn_dev_update_forward(Wp.data_ptr(), dWp.data_ptr() if dWp is not None else None, ld, rows, cols,
                                        delta.data_ptr(), xp.data_ptr(), y.data_ptr(), update, lr, alpha, act_kind,
                                        _s()), "hpnn_dev_update_forward")
    return y


def forward_batch(Wp: torch.Tensor, cols: int, X: torch.Tensor, act_kind: int = 0, precision: str = "fp64") -> torch.Tensor:
    """K-G / K-T: ``Y[B, N] = f(X[B, M] W^T)``.  ``precision`` "fp64" (FMA pipe, parity) or
    "bf16" (opt-in: tcgen05 tensor cores, fp32 accumulation in TMEM)."""
    lib = N.load()
    rows, ld = Wp.shape
    B = X.shape[0]
    if X.stride(1) != 1 or X.stride(0) % 2 or X.data_ptr() % 16:  # 16-byte aligned rows for cp.async / TMA
        Xa = torch.zeros((B, (X.shape[1] + 7) & ~7), dtype=torch.float64, device=Wp.device)
        Xa[:, : X.shape[1]].copy_(X)
        X = Xa
    Y = torch.empty((B, (rows + 7) & ~7), dtype=torch.float64, device=Wp.device)[:, :rows]
    if precision == "fp64":
        N.check(lib.hpnn_dev_forward_batch_f64(Wp.data_ptr(), ld, rows, cols, X.data_ptr(), X.stride(0), B,
                                               Y.data_ptr(), Y.stride(0), act_kind, _s()),
                "hpnn_dev_forward_batch_f64")
    elif precision == "bf16":
        Wb = cast_bf16(Wp, cols)
        Xb = cast_bf16(X, cols)
        gemm_bf16(Wb, Xb, cols, Y, act_kind)
    else:
        raise ValueError(f"unknown precision {precision!r}")
    return Y


def pad64(v: int) -> int:
    """Leading dimension of the bf16 operands of K-T: rows of 128-byte multiples, so that every 128-byte box row
    of a TMA load or store is one cache line (with pad8 rows a box row straddles two lines: +25 % L2 sectors)."""
    return (v + 63) & ~63


def cast_bf16(A: torch.Tensor, cols: int) -> torch.Tensor:
    """fp64 ``[R, >=cols]`` -> bf16 ``[R, pad64(cols)]`` (zero padded) with the native cast kernel."""
    lib = N.load()
    rows = A.shape[0]
    ld = pad64(cols)
    out = torch.empty((rows, ld), dtype=torch.bfloat16, device=A.device)
    N.check(lib.hpnn_dev_cast_bf16(A.data_ptr(), A.stride(0), rows, cols, out.data_ptr(), ld, _s()), "hpnn_dev_cast_bf16")
    return out


def gemm_bf16(Wb: torch.Tensor, Xb: torch.Tensor, cols: int, Y: Optional[torch.Tensor] = None, act_kind: int = 1) -> torch.Tensor:
    """K-T: ``Y = f(Xb Wb^T)`` on tcgen05 tensor cores (bf16 in, fp32 accumulate, fp64 out)."""
    lib = N.load()
    rows, B = Wb.shape[0], Xb.shape[0]
    if Y is None:
        Y = torch.empty((B, rows), dtype=torch.float64, device=Wb.device)
    N.check(lib.hpnn_dev_gemm_bf16(Wb.data_ptr(), Wb.stride(0), rows, cols, Xb.data_ptr(), Xb.stride(0), B, Y.data_ptr(),
                                   Y.stride(0), act_kind, _s()), "hpnn_dev_gemm_bf16")
    return Y


def gemm_bf16_chain(Wb: torch.Tensor, Xb: torch.Tensor, cols: int, act_kind: int = 0) -> torch.Tensor:
    """K-T for hidden layers: bf16 in, bf16 out ``[B, pad64(N)]`` (activation in fp32), written by TMA
    stores -- the result is the next layer's GEMM operand, no FP64 round trip and no cast kernel.  The pad
    columns are written by the kernel (zeros): no memset."""
    lib = N.load()
    rows, B = Wb.shape[0], Xb.shape[0]
    Yb = torch.empty((B, pad64(rows)), dtype=torch.bfloat16, device=Wb.device)
    N.check(lib.hpnn_dev_gemm_bf16_bf16(Wb.data_ptr(), Wb.stride(0), rows, cols, Xb.data_ptr(), Xb.stride(0), B,
                                        Yb.data_ptr(), Yb.stride(0), act_kind, _s()), "hpnn_dev_gemm_bf16_bf16")
    return Yb


def softmax_rows_(Z: torch.Tensor, stable: bool = False) -> torch.Tensor:
    lib = N.load()
    B, n = Z.shape
    N.check(lib.hpnn_dev_softmax_rows(Z.data_ptr(), Z.stride(0), B, n, int(stable), _s()), "hpnn_dev_softmax_rows")
    return Z


def grad_merge_(W: torch.Tensor, grad_ptrs: Sequence[int], scale: float = 1.0) -> torch.Tensor:
    """K-M: ``W += scale * sum_r G_r`` with the ranks summed in index order; the pointers may
    be peer-mapped buffers of other GPUs (the loads then cross NVLink inside the kernel)."""
    lib = N.load()
    n = W.numel()
    arr = (C.c_void_p * len(grad_ptrs))(*grad_ptrs)
    N.check(lib.hpnn_dev_grad_merge(W.data_ptr(), n, arr, len(grad_ptrs), scale, _s()), "hpnn_dev_grad_merge")
    return W


def launch_count(reset: bool = False) -> int:
    return int(N.load().hpnn_dev_launch_count(int(reset)))
