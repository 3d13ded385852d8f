"""Plain PyTorch fp64 statement of every hpnn operation.

This is the readable specification: it runs on CPU or GPU, is what the CUDA kernels
are tested against, and is the compute path of the collective-based (gloo / NCCL)
row-split baseline in :mod:`hpnn_b200.parallel.tp_collective`."""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import torch

TINY = 1e-14
TYPE_ANN, TYPE_SNN = 0, 2
TRAIN_BP, TRAIN_BPM = 0, 1


def act(x: torch.Tensor) -> torch.Tensor:
    """``2/(1+exp(-x))-1`` (src/ann.c:883-885)."""
    return 2.0 / (1.0 + torch.exp(-1.0 * x)) - 1.0


def dact(y: torch.Tensor) -> torch.Tensor:
    """Derivative expressed with the *output*: ``-0.5(y^2-1)`` (src/ann.c:886-888)."""
    return -0.5 * (y * y - 1.0)


def softmax_ref(z: torch.Tensor, stable: bool = False) -> torch.Tensor:
    """``exp(z-1)/(TINY+sum exp(z-1))`` -- the reference does not subtract the max
    (src/snn.c:282-334); ``stable=True`` shifts by the max instead of 1."""
    shift = z.max(dim=-1, keepdim=True).values if stable else 1.0
    e = torch.exp(z - shift)
    return e / (e.sum(dim=-1, keepdim=True) + TINY)


def forward(weights: Sequence[torch.Tensor], x: torch.Tensor, nn_type: int, stable: bool = False) -> List[torch.Tensor]:
    """Returns ``[x, h_0, ..., h_last, o]``; works for a vector or a ``[B, n_in]`` batch."""
    acts = [x]
    L = len(weights)
    for l, W in enumerate(weights):
        z = acts[-1] @ W.T
        if l == L - 1 and nn_type == TYPE_SNN:
            acts.append(softmax_ref(z, stable))
        else:
            acts.append(act(z))
    return acts


def error(o: torch.Tensor, t: torch.Tensor, nn_type: int) -> torch.Tensor:
    if nn_type == TYPE_SNN:
        term = torch.where(o > 0, t * torch.log(o + TINY), torch.zeros_like(o))
        return -term.sum(dim=-1) / o.shape[-1]
    return 0.5 * ((t - o) ** 2).sum(dim=-1)


def output_delta(o: torch.Tensor, t: torch.Tensor, nn_type: int) -> torch.Tensor:
    return (t - o) if nn_type == TYPE_SNN else (t - o) * dact(o)


def deltas(weights: Sequence[torch.Tensor], acts: Sequence[torch.Tensor], t: torch.Tensor, nn_type: int) -> List[torch.Tensor]:
    """All deltas from the *old* weights, top-down (src/ann.c:1279-1592)."""
    L = len(weights)
    d = [None] * L
    d[L - 1] = output_delta(acts[L], t, nn_type)
    for l in range(L - 1, 0, -1):
        d[l - 1] = (weights[l].T @ d[l]) * dact(acts[l])
    return d


def apply_update(weights: List[torch.Tensor], momentum: Optional[List[torch.Tensor]], d: Sequence[torch.Tensor],
                 acts: Sequence[torch.Tensor], lr: float, alpha: float) -> None:
    """BP ``W += lr d (x) h``; BPM ``dw += lr d (x) h; W += dw; dw *= alpha`` (in place)."""
    for l, W in enumerate(weights):
        g = torch.outer(lr * d[l], acts[l])
        if momentum is None:
            W += g
        else:
            momentum[l] += g
            W += momentum[l]
            momentum[l] *= alpha


def argmax_first(o: torch.Tensor) -> int:
    """Index the reference's scan ``probe=-1; if(probe<o[i])`` ends on (src/ann.c:2346-2351)."""
    best, idx = -1.0, 0
    for i, v in enumerate(o.tolist()):
        if best < v:
            best, idx = v, i
    return idx


def target_index(t: torch.Tensor) -> int:
    idx = 0
    for i, v in enumerate(t.tolist()):
        if v == 1.0:
            idx = i
    return idx


@dataclass
class TrainResult:
    init_err: float = 0.0
    final_de: float = 0.0
    last_err: float = 0.0
    n_iter: int = 0
    first_ok: bool = False
    success: bool = False


def default_lr(nn_type: int, train: int, cuda_defaults: bool) -> float:
    """the CUDA build trains with 0.01 everywhere, the CPU ANN build with the
    header constants 0.001 (BP) / 0.0005 (BPM), the CPU SNN build with 0.01."""
    if cuda_defaults or nn_type == TYPE_SNN:
        return 0.01
    return 0.0005 if train == TRAIN_BPM else 0.001


def train_sample(weights: List[torch.Tensor], x: torch.Tensor, t: torch.Tensor, nn_type: int, train: int,
                 lr: float, alpha: float = 0.2, delta: float = 1e-6, min_iter: Optional[int] = None,
                 max_iter: int = 102399, stable: bool = False) -> Tuple[TrainResult, torch.Tensor]:
    """Per-sample convergence loop (src/ann.c:2281-2467, src/snn.c:1417-1595)."""
    if min_iter is None:
        min_iter = 15 if train == TRAIN_BPM else 31
    mom = [torch.zeros_like(W) for W in weights] if train == TRAIN_BPM else None
    acts = forward(weights, x, nn_type, stable)
    E = float(error(acts[-1], t, nn_type))
    res = TrainResult(init_err=E)
    p_trg = target_index(t)
    it = 0
    while True:
        it += 1
        d = deltas(weights, acts, t, nn_type)
        apply_update(weights, mom, d, acts, lr, alpha)
        acts = forward(weights, x, nn_type, stable)
        En = float(error(acts[-1], t, nn_type))
        dE, E = E - En, En
        ok = argmax_first(acts[-1]) == p_trg
        if it == 1:
            res.first_ok = ok
        if it > max_iter:
            break
        ok = ok and it > min_iter
        if not (dE > delta or not ok):
            break
    res.n_iter, res.final_de, res.last_err, res.success = it, dE, E, ok
    return res, acts[-1]


def train_iterations_deferred(weights: List[torch.Tensor], x: torch.Tensor, t: torch.Tensor, nn_type: int, lr: float,
                              iterations: int, depth: int = 8, stable: bool = False) -> List[torch.Tensor]:
    """Plain PyTorch restatement of the deferred-update BP trainers (csrc/kernels/train_lazy.cu,
    train_lazy1.cu): ``iterations`` BP iterations on one sample, updating ``weights`` in place.

    Layer l >= 1 keeps its last m < ``depth`` rank-1 updates as factors ``u_j = lr*delta_l``,
    ``v_j = h_(l-1)`` instead of adding them to the stored matrix S:

        W_true^T d = S^T d + sum_j v_j (u_j . d)        W_true x = S x + sum_j u_j (v_j . x)

    and S is brought up to date every ``depth`` iterations (and at the end).  Layer 0 always sees
    the same input, so its updates collapse into one factor pair ``S0 + ubar (x) x``.  Returns the
    activations of the last forward pass.  Equal to ``apply_update`` after every iteration up to
    the rounding of the two correction sums (tests/test_deferred_reference.py)."""
    L = len(weights)
    U: List[List[torch.Tensor]] = [[] for _ in range(L)]  # pending factors per layer (l >= 1)
    V: List[List[torch.Tensor]] = [[] for _ in range(L)]
    ubar = torch.zeros(weights[0].shape[0], dtype=weights[0].dtype, device=weights[0].device)
    xx = x @ x

    def fwd() -> List[torch.Tensor]:
        acts = [x]
        for l, S in enumerate(weights):
            h = acts[-1]
            z = S @ h
            if l == 0:
                z = z + ubar * xx
            else:
                for u, v in zip(U[l], V[l]):
                    z = z + u * (v @ h)
            acts.append(softmax_ref(z, stable) if (l == L - 1 and nn_type == TYPE_SNN) else act(z))
        return acts

    def flush(l: int) -> None:
        for u, v in zip(U[l], V[l]):  # oldest first, like the immediate updates
            weights[l] += torch.outer(u, v)
        U[l].clear()
        V[l].clear()

    acts = fwd()
    for _ in range(iterations):
        d = [None] * L
        d[L - 1] = output_delta(acts[L], t, nn_type)
        for l in range(L - 1, 0, -1):
            p = weights[l].T @ d[l]
            for u, v in zip(U[l], V[l]):
                p = p + v * (u @ d[l])
            d[l - 1] = p * dact(acts[l])
        for l in range(1, L):
            U[l].append(lr * d[l])
            V[l].append(acts[l].clone())
            if len(U[l]) >= depth:
                flush(l)
        ubar += lr * d[0]
        acts = fwd()
    for l in range(1, L):
        flush(l)
    weights[0] += torch.outer(ubar, x)
    return acts
