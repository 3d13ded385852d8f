"""Every stand-alone sm_100a kernel against the plain PyTorch fp64 statement of the same op."""
import numpy as np
import pytest
import torch

import hpnn_b200 as hp
from hpnn_b200 import _native as N
from hpnn_b200.ops import reference as R

pytestmark = pytest.mark.gpu

SHAPES = [(10, 1000), (300, 784), (1000, 1000), (230, 851), (37, 13), (2048, 4096)]


def _dev():
    assert torch.cuda.is_available()
    return torch.device("cuda", 0)


def _rand(*shape, seed=0):
    g = torch.Generator().manual_seed(seed)
    return (torch.rand(*shape, generator=g, dtype=torch.float64) * 2 - 1).to(_dev())


@pytest.mark.parametrize("rows,cols", SHAPES)
@pytest.mark.parametrize("act_kind", [0, 1])
def test_forward_layer(rows, cols, act_kind):
    from hpnn_b200.ops import cuda as ops

    W = _rand(rows, cols, seed=1) / np.sqrt(cols)
    x = _rand(cols, seed=2)
    y = ops.forward_layer(ops.padded_weights(W), cols, ops.padded_vector(x), act_kind)
    ref = W @ x
    if act_kind == 0:
        ref = R.act(ref)
    torch.testing.assert_close(y, ref, atol=1e-13, rtol=0)


@pytest.mark.parametrize("nn_type", [N.TYPE_ANN, N.TYPE_SNN])
@pytest.mark.parametrize("n", [3, 10, 230, 5000])
@pytest.mark.parametrize("stable", [False, True])
def test_output_stats(nn_type, n, stable):
    from hpnn_b200.ops import cuda as ops

    z = _rand(n, seed=3) * (3.0 if nn_type == N.TYPE_SNN else 0.99)
    lo = 0.0 if nn_type == N.TYPE_SNN else -1.0
    t = torch.full((n,), lo, dtype=torch.float64, device=_dev())
    t[n // 2] = 1.0
    o, d, E, amax, trg = ops.output_stats(nn_type, z, t, stable)
    o_ref = R.softmax_ref(z, stable) if nn_type == N.TYPE_SNN else z
    torch.testing.assert_close(o, o_ref, atol=1e-14, rtol=0)
    torch.testing.assert_close(d, R.output_delta(o_ref, t, nn_type), atol=1e-14, rtol=0)
    assert abs(E - float(R.error(o_ref, t, nn_type))) < 1e-12 * max(1.0, abs(E))
    assert amax == R.argmax_first(o_ref.cpu()) and trg == n // 2


def test_output_stats_argmax_ties_and_floor():
    from hpnn_b200.ops import cuda as ops

    z = torch.tensor([0.2, 0.7, 0.7, -0.3], dtype=torch.float64, device=_dev())
    t = torch.tensor([-1.0, -1.0, 1.0, 1.0], dtype=torch.float64, device=_dev())
    _, _, _, amax, trg = ops.output_stats(N.TYPE_ANN, z, t)
    assert amax == 1 and trg == 3  # first maximum, last target==1.0 (src/ann.c:2346-2352)
    z = torch.full((4,), -1.0, dtype=torch.float64, device=_dev())
    assert ops.output_stats(N.TYPE_ANN, z, t)[3] == 0  # nothing beats probe=-1.0


@pytest.mark.parametrize("rows,cols", SHAPES)
@pytest.mark.parametrize("update", [0, 1, 2])
def test_backward_update(rows, cols, update):
    from hpnn_b200.ops import cuda as ops

    lr, alpha = 0.01, 0.2
    W = _rand(rows, cols, seed=4) / np.sqrt(cols)
    M = _rand(rows, cols, seed=5) * 1e-3
    d = _rand(rows, seed=6)
    h = _rand(cols, seed=7)
    Wp, Mp = ops.padded_weights(W), ops.padded_weights(M)
    part = ops.backward_update(Wp, cols, d, ops.padded_vector(h), update, lr, alpha, Mp if update == 2 else None)
    torch.testing.assert_close(part, W.T @ d, atol=1e-12, rtol=0)  # from the OLD weights
    g = torch.outer(lr * d, h)
    if update == 0:
        W_ref, M_ref = W, M
    elif update == 1:
        W_ref, M_ref = W + g, M
    else:
        W_ref, M_ref = W + (M + g), (M + g) * alpha
    torch.testing.assert_close(Wp[:, :cols], W_ref, atol=1e-15, rtol=0)
    torch.testing.assert_close(Mp[:, :cols], M_ref, atol=1e-15, rtol=0)
    assert (Wp[:, cols:] == 0).all() and (Mp[:, cols:] == 0).all()  # padding stays zero


def test_delta_finalize():
    from hpnn_b200.ops import cuda as ops

    parts = _rand(4, 1000, seed=8)
    h = _rand(1000, seed=9)
    d = ops.delta_finalize(parts, h)
    ref = ((parts[0] + parts[1]) + parts[2] + parts[3]) * R.dact(h)
    torch.testing.assert_close(d, ref, atol=1e-15, rtol=0)


@pytest.mark.parametrize("rows,cols", SHAPES[:5])
@pytest.mark.parametrize("update", [1, 2])
def test_update_forward(rows, cols, update):
    from hpnn_b200.ops import cuda as ops

    lr, alpha = 0.01, 0.2
    W = _rand(rows, cols, seed=10) / np.sqrt(cols)
    M = _rand(rows, cols, seed=11) * 1e-3
    d = _rand(rows, seed=12)
    x = _rand(cols, seed=13)
    Wp, Mp = ops.padded_weights(W), ops.padded_weights(M)
    y = ops.update_forward(Wp, cols, d, ops.padded_vector(x), update, lr, alpha, Mp if update == 2 else None)
    g = torch.outer(lr * d, x)
    W_ref = W + g if update == 1 else W + (M + g)
    torch.testing.assert_close(Wp[:, :cols], W_ref, atol=1e-15, rtol=0)
    torch.testing.assert_close(y, R.act(W_ref @ x), atol=1e-13, rtol=0)
    if update == 2:
        torch.testing.assert_close(Mp[:, :cols], (M + g) * alpha, atol=1e-15, rtol=0)


@pytest.mark.parametrize("batch", [1, 7, 64, 333])
def test_forward_batch_fp64_and_softmax(batch):
    from hpnn_b200.ops import cuda as ops

    rows, cols = 230, 851
    W = _rand(rows, cols, seed=14) / np.sqrt(cols)
    X = _rand(batch, cols, seed=15)
    Y = ops.forward_batch(ops.padded_weights(W), cols, X, act_kind=1)
    torch.testing.assert_close(Y, X @ W.T, atol=1e-12, rtol=0)
    Z = Y.clone()
    ops.softmax_rows_(Z)
    torch.testing.assert_close(Z, R.softmax_ref(Y), atol=1e-14, rtol=0)
    Ya = ops.forward_batch(ops.padded_weights(W), cols, X, act_kind=0)
    torch.testing.assert_close(Ya, R.act(X @ W.T), atol=1e-12, rtol=0)


def test_grad_merge_local():
    from hpnn_b200.ops import cuda as ops

    W = _rand(1001, seed=16)
    gs = [_rand(1001, seed=20 + i) for i in range(3)]
    ref = W + 0.5 * ((gs[0] + gs[1]) + gs[2])
    ops.grad_merge_(W, [g.data_ptr() for g in gs], 0.5)
    torch.testing.assert_close(W, ref, atol=1e-15, rtol=0)
