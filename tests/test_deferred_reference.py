"""The arithmetic behind the deferred-update GPU trainers (K-PL / K-PL1), checked on the CPU:
keeping the rank-1 BP updates as factors and writing them back every `depth` iterations must
give the weights of the immediate updates up to round-off, for ANN and SNN, any depth, and any
iteration count relative to the write-back period."""
import pytest
import torch

import hpnn_b200 as hp
from hpnn_b200 import _native as N
from hpnn_b200.ops import reference as R


@pytest.mark.parametrize("nn_type", [N.TYPE_ANN, N.TYPE_SNN])
@pytest.mark.parametrize("depth", [1, 2, 8])
@pytest.mark.parametrize("iterations", [1, 7, 16, 29])
def test_deferred_updates_equal_immediate_updates(nn_type, depth, iterations):
    k, _ = hp.generate_kernel(4711, 12, [17, 9, 11], 5)
    torch.manual_seed(3)
    x = torch.rand(12, dtype=torch.float64)
    t = torch.full((5,), 0.0 if nn_type == N.TYPE_SNN else -1.0, dtype=torch.float64)
    t[2] = 1.0
    lr = 0.01
    Wa = [torch.from_numpy(w.copy()) for w in k.weights]
    Wb = [torch.from_numpy(w.copy()) for w in k.weights]
    for _ in range(iterations):  # the reference's order: deltas from the old weights, update, forward
        acts = R.forward(Wa, x, nn_type)
        d = R.deltas(Wa, acts, t, nn_type)
        R.apply_update(Wa, None, d, acts, lr, 0.0)
    out_a = R.forward(Wa, x, nn_type)[-1]
    out_b = R.train_iterations_deferred(Wb, x, t, nn_type, lr, iterations, depth=depth)[-1]
    for a, b in zip(Wa, Wb):
        torch.testing.assert_close(a, b, atol=1e-13, rtol=0)
    torch.testing.assert_close(out_a, out_b, atol=1e-13, rtol=0)
