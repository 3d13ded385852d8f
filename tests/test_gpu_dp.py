"""Data-parallel gradient merge (extension): the device path in factored form on one GPU.

With one rank the merged update is the rank's own BP update, so three merged steps must equal three
plain BP steps of the PyTorch specification; the multi-rank equalities (factored == NVLS == peer
loads == NCCL == serial sum, replicas bit-identical) are in tests/mp_tp_check.py."""
import numpy as np
import pytest
import torch

import hpnn_b200 as hp
from hpnn_b200 import _native as N
from hpnn_b200.ops import reference as R

from conftest import make_problem

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("nn_type", [N.TYPE_ANN, N.TYPE_SNN])
@pytest.mark.parametrize("topo", [(50, (30, 20), 6), (37, (1100, 70), 10)])
def test_factored_merge_single_rank_equals_plain_bp_steps(nn_type, topo):
    from hpnn_b200.ops import cuda as ops
    from hpnn_b200.parallel.dp import DataParallelTrainer

    n_in, hid, n_out = topo
    k, _ = hp.generate_kernel(5, n_in, list(hid), n_out)
    X, T = make_problem(nn_type, n_in, hid, n_out, n_samples=3, seed=3)
    tr = DataParallelTrainer(k, nn_type, lr=0.01, merge="factored")
    W = [torch.from_numpy(w.copy()).cuda() for w in k.weights]
    launches0 = ops.launch_count()
    for x, t in zip(X, T):
        xs, ts = torch.from_numpy(x).cuda(), torch.from_numpy(t).cuda()
        acts = R.forward(W, xs, nn_type)
        E_ref = float(R.error(acts[-1], ts, nn_type))
        d = R.deltas(W, acts, ts, nn_type)
        for l in range(len(W)):
            W[l] += torch.outer(0.01 * d[l], acts[l])
        E = tr.step(x, t)
        assert abs(E - E_ref) < 1e-12
    tr.check()
    assert ops.launch_count() - launches0 >= 3 * (1 + len(W))  # publish + one apply per layer, per step
    for a, b in zip(tr.W, W):
        torch.testing.assert_close(a, b, atol=1e-13, rtol=0)
    tr.close()
