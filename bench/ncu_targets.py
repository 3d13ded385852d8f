"""Short, single-purpose scenarios for `ncu --set full` (one GPU, few launches).

    python bench/ncu_targets.py kt | kt_chain [batch ...] | kg | kp_wide | kpl1_wide | kp_snn_stream | kpr_snn | ops_wide"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpnn_b200 as hp  # noqa: E402
from hpnn_b200.ops import cuda as ops  # noqa: E402
from hpnn_b200.trainer import CudaEngine  # noqa: E402


def train_case(nn_type, train, n_in, hid, n_out, iters):
    k, _ = hp.generate_kernel(10958, n_in, hid, n_out)
    eng = CudaEngine(k, nn_type, train, max_iter=iters, min_iter=iters + 10, delta=1e-300)
    rng = np.random.default_rng(0)
    x = rng.random(n_in)
    t = np.full(n_out, 0.0 if nn_type == hp.TYPE_SNN else -1.0)
    t[3] = 1.0
    for _ in range(3):
        r = eng.train(x, t)
    print("iterations per launch:", r.n_iter)
    eng.close()


def main():
    which = sys.argv[1]
    if which == "kt_chain":  # bf16-in / bf16-out hidden-layer GEMM (persistent, two TMEM accumulators, TMA store)
        rows, cols = 1000, 1000
        W = torch.rand(rows, cols, dtype=torch.float64, device="cuda") - 0.5
        Wb = ops.cast_bf16(ops.padded_weights(W), cols)
        for B in ([int(a) for a in sys.argv[2:]] or [4096]):
            X = torch.rand(B, cols, dtype=torch.float64, device="cuda")
            Xb = ops.cast_bf16(X, cols)
            for _ in range(3):
                ops.gemm_bf16_chain(Wb, Xb, cols, 0)
        torch.cuda.synchronize()
    elif which in ("kt", "kg"):
        rows, cols, B = 1000, 1000, 4096
        W = torch.rand(rows, cols, dtype=torch.float64, device="cuda") - 0.5
        X = torch.rand(B, cols, dtype=torch.float64, device="cuda")
        Wp = ops.padded_weights(W)
        if which == "kt":
            Wb, Xb = ops.cast_bf16(Wp, cols), ops.cast_bf16(X, cols)
            for _ in range(4):
                ops.gemm_bf16(Wb, Xb, cols, act_kind=0)
        else:
            for _ in range(4):
                ops.forward_batch(Wp, cols, X, 0)
        torch.cuda.synchronize()
    elif which == "kp_wide":  # plain streaming trainer
        os.environ["HPNN_LAZY"] = "0"
        train_case(hp.TYPE_ANN, hp.TRAIN_BP, 784, [8192] * 4, 10, 4)
    elif which == "kpl1_wide":  # deferred-update row-panel trainer: 16 iterations = two write-back passes
        train_case(hp.TYPE_ANN, hp.TRAIN_BP, 784, [8192] * 4, 10, 16)
    elif which == "kp_snn_stream":
        os.environ["HPNN_NO_RESIDENT"] = "1"
        train_case(hp.TYPE_SNN, hp.TRAIN_BPM, 784, [1000, 1000], 10, 100)
    elif which == "kpr_snn":
        train_case(hp.TYPE_SNN, hp.TRAIN_BPM, 784, [1000, 1000], 10, 100)
    elif which == "ops_wide":
        rows = cols = 8192
        W = torch.rand(rows, cols, dtype=torch.float64, device="cuda") - 0.5
        x = torch.rand(cols, dtype=torch.float64, device="cuda")
        d = torch.rand(rows, dtype=torch.float64, device="cuda")
        Wp, xp = ops.padded_weights(W), ops.padded_vector(x)
        for _ in range(3):
            ops.forward_layer(Wp, cols, xp)
            ops.backward_update(Wp, cols, d, xp, 1, 1e-9)
            ops.update_forward(Wp, cols, d, xp, 1, 1e-9)
        torch.cuda.synchronize()


if __name__ == "__main__":
    main()
