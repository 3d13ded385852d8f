"""run_nn inference sweep (BASELINE.json config 5): SNN 784-1000-1000-10, batch 1..4096.

Rows: samples/s of
  * ours, per-sample persistent kernel (the reference-faithful batch-1 path),
  * ours, batched FP64 GEMM path (K-G), and opt-in bf16 tensor-core path (K-T),
  * torch/cuBLAS fp64 and bf16 matmul of the same layers (library roofline reference),
  * the unmodified reference's nn_run_kernel (batch 1 only; batch B = B x single-sample time)."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpnn_b200 as hp  # noqa: E402
from hpnn_b200.models.net import HpnnNet  # noqa: E402
from hpnn_b200.ops import cuda as ops  # noqa: E402
from hpnn_b200.ops import reference as R  # noqa: E402


def timed(fn, n=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def main():
    dims = [784, 1000, 1000, 10]
    k, _ = hp.generate_kernel(10958, dims[0], dims[1:-1], dims[-1])
    net = HpnnNet(k, hp.TYPE_SNN, momentum=False)
    Wt = [w.clone() for w in net.weights]
    Wb16 = [w.to(torch.bfloat16) for w in Wt]
    Wbf = [ops.cast_bf16(net.W_full[l], dims[l]) for l in range(3)]
    out = []
    rng = np.random.default_rng(0)
    x1 = torch.from_numpy(rng.random(784)).pin_memory()
    ms1 = timed(lambda: net.run(x1), n=200)
    out.append({"path": "ours K-P run (batch 1, fp64, H2D per sample)", "batch": 1, "samples_per_s": 1e3 / ms1})
    print(out[-1], flush=True)
    for B in (1, 8, 64, 256, 1024, 4096, 16384):
        X = torch.from_numpy(rng.random((B, 784))).cuda()
        row = {"batch": B}
        row["ours_fp64_KG"] = B * 1e3 / timed(lambda: net.forward_batch(X))
        row["ours_bf16_KT"] = B * 1e3 / timed(lambda: net.forward_batch(X, precision="bf16"))

        def kt_cached():
            H = X
            for l in range(3):
                Xb = ops.cast_bf16(H, dims[l])
                H = ops.gemm_bf16(Wbf[l], Xb, dims[l], act_kind=1 if l == 2 else 0)
            return ops.softmax_rows_(H)

        row["ours_bf16_KT_cached_weights"] = B * 1e3 / timed(kt_cached)
        row["torch_cublas_fp64"] = B * 1e3 / timed(lambda: R.forward(Wt, X, hp.TYPE_SNN)[-1])
        Xh = X.to(torch.bfloat16)

        def tb():
            H = Xh
            for l, W in enumerate(Wb16):
                H = H @ W.T
                H = torch.softmax(H.float(), 1) if l == 2 else (2.0 / (1.0 + torch.exp(-H.float())) - 1.0).to(torch.bfloat16)
            return H

        row["torch_cublas_bf16"] = B * 1e3 / timed(tb)
        out.append(row)
        print({k_: (round(v, 1) if isinstance(v, float) else v) for k_, v in row.items()}, flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/infer_sweep.json", "w"), indent=1)
    net.close()


if __name__ == "__main__":
    main()
