"""K-T with and without the CTA-pair / TMA-multicast tile mapping: CUDA-event times of the hidden-layer GEMM
(bf16 in, bf16 out) and of the whole 784-1000-1000-10 chain, inputs rotated through buffers larger than L2.

    python bench/kt_ab.py [batch ...]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hpnn_b200 import _native as N  # noqa: E402
from hpnn_b200.ops import cuda as ops  # noqa: E402


def timed(fn, reps=40, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / reps  # us


def main():
    lib = N.load()
    batches = [int(a) for a in sys.argv[1:]] or [4096, 16384]
    dims = [784, 1000, 1000, 10]
    Ws = [torch.rand(dims[l + 1], dims[l], dtype=torch.float64, device="cuda") - 0.5 for l in range(3)]
    Wb = [ops.cast_bf16(ops.padded_weights(W), W.shape[1]) for W in Ws]
    for B in batches:
        nbuf = max(2, int(160e6 // (B * 784 * 2)) + 1)  # rotate inputs: > 126 MB of L2 between reuses
        Xs = [ops.cast_bf16(torch.rand(B, 784, dtype=torch.float64, device="cuda"), 784) for _ in range(nbuf)]
        H1 = ops.gemm_bf16_chain(Wb[0], Xs[0], 784, 0)
        out = {"batch": B, "input_buffers": nbuf}
        ref = None
        for mode, name in ((0, "single"), (1, "pair_multicast"), (2, "pair_mma")):
            lib.hpnn_dev_set_kt_cluster(mode)
            t_hidden = timed(lambda: ops.gemm_bf16_chain(Wb[1], H1, 1000, 0))
            k = [0]

            def chain():
                k[0] = (k[0] + 1) % nbuf
                h = ops.gemm_bf16_chain(Wb[0], Xs[k[0]], 784, 0)
                h = ops.gemm_bf16_chain(Wb[1], h, 1000, 0)
                return ops.gemm_bf16(Wb[2], h, 1000, None, 1)

            t_chain = timed(chain)
            y = chain()
            if ref is None:
                ref = y
            else:
                out["max_abs_diff_between_mappings"] = float((y - ref).abs().max())
            flop = 2.0 * B * 1000 * 1000
            out[name] = {"hidden_gemm_us": round(t_hidden, 2), "hidden_gemm_tflops": round(flop / t_hidden / 1e6, 1),
                         "chain_us": round(t_chain, 2), "chain_samples_per_s": round(B / t_chain * 1e6)}
        lib.hpnn_dev_set_kt_cluster(-1)
        # library GEMM on the same operands for scale (no activation, so less work): torch.matmul -> cuBLAS
        Wt = Wb[1][:, :1000].t()
        t_lib = timed(lambda: torch.matmul(H1[:, :1000], Wt))
        out["torch_matmul_bf16_no_activation"] = {"hidden_gemm_us": round(t_lib, 2),
                                                  "hidden_gemm_tflops": round(2.0 * B * 1e6 / t_lib / 1e6, 1)}
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
