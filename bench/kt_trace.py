"""Time line of CTA 0 of one K-T launch (clock64 stamps written by the kernel itself): set-up, the
producer's and the MMA thread's progress per k-step, the epilogue of each tile.

    python bench/kt_trace.py [batch] """
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hpnn_b200 import _native as N  # noqa: E402
from hpnn_b200.ops import cuda as ops  # noqa: E402


def main():
    lib = N.load()
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    W = torch.rand(1000, 1000, dtype=torch.float64, device="cuda") - 0.5
    X = torch.rand(B, 1000, dtype=torch.float64, device="cuda")
    Wb, Xb = ops.cast_bf16(ops.padded_weights(W), 1000), ops.cast_bf16(X, 1000)
    trace = torch.zeros(160, dtype=torch.int64, device="cuda")
    for mode, name in ((0, "single"), (1, "pair_multicast"), (2, "pair_mma")):
        lib.hpnn_dev_set_kt_cluster(mode)
        for _ in range(3):
            ops.gemm_bf16_chain(Wb, Xb, 1000, 0)
        torch.cuda.synchronize()
        trace.zero_()
        lib.hpnn_dev_set_kt_trace(trace.data_ptr())
        ops.gemm_bf16_chain(Wb, Xb, 1000, 0)
        torch.cuda.synchronize()
        lib.hpnn_dev_set_kt_trace(None)
        t = trace.cpu().tolist()
        t0 = t[0]
        rel = lambda v: (v - t0) if v else None
        print(json.dumps({"mapping": name, "batch": B, "setup_done": rel(t[1]),
                          "producer_past_ring_wait": [rel(v) for v in t[8:72] if v],
                          "mma_past_full_wait": [rel(v) for v in t[72:136] if v],
                          "epilogue_start_end": [[rel(t[136 + 2 * i]), rel(t[137 + 2 * i])] for i in range(8) if t[136 + 2 * i]]}),
              flush=True)
    lib.hpnn_dev_set_kt_cluster(-1)


if __name__ == "__main__":
    main()
