"""Per-kernel bandwidth of the stand-alone phase kernels on a large layer (roofline check).

Reports achieved GB/s of algorithmic bytes against MEASURED_PEAKS.json ``hbm_gbs``."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hpnn_b200.ops import cuda as ops  # noqa: E402


def timeit(fn, n=10, warm=3):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    peak = 6539.5
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = json.load(open(p)).get("hbm_gbs", peak)
    out = []
    for rows, cols in ((8192, 8192), (1024, 8192), (8192, 784)):
        W = torch.rand(rows, cols, dtype=torch.float64, device="cuda") - 0.5
        M = torch.zeros_like(W)
        x = torch.rand(cols, dtype=torch.float64, device="cuda")
        d = torch.rand(rows, dtype=torch.float64, device="cuda")
        Wp, Mp, xp = ops.padded_weights(W), ops.padded_weights(M), ops.padded_vector(x)
        nb = rows * cols * 8
        cases = [
            ("forward_layer (K-A)", lambda: ops.forward_layer(Wp, cols, xp), nb),
            ("backward only (K-C)", lambda: ops.backward_update(Wp, cols, d, xp, 0), nb),
            ("backward+BP update (K-C+K-D)", lambda: ops.backward_update(Wp, cols, d, xp, 1, 1e-9), 2 * nb),
            ("backward+BPM update (K-C+K-E)", lambda: ops.backward_update(Wp, cols, d, xp, 2, 1e-9, 0.2, Mp), 4 * nb),
            ("update+forward BP (K-D)", lambda: ops.update_forward(Wp, cols, d, xp, 1, 1e-9), 2 * nb),
            ("update+forward BPM (K-E)", lambda: ops.update_forward(Wp, cols, d, xp, 2, 1e-9, 0.2, Mp), 4 * nb),
        ]
        for name, fn, bytes_ in cases:
            ms = timeit(fn)
            gbs = bytes_ / ms / 1e6
            out.append({"kernel": name, "rows": rows, "cols": cols, "ms": round(ms, 4), "GB/s": round(gbs, 1),
                        "frac_of_measured_hbm": round(gbs / peak, 3)})
            print(out[-1], flush=True)
    json.dump(out, open(os.path.join("gpurun_out", "op_bw.json"), "w"), indent=1)


if __name__ == "__main__":
    os.makedirs("gpurun_out", exist_ok=True)
    main()
