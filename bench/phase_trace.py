"""Per-phase timing of the persistent trainer from its on-device trace (clock64 of CTA 0)."""
import os
import sys
from collections import defaultdict

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpnn_b200 as hp  # noqa: E402
from hpnn_b200 import _native as N  # noqa: E402
from hpnn_b200.trainer import CudaEngine  # noqa: E402

CFG = {"snn_mnist": (hp.TYPE_SNN, hp.TRAIN_BPM, 784, [1000, 1000], 10),
       "ann_mnist": (hp.TYPE_ANN, hp.TRAIN_BP, 784, [300], 10),
       "xrd_snn": (hp.TYPE_SNN, hp.TRAIN_BPM, 851, [230], 230),
       "tiny": (hp.TYPE_ANN, hp.TRAIN_BP, 8, [8, 8], 8),
       "wide_ann": (hp.TYPE_ANN, hp.TRAIN_BP, 784, [8192] * 4, 10)}


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "snn_mnist"
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 200
    nn_type, train, n_in, hid, n_out = CFG[name]
    k, _ = hp.generate_kernel(10958, n_in, hid, n_out)
    eng = CudaEngine(k, nn_type, train, max_iter=iters, min_iter=iters + 10, delta=1e-300)
    rng = np.random.default_rng(0)
    x = rng.random(n_in)
    t = np.full(n_out, 0.0 if nn_type == hp.TYPE_SNN else -1.0)
    t[3 % n_out] = 1.0
    eng.train(x, t)  # warm
    cap = 64 * (iters + 2)
    buf = torch.zeros(2 * cap, dtype=torch.int64, device="cuda")
    lib = N.load()
    lib.hpnn_dev_set_phase_trace(buf.data_ptr(), cap)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    r = eng.train(x, t)
    b.record()
    torch.cuda.synchronize()
    lib.hpnn_dev_set_phase_trace(None, 0)
    ms = a.elapsed_time(b)
    tr = buf.cpu().numpy().reshape(-1, 2)
    tr = tr[tr[:, 0] != 0]
    print(f"{name}: {r.n_iter} iterations in {ms:.3f} ms -> {ms * 1e3 / r.n_iter:.2f} us/iteration")
    # skip the first 10 iterations (cold L2), average the cycle deltas per (previous tag -> tag)
    starts = np.where(tr[:, 0] == 1)[0]
    seg = defaultdict(list)
    for s0, s1 in zip(starts[10:-1], starts[11:]):
        for j in range(s0 + 1, s1 + 1):
            seg[(int(tr[j - 1, 0]), int(tr[j, 0]))].append(int(tr[j, 1] - tr[j - 1, 1]))
    tot = 0.0
    for (a_, b_), v in seg.items():
        m = float(np.mean(v))
        tot += m
        print(f"  {a_:4d} -> {b_:4d}: {m:9.0f} cycles   (min {min(v)}, max {max(v)})")
    print(f"  total {tot:9.0f} cycles/iteration (SM clock)")
    eng.close()


if __name__ == "__main__":
    main()
