"""Per-phase cycles of K-P on a row-split net (run under torchrun; rank 0 prints its CTA-0 trace)."""
import os
import sys
from collections import defaultdict

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpnn_b200 as hp  # noqa: E402
from hpnn_b200 import _native as N  # noqa: E402
from hpnn_b200.parallel.dist import init_world  # noqa: E402
from hpnn_b200.trainer import CudaEngine  # noqa: E402
from phase_trace import CFG  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "wide_ann"
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    group = init_world("nccl")
    rank, world = dist.get_rank(), dist.get_world_size()
    nn_type, train, n_in, hid, n_out = CFG[name]
    k, _ = hp.generate_kernel(10958, n_in, hid, n_out)
    eng = CudaEngine(k, nn_type, train, group=group, split="always", max_iter=iters, min_iter=iters + 10, delta=1e-300)
    rng = np.random.default_rng(0)
    x = rng.random(n_in)
    t = np.full(n_out, 0.0 if nn_type == hp.TYPE_SNN else -1.0)
    t[3 % n_out] = 1.0
    eng.train(x, t)
    cap = 64 * (iters + 2)
    buf = torch.zeros(2 * cap, dtype=torch.int64, device="cuda")
    lib = N.load()
    lib.hpnn_dev_set_phase_trace(buf.data_ptr(), cap)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); dist.barrier()
    a.record()
    r = eng.train(x, t)
    b.record()
    torch.cuda.synchronize()
    lib.hpnn_dev_set_phase_trace(None, 0)
    if rank == 0:
        ms = a.elapsed_time(b)
        tr = buf.cpu().numpy().reshape(-1, 2)
        tr = tr[tr[:, 0] != 0]
        print(f"{name} world={world}: {r.n_iter} iterations in {ms:.3f} ms -> {ms * 1e3 / r.n_iter:.2f} us/iteration")
        starts = np.where(tr[:, 0] == 1)[0]
        seg = defaultdict(list)
        for s0, s1 in zip(starts[3:-1], starts[4:]):
            for j in range(s0 + 1, s1 + 1):
                seg[(int(tr[j - 1, 0]), int(tr[j, 0]))].append(int(tr[j, 1] - tr[j - 1, 1]))
        tot = 0.0
        for (a_, b_), v in seg.items():
            m = float(np.mean(v)); tot += m
            print(f"  {a_:4d} -> {b_:4d}: {m:9.0f} cycles")
        print(f"  total {tot:9.0f} cycles/iteration")
    eng.close()
    dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
