"""One torchrun session on an N-GPU box that measures the row-split trainer at several world sizes and
hand-over protocols (sub-groups of the ranks; ranks outside a sub-group idle), with the per-phase
cycle trace of rank 0's CTA 0.  Per-iteration device time, max over the ranks of the sub-group.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29711 bench/scale_probe.py [config] [iterations] [worlds e.g. 8,4,2] [protocols]

HPNN_TRACE_NS=1: the phase stamps are nanoseconds (%globaltimer) instead of SM cycles (the SM clock follows the power
draw, so cycle counts of phases with different load are not comparable as times).
"""
import json
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

CFG = {
    "wide_ann": (hp.TYPE_ANN, hp.TRAIN_BP, 784, [8192, 8192, 8192, 8192], 10),
    "wide_ann_bpm": (hp.TYPE_ANN, hp.TRAIN_BPM, 784, [8192, 8192, 8192, 8192], 10),
    # one GPU's view of the 8-rank job: 1024-row layers of 8192 columns = 7-row panels per CTA, no communication
    "short_panels": (hp.TYPE_ANN, hp.TRAIN_BP, 784, [8192, 1024, 8192, 1024], 10),
    "short_panels6": (hp.TYPE_ANN, hp.TRAIN_BP, 784, [8192, 1024, 8192, 1024, 8192, 1024], 10),
    "wide6": (hp.TYPE_ANN, hp.TRAIN_BP, 784, [4096] * 6, 10),
    "snn_mnist": (hp.TYPE_SNN, hp.TRAIN_BPM, 784, [1000, 1000], 10),
    "xrd_snn": (hp.TYPE_SNN, hp.TRAIN_BPM, 851, [230], 230),
}


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "wide_ann"
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 200
    worlds = [int(w) for w in (sys.argv[3] if len(sys.argv) > 3 else "8,4,2").split(",")]
    protos = (sys.argv[4] if len(sys.argv) > 4 else "tagged,barrier").split(",")
    init_world("nccl")
    rank, total = dist.get_rank(), dist.get_world_size()
    nn_type, train, n_in, hid, n_out = CFG[name]
    k, _ = hp.generate_kernel(10958, n_in, hid, n_out)
    rng = np.random.default_rng(0)
    x = rng.random(n_in)
    t = np.full(n_out, 0.0 if nn_type == hp.TYPE_SNN else -1.0)
    t[3 % n_out] = 1.0
    lib = N.load()
    for world in worlds:
        if world > total:
            continue
        group = dist.new_group(list(range(world)))  # collective over ALL ranks
        if rank < world:
            eng = CudaEngine(k, nn_type, train, group=group if world > 1 else None, split="always", max_iter=iters, min_iter=iters + 10,
                             delta=1e-300)
            for proto in protos if world > 1 else ["single"]:
                if world > 1:
                    os.environ["HPNN_TP_PROTOCOL"] = proto
                eng.net.load_weights(k)
                eng.train(x, t)  # warm
                eng.net.load_weights(k)
                cap = 64 * (iters + 2)
                buf = torch.zeros(2 * cap, dtype=torch.int64, device="cuda")
                lib.hpnn_dev_set_phase_trace(buf.data_ptr(), cap)
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                if world > 1:
                    dist.barrier(group=group)
                a.record()
                r = eng.train(x, t)
                b.record()
                torch.cuda.synchronize()
                lib.hpnn_dev_set_phase_trace(None, 0)
                ms = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device="cuda")
                if world > 1:
                    dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=group)
                tr = buf.cpu().numpy().reshape(-1, 2)
                tr = tr[tr[:, 0] != 0]
                starts = np.where(tr[:, 0] == 1)[0]
                seg = defaultdict(list)
                for s0, s1 in zip(starts[3:-1], starts[4:]):
                    for j in range(s0 + 1, s1 + 1):
                        seg[(int(tr[j - 1, 0]), int(tr[j, 0]))].append(int(tr[j, 1] - tr[j - 1, 1]))
                phases = {f"{a_}->{b_}": round(float(np.mean(v))) for (a_, b_), v in seg.items()}
                by_rank = [phases]
                if world > 1:  # CTA 0 of EVERY rank: who waits for whom
                    by_rank = [None] * world
                    dist.all_gather_object(by_rank, phases, group=group)
                if rank == 0:
                    us = float(ms.item()) * 1e3 / r.n_iter
                    print(json.dumps({"config": name, "world": world, "protocol": proto, "iterations": r.n_iter,
                                      "us_per_iteration": round(us, 2), "iterations_per_s": round(1e6 / us, 1),
                                      "last_err": r.last_err, "phase_cycles": phases,
                                      "phase_cycles_by_rank": by_rank}), flush=True)
            eng.close()
        dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
