"""run_nn inference sweep on N GPUs (BASELINE.json config 5), under torchrun, one rank per GPU:
SNN 784-1000-1000-10, FP64, batch 1..4096, samples/s (device-timed, max over ranks) of
  fused  row-split net, GEMM epilogue stores into every rank's activation matrix (GEMM + all-gather),
  nccl   same partition, dist.all_gather after every layer (the baseline),
  batch  full replica per rank, the batch is split (no communication until the final gather),
  kp     batch 1 only: the persistent per-sample kernel on the row-split net (run mode)."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpnn_b200 as hp  # noqa: E402
from hpnn_b200.models.net import HpnnNet  # noqa: E402
from hpnn_b200.parallel.dist import init_world  # noqa: E402
from hpnn_b200.parallel.tp import TPBatchRunner, forward_batch_parallel  # noqa: E402


def timed(fn, n, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / n], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    group = init_world("nccl")
    rank, world = dist.get_rank(), dist.get_world_size()
    k, _ = hp.generate_kernel(10958, 784, [1000, 1000], 10)
    tpnet = HpnnNet(k, hp.TYPE_SNN, momentum=False, group=group)
    full = HpnnNet(k, hp.TYPE_SNN, momentum=False)
    runner = TPBatchRunner(tpnet, max_batch=4096)
    rng = np.random.default_rng(1)
    rows = []
    for B in (1, 64, 1024, 4096):
        X = torch.from_numpy(rng.random((B, 784))).cuda()
        n = 50 if B <= 64 else 20
        res = {"batch": B, "n_gpus": world}
        for mode in ("fused", "nccl"):
            ms = timed(lambda: runner.forward(X, mode), n)
            res[mode + "_samples_per_s"] = B / (ms * 1e-3)
        if B >= world:
            ms = timed(lambda: forward_batch_parallel(full, X, group=group), n)
            res["batch_parallel_samples_per_s"] = B / (ms * 1e-3)
        if B == 1:
            x = X[0].contiguous()
            ms = timed(lambda: tpnet.run(x), 50)
            res["kp_samples_per_s"] = 1.0 / (ms * 1e-3)
        rows.append(res)
        if rank == 0:
            print(json.dumps(res), flush=True)
    runner.close(); tpnet.close(); full.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
