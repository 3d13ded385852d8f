"""Sample-parallel gradient merge (extension, BASELINE.json "MPI sample-parallel gradient merge"):
merged BP steps per second for the four ways the per-rank updates can meet (hpnn_b200/parallel/dp.py).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29721 bench/dp_merge.py [config] [steps] [modes]
    python bench/dp_merge.py wide_ann 20            # one GPU

One JSON line per mode: samples/s = world * steps/s (device-timed, max over ranks), the bytes every rank
puts on NVLink per step and the roofline fraction of the merge: slower of (replica bytes streamed by the
step / measured HBM copy bandwidth) and (NVLink bytes / 770 GB/s measured peer bandwidth)."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpnn_b200 as hp  # noqa: E402
from hpnn_b200.parallel.dist import init_world  # noqa: E402
from hpnn_b200.parallel.dp import DataParallelTrainer  # noqa: E402

CFG = {
    "wide_ann": (hp.TYPE_ANN, 784, [8192, 8192, 8192, 8192], 10),
    "snn_mnist": (hp.TYPE_SNN, 784, [1000, 1000], 10),
}
HBM_GBS, NVLINK_GBS = 6539.5, 770.0  # MEASURED_PEAKS.json copy bandwidth; measured peer copy (profiling recipe)


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "wide_ann"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    modes = (sys.argv[3] if len(sys.argv) > 3 else "factored,nvls,fused,nccl").split(",")
    group = init_world("nccl")
    rank = dist.get_rank() if group is not None else 0
    world = dist.get_world_size() if group is not None else 1
    nn_type, n_in, hid, n_out = CFG[name]
    k, _ = hp.generate_kernel(10958, n_in, hid, n_out)
    rng = np.random.default_rng(100 + rank)
    X = rng.random((steps + 3, n_in))
    T = np.full((steps + 3, n_out), 0.0 if nn_type == hp.TYPE_SNN else -1.0)
    T[np.arange(steps + 3), rng.integers(n_out, size=steps + 3)] = 1.0
    wbytes = 8 * sum(a * b for a, b in zip([n_in, *hid], [*hid, n_out]))
    fbytes = 8 * sum(a + b for a, b in zip([n_in, *hid], [*hid, n_out]))
    for mode in modes:
        try:
            tr = DataParallelTrainer(k, nn_type, lr=0.001, group=group, merge=mode)
            if mode == "nvls" and not (tr.grads.mc_ptr and tr._wbuf.mc_ptr):
                raise RuntimeError("no NVLS multicast mapping")
        except Exception as e:  # a mode that cannot run here is reported, not fatal
            if rank == 0:
                print(json.dumps({"bench": "dp_merge", "config": name, "merge": mode, "n_gpus": world,
                                  "unavailable": str(e)[:200]}), flush=True)
            continue
        for i in range(3):
            tr.step(X[i], T[i], want_error=False)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(3, steps + 3):
            tr.step(X[i], T[i], want_error=False)
        b.record()
        torch.cuda.synchronize()
        if mode == "factored":
            tr.check()
        ms = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        per_step = float(ms.item()) / steps
        # replica bytes streamed per step: forward + backward read the weights once each; the factored
        # apply reads and writes them once; the matrix forms also write and read a gradient matrix
        hbm = wbytes * (4 if mode == "factored" else 7)
        nvl = fbytes * (1 if mode == "factored" else 0) + (0 if mode == "factored" else wbytes * max(world - 1, 0) // (1 if mode != "nvls" else world) * (2 if mode == "nvls" else 1))
        floor_ms = max(hbm / (HBM_GBS * 1e6), nvl / (NVLINK_GBS * 1e6))
        if rank == 0:
            print(json.dumps({"bench": "dp_merge", "config": name, "merge": mode, "n_gpus": world, "steps": steps,
                              "ms_per_step": round(per_step, 4), "samples_per_s": round(world * 1e3 / per_step, 1),
                              "nvlink_bytes_per_rank_per_step": int(nvl), "hbm_bytes_per_rank_per_step": int(hbm),
                              "roofline_ms": round(floor_ms, 4), "fraction_of_roofline": round(floor_ms / per_step, 3),
                              "multicast": bool(getattr(tr, "fbuf", None) and tr.fbuf.mc_ptr) if mode == "factored" else None}),
                  flush=True)
        tr.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
