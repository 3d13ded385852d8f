"""Multi-GPU check of the fused row-split trainer; run under torchrun, one rank per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tests/mp_tp_check.py

Every rank trains the same samples with K-P over NVLink peer stores; rank 0 compares N_ITER,
errors and the gathered weights with the CPU engine, and with the NCCL-collective baseline."""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import hpnn_b200 as hp  # noqa: E402
from hpnn_b200 import _native as N  # noqa: E402
from hpnn_b200.parallel.dist import init_world  # noqa: E402
from hpnn_b200.parallel.tp_collective import CollectiveTPEngine  # noqa: E402
from hpnn_b200.trainer import CpuEngine, CudaEngine  # noqa: E402


def problem(nn_type, n_in, n_out, n, seed):
    rng = np.random.default_rng(seed)
    lo = 0.0 if nn_type == N.TYPE_SNN else -1.0
    X = rng.random((n, n_in))
    T = np.full((n, n_out), lo)
    for i in range(n):
        T[i, int(rng.integers(n_out))] = 1.0
    return X, T


def main():
    group = init_world("nccl")
    rank, world = dist.get_rank(), dist.get_world_size()
    fails = 0
    cases = [(N.TYPE_ANN, N.TRAIN_BP, 64, [48, 40], 10), (N.TYPE_SNN, N.TRAIN_BPM, 100, [37, 29, 23], 7),
             (N.TYPE_SNN, N.TRAIN_BPM, 784, [300, 200], 10), (N.TYPE_ANN, N.TRAIN_BPM, 33, [5], 3)]
    for nn_type, train, n_in, hid, n_out in cases:
        k, _ = hp.generate_kernel(77, n_in, hid, n_out)
        X, T = problem(nn_type, n_in, n_out, 3, 11)
        kw = dict(max_iter=400)
        fused = CudaEngine(k, nn_type, train, group=group, split="always", **kw)
        if rank == 0 and n_in == 64:
            sy = fused.net._symm
            print(f"case symmetric-memory backend={sy.backend} multicast={'yes' if sy.mc_ptr else 'no'} world={world}", flush=True)
        coll = CollectiveTPEngine(k, nn_type, train, group=group, device=f"cuda:{torch.cuda.current_device()}", **kw)
        cpu = CpuEngine(k, nn_type, train, cuda_defaults=True, **kw) if rank == 0 else None
        t_f = t_c = 0.0
        for x, t in zip(X, T):
            torch.cuda.synchronize(); dist.barrier(); t0 = time.perf_counter()
            rf = fused.train(x, t)
            torch.cuda.synchronize(); t_f += time.perf_counter() - t0
            dist.barrier(); t0 = time.perf_counter()
            rc = coll.train(x, t)
            torch.cuda.synchronize(); t_c += time.perf_counter() - t0
            if rank == 0:
                r0 = cpu.train(x, t)
                ok = (abs(rf.n_iter - r0.n_iter) <= max(3, r0.n_iter // 100) and abs(rf.init_err - r0.init_err) < 1e-11
                      and rf.first_ok == r0.first_ok and rf.success == r0.success
                      and abs(rc.n_iter - r0.n_iter) <= max(3, r0.n_iter // 100))
                if not ok:
                    fails += 1
                    print(f"MISMATCH type={nn_type} train={train}: fused={rf} coll={rc} cpu={r0}", flush=True)
        kf, kc = fused.kernel(), coll.kernel()
        of = fused.run(X[0])
        if rank == 0:
            k0 = cpu.kernel()
            e_f = max(float(np.abs(a - b).max()) for a, b in zip(kf.weights, k0.weights))
            e_c = max(float(np.abs(a - b).max()) for a, b in zip(kc.weights, k0.weights))
            e_o = float(np.abs(of - cpu.run(X[0])).max())
            print(f"case type={nn_type} train={train} {n_in}-{hid}-{n_out} world={world}: max|dW| fused={e_f:.2e} "
                  f"collective={e_c:.2e} max|do|={e_o:.2e}  time fused={t_f * 1e3:.1f} ms collective={t_c * 1e3:.1f} ms",
                  flush=True)
            if not (e_f < 1e-6 and e_c < 1e-6 and e_o < 1e-6):
                fails += 1
        fused.close()
    # ---- deferred-update trainer (train_lazy.cu) on the row-split net: fixed iteration counts,
    # weights must match the CPU engine to round-off
    lib = N.load()
    os.environ["HPNN_TP_PROTOCOL"] = "barrier"  # the variant uses the fence + flag barrier protocol
    lib.hpnn_dev_set_lazy(8, 0)
    # (panel, protocol): row panels (train_lazy1.cu) with fence + flag barriers or tagged pairs,
    # column tasks (train_lazy.cu, barrier protocol only)
    for panel, proto in ((1, "barrier"), (1, "tagged"), (0, "barrier")):
        lib.hpnn_dev_set_lazy_panel(panel)
        os.environ["HPNN_TP_PROTOCOL"] = proto
        for nn_type in (N.TYPE_ANN, N.TYPE_SNN):
            k, _ = hp.generate_kernel(31, 72, [520, 264, 130], 10)
            X, T = problem(nn_type, 72, 10, 2, 17)
            for n_it in (5, 19):
                kw = dict(max_iter=n_it - 1, min_iter=n_it + 10, delta=1e-300)
                lazy = CudaEngine(k, nn_type, N.TRAIN_BP, group=group, split="always", **kw)
                assert lazy.net.net.hist_depth == 8
                cpu = CpuEngine(k, nn_type, N.TRAIN_BP, cuda_defaults=True, **kw) if rank == 0 else None
                e_err = 0.0
                for x, t in zip(X, T):
                    rl = lazy.train(x, t)
                    if rank == 0:
                        r0 = cpu.train(x, t)
                        e_err = max(e_err, abs(rl.last_err - r0.last_err), float(abs(rl.n_iter - r0.n_iter)))
                kl = lazy.kernel()
                if rank == 0:
                    e_w = max(float(np.abs(a - b).max()) for a, b in zip(kl.weights, cpu.kernel().weights))
                    print(f"case deferred-update panel={panel} {proto} type={nn_type} iters={n_it} world={world}: "
                          f"max|dW|={e_w:.2e} max|dE|={e_err:.2e}", flush=True)
                    if not (e_w < 1e-11 and e_err < 1e-11):
                        fails += 1
                lazy.close()
    lib.hpnn_dev_set_lazy_panel(-1)
    lib.hpnn_dev_set_lazy(-1, -1)
    del os.environ["HPNN_TP_PROTOCOL"]
    # ---- data-parallel gradient merge (extension): fused peer-load merge == NCCL all_reduce == serial sum
    from hpnn_b200.parallel.dp import DataParallelTrainer
    from hpnn_b200.ops import reference as R

    k, _ = hp.generate_kernel(5, 50, [30, 20], 6)
    X, T = problem(N.TYPE_ANN, 50, 6, world * 3, 3)
    modes = ["factored", "fused", "nccl"]
    probe = DataParallelTrainer(k, N.TYPE_ANN, lr=0.01, group=group, merge="factored")
    have_mc = bool(probe.fbuf.mc_ptr)
    probe.close()
    if have_mc:
        modes.append("nvls")
    trainers = {m: DataParallelTrainer(k, N.TYPE_ANN, lr=0.01, group=group, merge=m) for m in modes}
    Wser = [torch.from_numpy(w.copy()).cuda() for w in k.weights]
    for s in range(3):
        for tr in trainers.values():
            tr.step(X[s * world + rank], T[s * world + rank])
        upd = [torch.zeros_like(w) for w in Wser]
        for r in range(world):  # serial reference: sum of the per-sample updates from the same weights
            xs = torch.from_numpy(X[s * world + r]).cuda(); ts = torch.from_numpy(T[s * world + r]).cuda()
            acts = R.forward(Wser, xs, N.TYPE_ANN); d = R.deltas(Wser, acts, ts, N.TYPE_ANN)
            for l in range(len(Wser)):
                upd[l] += torch.outer(0.01 * d[l], acts[l])
        for l in range(len(Wser)):
            Wser[l] += upd[l]
    trainers["factored"].check()
    errs = {m: max(float((x - y).abs().max()) for x, y in zip(tr.W, Wser)) for m, tr in trainers.items()}
    # replicas must be bit-identical across ranks (no broadcast ever happens)
    flat = trainers["factored"].flat.clone()
    ref0 = flat.clone()
    dist.broadcast(ref0, src=0)
    same = float((flat - ref0).abs().max())
    if rank == 0:
        print(f"case dp-merge world={world} multicast={'yes' if have_mc else 'no'}: " +
              " ".join(f"max|{m}-serial|={e:.2e}" for m, e in errs.items()) + f" replica spread={same:.1e}", flush=True)
    if not (max(errs.values()) < 1e-13 and same == 0.0):
        fails += 1
    for tr in trainers.values():
        tr.close()
    # ---- batched inference on the row-split net: fused GEMM+all-gather == NCCL == single-rank
    from hpnn_b200.models.net import HpnnNet
    from hpnn_b200.parallel.tp import TPBatchRunner, forward_batch_parallel

    for nn_type in (N.TYPE_ANN, N.TYPE_SNN):
        k, _ = hp.generate_kernel(9, 101, [67, 45], 11)
        Xb = torch.from_numpy(np.random.default_rng(4).random((300, 101))).cuda()
        tpnet = HpnnNet(k, nn_type, momentum=False, group=group)
        full = HpnnNet(k, nn_type, momentum=False)
        runner = TPBatchRunner(tpnet, max_batch=512)
        ref_out = full.forward_batch(Xb)
        yf = runner.forward(Xb, "fused")
        yn = runner.forward(Xb, "nccl")
        yf2 = runner.forward(Xb, "fused")
        yb = forward_batch_parallel(full, Xb, group=group)
        errs = [float((y - ref_out).abs().max()) for y in (yf, yn, yf2, yb)]
        if rank == 0:
            print(f"case tp-batch type={nn_type} world={world}: max|fused-ref|={errs[0]:.2e} nccl={errs[1]:.2e} "
                  f"fused again={errs[2]:.2e} batch-parallel={errs[3]:.2e}", flush=True)
        if max(errs) > 1e-12:
            fails += 1
        runner.close(); tpnet.close(); full.close()
    flag = torch.tensor([fails], device="cuda")
    dist.all_reduce(flag)
    if rank == 0:
        print("TP_CHECK_OK" if int(flag.item()) == 0 else "TP_CHECK_FAILED", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 0 else 1)


if __name__ == "__main__":
    main()
