"""Multi-process host logic on CPU (gloo, world_size 2 and 3): the collective row split must
reproduce the single-rank result, and kernels must broadcast intact."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import hpnn_b200 as hp
from hpnn_b200 import _native as N

from conftest import make_problem


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, nn_type, train, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from hpnn_b200.parallel.dist import broadcast_kernel
    from hpnn_b200.parallel.tp_collective import CollectiveTPEngine

    k = hp.generate_kernel(31, 12, [10, 7], 5)[0] if rank == 0 else None
    k = broadcast_kernel(k)
    X, T = make_problem(nn_type, 12, (10, 7), 5, n_samples=2, seed=2)
    eng = CollectiveTPEngine(k, nn_type, train, max_iter=150)
    res = [eng.train(x, t) for x, t in zip(X, T)]
    out = eng.run(X[0])
    full = eng.kernel()
    if rank == 0:
        q.put(([(r.n_iter, r.init_err, r.final_de, r.first_ok, r.success) for r in res], out,
               [w.copy() for w in full.weights], eng.n_collectives))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("nn_type,train", [(N.TYPE_ANN, N.TRAIN_BP), (N.TYPE_SNN, N.TRAIN_BPM)])
def test_collective_row_split_equals_single_rank(world, nn_type, train):
    from hpnn_b200.trainer import TorchEngine

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, nn_type, train, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    res, out, weights, n_coll = got
    k = hp.generate_kernel(31, 12, [10, 7], 5)[0]
    X, T = make_problem(nn_type, 12, (10, 7), 5, n_samples=2, seed=2)
    ref = TorchEngine(k, nn_type, train, max_iter=150, cuda_defaults=True)
    for (n_iter, init, final, first_ok, success), (x, t) in zip(res, zip(X, T)):
        r = ref.train(x, t)
        assert n_iter == r.n_iter and first_ok == r.first_ok and success == r.success
        assert abs(init - r.init_err) < 1e-13 and abs(final - r.final_de) < 1e-12
    for a, b in zip(weights, ref.kernel().weights):
        np.testing.assert_allclose(a, b, atol=1e-12)
    np.testing.assert_allclose(out, ref.run(X[0]), atol=1e-12)
    assert n_coll > 0


def test_split_rows_is_a_partition():
    from hpnn_b200.models.net import split_rows

    for n in (1, 7, 10, 230, 1000, 8192):
        for world in (1, 2, 3, 4, 8):
            parts = [split_rows(n, g, world) for g in range(world)]
            assert parts[0][0] == 0 and sum(nr for _, nr in parts) == n
            assert all(parts[g][0] + parts[g][1] == parts[g + 1][0] for g in range(world - 1))
            assert max(nr for _, nr in parts) - min(nr for _, nr in parts) <= 1


def _fd_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from hpnn_b200.parallel.symm import _all_ok, _pass_fds

    import tempfile

    # every rank owns an (unlinked) file; its descriptor travels to every peer
    f = tempfile.TemporaryFile()
    f.write(b"from%d" % rank)
    f.flush()
    got = _pass_fds(f.fileno(), list(range(world)), rank, world, None)
    seen = {src: os.pread(fd, 16, 0).decode() for src, fd in got.items()}
    # one-to-all variant (how the multicast object of rank 0 is shared)
    f2 = tempfile.TemporaryFile()
    f2.write(b"mc")
    f2.flush()
    got2 = _pass_fds(f2.fileno() if rank == 0 else None, [0], rank, world, None)
    seen2 = {src: os.pread(fd, 16, 0).decode() for src, fd in got2.items()}
    ok = _all_ok(rank != 1, None)
    q.put((rank, seen, seen2, ok))
    dist.barrier()
    dist.destroy_process_group()


def test_file_descriptors_cross_processes():
    """The plumbing of the VMM symmetric memory (hpnn_b200/parallel/symm.py): POSIX file descriptors
    are handed from every rank to every other rank (allocation handles) and from rank 0 to all
    (multicast object) over unix sockets."""
    world = 3
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_fd_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, seen, seen2, ok in res:
        assert seen == {s: "from%d" % s for s in range(world) if s != rank}
        assert seen2 == ({} if rank == 0 else {0: "mc"})
        assert ok is False  # rank 1 reported a failure: everybody must see it
