"""Multi-rank protocols of the persistent trainers on ONE GPU: P virtual ranks inside one
cooperative launch (hpnn_dev_train_sample_vranks), checked against the CPU engine.

Covers what tests/test_gpu_multi.py can only do on a multi-GPU box: the row partition, the
all-gather / reduce-scatter by stores into the other ranks' arenas, the fence + flag barrier and
the tagged {value, tag} hand-over, for the streaming kernel (K-P) and the row-panel
deferred-update kernel (K-PL1)."""
import numpy as np
import pytest

import hpnn_b200 as hp
from hpnn_b200 import _native as N
from hpnn_b200.trainer import CpuEngine, make_engine

from conftest import make_problem

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("proto", ["barrier", "tagged"])
@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("nn_type,train", [(N.TYPE_ANN, N.TRAIN_BP), (N.TYPE_SNN, N.TRAIN_BPM)])
def test_streaming_trainer_on_virtual_ranks(monkeypatch, world, proto, nn_type, train):
    monkeypatch.setenv("HPNN_TP_PROTOCOL", proto)
    k, _ = hp.generate_kernel(21, 40, [37, 29], 6)
    X, T = make_problem(nn_type, 40, (37, 29), 6, n_samples=3, seed=4)
    kw = dict(max_iter=60, min_iter=70, delta=1e-300)
    eng = make_engine(f"virtual:{world}", k, nn_type, train, **kw)
    cpu = CpuEngine(k, nn_type, train, cuda_defaults=True, **kw)
    for x, t in zip(X, T):
        a, b = eng.train(x, t), cpu.train(x, t)
        assert a.n_iter == b.n_iter == 61
        assert abs(a.init_err - b.init_err) < 1e-12 and abs(a.last_err - b.last_err) < 1e-12
    for wa, wb in zip(eng.kernel().weights, cpu.kernel().weights):
        np.testing.assert_allclose(wa, wb, atol=1e-12)
    np.testing.assert_allclose(eng.run(X[0]), cpu.run(X[0]), atol=1e-12)
    eng.close()
    cpu.close()


@pytest.mark.parametrize("train", [N.TRAIN_BP, N.TRAIN_BPM])
@pytest.mark.parametrize("proto", ["barrier", "tagged", "barrier-slicewise"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_row_panel_trainer_on_virtual_ranks(monkeypatch, world, proto, train):
    """K-PL1 with several ranks: per-rank column-slice sums scattered to the owning rank, forward
    results and partial dot products stored into every rank's arena; 19 iterations = two
    write-back periods of 8 plus a partial one."""
    # "barrier-slicewise": the opt-in forward pass that waits per producer rank (HPNN_SLICEWISE is read once
    # per process, so the variant is selected through the setter)
    monkeypatch.setenv("HPNN_TP_PROTOCOL", proto.split("-")[0])
    lib = N.load()
    lib.hpnn_dev_set_slicewise(1 if proto.endswith("slicewise") else 0)
    lib.hpnn_dev_set_lazy(8, 0)
    lib.hpnn_dev_set_lazy_panel(1)
    try:
        # the second shape has a 2600-wide input to its second layer: three 1024-column chunks in the
        # slice-by-slice forward pass, whose columns come from several producer ranks each
        for nn_type, (n_in, hid, n_out) in ((N.TYPE_ANN, (40, (600, 300, 150), 12)), (N.TYPE_SNN, (40, (600, 300, 150), 12)),
                                            (N.TYPE_ANN, (24, (2600, 40), 10))):
            k, _ = hp.generate_kernel(77, n_in, list(hid), n_out)
            X, T = make_problem(nn_type, n_in, hid, n_out, n_samples=2, seed=3)
            n_it = 19
            kw = dict(max_iter=n_it - 1, min_iter=n_it + 10, delta=1e-300)
            eng = make_engine(f"virtual:{world}", k, nn_type, train, **kw)
            assert eng.nets[0].net.hist_depth == 8 and eng.nets[0].net.panel_stride > 0
            cpu = CpuEngine(k, nn_type, train, cuda_defaults=True, **kw)
            for x, t in zip(X, T):
                a, b = eng.train(x, t), cpu.train(x, t)
                assert a.n_iter == b.n_iter == n_it
                assert abs(a.init_err - b.init_err) < 1e-12 and abs(a.last_err - b.last_err) < 1e-11
            for wa, wb in zip(eng.kernel().weights, cpu.kernel().weights):
                np.testing.assert_allclose(wa, wb, atol=1e-11)
            eng.close()
            cpu.close()
    finally:
        lib.hpnn_dev_set_slicewise(-1)
        lib.hpnn_dev_set_lazy(-1, -1)
        lib.hpnn_dev_set_lazy_panel(-1)


def test_convergent_loop_on_virtual_ranks_takes_the_same_exit():
    """The data-dependent exit (dE <= 1e-6 and argmax == target) is taken redundantly by every
    rank; VirtualRanksEngine.train raises if they disagree."""
    k, _ = hp.generate_kernel(4242, 64, [48], 10)
    X, T = make_problem(N.TYPE_ANN, 64, (48,), 10, n_samples=2, seed=5)
    eng = make_engine("virtual:4", k, N.TYPE_ANN, N.TRAIN_BP, max_iter=3000)
    cpu = CpuEngine(k, N.TYPE_ANN, N.TRAIN_BP, cuda_defaults=True, max_iter=3000)
    for x, t in zip(X, T):
        a, b = eng.train(x, t), cpu.train(x, t)
        assert abs(a.n_iter - b.n_iter) <= max(3, b.n_iter // 100) and a.success == b.success
    eng.close()
    cpu.close()


@pytest.mark.parametrize("proto", ["barrier", "tagged"])
def test_missing_rank_is_reported_not_waited_for(monkeypatch, proto):
    """Fault injection: one rank of four never shows up (its CTAs leave at kernel entry, as after a
    failed launch).  Every other rank must give up within the spin time-out and say so in
    hpnn_train_result.status; the call returns an error in well under two seconds, and the same
    engine works again afterwards."""
    import time

    monkeypatch.setenv("HPNN_TP_PROTOCOL", proto)
    lib = N.load()
    k, _ = hp.generate_kernel(21, 40, [37, 29], 6)
    X, T = make_problem(N.TYPE_ANN, 40, (37, 29), 6, n_samples=1, seed=4)
    kw = dict(max_iter=60, min_iter=70, delta=1e-300)
    eng = make_engine("virtual:4", k, N.TYPE_ANN, N.TRAIN_BP, **kw)
    eng.train(X[0], T[0])  # warm: module load, first launch
    lib.hpnn_dev_set_spin_timeout_ms(200)
    lib.hpnn_dev_set_fault(1, 2)
    try:
        t0 = time.perf_counter()
        with pytest.raises(RuntimeError, match="abandoned"):
            eng.train(X[0], T[0])
        assert time.perf_counter() - t0 < 2.0
        # ranks 0, 1 and 3 ran and gave up; rank 2 never wrote its result slot
        assert [eng.last_status[r] for r in (0, 1, 3)] == [1, 1, 1]
    finally:
        lib.hpnn_dev_set_fault(0, -1)
        lib.hpnn_dev_set_spin_timeout_ms(-1)
    # the engine is usable again (weights were garbage-updated by the abandoned sample: reload them)
    for n in eng.nets:
        n.load_weights(k)
    cpu = CpuEngine(k, N.TYPE_ANN, N.TRAIN_BP, cuda_defaults=True, **kw)
    a, b = eng.train(X[0], T[0]), cpu.train(X[0], T[0])
    assert a.n_iter == b.n_iter and abs(a.last_err - b.last_err) < 1e-12
    eng.close()
    cpu.close()


@pytest.mark.parametrize("backend,proto,panel", [("virtual:4", "barrier", False), ("virtual:4", "tagged", False),
                                                 ("virtual:4", "barrier", True), ("virtual:3", "tagged", True),
                                                 ("cuda", "barrier", False)])
def test_schedule_perturbation_changes_no_bit(monkeypatch, backend, proto, panel):
    """The stand-in for a race detector (compute-sanitizer is closed on the pool): with pseudo-random delays of up to
    3 us before every barrier arrival, cross-rank signal and tagged store (hpnn_dev_set_jitter_ns) the trainers
    must reproduce the undisturbed run bit for bit -- all summation orders are fixed, so any difference would be
    a hand-over that only works because the other side happens to be later.  Virtual ranks: K-P with both
    protocols, the row-panel trainer K-PL1 with both; "cuda": the shared-memory-resident K-PR (tagged pairs only)."""
    monkeypatch.setenv("HPNN_TP_PROTOCOL", proto)
    lib = N.load()
    if panel:
        lib.hpnn_dev_set_lazy(8, 0)
        lib.hpnn_dev_set_lazy_panel(1)
    try:
        dims = (40, (600, 300, 150), 12) if panel else (40, (37, 29), 6)
        n_it = 21 if panel else 45
        kw = dict(max_iter=n_it - 1, min_iter=n_it + 10, delta=1e-300)
        runs = []
        for jitter in (0, 3000, 700):
            lib.hpnn_dev_set_jitter_ns(jitter)
            k, _ = hp.generate_kernel(9, dims[0], list(dims[1]), dims[2])
            X, T = make_problem(N.TYPE_SNN, dims[0], dims[1], dims[2], n_samples=2, seed=8)
            eng = make_engine(backend, k, N.TYPE_SNN, N.TRAIN_BPM if not panel else N.TRAIN_BP, **kw)
            res = [eng.train(x, t) for x, t in zip(X, T)]
            runs.append(([r.n_iter for r in res], [r.last_err for r in res], [w.copy() for w in eng.kernel().weights]))
            eng.close()
        for iters, errs, ws in runs[1:]:
            assert iters == runs[0][0] == [n_it, n_it]
            assert errs == runs[0][1]
            for a, b in zip(ws, runs[0][2]):
                assert np.array_equal(a, b)
    finally:
        lib.hpnn_dev_set_jitter_ns(0)
        lib.hpnn_dev_set_lazy(-1, -1)
        lib.hpnn_dev_set_lazy_panel(-1)
