"""The persistent per-sample trainer (K-P) against the CPU engine (itself bit-identical to the
reference binary, see test_cpu_engine.py) and the multi-launch path built from K-A..K-E."""
import numpy as np
import pytest
import torch

import hpnn_b200 as hp
from hpnn_b200 import _native as N
from hpnn_b200.trainer import CpuEngine, CudaEngine

from conftest import make_problem

pytestmark = pytest.mark.gpu

CASES = [(N.TYPE_ANN, N.TRAIN_BP), (N.TYPE_ANN, N.TRAIN_BPM), (N.TYPE_SNN, N.TRAIN_BP), (N.TYPE_SNN, N.TRAIN_BPM)]


@pytest.mark.parametrize("nn_type,train", CASES)
@pytest.mark.parametrize("topo", [(8, (5, 4), 3), (64, (48,), 10), (100, (37, 29, 23), 7)])
def test_persistent_trainer_matches_cpu_engine(nn_type, train, topo):
    n_in, hid, n_out = topo
    k, _ = hp.generate_kernel(4242, n_in, list(hid), n_out)
    X, T = make_problem(nn_type, n_in, hid, n_out, n_samples=3, seed=5)
    gpu = CudaEngine(k, nn_type, train, max_iter=2000)
    cpu = CpuEngine(k, nn_type, train, max_iter=2000, cuda_defaults=True)
    same_work = True
    for x, t in zip(X, T):
        rg, rc = gpu.train(x, t), cpu.train(x, t)
        assert abs(rg.init_err - rc.init_err) < 1e-11
        assert rg.first_ok == rc.first_ok and rg.success == rc.success
        # different fp64 summation orders may move the exit by a few iterations
        assert abs(rg.n_iter - rc.n_iter) <= max(3, rc.n_iter // 100)
        if rg.n_iter == rc.n_iter:
            assert abs(rg.final_de - rc.final_de) < 1e-10
        else:
            same_work = False
    # equal iteration counts on every sample = the same arithmetic up to summation order: round-off
    # only.  An exit that moved by an iteration leaves one more / one fewer update of size ~lr*delta.
    tol = 1e-11 if same_work else 1e-6
    wg, wc = gpu.kernel(), cpu.kernel()
    for a, b in zip(wg.weights, wc.weights):
        np.testing.assert_allclose(a, b, atol=tol)
    np.testing.assert_allclose(gpu.run(X[0]), cpu.run(X[0]), atol=tol)
    gpu.close()
    cpu.close()


@pytest.mark.parametrize("nn_type,train", CASES)
def test_fixed_iteration_count_is_tight(nn_type, train):
    """With min_iter == max_iter both engines do exactly the same number of iterations, so the
    weights must agree to round-off (ChangeLog:33-38 asks for < 1e-12 between back-ends)."""
    k, _ = hp.generate_kernel(99, 32, [24, 16], 5)
    X, T = make_problem(nn_type, 32, (24, 16), 5, n_samples=2, seed=9)
    kw = dict(max_iter=50, min_iter=60, delta=1e-300)
    gpu = CudaEngine(k, nn_type, train, **kw)
    cpu = CpuEngine(k, nn_type, train, cuda_defaults=True, **kw)
    for x, t in zip(X, T):
        rg, rc = gpu.train(x, t), cpu.train(x, t)
        assert rg.n_iter == rc.n_iter == 51
        assert abs(rg.last_err - rc.last_err) < 1e-12
    for a, b in zip(gpu.kernel().weights, cpu.kernel().weights):
        np.testing.assert_allclose(a, b, atol=1e-12)
    gpu.close()
    cpu.close()


def test_run_only_and_pipelined_pass():
    k, _ = hp.generate_kernel(7, 20, [16], 4)
    X, T = make_problem(N.TYPE_SNN, 20, (16,), 4, n_samples=5, seed=3)
    gpu = CudaEngine(k, N.TYPE_SNN, N.TRAIN_BPM, max_iter=300)
    cpu = CpuEngine(k, N.TYPE_SNN, N.TRAIN_BPM, max_iter=300, cuda_defaults=True)
    np.testing.assert_allclose(gpu.run(X[0]), cpu.run(X[0]), atol=1e-13)
    Xp, Tp = torch.from_numpy(X).pin_memory(), torch.from_numpy(T).pin_memory()
    res = gpu.train_many(Xp, Tp)
    ref = [cpu.train(x, t) for x, t in zip(X, T)]
    assert [abs(a.n_iter - b.n_iter) <= 3 for a, b in zip(res, ref)] == [True] * 5
    for a, b in zip(gpu.kernel().weights, cpu.kernel().weights):
        np.testing.assert_allclose(a, b, atol=1e-6)
    gpu.close()


def test_native_library_is_the_one_in_tree():
    from hpnn_b200.ops import cuda as ops

    assert N.lib_path().exists() and "hpnn_b200/lib" in str(N.lib_path())
    before = ops.launch_count()
    k, _ = hp.generate_kernel(7, 20, [16], 4)
    gpu = CudaEngine(k, N.TYPE_ANN, N.TRAIN_BP, max_iter=5)
    gpu.train(np.zeros(20), np.array([1.0, -1, -1, -1]))
    assert ops.launch_count() == before + 1  # one cooperative launch per sample
    gpu.close()


@pytest.mark.parametrize("world", [1, 3, 8])
@pytest.mark.parametrize("nn_type,train", [(N.TYPE_ANN, N.TRAIN_BP), (N.TYPE_SNN, N.TRAIN_BPM)])
def test_multi_launch_path_with_logical_ranks(world, nn_type, train):
    """K-A..K-E driven from the host over P logical ranks on one GPU (the successor of the
    reference's DEBUG fake-multi-GPU build) equals the single-rank CPU engine."""
    from hpnn_b200.trainer import make_engine

    k, _ = hp.generate_kernel(21, 40, [17, 13], 6)
    X, T = make_problem(nn_type, 40, (17, 13), 6, n_samples=2, seed=4)
    kw = dict(max_iter=60, min_iter=70, delta=1e-300)
    eng = make_engine(f"logical:{world}", k, nn_type, train, **kw)
    cpu = CpuEngine(k, nn_type, train, cuda_defaults=True, **kw)
    for x, t in zip(X, T):
        a, b = eng.train(x, t), cpu.train(x, t)
        assert a.n_iter == b.n_iter == 61 and abs(a.last_err - b.last_err) < 1e-12
    for wa, wb in zip(eng.kernel().weights, cpu.kernel().weights):
        np.testing.assert_allclose(wa, wb, atol=1e-12)
    np.testing.assert_allclose(eng.run(X[0]), cpu.run(X[0]), atol=1e-12)


@pytest.mark.parametrize("panel", [0, 1])
@pytest.mark.parametrize("depth", [2, 3, 8])
@pytest.mark.parametrize("nn_type", [N.TYPE_ANN, N.TYPE_SNN])
def test_deferred_update_trainer_matches_cpu_engine(depth, nn_type, panel):
    """K-PL (train_lazy.cu) keeps the rank-1 updates of the layers >= 1 in factored form and writes
    them back every `depth` iterations; a fixed iteration count must reproduce the CPU engine's
    weights to round-off, whatever the count is relative to the write-back period."""
    lib = N.load()
    lib.hpnn_dev_set_lazy(depth, 0)  # reserve the factor history for every shape
    lib.hpnn_dev_set_lazy_panel(panel)  # 1: row panels (train_lazy1.cu), 0: column tasks (train_lazy.cu)
    # the second shape has a layer wider than 4096 columns (streaming forward pass of the panel
    # variant, with a partially filled second column group) and more rows than CTAs
    shapes = [(40, (600, 300, 150), 12)] if depth != 8 else [(40, (600, 300, 150), 12), (24, (4200, 310), 10)]
    try:
        for n_in, hid, n_out in shapes:
            k, _ = hp.generate_kernel(77, n_in, list(hid), n_out)
            X, T = make_problem(nn_type, n_in, hid, n_out, n_samples=3, seed=3)
            for n_it in (1, 7, 16, 37):
                kw = dict(max_iter=n_it - 1, min_iter=n_it + 10, delta=1e-300)
                gpu = CudaEngine(k, nn_type, N.TRAIN_BP, **kw)
                assert gpu.net.net.hist_depth == 8
                cpu = CpuEngine(k, nn_type, N.TRAIN_BP, cuda_defaults=True, **kw)
                for x, t in zip(X, T):
                    rg, rc = gpu.train(x, t), cpu.train(x, t)
                    assert rg.n_iter == rc.n_iter == n_it
                    assert abs(rg.init_err - rc.init_err) < 1e-12
                    assert abs(rg.last_err - rc.last_err) < 1e-11
                for a, b in zip(gpu.kernel().weights, cpu.kernel().weights):
                    np.testing.assert_allclose(a, b, atol=1e-11)
                np.testing.assert_allclose(gpu.run(X[0]), cpu.run(X[0]), atol=1e-10)
                gpu.close()
                cpu.close()
    finally:
        lib.hpnn_dev_set_lazy(-1, -1)
        lib.hpnn_dev_set_lazy_panel(-1)


def test_deferred_update_trainer_on_the_wide_bench_width():
    """The 8192-wide benchmark width: 55 rows per CTA panel (> 8), two column groups per thread in
    the streaming passes, write-back depth 8 and 26 iterations = three write-back periods plus a
    partial one.  Weights must equal the CPU engine's to 1e-11."""
    n_in, hid, n_out = 784, (8192, 8192), 10
    k, _ = hp.generate_kernel(10958, n_in, list(hid), n_out)
    X, T = make_problem(N.TYPE_ANN, n_in, hid, n_out, n_samples=1, seed=11)
    n_it = 26
    kw = dict(max_iter=n_it - 1, min_iter=n_it + 10, delta=1e-300)
    gpu = CudaEngine(k, N.TYPE_ANN, N.TRAIN_BP, **kw)
    assert gpu.net.net.hist_depth == 8 and gpu.net.net.panel_stride > 0  # K-PL1 is the trainer in use
    cpu = CpuEngine(k, N.TYPE_ANN, N.TRAIN_BP, cuda_defaults=True, **kw)
    rg, rc = gpu.train(X[0], T[0]), cpu.train(X[0], T[0])
    assert rg.n_iter == rc.n_iter == n_it
    assert abs(rg.init_err - rc.init_err) < 1e-12 and abs(rg.last_err - rc.last_err) < 1e-11
    for a, b in zip(gpu.kernel().weights, cpu.kernel().weights):
        np.testing.assert_allclose(a, b, atol=1e-11)
    gpu.close()
    cpu.close()


@pytest.mark.parametrize("nn_type", [N.TYPE_ANN, N.TYPE_SNN])
def test_row_panel_momentum_trainer_matches_cpu_engine(nn_type):
    """BPM on the row-panel trainer (K-PL1 with immediate updates): the momentum triple
    dw += lr*delta (x) h; W += dw; dw *= alpha (src/ann.c:1996-1999) applied by the CTA that owns the
    panel, layer 0 in factored form.  Fixed iteration counts, weights equal to the CPU engine's."""
    lib = N.load()
    lib.hpnn_dev_set_lazy(8, 0)  # reserve the panel slots for every shape
    lib.hpnn_dev_set_lazy_panel(1)
    try:
        for n_in, hid, n_out in [(40, (600, 300, 150), 12), (24, (4200, 310), 10)]:
            k, _ = hp.generate_kernel(77, n_in, list(hid), n_out)
            X, T = make_problem(nn_type, n_in, hid, n_out, n_samples=2, seed=3)
            for n_it in (1, 9, 22):
                kw = dict(max_iter=n_it - 1, min_iter=n_it + 10, delta=1e-300)
                gpu = CudaEngine(k, nn_type, N.TRAIN_BPM, **kw)
                assert gpu.net.net.panel_stride > 0
                cpu = CpuEngine(k, nn_type, N.TRAIN_BPM, cuda_defaults=True, **kw)
                for x, t in zip(X, T):
                    rg, rc = gpu.train(x, t), cpu.train(x, t)
                    assert rg.n_iter == rc.n_iter == n_it
                    assert abs(rg.init_err - rc.init_err) < 1e-12 and abs(rg.last_err - rc.last_err) < 1e-11
                for a, b in zip(gpu.kernel().weights, cpu.kernel().weights):
                    np.testing.assert_allclose(a, b, atol=1e-11)
                gpu.close()
                cpu.close()
    finally:
        lib.hpnn_dev_set_lazy(-1, -1)
        lib.hpnn_dev_set_lazy_panel(-1)


def test_hpnn_net_is_a_torch_module():
    """parameters(), state_dict() / load_state_dict() and forward() of the device-resident net."""
    from hpnn_b200.models.net import HpnnNet

    k, _ = hp.generate_kernel(7, 20, [16, 12], 4)
    k2, _ = hp.generate_kernel(8, 20, [16, 12], 4)
    a, b = HpnnNet(k, N.TYPE_SNN), HpnnNet(k2, N.TYPE_SNN)
    assert isinstance(a, torch.nn.Module)
    ps = list(a.parameters())
    assert len(ps) == 3 and not any(p.requires_grad for p in ps) and ps[0].shape == (16, 24)  # padded to 8 columns
    assert sum(1 for _ in a.buffers()) == 3  # momenta
    x = torch.rand(20, dtype=torch.float64, device="cuda")
    X = torch.rand(5, 20, dtype=torch.float64, device="cuda")
    ya = a(x)
    assert not torch.allclose(ya, b(x))
    b.load_state_dict(a.state_dict())  # copies in place: the kernels' pointers stay valid
    torch.testing.assert_close(b(x), ya, atol=0, rtol=0)
    torch.testing.assert_close(a(X)[0], a(X[0]), atol=1e-13, rtol=0)
    a.double().to("cuda")  # same device and dtype: allowed, pointers refreshed
    torch.testing.assert_close(a(x), ya, atol=0, rtol=0)
    with pytest.raises(RuntimeError):
        a.float()
    a.close()
    b.close()
