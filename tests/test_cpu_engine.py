"""CPU/OpenMP engine vs the PyTorch specification and vs golden numbers recorded from the
reference binary (gcc -O2 -fopenmp -D_OMP build of /root/reference, toy problem below)."""
import numpy as np
import pytest
import torch

import hpnn_b200 as hp
from hpnn_b200 import _native as N
from hpnn_b200.ops import reference as R
from hpnn_b200.trainer import CpuEngine, TorchEngine

from conftest import make_problem

# reference train_nn output for: ANN 8-5-4-3, BP, seed 10958, samples = make_problem(ANN)
# written with write_sample (%7.5f) and visited in this order
REF_ORDER = [2, 4, 1, 0, 5, 3]
REF_NITER = [31403, 24208, 9146, 25292, 18347, 21753]
REF_INIT = [1.5139345432, 3.2261434563, 0.1447097800, 3.7106713300, 3.4675286205, 1.5578267625]


def _rounded(X):
    return np.array([[float("%7.5f" % v) for v in row] for row in X])


def test_matches_reference_binary_golden():
    k, _ = hp.generate_kernel(10958, 8, [5, 4], 3)
    X, T = make_problem(N.TYPE_ANN)
    X = _rounded(X)
    eng = CpuEngine(k, N.TYPE_ANN, N.TRAIN_BP)
    for j, i in enumerate(REF_ORDER):
        r = eng.train(X[i], T[i])
        assert r.n_iter == REF_NITER[j]
        assert abs(r.init_err - REF_INIT[j]) < 1e-9
        assert r.success and r.final_de <= 1e-6
    eng.close()


@pytest.mark.parametrize("nn_type,train", [(N.TYPE_ANN, N.TRAIN_BP), (N.TYPE_ANN, N.TRAIN_BPM),
                                           (N.TYPE_SNN, N.TRAIN_BP), (N.TYPE_SNN, N.TRAIN_BPM)])
def test_cpu_engine_equals_torch_spec(nn_type, train):
    k, _ = hp.generate_kernel(77, 8, [6, 5], 3)
    X, T = make_problem(nn_type, n_samples=2)
    a = CpuEngine(k, nn_type, train, max_iter=300)
    b = TorchEngine(k, nn_type, train, max_iter=300)
    for x, t in zip(X, T):
        ra, rb = a.train(x, t), b.train(x, t)
        assert ra.n_iter == rb.n_iter and ra.first_ok == rb.first_ok and ra.success == rb.success
        assert abs(ra.init_err - rb.init_err) < 1e-12 and abs(ra.final_de - rb.final_de) < 1e-12
    for wa, wb in zip(a.kernel().weights, b.kernel().weights):
        np.testing.assert_allclose(wa, wb, atol=1e-12)
    np.testing.assert_allclose(a.run(X[0]), b.run(X[0]), atol=1e-13)
    a.close()


def test_default_learning_rates():
    o = N.TrainOpts()
    lib = N.load()
    import ctypes as C
    for (ty, tr, cuda), lr in {(N.TYPE_ANN, N.TRAIN_BP, 0): 0.001, (N.TYPE_ANN, N.TRAIN_BPM, 0): 0.0005,
                               (N.TYPE_SNN, N.TRAIN_BP, 0): 0.01, (N.TYPE_ANN, N.TRAIN_BP, 1): 0.01,
                               (N.TYPE_ANN, N.TRAIN_BPM, 1): 0.01}.items():
        lib.hpnn_train_opts_default(C.byref(o), ty, tr, cuda)
        assert o.lr == lr and o.alpha == 0.2 and o.delta == 1e-6 and o.max_iter == 102399
        assert o.min_iter == (15 if tr == N.TRAIN_BPM else 31)


def test_softmax_reference_formula_and_stable_mode():
    z = torch.tensor([0.3, -1.2, 2.0], dtype=torch.float64)
    o = R.softmax_ref(z)
    e = torch.exp(z - 1.0)
    torch.testing.assert_close(o, e / (e.sum() + 1e-14))
    torch.testing.assert_close(R.softmax_ref(z, stable=True), torch.softmax(z, 0), atol=1e-13, rtol=0)
    big = torch.tensor([800.0, 0.0], dtype=torch.float64)  # overflows without the max shift
    assert not torch.isfinite(R.softmax_ref(big)).all()
    assert torch.isfinite(R.softmax_ref(big, stable=True)).all()


def test_loop_bounds_min_and_max_iter():
    k, _ = hp.generate_kernel(5, 8, [5], 3)
    X, T = make_problem(N.TYPE_ANN, n_samples=1)
    e = CpuEngine(k, N.TYPE_ANN, N.TRAIN_BP, max_iter=10)
    r = e.train(X[0], T[0])
    assert r.n_iter == 11  # `if(iter>MAX) break` fires at MAX+1 (src/ann.c:2360)
    e2 = CpuEngine(k, N.TYPE_ANN, N.TRAIN_BP, delta=1e9, min_iter=40)
    r2 = e2.train(X[0], T[0])
    assert r2.n_iter >= 41  # at least MIN+1 iterations even when dE is already small


@pytest.mark.parametrize("threads", [1, 3])
@pytest.mark.parametrize("nn_type,train", [(N.TYPE_ANN, N.TRAIN_BP), (N.TYPE_ANN, N.TRAIN_BPM), (N.TYPE_SNN, N.TRAIN_BP),
                                           (N.TYPE_SNN, N.TRAIN_BPM)])
def test_packed_sample_loop_equals_single_step_path_bitwise(nn_type, train, threads):
    """hpnn_cpu_train_sample runs on weights packed 8 rows per column with the update fused into
    the forward pass; the single-step entry points (what ann_kernel_train / snn_kernel_train of the
    C API call) keep the plain row-major loops.  Same operations in the same order per row, so after
    the same number of iterations every weight must be equal to the last bit -- for row counts that
    are not multiples of 8, layers wide enough for a thread team and several hidden layers."""
    import ctypes as C
    lib = N.load()
    dp = C.POINTER(C.c_double)
    lib.hpnn_set_omp_threads(threads)
    try:
        for n_in, hid, n_out in [(13, [9, 7], 3), (70, [300, 41], 10), (5, [8], 8)]:
            k, _ = hp.generate_kernel(321, n_in, hid, n_out)
            rng = np.random.default_rng(7)
            x = rng.random(n_in)
            t = np.full(n_out, 0.0 if nn_type == N.TYPE_SNN else -1.0)
            t[1] = 1.0
            n_it = 23
            a = CpuEngine(k, nn_type, train, max_iter=n_it - 1, min_iter=n_it + 5, delta=1e-300)
            assert a.train(x, t).n_iter == n_it
            b = CpuEngine(k, nn_type, train)
            if train == N.TRAIN_BPM:
                lib.hpnn_cpu_momentum_init(b.h)
                lib.hpnn_cpu_momentum_zero(b.h)
            inp = lib.hpnn_cpu_act(b.h, 0)
            for i in range(n_in):
                inp[i] = x[i]
            lib.hpnn_cpu_forward_inplace(b.h, nn_type, 0)
            tt = np.ascontiguousarray(t)
            for _ in range(n_it):
                lib.hpnn_cpu_train_iteration(b.h, C.byref(b.opts), tt.ctypes.data_as(dp))
            for wa, wb in zip(a.kernel().weights, b.kernel().weights):
                assert np.array_equal(wa, wb)
            assert np.array_equal(a.run(x), b.run(x))
            a.close()
            b.close()
    finally:
        lib.hpnn_set_omp_threads(0)


def test_deep_network_beyond_the_device_cap(tmp_path):
    """The reference accepts any number of hidden layers (src/ann.c:206-766); only the device
    descriptor is capped at HPNN_MAX_LAYERS, so host kernels, kernel files and the CPU engine
    must take a deeper net."""
    import hpnn_b200 as hp
    from hpnn_b200 import _native as N
    from hpnn_b200.trainer import CpuEngine

    hid = [6] * 20  # 21 weight layers > HPNN_MAX_LAYERS
    k, _ = hp.generate_kernel(5, 7, hid, 3)
    path = tmp_path / "deep.kernel"
    hp.dump_kernel(k, str(path))
    k2 = hp.load_kernel(str(path))
    assert list(k2.dims) == [7] + hid + [3]
    cpu = CpuEngine(k2, N.TYPE_ANN, N.TRAIN_BP, max_iter=40)
    r = cpu.train(np.linspace(0.1, 0.9, 7), np.array([1.0, -1.0, -1.0]))
    assert r.n_iter >= 1 and np.isfinite(r.last_err)
    assert cpu.run(np.linspace(0.1, 0.9, 7)).shape == (3,)
    cpu.close()


@pytest.mark.parametrize("train", [N.TRAIN_BP, N.TRAIN_BPM])
def test_simd_paths_are_bit_identical(train, monkeypatch):
    """HPNN_CPU_ISA = generic | avx2 | avx512 (the wider ones only where the CPU has them): every path performs the
    reference's operations in the reference's order, products and sums rounded separately (no FMA contraction),
    so the trained weights agree to the last bit.  150 rows exercise the 8-, 4-, 2- and 1-block groups."""
    results = []
    for isa in ("generic", "avx2", "avx512"):
        monkeypatch.setenv("HPNN_CPU_ISA", isa)
        k, _ = hp.generate_kernel(5, 33, [150, 21], 7)
        X, T = make_problem(N.TYPE_ANN, 33, (150, 21), 7, n_samples=2, seed=4)
        eng = CpuEngine(k, N.TYPE_ANN, train, max_iter=40, min_iter=40)
        iters = [eng.train(x, t).n_iter for x, t in zip(X, T)]
        results.append((iters, [w.copy() for w in eng.kernel().weights]))
        eng.close()
    for iters, ws in results[1:]:
        assert iters == results[0][0]
        for a, b in zip(ws, results[0][1]):
            assert np.array_equal(a, b)
