"""Does the schedule perturbation (hpnn_dev_set_jitter_ns) reach the kernels?  Time of 400 iterations with and without it."""
import time, numpy as np, torch, hpnn_b200 as hp
from hpnn_b200 import _native as N
from hpnn_b200.trainer import make_engine
import sys
sys.path.insert(0, "tests")
from conftest import make_problem
lib = N.load()
for backend in ("virtual:4", "cuda"):
    for jitter in (0, 3000):
        lib.hpnn_dev_set_jitter_ns(jitter)
        k, _ = hp.generate_kernel(9, 40, [37, 29], 6)
        X, T = make_problem(N.TYPE_SNN, 40, (37, 29), 6, n_samples=2, seed=8)
        eng = make_engine(backend, k, N.TYPE_SNN, N.TRAIN_BPM, max_iter=399, min_iter=500, delta=1e-300)
        eng.train(X[0], T[0]); torch.cuda.synchronize()
        t0 = time.perf_counter(); eng.train(X[1], T[1]); torch.cuda.synchronize(); dt = time.perf_counter() - t0
        print(backend, "jitter", jitter, "ms for 400 iterations: %.2f" % (dt * 1e3), flush=True)
        eng.close()
lib.hpnn_dev_set_jitter_ns(0)
