import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")


@pytest.fixture(scope="session", autouse=True)
def _native_built():
    from hpnn_b200 import _native

    _native.load()  # builds in-tree on first use; raises if nvcc is unavailable
    yield


def make_problem(nn_type, n_in=8, hid=(5, 4), n_out=3, n_samples=6, seed=1):
    """Synthetic classification samples with the target convention of the type."""
    from hpnn_b200 import TYPE_SNN

    rng = np.random.default_rng(seed)
    lo = 0.0 if nn_type == TYPE_SNN else -1.0
    xs, ts = [], []
    for i in range(n_samples):
        x = rng.random(n_in)
        t = np.full(n_out, lo)
        t[i % n_out] = 1.0
        xs.append(x)
        ts.append(t)
    return np.stack(xs), np.stack(ts)


@pytest.fixture
def toy_dir(tmp_path):
    """A conf + sample/test directories for the 8-5-4-3 toy net."""
    from hpnn_b200 import TYPE_ANN, write_sample

    X, T = make_problem(TYPE_ANN)
    (tmp_path / "samples").mkdir()
    (tmp_path / "tests").mkdir()
    for i, (x, t) in enumerate(zip(X, T)):
        write_sample(str(tmp_path / "samples" / f"s{i:03d}.txt"), x, t, str(i % 3))
        write_sample(str(tmp_path / "tests" / f"s{i:03d}.txt"), x, t, str(i % 3))
    return tmp_path


def write_conf(path, nn_type="ANN", train="BP", init="generate", seed=10958, n_in=8, hid="5 4", n_out=3,
               samples="./samples", tests="./tests"):
    path.write_text(f"# test conf\n[name] toy\n[type] {nn_type}\n[init] {init}\n[seed] {seed}\n[input] {n_in}\n"
                    f"[hidden] {hid}\n[output] {n_out}\n[train] {train}\n[sample_dir] {samples}\n[test_dir] {tests}\n")
