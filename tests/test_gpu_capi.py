"""The C API on GPUs: libhpnn.so drives the device engine (and every visible GPU) by itself."""
import os
import subprocess

import numpy as np
import pytest
import torch

import hpnn_b200 as hp
from hpnn_b200 import build as B

from conftest import ROOT, make_problem, write_conf

pytestmark = pytest.mark.gpu


def _setup(tmp_path, nn_type="ANN", train="BP", hid="48 40", n_in=64, n_out=10):
    t = hp.TYPE_SNN if nn_type == "SNN" else hp.TYPE_ANN
    X, T = make_problem(t, n_in, (48, 40), n_out, n_samples=5, seed=8)
    (tmp_path / "samples").mkdir()
    (tmp_path / "tests").mkdir()
    for i, (x, tt) in enumerate(zip(X, T)):
        hp.write_sample(str(tmp_path / "samples" / f"s{i:03d}.txt"), x, tt, str(i))
        hp.write_sample(str(tmp_path / "tests" / f"s{i:03d}.txt"), x, tt, str(i))
    write_conf(tmp_path / "nn.conf", nn_type=nn_type, train=train, n_in=n_in, hid=hid, n_out=n_out)
    return tmp_path


def _train(exe, d, env_extra):
    env = dict(os.environ)
    env.update(env_extra)
    r = subprocess.run([str(exe), "-v", "-v", "-v", "nn.conf"], cwd=d, capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    rows = [(l.split("TRAINING FILE:")[1].split()[0], int(l.split("N_ITER=")[1].split()[0]))
            for l in r.stdout.splitlines() if "TRAINING FILE" in l]
    return r.stdout, rows, hp.load_kernel(str(d / "kernel.opt"))


@pytest.mark.parametrize("nn_type,train", [("ANN", "BP"), ("SNN", "BPM")])
def test_c_driver_uses_the_device_engine(tmp_path, nn_type, train):
    arts = B.build(verbose=False)
    d = _setup(tmp_path, nn_type, train)
    out, rows_gpu, k_gpu = _train(arts["train_nn"], d, {"CUDA_VISIBLE_DEVICES": "0", "HPNN_LR": "0.01"})
    assert "[GPU] ANN total allocation" in out and "CUDA started, found 1 GPU(s)" in out
    _, rows_cpu, k_cpu = _train(arts["train_nn"], d, {"CUDA_VISIBLE_DEVICES": "", "HPNN_LR": "0.01"})
    assert [n for n, _ in rows_gpu] == [n for n, _ in rows_cpu]
    for (_, a), (_, b) in zip(rows_gpu, rows_cpu):
        assert abs(a - b) <= max(3, b // 100)
    for a, b in zip(k_gpu.weights, k_cpu.weights):
        np.testing.assert_allclose(a, b, atol=1e-6)
    write_conf(d / "nn.conf", nn_type=nn_type, train=train, n_in=64, hid="48 40", n_out=10, init="kernel.opt")
    outs = []
    for vis in ("0", ""):  # the device forward and the CPU forward classify every file alike
        r = subprocess.run([str(arts["run_nn"]), "-vv", "nn.conf"], cwd=d, capture_output=True, text=True,
                           env=dict(os.environ, CUDA_VISIBLE_DEVICES=vis), timeout=300)
        outs.append([l.split("P=")[0] + l[l.index("["):] if "P=" in l else l for l in r.stdout.splitlines() if "TESTING FILE" in l])
    assert len(outs[0]) == 5 and outs[0] == outs[1]


@pytest.mark.parametrize("n_gpus", [2, 4, 8])
def test_c_driver_row_splits_over_all_gpus(tmp_path, n_gpus):
    if torch.cuda.device_count() < n_gpus:
        pytest.skip(f"needs {n_gpus} GPUs")
    arts = B.build(verbose=False)
    d = _setup(tmp_path, "SNN", "BPM")
    vis = ",".join(str(i) for i in range(n_gpus))
    out, rows_n, k_n = _train(arts["train_nn"], d, {"CUDA_VISIBLE_DEVICES": vis, "HPNN_TP_MIN_MB": "0"})  # force the split
    assert f"CUDA started, found {n_gpus} GPU(s)" in out
    _, rows_1, k_1 = _train(arts["train_nn"], d, {"CUDA_VISIBLE_DEVICES": "0"})
    for (_, a), (_, b) in zip(rows_n, rows_1):
        assert abs(a - b) <= max(3, b // 100)
    for a, b in zip(k_n.weights, k_1.weights):
        np.testing.assert_allclose(a, b, atol=1e-6)
