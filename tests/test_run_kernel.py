"""``run_kernel`` (nn_run_kernel, src/libhpnn.c:1306-1536): the vectorised verdicts of the batched
default path must be the per-file ones, line for line."""
import io

import numpy as np
import pytest

import hpnn_b200 as hp
from hpnn_b200 import _native as N
from hpnn_b200.trainer import classify, classify_batch
from hpnn_b200.utils import log

from conftest import make_problem


@pytest.mark.parametrize("nn_type", [N.TYPE_ANN, N.TYPE_SNN])
def test_classify_batch_equals_classify(nn_type):
    rng = np.random.default_rng(0)
    Y = rng.random((300, 7)) * 2 - 1
    Y[3] = -2.0          # nothing above the scan's starting value
    Y[5, :] = Y[5, 0]    # ties: the first maximum wins
    Y[7, 2] = np.nan
    lo = -1.0 if nn_type == N.TYPE_ANN else 0.0
    T = np.where(rng.random((300, 7)) > 0.8, 1.0, lo)  # several / no targets set: the LAST one counts
    T[11] = lo
    g, w, r = classify_batch(Y, T, nn_type)
    for i in range(300):
        a = classify(Y[i], T[i], nn_type)
        assert (a[0], a[1], a[2]) == (g[i], w[i], r[i])


@pytest.mark.gpu
@pytest.mark.parametrize("nn_type,name", [(N.TYPE_ANN, "ANN"), (N.TYPE_SNN, "SNN")])
def test_batched_default_prints_the_per_file_lines(tmp_path, nn_type, name):
    from conftest import write_conf

    n_in, hid, n_out = 20, (16,), 4
    X, T = make_problem(nn_type, n_in, hid, n_out, n_samples=37, seed=3)
    (tmp_path / "samples").mkdir()
    (tmp_path / "tests").mkdir()
    for i, (x, t) in enumerate(zip(X, T)):
        hp.write_sample(str(tmp_path / "tests" / f"s{i:03d}.txt"), x, t, str(i % n_out))
    write_conf(tmp_path / "nn.conf", nn_type=name, n_in=n_in, hid="16", n_out=n_out, samples=str(tmp_path / "samples"),
               tests=str(tmp_path / "tests"))
    conf = hp.load_conf(str(tmp_path / "nn.conf"))
    log.set_verbose(2)
    try:
        a, b = io.StringIO(), io.StringIO()
        sa = hp.run_kernel(conf, backend="cuda", stream=a)            # default: whole directory batched
        sb = hp.run_kernel(conf, backend="cuda", stream=b, batch=0)   # the reference's one file per launch
    finally:
        log.set_verbose(0)
    assert sa == sb and sa["total"] == 37
    la, lb = a.getvalue().splitlines(), b.getvalue().splitlines()
    assert len(la) == 37
    if nn_type == N.TYPE_SNN:  # probabilities are printed with 10 digits: the two GEMV/GEMM paths may differ in the last
        strip = lambda l: l.split(" P=")[0] + l[l.index(" ["):]
        assert [strip(l) for l in la] == [strip(l) for l in lb]
        pa = [float(l.split(" P=")[1].split()[0]) for l in la]
        pb = [float(l.split(" P=")[1].split()[0]) for l in lb]
        assert np.allclose(pa, pb, atol=1e-8)
    else:
        assert la == lb


@pytest.mark.gpu
def test_bf16_precision_option_agrees_with_fp64(tmp_path):
    """run_kernel(precision="bf16"): the tensor-core chain gives the FP64 path's probabilities to ~1e-3 and the
    same verdicts wherever the best two classes are further apart than that; refused off the batched CUDA path."""
    from conftest import write_conf

    nn_type, n_in, n_out = N.TYPE_SNN, 70, 5
    X, T = make_problem(nn_type, n_in, (48, 40), n_out, n_samples=200, seed=5)
    (tmp_path / "samples").mkdir()
    (tmp_path / "tests").mkdir()
    for i, (x, t) in enumerate(zip(X, T)):
        hp.write_sample(str(tmp_path / "tests" / f"s{i:03d}.txt"), x, t, str(i % n_out))
    write_conf(tmp_path / "nn.conf", nn_type="SNN", n_in=n_in, hid="48 40", n_out=n_out, samples=str(tmp_path / "samples"),
               tests=str(tmp_path / "tests"))
    conf = hp.load_conf(str(tmp_path / "nn.conf"))
    log.set_verbose(2)
    try:
        a, b = io.StringIO(), io.StringIO()
        sa = hp.run_kernel(conf, backend="cuda", stream=a)
        sb = hp.run_kernel(conf, backend="cuda", stream=b, precision="bf16")
    finally:
        log.set_verbose(0)
    assert sa["total"] == sb["total"] == 200
    la, lb = a.getvalue().splitlines(), b.getvalue().splitlines()
    pa = np.array([float(l.split(" P=")[1].split()[0]) for l in la])
    pb = np.array([float(l.split(" P=")[1].split()[0]) for l in lb])
    assert np.allclose(pa, pb, atol=0.5)  # percent: 5e-3 in probability
    same = sum(x.split(" P=")[0] == y.split(" P=")[0] for x, y in zip(la, lb))
    assert same >= 190 and abs(sa["pass"] - sb["pass"]) <= 10
    with pytest.raises(ValueError):
        hp.run_kernel(conf, backend="cuda", batch=0, precision="bf16")
    with pytest.raises(ValueError):
        hp.run_kernel(conf, backend="cpu", precision="bf16")


def test_cli_accepts_precision_only_for_run_nn():
    from hpnn_b200.cli._common import parse

    o = parse(["run_nn", "--precision", "bf16", "--batch", "64", "nn.conf"], "run_nn", False)
    assert o["precision"] == "bf16" and o["batch"] == 64 and o["conf"] == "nn.conf"
    with pytest.raises(SystemExit):
        parse(["train_nn", "--precision", "bf16", "nn.conf"], "train_nn", True)
