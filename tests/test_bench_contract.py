"""bench.py contract checks that need no GPU: both arms describe the same config, and the
reference arm creates its data without touching this repository's native code."""
import argparse
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_sample_writer_matches_native(tmp_path):
    import bench
    from hpnn_b200.samples import write_sample

    rng = np.random.default_rng(7)
    x = np.concatenate([rng.random(37), [0.0, 1.0, 0.999995, 1e-6, 0.5]])
    t = np.full(10, -1.0)
    t[4] = 1.0
    a, b = tmp_path / "a.txt", tmp_path / "b.txt"
    bench.write_sample_text(str(a), x, t, "4")
    write_sample(str(b), x, t, "4")
    assert a.read_bytes() == b.read_bytes()


def test_config_block_is_arm_independent():
    import bench

    args = argparse.Namespace(config="wide_ann", ref_streams=1)
    for n in (1, 2, 8):
        c = bench.config_block(args, bench.CONFIGS["wide_ann"], n)
        assert c == bench.config_block(args, bench.CONFIGS["wide_ann"], n)
        assert set(c) == {"model", "name", "baseline_config", "global_batch", "seq_len", "parallelism", "l2"}
        assert c["model"] == "ANN 784-8192-8192-8192-8192-10"


def test_reference_data_path_loads_no_repo_library(tmp_path):
    code = (
        "import sys; sys.path.insert(0, %r); import bench\n"
        "bench.make_dataset(bench.CONFIGS['ann_mnist'], 1, 2, %r)\n"
        "assert not any(m.startswith('hpnn_b200') for m in sys.modules), 'hpnn_b200 imported'\n"
        "assert 'hpnn_b200/lib' not in open('/proc/self/maps').read()\n" % (ROOT, str(tmp_path)))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert len(os.listdir(tmp_path / "timed")) == 2
