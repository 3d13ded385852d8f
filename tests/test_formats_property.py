"""Property tests of the file formats (hypothesis): whatever topology, values and spacing, what is
written is read back, and the native (C++) and Python parsers agree."""
import os
import tempfile

import numpy as np
from hypothesis import given, settings
from hypothesis import strategies as st

import hpnn_b200 as hp
from hpnn_b200 import _native as N
from hpnn_b200.config import dump_conf, parse_conf

dims = st.lists(st.integers(min_value=1, max_value=12), min_size=3, max_size=6)
finite = st.floats(min_value=-1e3, max_value=1e3, allow_nan=False, allow_infinity=False)


@settings(max_examples=40, deadline=None)
@given(d=dims, seed=st.integers(min_value=1, max_value=2**31 - 1))
def test_kernel_text_roundtrip_any_topology(d, seed):
    k, _ = hp.generate_kernel(seed, d[0], d[1:-1], d[-1])
    with tempfile.TemporaryDirectory() as tmp:
        p = os.path.join(tmp, "k.opt")
        hp.dump_kernel(k, p)
        k2 = hp.load_kernel(p)
        assert k2.dims == k.dims
        for a, b in zip(k.weights, k2.weights):
            np.testing.assert_allclose(a, b, atol=5e-16)  # %17.15f keeps 15 decimals
        # a second dump of the reloaded kernel is byte-identical (the text form is a fixed point)
        p2 = os.path.join(tmp, "k2.opt")
        hp.dump_kernel(k2, p2)
        assert open(p).read() == open(p2).read()


@settings(max_examples=40, deadline=None)
@given(x=st.lists(finite, min_size=1, max_size=40), hot=st.integers(min_value=0, max_value=9),
       n_out=st.integers(min_value=1, max_value=10), snn=st.booleans())
def test_sample_roundtrip_any_values(x, hot, n_out, snn):
    t = np.full(n_out, 0.0 if snn else -1.0)
    t[hot % n_out] = 1.0
    xs = np.array(x)
    with tempfile.TemporaryDirectory() as tmp:
        p = os.path.join(tmp, "s.txt")
        hp.write_sample(p, xs, t, "c")
        x2, t2 = hp.read_sample(p)
        np.testing.assert_allclose(x2, xs, atol=5.1e-6)  # inputs are written with 5 decimals
        np.testing.assert_array_equal(t2, t)


@settings(max_examples=30, deadline=None)
@given(d=dims, typ=st.sampled_from(["ANN", "SNN"]), train=st.sampled_from(["BP", "BPM"]),
       seed=st.integers(min_value=0, max_value=10**6), pad=st.sampled_from(["", " ", "\t", "   # trailing comment"]))
def test_conf_roundtrip_any_topology(d, typ, train, seed, pad):
    text = (f"[name] prop{pad}\n[type] {typ}\n[init] generate\n[seed] {seed}\n[input] {d[0]}\n"
            f"[hidden] {' '.join(str(v) for v in d[1:-1])}{pad}\n[output] {d[-1]}\n[train] {train}\n"
            f"[sample_dir] ./s\n[test_dir] ./t\n")
    with tempfile.TemporaryDirectory() as tmp:
        p = os.path.join(tmp, "nn.conf")
        open(p, "w").write(text)
        c = parse_conf(p)
        assert c.n_inputs == d[0] and c.hiddens == d[1:-1] and c.n_outputs == d[-1] and c.seed == seed
        assert c.type == (N.TYPE_SNN if typ == "SNN" else N.TYPE_ANN)
        assert c.train == (N.TRAIN_BPM if train == "BPM" else N.TRAIN_BP)
        p2 = os.path.join(tmp, "nn2.conf")
        open(p2, "w").write(dump_conf(c))
        c2 = parse_conf(p2)
        assert (c2.n_inputs, c2.hiddens, c2.n_outputs, c2.type, c2.train) == (c.n_inputs, c.hiddens, c.n_outputs, c.type, c.train)
