"""File formats, RNG parity and directory walk."""
import io
import os

import numpy as np
import pytest

import hpnn_b200 as hp
from hpnn_b200 import _native as N
from hpnn_b200.config import dump_conf, parse_conf
from hpnn_b200.kernel_io import dump_kernel_binary, load_kernel_binary
from hpnn_b200.samples import list_dir

from conftest import write_conf

# first weights of `[seed] 10958`, 8-5-4-3, as written by the reference binary's kernel.tmp
REF_FIRST_ROW = [-0.281305454478708, -0.027710347937003, -0.077296293698722, 0.154056015704165,
                 0.042771657614647, -0.165060284230241, 0.024460498947952, 0.243725484355540]


def test_generate_matches_reference_binary():
    k, seed = hp.generate_kernel(10958, 8, [5, 4], 3)
    assert seed == 10958 and k.dims == [8, 5, 4, 3] and k.name is None
    np.testing.assert_allclose(k.weights[0][0], REF_FIRST_ROW, atol=5e-16)
    for l, w in enumerate(k.weights):  # |w| <= 1/sqrt(fan_in)
        assert np.abs(w).max() <= 1.0 / np.sqrt(k.dims[l]) + 1e-15


def test_seed_zero_uses_time():
    _, seed = hp.generate_kernel(0, 4, [3], 2)
    assert seed != 0


def test_shuffle_is_glibc_random():
    assert hp.shuffle_order(10958, 8) == [0, 3, 5, 4, 2, 6, 1, 7]
    assert sorted(hp.shuffle_order(1, 100)) == list(range(100))


def test_kernel_text_roundtrip(tmp_path):
    k, _ = hp.generate_kernel(42, 7, [6, 5], 4)
    p = tmp_path / "kernel.tmp"
    hp.dump_kernel(k, str(p))
    text = p.read_text().splitlines()
    assert text[0] == "[name] (null)"  # generated kernels are unnamed
    assert text[1] == "[param] 7 6 5 4" and text[2] == "[input] 7" and text[3] == "[hidden 1] 6"
    assert text[4] == "[neuron 1] 7"
    assert all(len(tok.split(".")[1]) == 15 for tok in text[5].split())
    k2 = hp.load_kernel(str(p))
    assert k2.dims == k.dims and k2.name == "(null)"
    for a, b in zip(k.weights, k2.weights):
        np.testing.assert_allclose(a, b, atol=1e-15)
    # dump of the reloaded kernel is byte-identical except for nothing: name line equal too
    p2 = tmp_path / "kernel2"
    hp.dump_kernel(k2, str(p2))
    assert p2.read_text() == p.read_text()


def test_kernel_binary_sidecar_is_exact(tmp_path):
    k, _ = hp.generate_kernel(5, 9, [4], 3)
    k.name = "exact"
    dump_kernel_binary(k, str(tmp_path / "k.bin"))
    k2 = load_kernel_binary(str(tmp_path / "k.bin"))
    assert k2.name == "exact" and all((a == b).all() for a, b in zip(k.weights, k2.weights))


def test_kernel_load_errors(tmp_path):
    p = tmp_path / "bad"
    p.write_text("[nope]\n")
    with pytest.raises(IOError):
        hp.load_kernel(str(p))
    p.write_text("[name] x\n[param] 3 0 2\n")
    with pytest.raises(IOError):
        hp.load_kernel(str(p))
    with pytest.raises(IOError):
        hp.load_kernel(str(tmp_path / "missing"))


def test_conf_parse_and_dump(tmp_path):
    c = tmp_path / "nn.conf"
    write_conf(c, nn_type="SNN", train="BPM", hid="30 20 10", n_in=12, n_out=5, samples="/a/b  # comment")
    conf = parse_conf(str(c))
    assert conf.name == "toy" and conf.type == N.TYPE_SNN and conf.train == N.TRAIN_BPM
    assert conf.need_init and conf.seed == 10958 and conf.n_inputs == 12 and conf.hiddens == [30, 20, 10]
    assert conf.n_outputs == 5 and conf.samples == "/a/b" and conf.tests == "./tests"
    text = dump_conf(conf)
    assert "[type] SNN" in text and "[train] BPM" in text and "[hiddens] 30 20 10 " in text
    # the dumped form ([inputs]/[hiddens]/[outputs]) is accepted back
    c2 = tmp_path / "nn2.conf"
    c2.write_text(text)
    conf2 = parse_conf(str(c2))
    assert conf2.hiddens == [30, 20, 10] and conf2.n_inputs == 12 and conf2.n_outputs == 5


@pytest.mark.parametrize("train,expect", [("BP", N.TRAIN_BP), ("BPM", N.TRAIN_BPM), ("CG", N.TRAIN_CG),
                                          ("SPLX", N.TRAIN_SPLX), ("zzz", N.TRAIN_UKN)])
def test_conf_train_keywords(tmp_path, train, expect):
    c = tmp_path / "nn.conf"
    write_conf(c, train=train)
    assert parse_conf(str(c)).train == expect


def test_conf_errors(tmp_path):
    c = tmp_path / "nn.conf"
    c.write_text("# nothing\n[name] x\n")
    with pytest.raises(ValueError):
        parse_conf(str(c))  # missing [type]
    write_conf(c, hid="5 0")
    with pytest.raises(ValueError):
        parse_conf(str(c))
    c.write_text("#\n[type] ANN\n[init] generate\n[hidden] 3\n[output] 2\n")
    with pytest.raises(ValueError):
        parse_conf(str(c))  # missing [input]
    with pytest.raises(ValueError):
        parse_conf(str(tmp_path / "nope.conf"))


def test_conf_init_from_file(tmp_path):
    k, _ = hp.generate_kernel(3, 8, [5, 4], 3)
    hp.dump_kernel(k, str(tmp_path / "kernel.opt"))
    c = tmp_path / "nn.conf"
    write_conf(c, init=str(tmp_path / "kernel.opt"))
    conf = hp.load_conf(str(c))
    assert not conf.need_init and conf.kernel.dims == [8, 5, 4, 3]


def test_sample_roundtrip_and_comment(tmp_path):
    x = np.array([0.5, 255.0, 0.12345678, 3.0])
    t = np.array([-1.0, 1.0, -1.0])
    p = tmp_path / "s00001.txt"
    hp.write_sample(str(p), x, t, "1")
    lines = p.read_text().splitlines()
    assert lines[0] == "[input] 4" and lines[2] == "[output] 3  #1" and lines[3] == "-1.0 1.0 -1.0"
    assert lines[1] == "0.50000 255.00000 0.12346 3.00000"
    x2, t2 = hp.read_sample(str(p))
    np.testing.assert_allclose(x2, [0.5, 255.0, 0.12346, 3.0])
    np.testing.assert_array_equal(t2, t)
    with pytest.raises(IOError):
        hp.read_sample(str(tmp_path / "missing.txt"))


def test_dir_walk_skips_dotfiles_and_covers_all(toy_dir):
    (toy_dir / "samples" / ".hidden").write_text("x")
    names = list_dir(str(toy_dir / "samples"))
    assert ".hidden" not in names and sorted(names) == [f"s{i:03d}.txt" for i in range(6)]
    files = hp.walk(str(toy_dir / "samples"), 10958)
    assert sorted(os.path.basename(f) for f in files) == sorted(names)
    assert files == hp.walk(str(toy_dir / "samples"), 10958)  # deterministic for a seed


def test_arena_layout_regions_are_disjoint_and_cover_the_variants():
    """hpnn_dev_arena_layout is host code: every region the trainers index must lie inside the arena
    and overlap no other region, for one rank and for a row-split job, small and wide shapes."""
    import ctypes as C

    from hpnn_b200 import _native as N

    lib = N.load()

    def layout(dims, rank, world):
        net = N.DevNet()
        net.n_layers = len(dims) - 1
        for i, d in enumerate(dims):
            net.dims[i] = d
        net.rank, net.world = rank, world
        total = lib.hpnn_dev_arena_layout(C.byref(net))
        assert total > 0 and total == net.arena_doubles
        return net, total

    def regions(net, total):
        NL, W = net.n_layers, net.world
        pad = lambda v: (v + 7) & ~7
        regs = [("act%d" % i, net.off_act[i], pad(net.dims[i]) * (1 if i == 0 else 4)) for i in range(NL + 1)]
        regs += [("part%d" % l, net.off_part[l], 2 * W * 4 * net.part_stride[l]) for l in range(NL)]
        regs += [("target", net.off_target, pad(net.dims[NL])), ("out", net.off_out, pad(net.dims[NL])),
                 ("flags", net.off_flags, 2 * N.MAX_WORLD)]
        if net.cta_slots:
            regs += [("cta%d" % l, net.off_cta_part[l], 2 * net.cta_slots * net.ld[l]) for l in range(1, NL)]
        if net.hist_depth:
            for l in range(1, NL):
                regs += [("hu%d" % l, net.off_hist_u[l], 8 * net.part_stride[l]), ("hv%d" % l, net.off_hist_v[l], 8 * net.ld[l]),
                         ("hd%d" % l, net.off_hist_dot[l], 2 * 8 * W * 160)]
            if net.panel_stride:
                regs.append(("panel", net.off_panel, 2 * 160 * net.panel_stride))
        if net.off_zpart:  # folded output layer: working copies of the layer and its momentum, partial output sums
            wl = pad(net.dims[NL] * net.ld[NL - 1])
            regs += [("wlast", net.off_wlast, wl), ("mlast", net.off_mlast, wl), ("zpart", net.off_zpart, W * 160 * 32)]
        return regs

    def check(net, total):
        regs = sorted(regions(net, total), key=lambda r: r[1])
        for (na, oa, sa), (nb, ob, _) in zip(regs, regs[1:]):
            assert oa >= 0 and oa + sa <= ob, (na, nb)
        assert regs[-1][1] + regs[-1][2] <= total

    small, _ = layout([784, 300, 10], 0, 1)
    assert small.cta_slots > 0 and small.hist_depth == 0 and small.panel_stride == 0
    check(small, small.arena_doubles)
    for world in (1, 2, 8):
        for rank in (0, world - 1):
            wide, total = layout([784, 8192, 8192, 8192, 8192, 10], rank, world)
            assert wide.hist_depth == 8 and wide.panel_stride == 8192 + 8 and wide.cta_slots == 0
            assert wide.rows[1] == 8192 // world and wide.row0[1] == rank * (8192 // world)
            assert wide.off_wlast > 0 and wide.off_mlast > wide.off_wlast and wide.off_zpart > wide.off_mlast
            check(wide, total)
    many_out, total = layout([784, 8192, 8192, 64], 0, 2)  # more than 32 outputs: the output layer is not folded
    assert many_out.hist_depth == 8 and many_out.off_zpart == 0
    check(many_out, total)
    # the history threshold is a setting (tests use it to exercise the deferred-update trainers on small nets)
    lib.hpnn_dev_set_lazy(-1, 0)
    try:
        forced, total = layout([40, 600, 300, 12], 0, 1)
        assert forced.hist_depth == 8 and forced.cta_slots > 0
        check(forced, total)
    finally:
        lib.hpnn_dev_set_lazy(-1, -1)
