"""``nn.conf`` handling (reference: src/libhpnn.c:658-937, tutorials/README.md:113-137)."""
from __future__ import annotations

import ctypes as C
import io
from dataclasses import dataclass, field
from typing import List, Optional

from . import _native as N

TYPE_NAMES = {N.TYPE_ANN: "ANN", N.TYPE_LNN: "LNN", N.TYPE_SNN: "SNN"}
TRAIN_NAMES = {N.TRAIN_BP: "BP", N.TRAIN_BPM: "BPM", N.TRAIN_CG: "CG", N.TRAIN_SPLX: "SPLX"}


@dataclass
class NNConf:
    """Python twin of ``nn_def`` (include/libhpnn.h:78-89)."""

    name: Optional[str] = None
    type: int = N.TYPE_UKN
    need_init: bool = False
    seed: int = 0
    f_kernel: Optional[str] = None
    train: int = N.TRAIN_UKN
    samples: Optional[str] = None
    tests: Optional[str] = None
    n_inputs: int = 0
    hiddens: List[int] = field(default_factory=list)
    n_outputs: int = 0
    kernel: object = None  # hpnn_b200.kernel_io.Kernel once generated / loaded

    @property
    def type_name(self) -> str:
        return TYPE_NAMES.get(self.type, "UKN")

    @property
    def train_name(self) -> str:
        return TRAIN_NAMES.get(self.train, "none")


def parse_conf(path: str) -> NNConf:
    """Parse the file only (no kernel generation / loading)."""
    lib = N.load()
    c = N.Conf()
    if lib.hpnn_conf_load(str(path).encode(), C.byref(c)) != 0:
        raise ValueError(f"malformed or unreadable NN configuration file: {path}")
    try:
        dec = lambda b: b.decode() if b is not None else None
        return NNConf(name=dec(c.name), type=c.type, need_init=bool(c.need_init), seed=int(c.seed),
                      f_kernel=dec(c.f_kernel), train=c.train, samples=dec(c.samples), tests=dec(c.tests),
                      n_inputs=c.n_in, hiddens=[c.hid[i] for i in range(c.n_hid)], n_outputs=c.n_out)
    finally:
        lib.hpnn_conf_free(C.byref(c))


def load_conf(path: str) -> NNConf:
    """``nn_load_conf``: parse, then generate or load the kernel (src/libhpnn.c:830-865)."""
    from .kernel_io import generate_kernel, load_kernel
    from .utils.log import nn_out

    conf = parse_conf(path)
    if conf.type not in (N.TYPE_ANN, N.TYPE_SNN):
        raise ValueError("FAILED to generate NN kernel! [type] unsupported...")
    if conf.need_init:
        nn_out("generating kernel!\n")
        conf.kernel, conf.seed = generate_kernel(conf.seed, conf.n_inputs, conf.hiddens, conf.n_outputs)
    else:
        nn_out("loading kernel!\n")
        conf.kernel = load_kernel(conf.f_kernel)
    return conf


def dump_conf(conf: NNConf, fp=None) -> str:
    """``nn_dump_conf`` text (src/libhpnn.c:885-937).  Unlike the reference the sizes of a
    freshly generated kernel are printed."""
    k = conf.kernel
    out = io.StringIO()
    out.write("# NN configuration\n")
    out.write(f"[name] {conf.name if conf.name is not None else '(null)'}\n")
    out.write(f"[type] {conf.type_name if conf.type in TYPE_NAMES else 'ANN'}\n")
    if conf.need_init:
        out.write("[init] generate\n")
    elif conf.f_kernel is not None:
        out.write(f"[init] {conf.f_kernel}\n")
    else:
        out.write("[init] INVALID <- this should trigger an error\n")
    out.write(f"[seed] {conf.seed}\n")
    out.write(f"[inputs] {k.n_inputs if k is not None else conf.n_inputs}\n")
    out.write("[hiddens] " + "".join(f"{h} " for h in (k.hiddens if k is not None else conf.hiddens)) + "\n")
    out.write(f"[outputs] {k.n_outputs if k is not None else conf.n_outputs}\n")
    out.write(f"[train] {conf.train_name}\n")
    out.write(f"[sample_dir] {conf.samples}\n" if conf.samples is not None
              else "[sample_dir] INVALID <- this should trigger an error\n")
    out.write(f"[test_dir] {conf.tests}\n" if conf.tests is not None
              else "[test_dir] INVALID <- this should trigger an error\n")
    text = out.getvalue()
    if fp is not None:
        fp.write(text)
    return text
