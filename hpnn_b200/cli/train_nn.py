"""``train_nn`` (tests/train_nn.c): load nn.conf, dump ``kernel.tmp``, train one pass over
the sample directory, dump ``kernel.opt``."""
from __future__ import annotations

import sys

from .. import _native as N
from ..config import load_conf
from ..kernel_io import dump_kernel
from ..trainer import train_kernel
from ..utils import log
from ._common import help_text, parse


def main(argv=None) -> int:
    argv = list(sys.argv if argv is None else argv)
    log.set_verbose(0)
    o = parse(argv, "train_nn", True)
    if o["help"]:
        sys.stdout.write(help_text("train_nn", True))
        return 0
    if o["omp"]:
        N.load().hpnn_set_omp_threads(o["omp"])
    try:
        conf = load_conf(o["conf"])
    except (ValueError, IOError) as e:
        log.nn_error(str(e) + "\n")
        sys.stderr.write("FAILED to read NN configuration file! (ABORTING)\n")
        return 255
    try:
        dump_kernel(conf.kernel, "kernel.tmp")
    except IOError:
        sys.stderr.write("FAILED to open kernel.tmp for WRITE!\n")
        return 255
    kw = {}
    if o["lr"]:
        kw["lr"] = o["lr"]
    if not train_kernel(conf, backend=o["backend"], **kw):
        sys.stderr.write("FAILED to train kernel!\n")
        return 255
    try:
        dump_kernel(conf.kernel, "kernel.opt")
    except IOError:
        sys.stderr.write("FAILED to open kernel.opt for WRITE!\n")
        return 255
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
