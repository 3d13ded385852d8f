"""``run_nn`` (tests/run_nn.c): load nn.conf (and its kernel), evaluate the test directory."""
from __future__ import annotations

import sys

from .. import _native as N
from ..config import load_conf
from ..trainer import run_kernel
from ..utils import log
from ._common import help_text, parse


def main(argv=None) -> int:
    argv = list(sys.argv if argv is None else argv)
    log.set_verbose(0)
    o = parse(argv, "run_nn", False)
    if o["help"]:
        sys.stdout.write(help_text("run_nn", False))
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
        run_kernel(conf, backend=o["backend"], batch=o["batch"], precision=o["precision"])
    except ValueError as e:  # an extension option that does not fit the chosen back-end
        log.nn_error(str(e) + "\n")
        return 255
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
