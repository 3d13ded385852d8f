"""Argument handling shared by ``train_nn`` and ``run_nn``.

The flags are the reference's (tests/train_nn.c:33-58,83-210): ``-h -v -x -O n -B n -S n``
with ``-On`` / ``-O n`` both accepted, switches combinable (``-vvv``) except the numeric
ones, and the first non-switch argument taken as the configuration file."""
from __future__ import annotations

import sys
from typing import List, Tuple

from ..utils import log


def help_text(prog: str, train: bool) -> str:
    bar = "***********************************\n"
    s = bar + f"usage:  {prog} [-options] [input]\n" + bar + "options:\n"
    s += "-h \tdisplay this help;\n-v \tincrease verbosity;\n"
    if train:
        s += "-x \tdiscard results.\n"
    s += "-O \tnumber of openMP threads.\n-B \tnumber of BLAS threads (MKL).\n-S \tnumber of CUDA streams.\n"
    s += "--backend cuda|cpu|torch \tengine (extension).\n--lr x \tlearning rate override (extension).\n"
    if not train:
        s += "--batch n \tsamples per launch, batched GEMM path (extension).\n"
        s += "--precision fp64|bf16 \tbatched path: FP64 (default) or tensor-core bf16 (extension).\n"
    s += bar + "input:     neural network .def file\ncontains the network definition and\n"
    s += "topology. May contain weight values\nor context for a random generation.\n" + bar
    return s


def parse(argv: List[str], prog: str, train: bool):
    opts = {"omp": None, "blas": None, "streams": 1, "dry": False, "backend": None, "lr": None, "batch": -1,
            "precision": "fp64", "conf": None, "help": False}
    i = 1

    def bad(flag):
        sys.stderr.write(f"syntax error: bad -{flag} parameter!\n")
        sys.stdout.write(help_text(prog, train))
        raise SystemExit(255)

    while i < len(argv):
        a = argv[i]
        if a.startswith("--"):
            key = a[2:]
            if key in ("backend", "lr", "batch", "precision") and i + 1 < len(argv) and (train is False or key != "precision"):
                val = argv[i + 1]
                opts[key] = val if key in ("backend", "precision") else (float(val) if key == "lr" else int(val))
                i += 2
                continue
            sys.stderr.write("syntax error: unrecognized option!\n")
            sys.stdout.write(help_text(prog, train))
            raise SystemExit(255)
        if a.startswith("-"):
            j = 1
            while j < len(a):
                ch = a[j]
                if ch == "h":
                    opts["help"] = True
                    return opts
                if ch == "v":
                    log.inc_verbose()
                    j += 1
                elif ch == "x" and train:
                    opts["dry"] = True  # nn_toggle_dry is a no-op in the reference
                    j += 1
                elif ch in "OBS":
                    rest = a[j + 1:]
                    if rest == "":
                        i += 1
                        rest = argv[i] if i < len(argv) else ""
                    digits = ""
                    for c in rest.strip():
                        if c.isdigit():
                            digits += c
                        else:
                            break
                    if not digits or int(digits) == 0:
                        bad(ch)
                    opts[{"O": "omp", "B": "blas", "S": "streams"}[ch]] = int(digits)
                    break  # no combination allowed after a numeric switch
                else:
                    sys.stderr.write("syntax error: unrecognized option!\n")
                    sys.stdout.write(help_text(prog, train))
                    raise SystemExit(255)
        else:
            if opts["conf"] is not None:
                raise SystemExit(255)
            opts["conf"] = a
        i += 1
    if opts["conf"] is None:
        opts["conf"] = "./nn.conf"
    return opts
