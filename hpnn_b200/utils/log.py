"""Verbosity-gated logging with the reference's prefixes (include/libhpnn.h:95-122).

Level 0..3: ``NN(WARN)`` needs >0, ``NN:`` / continuation output needs >1,
``NN(DBG)`` needs >2.  Under ``torch.distributed`` only rank 0 prints, like the
reference's ``_OUT`` under MPI (include/libhpnn/common.h:81-91)."""
from __future__ import annotations

import os
import sys

_verbose = 0


def set_verbose(v: int) -> None:
    global _verbose
    _verbose = max(0, min(3, int(v)))


def inc_verbose() -> None:
    set_verbose(_verbose + 1)


def dec_verbose() -> None:
    set_verbose(_verbose - 1)


def get_verbose() -> int:
    return _verbose


def _rank0() -> bool:
    return int(os.environ.get("RANK", "0")) == 0


def _emit(stream, text: str) -> None:
    if _rank0():
        stream.write(text)


def nn_dbg(text: str, stream=None) -> None:
    if _verbose > 2:
        _emit(stream or sys.stdout, "NN(DBG): " + text)


def nn_out(text: str, stream=None) -> None:
    if _verbose > 1:
        _emit(stream or sys.stdout, "NN: " + text)


def nn_cout(text: str, stream=None) -> None:
    if _verbose > 1:
        _emit(stream or sys.stdout, text)


def nn_warn(text: str, stream=None) -> None:
    if _verbose > 0:
        _emit(stream or sys.stdout, "NN(WARN): " + text)


def nn_error(text: str, stream=None) -> None:
    _emit(stream or sys.stderr, "NN(ERR): " + text)
