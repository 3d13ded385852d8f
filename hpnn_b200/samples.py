"""Sample files and the reference's directory walk.

Reference: ``nn_read_sample`` src/libhpnn.c:1070-1145; listing + shuffle
src/libhpnn.c:1184-1229 (``readdir`` order, dot-files skipped, then
``srandom(seed)`` and rejection draws without replacement)."""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import _native as N


def read_sample(path: str) -> Tuple[np.ndarray, np.ndarray]:
    """Return ``(input, target)`` fp64 vectors of one sample file."""
    lib = N.load()
    pin, pout = C.POINTER(C.c_double)(), C.POINTER(C.c_double)()
    nin, nout = C.c_int(0), C.c_int(0)
    rc = lib.hpnn_sample_read(str(path).encode(), C.byref(pin), C.byref(nin), C.byref(pout), C.byref(nout))
    if rc != 0 or not pin or not pout:
        if pin:
            lib.hpnn_free(pin)
        if pout:
            lib.hpnn_free(pout)
        raise IOError(f"sample {path} read failed!")
    try:
        x = np.ctypeslib.as_array(pin, shape=(nin.value,)).copy()
        t = np.ctypeslib.as_array(pout, shape=(nout.value,)).copy()
    finally:
        lib.hpnn_free(pin)
        lib.hpnn_free(pout)
    return x, t


def write_sample(path: str, x: Sequence[float], t: Sequence[float], comment: Optional[str] = None) -> None:
    """Writer compatible with pmnist/pdif output (``%7.5f`` inputs, ``%.1f`` targets)."""
    lib = N.load()
    xa = np.ascontiguousarray(x, dtype=np.float64)
    ta = np.ascontiguousarray(t, dtype=np.float64)
    dp = C.POINTER(C.c_double)
    rc = lib.hpnn_sample_write(str(path).encode(), xa.ctypes.data_as(dp), xa.size, ta.ctypes.data_as(dp), ta.size,
                               comment.encode() if comment is not None else None)
    if rc != 0:
        raise IOError(f"cannot write sample {path}")


def list_dir(directory: str) -> List[str]:
    """File names in raw ``readdir`` order (never sorted: the order feeds the shuffle)."""
    lib = N.load()
    names = C.POINTER(C.c_char_p)()
    n = C.c_int(0)
    if lib.hpnn_dir_list(str(directory).encode(), C.byref(names), C.byref(n)) != 0:
        raise IOError(f"can't open sample directory: {directory}")
    try:
        return [names[i].decode() for i in range(n.value)]
    finally:
        lib.hpnn_dir_list_free(names, n.value)


def shuffle_order(seed: int, n: int) -> List[int]:
    """Visit order of ``n`` files for ``seed`` -- the reference's glibc-random draw."""
    if n <= 0:
        return []
    lib = N.load()
    order = (C.c_int * n)()
    lib.hpnn_shuffle_order(C.c_uint(int(seed) & 0xFFFFFFFF), n, order)
    return list(order)


def walk(directory: str, seed: int) -> List[str]:
    """Full paths of the files of ``directory`` in the order the reference visits them."""
    names = list_dir(directory)
    return [os.path.join(directory, names[i]) for i in shuffle_order(seed, len(names))]


class SampleSet:
    """Samples of a directory parsed once into two dense fp64 matrices (pinned when CUDA
    is available) so that a training pass streams them to the GPU without re-parsing."""

    def __init__(self, paths: Sequence[str]):
        import torch

        paths = list(paths)
        self.paths: List[str] = []
        self.inputs = torch.empty(0, 0, dtype=torch.float64)
        self.targets = torch.empty(0, 0, dtype=torch.float64)
        first = None
        for p in paths:  # the first readable file fixes the shape
            try:
                first = read_sample(p)
                break
            except IOError:
                continue
        if first is None:
            return
        n_in, n_out = first[0].size, first[1].size
        pin = torch.cuda.is_available()
        X = torch.zeros((len(paths), n_in), dtype=torch.float64, pin_memory=pin)
        T = torch.zeros((len(paths), n_out), dtype=torch.float64, pin_memory=pin)
        ok = np.zeros(len(paths), dtype=np.uint8)
        arr = (C.c_char_p * len(paths))(*[p.encode() for p in paths])
        N.load().hpnn_sample_read_many(arr, len(paths), n_in, n_out, X.data_ptr(), T.data_ptr(), ok.ctypes.data)
        if ok.all():
            self.paths, self.inputs, self.targets = paths, X, T
        else:  # the reference skips unreadable samples silently
            keep = torch.from_numpy(ok.astype(bool))
            self.paths = [p for p, k in zip(paths, ok) if k]
            self.inputs, self.targets = X[keep], T[keep]
            if pin:
                self.inputs, self.targets = self.inputs.pin_memory(), self.targets.pin_memory()

    def __len__(self):
        return len(self.paths)
