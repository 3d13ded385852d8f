"""Network ("kernel") objects and the ``kernel.tmp`` / ``kernel.opt`` text format.

Parsing, dumping and the glibc-``random()`` weight initialisation are done by the
native core (``csrc/core/io.cpp``) so that Python, the C shim and the C tools share
one implementation.  Reference behaviour: src/ann.c:206-631 (load), 632-766
(generate), 770-857 (dump).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np

from . import _native as N


@dataclass
class Kernel:
    """Topology + fp64 weights on the host.  ``weights[l]`` is ``[dims[l+1], dims[l]]``."""

    dims: List[int]
    weights: List[np.ndarray] = field(default_factory=list)
    name: Optional[str] = None

    def __post_init__(self):
        if not self.weights:
            self.weights = [np.zeros((self.dims[l + 1], self.dims[l]), dtype=np.float64)
                            for l in range(len(self.dims) - 1)]

    @property
    def n_layers(self) -> int:
        return len(self.dims) - 1

    @property
    def n_inputs(self) -> int:
        return self.dims[0]

    @property
    def n_outputs(self) -> int:
        return self.dims[-1]

    @property
    def hiddens(self) -> List[int]:
        return list(self.dims[1:-1])

    def copy(self) -> "Kernel":
        return Kernel(list(self.dims), [w.copy() for w in self.weights], self.name)

    # ---- native handle round trip -------------------------------------------------
    def to_native(self):
        lib = N.load()
        dims = (C.c_int * len(self.dims))(*self.dims)
        h = lib.hpnn_kernel_create(len(self.dims), dims)
        if not h:
            raise ValueError(f"invalid topology {self.dims}")
        for l, w in enumerate(self.weights):
            ptr = lib.hpnn_kernel_weights(h, l)
            dst = np.ctypeslib.as_array(ptr, shape=(w.shape[0] * w.shape[1],))
            dst[:] = np.ascontiguousarray(w, dtype=np.float64).reshape(-1)
        if self.name is not None:
            lib.hpnn_kernel_set_name(h, self.name.encode())
        return h

    @staticmethod
    def from_native(h, destroy: bool = True) -> "Kernel":
        lib = N.load()
        nl = lib.hpnn_kernel_n_layers(h)
        dims = [lib.hpnn_kernel_dim(h, i) for i in range(nl + 1)]
        ws = []
        for l in range(nl):
            ptr = lib.hpnn_kernel_weights(h, l)
            a = np.ctypeslib.as_array(ptr, shape=(dims[l + 1] * dims[l],))
            ws.append(a.reshape(dims[l + 1], dims[l]).copy())
        nm = lib.hpnn_kernel_name(h)
        k = Kernel(dims, ws, nm.decode() if nm is not None else None)
        if destroy:
            lib.hpnn_kernel_destroy(h)
        return k


def generate_kernel(seed: int, n_inputs: int, hiddens: Sequence[int], n_outputs: int):
    """Random kernel, bit-identical to ``ann_generate`` for the same seed (glibc ``random()``).

    Returns ``(kernel, seed_used)``; seed 0 means ``time(NULL)`` as in the reference."""
    lib = N.load()
    dims = [int(n_inputs), *[int(h) for h in hiddens], int(n_outputs)]
    cdims = (C.c_int * len(dims))(*dims)
    s = C.c_uint(int(seed))
    h = lib.hpnn_kernel_generate(C.byref(s), len(dims), cdims)
    if not h:
        raise ValueError(f"invalid topology {dims}")
    return Kernel.from_native(h), int(s.value)


def load_kernel(path: str) -> Kernel:
    h = N.load().hpnn_kernel_load(str(path).encode())
    if not h:
        raise IOError(f"cannot load kernel file {path}")
    return Kernel.from_native(h)


def dump_kernel(kernel: Kernel, path: str) -> None:
    lib = N.load()
    h = kernel.to_native()
    try:
        if lib.hpnn_kernel_dump_path(h, str(path).encode()) != 0:
            raise IOError(f"cannot write kernel file {path}")
    finally:
        lib.hpnn_kernel_destroy(h)


def dump_kernel_binary(kernel: Kernel, path: str) -> None:
    """Exact fp64 sidecar (the text format rounds to 15 decimals)."""
    lib = N.load()
    h = kernel.to_native()
    try:
        if lib.hpnn_kernel_dump_binary(h, str(path).encode()) != 0:
            raise IOError(f"cannot write {path}")
    finally:
        lib.hpnn_kernel_destroy(h)


def load_kernel_binary(path: str) -> Kernel:
    h = N.load().hpnn_kernel_load_binary(str(path).encode())
    if not h:
        raise IOError(f"cannot load binary kernel {path}")
    return Kernel.from_native(h)
