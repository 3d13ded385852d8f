"""Training / inference drivers: ``nn_train_kernel`` and ``nn_run_kernel`` re-done on top
of three interchangeable engines.

  * ``cuda``  -- :class:`HpnnNet` + the persistent sm_100a kernel (default when a GPU exists)
  * ``cpu``   -- the native C++/OpenMP engine (fallback and oracle)
  * ``torch`` -- the plain-PyTorch specification in :mod:`hpnn_b200.ops.reference`

The stdout lines are the reference's de-facto API (the tutorials ``grep`` them):
src/libhpnn.c:1230, src/ann.c:2322-2366, src/snn.c:1457-1498, src/libhpnn.c:1388-1514.
"""
from __future__ import annotations

import ctypes as C
import os
import sys
import time
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np
import torch

from . import _native as N
from .config import NNConf
from .kernel_io import Kernel
from .ops import reference as R
from .samples import SampleSet, read_sample, walk
from .utils import log


@dataclass
class SampleResult:
    init_err: float
    final_de: float
    last_err: float
    n_iter: int
    first_ok: bool
    success: bool

    @staticmethod
    def from_native(r) -> "SampleResult":
        return SampleResult(r.init_err, r.final_de, r.last_err, r.n_iter, bool(r.first_ok), bool(r.success))


def default_backend() -> str:
    env = os.environ.get("HPNN_BACKEND")
    if env:
        return env
    return "cuda" if torch.cuda.is_available() else "cpu"


# --------------------------------------------------------------------------- engines
class CpuEngine:
    """Native C++/OpenMP engine (csrc/cpu/engine.cpp); learning rates default to the
    reference's CPU values."""

    name = "cpu"

    def __init__(self, kernel: Kernel, nn_type: int, train: int, lr: Optional[float] = None, alpha: float = 0.2,
                 delta: float = -1.0, min_iter: int = -1, max_iter: int = -1, stable_softmax: bool = False,
                 cuda_defaults: bool = False, **_):
        self.lib = N.load()
        self.type, self.train_kind = nn_type, train
        self.h = kernel.to_native()
        self.name_ = kernel.name
        self.dims = list(kernel.dims)
        self.opts = N.TrainOpts()
        self.lib.hpnn_train_opts_default(C.byref(self.opts), nn_type, train, int(cuda_defaults))
        if lr is not None and lr > 0:
            self.opts.lr = lr
        self.opts.alpha = alpha
        if delta > 0:
            self.opts.delta = delta
        if min_iter >= 0:
            self.opts.min_iter = min_iter
        if max_iter >= 0:
            self.opts.max_iter = max_iter
        self.opts.stable_softmax = int(stable_softmax)

    def train(self, x: np.ndarray, t: np.ndarray) -> SampleResult:
        dp = C.POINTER(C.c_double)
        x = np.ascontiguousarray(x, dtype=np.float64)
        t = np.ascontiguousarray(t, dtype=np.float64)
        res = N.TrainResult()
        rc = self.lib.hpnn_cpu_train_sample(self.h, C.byref(self.opts), x.ctypes.data_as(dp), t.ctypes.data_as(dp),
                                            C.byref(res), None)
        if rc != 0:
            raise RuntimeError("hpnn_cpu_train_sample failed")
        return SampleResult.from_native(res)

    def run(self, x: np.ndarray) -> np.ndarray:
        dp = C.POINTER(C.c_double)
        x = np.ascontiguousarray(x, dtype=np.float64)
        out = np.zeros(self.dims[-1], dtype=np.float64)
        self.lib.hpnn_cpu_forward(self.h, self.type, self.opts.stable_softmax, x.ctypes.data_as(dp),
                                  out.ctypes.data_as(dp))
        return out

    def kernel(self) -> Kernel:
        k = Kernel.from_native(self.h, destroy=False)
        k.name = self.name_
        return k

    def close(self):
        if self.h:
            self.lib.hpnn_kernel_destroy(self.h)
            self.h = None


class TorchEngine:
    """The PyTorch specification as an engine (any device; slow, used for cross-checks)."""

    name = "torch"

    def __init__(self, kernel: Kernel, nn_type: int, train: int, lr: Optional[float] = None, alpha: float = 0.2,
                 delta: float = -1.0, min_iter: int = -1, max_iter: int = -1, stable_softmax: bool = False,
                 device: str = "cpu", cuda_defaults: bool = False, **_):
        self.type, self.train_kind = nn_type, train
        self.dev = torch.device(device)
        self.W = [torch.from_numpy(w.copy()).to(self.dev) for w in kernel.weights]
        self.dims, self.name_ = list(kernel.dims), kernel.name
        self.lr = lr if lr and lr > 0 else R.default_lr(nn_type, train, cuda_defaults)
        self.alpha = alpha
        self.delta = delta if delta > 0 else 1e-6
        self.min_iter = min_iter if min_iter >= 0 else None
        self.max_iter = max_iter if max_iter >= 0 else 102399
        self.stable = stable_softmax

    def train(self, x, t) -> SampleResult:
        xt = torch.as_tensor(np.asarray(x), dtype=torch.float64, device=self.dev)
        tt = torch.as_tensor(np.asarray(t), dtype=torch.float64, device=self.dev)
        r, _ = R.train_sample(self.W, xt, tt, self.type, self.train_kind, self.lr, self.alpha, self.delta,
                              self.min_iter, self.max_iter, self.stable)
        return SampleResult(r.init_err, r.final_de, r.last_err, r.n_iter, r.first_ok, r.success)

    def run(self, x) -> np.ndarray:
        xt = torch.as_tensor(np.asarray(x), dtype=torch.float64, device=self.dev)
        return R.forward(self.W, xt, self.type, self.stable)[-1].cpu().numpy()

    def kernel(self) -> Kernel:
        return Kernel(list(self.dims), [w.cpu().numpy().copy() for w in self.W], self.name_)

    def close(self):
        pass


class CudaEngine:
    """:class:`HpnnNet` + K-P.  One cooperative launch per sample; samples are streamed from
    pinned host memory and the per-sample results are read back in one small D2H copy."""

    name = "cuda"

    def __init__(self, kernel: Kernel, nn_type: int, train: int, lr: Optional[float] = None, alpha: float = 0.2,
                 delta: float = -1.0, min_iter: int = -1, max_iter: int = -1, stable_softmax: bool = False,
                 group=None, device=None, split: str = "auto", **_):
        from .models.net import HpnnNet

        self.type, self.train_kind = nn_type, train
        # Row-splitting pays only when streaming the weights dominates an iteration (1.6 GB net: 3.3x on 8
        # GPUs; 14 MB and 1.6 MB nets: ~2x SLOWER on 2..8 GPUs than on one, every hand-over that crosses
        # NVLink costs several L2 hand-overs, profiles/r2_probe8_small.jsonl).  split="auto": below
        # HPNN_TP_MIN_MB (default 128 MB of weights) every rank trains a full replica on its own GPU (all
        # ranks see the same samples, so the replicas stay identical); "always" / "never" force it.
        if group is not None and split != "always":
            nbytes = 8 * sum(a * b for a, b in zip(kernel.dims[:-1], kernel.dims[1:]))
            if split == "never" or nbytes < (int(os.environ.get("HPNN_TP_MIN_MB", "128")) << 20):
                group = None
        self.net = HpnnNet(kernel, nn_type, device=device, momentum=(train == N.TRAIN_BPM), group=group)
        self.opts = self.net.make_opts(train, lr, alpha, delta, min_iter, max_iter, stable_softmax)
        self.stable = stable_softmax
        n_in, n_out = kernel.dims[0], kernel.dims[-1]
        self._x = torch.zeros(n_in, dtype=torch.float64).pin_memory()
        self._t = torch.zeros(n_out, dtype=torch.float64).pin_memory()

    def train(self, x, t) -> SampleResult:
        self._x.copy_(torch.as_tensor(np.asarray(x), dtype=torch.float64))
        self._t.copy_(torch.as_tensor(np.asarray(t), dtype=torch.float64))
        return SampleResult.from_native(self.net.train_sample(self._x, self._t, self.opts))

    def train_many(self, X: torch.Tensor, T: torch.Tensor, sync_each: bool = False) -> List[SampleResult]:
        """Train on the rows of pinned ``X``/``T`` in order.  Without ``sync_each`` the whole
        pass is enqueued back to back (the host never waits for the device in between)."""
        out: List[SampleResult] = []
        n = X.shape[0]
        slots = self.net._n_slots
        for base in range(0, n, slots):
            m = min(slots, n - base)
            for i in range(m):
                self.net.load_sample(X[base + i], T[base + i])
                self.net.train_sample_async(self.opts, i)
                if sync_each:
                    torch.cuda.current_stream(self.net.device).synchronize()
            out += [SampleResult.from_native(r) for r in self.net.fetch_results(m)]
        return out

    def run(self, x) -> np.ndarray:
        self._x.copy_(torch.as_tensor(np.asarray(x), dtype=torch.float64))
        o = self.net.run(self._x, None, self.stable)
        return o.cpu().numpy()

    def kernel(self) -> Kernel:
        return self.net.to_kernel()

    def close(self):
        self.net.close()


def make_engine(backend: str, kernel: Kernel, nn_type: int, train: int, **kw):
    backend = backend or default_backend()
    if backend == "cuda":
        return CudaEngine(kernel, nn_type, train, **kw)
    if backend == "cpu":
        return CpuEngine(kernel, nn_type, train, **kw)
    if backend.startswith("logical"):  # "logical" or "logical:P": multi-launch path, P logical ranks
        from .parallel.logical import LogicalRanksEngine
        world = int(backend.split(":")[1]) if ":" in backend else kw.pop("world", 1)
        return LogicalRanksEngine(kernel, nn_type, train, world=world, **kw)
    if backend.startswith("virtual"):  # "virtual:P": P ranks of the persistent kernel in one cooperative launch
        from .parallel.virtual import VirtualRanksEngine
        world = int(backend.split(":")[1]) if ":" in backend else kw.pop("world", 2)
        return VirtualRanksEngine(kernel, nn_type, train, world=world, **kw)
    if backend.startswith("torch"):
        dev = kw.pop("device", None) or ("cuda" if backend == "torch-cuda" else "cpu")
        return TorchEngine(kernel, nn_type, train, device=dev, **kw)
    raise ValueError(f"unknown backend {backend!r}")


# --------------------------------------------------------------------------- log lines
def _line_training_file(path: str) -> str:
    return "TRAINING FILE: %16.16s\t" % os.path.basename(path)


def format_train_result(r: SampleResult, nn_type: int, train: int) -> str:
    s = " init=%15.10f" % r.init_err
    s += " OK" if r.first_ok else " NO"
    s += " N_ITER=%8i" % r.n_iter
    s += " final=%15.10f" % r.final_de
    if nn_type == N.TYPE_SNN and train == N.TRAIN_BP:
        return s + "\n"  # snn_train_BP prints no verdict (src/snn.c:1496-1497)
    return s + (" SUCCESS!\n" if r.success else " FAIL!\n")


def classify(out: np.ndarray, target: np.ndarray, nn_type: int):
    """``(guess, wanted)`` exactly as ``nn_run_kernel`` derives them (src/libhpnn.c:1443-1509)."""
    if nn_type == N.TYPE_SNN:
        res, guess, want = 0.0, 0, 0
        for i, v in enumerate(out.tolist()):
            if v > res:
                res, guess = v, i
            if target[i] > 0.1:
                want = i
        return guess, want, res
    res, guess, want = -1.0, len(out), 1  # is_ok starts as TRUE (=1) in the reference
    for i, v in enumerate(out.tolist()):
        if res < v:
            res, guess = v, i
        if target[i] > 0.5:
            want = i
    return guess, want, res


# --------------------------------------------------------------------------- drivers
class _Nvtx:
    """NVTX range when CUDA is up (shows the per-sample / per-pass structure in Nsight)."""

    def __init__(self, name: str):
        self.name, self.on = name, torch.cuda.is_available()

    def __enter__(self):
        if self.on:
            torch.cuda.nvtx.range_push(self.name)

    def __exit__(self, *a):
        if self.on:
            torch.cuda.nvtx.range_pop()


def _write_metrics(path: Optional[str], record: dict) -> None:
    if path:
        import json

        with open(path, "a") as f:
            f.write(json.dumps(record) + "\n")


def train_kernel(conf: NNConf, backend: Optional[str] = None, stream=None, pipeline: bool = True,
                 checkpoint_every: int = 0, checkpoint_path: str = "kernel.ckpt", metrics_path: Optional[str] = None,
                 **engine_kw) -> bool:
    """``nn_train_kernel`` (src/libhpnn.c:1149-1305): one pass over ``conf.samples`` in the
    reference's pseudo-random order; ``conf.kernel`` is replaced by the trained kernel.

    Beyond the reference (which loses the whole pass on a crash): ``checkpoint_every=n``
    dumps the kernel (text + exact binary sidecar) every n samples, ``metrics_path`` appends one
    JSON record per sample (``HPNN_METRICS`` env var as default)."""
    metrics_path = metrics_path or os.environ.get("HPNN_METRICS")
    out = stream or sys.stdout
    if conf.kernel is None or conf.samples is None or conf.type == N.TYPE_UKN:
        return False
    if conf.type not in (N.TYPE_ANN, N.TYPE_SNN):
        log.nn_error("unimplemented NN type!\n", out)
        return False
    if not os.path.isdir(conf.samples):
        log.nn_error("can't open sample directory: %s\n" % conf.samples)
        return False
    if conf.seed == 0:
        conf.seed = int(time.time())
    files = walk(conf.samples, conf.seed)
    if conf.train not in (N.TRAIN_BP, N.TRAIN_BPM):
        for f in files:  # CG / SPLX are enum values only in the reference: nothing is trained
            log.nn_out(_line_training_file(f), out)
        return True
    backend = backend or default_backend()
    eng = make_engine(backend, conf.kernel, conf.type, conf.train, **engine_kw)
    try:
        if backend == "cuda" and pipeline and not checkpoint_every:
            sset = SampleSet(files)
            results = eng.train_many(sset.inputs, sset.targets) if len(sset) else []
            by_path = dict(zip(sset.paths, results))
            for f in files:
                log.nn_out(_line_training_file(f), out)
                if f in by_path:
                    log.nn_cout(format_train_result(by_path[f], conf.type, conf.train), out)
                    _write_metrics(metrics_path, {"file": os.path.basename(f), **by_path[f].__dict__})
                    if by_path[f].final_de > 0.1:
                        log.nn_dbg("bad optimization!\n", out)
        else:
            n_done = 0
            for f in files:
                log.nn_out(_line_training_file(f), out)
                try:
                    x, t = read_sample(f)
                except IOError:
                    continue
                with _Nvtx("train_sample " + os.path.basename(f)):
                    r = eng.train(x, t)
                log.nn_cout(format_train_result(r, conf.type, conf.train), out)
                _write_metrics(metrics_path, {"file": os.path.basename(f), **r.__dict__})
                if r.final_de > 0.1:
                    log.nn_dbg("bad optimization!\n", out)
                out.flush()
                n_done += 1
                if checkpoint_every and n_done % checkpoint_every == 0:
                    from .kernel_io import dump_kernel, dump_kernel_binary

                    ck = eng.kernel()
                    if int(os.environ.get("RANK", "0")) == 0:
                        dump_kernel(ck, checkpoint_path)
                        dump_kernel_binary(ck, checkpoint_path + ".bin")
        conf.kernel = eng.kernel()
    finally:
        eng.close()
    return True


def classify_batch(Y: np.ndarray, T: np.ndarray, nn_type: int):
    """:func:`classify` for a whole matrix of outputs at once: ``(guess, want, res)`` arrays."""
    n = Y.shape[1]
    Yc = np.where(np.isnan(Y), -np.inf, Y)
    am = Yc.argmax(axis=1)  # first maximum, like the reference's strict '<' scan
    mx = Yc[np.arange(Y.shape[0]), am]
    if nn_type == N.TYPE_SNN:
        hit = mx > 0.0
        guess, res = np.where(hit, am, 0), np.where(hit, mx, 0.0)
        mask, default = T > 0.1, 0
    else:
        hit = mx > -1.0
        guess, res = np.where(hit, am, n), np.where(hit, mx, -1.0)
        mask, default = T > 0.5, 1  # is_ok starts as TRUE (=1) in the reference
    last = n - 1 - mask[:, ::-1].argmax(axis=1)  # the scan keeps the LAST index above the threshold
    want = np.where(mask.any(axis=1), last, default)
    return guess, want, res


def run_kernel(conf: NNConf, backend: Optional[str] = None, stream=None, batch: int = -1, precision: str = "fp64",
               **engine_kw) -> dict:
    """``nn_run_kernel`` (src/libhpnn.c:1306-1536).  Returns ``{"pass": n, "total": n}``.

    The reference evaluates one file per launch (and so does ``batch=0``).  By default
    (``batch=-1``) the CUDA back-end parses the whole directory with the native parallel reader and
    pushes it through the batched FP64 forward ``HPNN_RUN_BATCH`` (1024) samples at a time; verdicts
    and output lines are identical, only produced ~20x faster.  Verbosity 3 (per-class tables) and
    the other back-ends keep the per-file loop.  ``precision="bf16"`` (extension, batched CUDA path only)
    evaluates the layers on the tcgen05 tensor cores (bf16 operands, fp32 accumulation, FP64 outputs): the
    probabilities agree with FP64 to about 1e-3, a verdict can differ where two classes are that close."""
    if precision not in ("fp64", "bf16"):
        raise ValueError(f"unknown precision {precision!r}")
    out = stream or sys.stdout
    stats = {"pass": 0, "total": 0}
    if conf.kernel is None or conf.tests is None or conf.type not in (N.TYPE_ANN, N.TYPE_SNN):
        return stats
    if not os.path.isdir(conf.tests):
        log.nn_error("can't open test directory: %s\n" % conf.tests)
        return stats
    if conf.seed == 0:
        conf.seed = int(time.time())
    files = walk(conf.tests, conf.seed)
    backend = backend or default_backend()
    if batch < 0:
        batch = int(os.environ.get("HPNN_RUN_BATCH", "1024")) if (backend == "cuda" and len(files) > 1) else 0
    if precision != "fp64" and not (batch > 0 and backend == "cuda"):
        raise ValueError("precision='bf16' needs the batched CUDA path (backend cuda, batch > 0)")
    eng = make_engine(backend, conf.kernel, conf.type, N.TRAIN_BP, **engine_kw)
    try:
        outputs = {}
        if batch > 0 and backend == "cuda" and getattr(eng.net, "world", 1) == 1:
            sset = SampleSet(files)
            Ys = []
            for b0 in range(0, len(sset), batch):
                X = sset.inputs[b0:b0 + batch].to(eng.net.device, non_blocking=True)
                Ys.append(eng.net.forward_batch(X, engine_kw.get("stable_softmax", False), precision=precision))
            if Ys and log.get_verbose() < 3 and len(sset) == len(files):
                # whole-directory fast path: vectorised verdicts, one write for all lines
                Y = torch.cat(Ys).cpu().numpy()
                T = sset.targets.numpy()
                guess, want, res = classify_batch(Y, T, conf.type)
                ok = guess == want
                stats["total"], stats["pass"] = len(files), int(ok.sum())
                if log.get_verbose() > 1:
                    snn = conf.type == N.TYPE_SNN
                    lines = []
                    for i, p in enumerate(sset.paths):
                        line = "NN: TESTING FILE: %16.16s\t" % os.path.basename(p)
                        if snn:
                            line += " BEST CLASS idx=%i P=%15.10f" % (guess[i] + 1, res[i] * 100.0)
                        line += " [PASS]\n" if ok[i] else " [FAIL idx=%i]\n" % (want[i] + 1)
                        lines.append(line)
                    log.nn_cout("".join(lines), out)
                    out.flush()
                return stats
            if Ys:
                Y = torch.cat(Ys).cpu().numpy()
                for i, p in enumerate(sset.paths):
                    outputs[p] = (Y[i], sset.targets[i].numpy())
        for f in files:
            log.nn_out("TESTING FILE: %16.16s\t" % os.path.basename(f), out)
            if f in outputs:
                o, t = outputs[f]
            else:
                try:
                    x, t = read_sample(f)
                except IOError:
                    continue
                o = eng.run(x)
            guess, want, res = classify(o, t, conf.type)
            if conf.type == N.TYPE_SNN:
                log.nn_dbg(" CLASS | PROBABILITY (%)\n", out)
                log.nn_dbg("-------|----------------\n", out)
                for i, v in enumerate(o.tolist()):
                    log.nn_dbg(" %5i | %15.10f\n" % (i + 1, v * 100.0), out)
                log.nn_dbg("-------|----------------\n", out)
                log.nn_cout(" BEST CLASS idx=%i P=%15.10f" % (guess + 1, res * 100.0), out)
            stats["total"] += 1
            if guess == want:
                stats["pass"] += 1
                log.nn_cout(" [PASS]\n", out)
            else:
                log.nn_cout(" [FAIL idx=%i]\n" % (want + 1), out)
            out.flush()
    finally:
        eng.close()
    return stats
