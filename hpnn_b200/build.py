"""In-tree build of the native core (``libhpnn_core.so``), the libhpnn.h shim
(``libhpnn.so``) and the C tools, all for sm_100a.

The reference drives its build with autotools and passes no ``-arch`` at all
(configure.ac:535); here one ``nvcc`` invocation per translation unit produces
objects for ``compute_100a/sm_100a`` only, linked into shared objects that live
inside the package so they travel with the source tree.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIBDIR = ROOT / "hpnn_b200" / "lib"
OBJDIR = ROOT / "build" / "obj"
BINDIR = ROOT / "build" / "bin"

NVCC = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-fopenmp,-Wall,-Wno-unused-function,-ffp-contract=off",
          f"-I{ROOT / 'include'}", f"-I{ROOT / 'csrc' / 'core'}", f"-I{ROOT / 'csrc' / 'kernels'}"]
COMMON += os.environ.get("HPNN_EXTRA_NVCC_FLAGS", "").split()  # experiments, e.g. -DHPNN_W_STREAM

CORE_SOURCES = [
    "csrc/core/io.cpp",
    "csrc/cpu/engine.cpp",
    "csrc/core/device_net.cu",
    "csrc/core/symm.cu",
    "csrc/kernels/train_persistent.cu",
    "csrc/kernels/train_resident.cu",
    "csrc/kernels/train_lazy.cu",
    "csrc/kernels/train_lazy1.cu",
    "csrc/kernels/ops.cu",
    "csrc/kernels/gemm_umma.cu",
    "csrc/kernels/dp_merge.cu",
    "csrc/kernels/noalloc.cu",
]
SHIM_SOURCES = ["csrc/capi/libhpnn.cpp"]
HEADERS = ["include/hpnn_core.h", "csrc/core/kernel_host.hpp", "csrc/kernels/dev_common.cuh", "csrc/kernels/kp_common.cuh",
           "csrc/kernels/lazy_common.cuh",
           "include/libhpnn.h", "include/libhpnn/common.h", "include/libhpnn/ann.h", "include/libhpnn/snn.h"]


def _digest(paths, extra=""):
    h = hashlib.sha256(extra.encode())
    for p in paths:
        p = ROOT / p
        if p.exists():
            h.update(p.read_bytes())
    return h.hexdigest()[:16]


def _run(cmd, verbose):
    if verbose:
        print("+", " ".join(str(c) for c in cmd), flush=True)
    res = subprocess.run([str(c) for c in cmd], capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError(f"build step failed: {' '.join(str(c) for c in cmd)}")
    if verbose and (res.stdout or res.stderr):
        sys.stderr.write(res.stdout + res.stderr)


def _compile(src, verbose, ptxas_v=False):
    src_path = ROOT / src
    if not src_path.exists():
        return None
    OBJDIR.mkdir(parents=True, exist_ok=True)
    tag = _digest([src] + HEADERS, extra=" ".join(COMMON + ARCH))
    stem = src.replace("/", "_")
    obj = OBJDIR / f"{stem}.{tag}.o"
    for old in OBJDIR.glob(f"{stem}.*.o"):  # objects of earlier versions of this source
        if old != obj:
            old.unlink()
    if obj.exists():
        return obj
    cmd = [NVCC, *ARCH, *COMMON, "-c", src_path, "-o", obj]
    if ptxas_v:
        cmd[1:1] = ["-Xptxas", "-v"]
    _run(cmd, verbose)
    return obj


def build(verbose: bool = False, force: bool = False, ptxas_v: bool = False) -> dict:
    """Compile everything that is out of date.  Returns the artefact paths."""
    if force and OBJDIR.exists():
        shutil.rmtree(OBJDIR)
    LIBDIR.mkdir(parents=True, exist_ok=True)
    BINDIR.mkdir(parents=True, exist_ok=True)
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        core_objs = [o for o in ex.map(lambda s: _compile(s, verbose, ptxas_v), CORE_SOURCES) if o]
        shim_objs = [o for o in ex.map(lambda s: _compile(s, verbose, ptxas_v), SHIM_SOURCES) if o]
    out = {}
    core = LIBDIR / "libhpnn_core.so"
    stamp = LIBDIR / ".core.stamp"
    want = " ".join(sorted(o.name for o in core_objs))
    if force or not core.exists() or not stamp.exists() or stamp.read_text() != want:
        _run([NVCC, *ARCH, "-shared", "-Xcompiler", "-fPIC", "-o", core, *core_objs, "-lgomp", "-lcuda"], verbose)
        stamp.write_text(want)
    out["core"] = core
    if shim_objs:
        shim = LIBDIR / "libhpnn.so"
        sstamp = LIBDIR / ".shim.stamp"
        want2 = want + " " + " ".join(sorted(o.name for o in shim_objs))
        if force or not shim.exists() or not sstamp.exists() or sstamp.read_text() != want2:
            _run([NVCC, *ARCH, "-shared", "-Xcompiler", "-fPIC", "-o", shim, *shim_objs, *core_objs,
                  "-lgomp", "-lcuda"], verbose)
            sstamp.write_text(want2)
        out["shim"] = shim
    out.update(build_tools(verbose))
    return out


def build_tools(verbose: bool = False) -> dict:
    """C drivers (train_nn / run_nn against the shim) and the dataset converters."""
    out = {}
    cc = shutil.which("gcc") or "gcc"
    cxx = shutil.which("g++") or "g++"
    shim = LIBDIR / "libhpnn.so"
    for name in ("train_nn", "run_nn"):
        src = ROOT / "tools" / f"{name}.c"
        if src.exists() and shim.exists():
            exe = BINDIR / name
            if not exe.exists() or exe.stat().st_mtime < max(src.stat().st_mtime, shim.stat().st_mtime):
                _run([cc, "-O2", "-D_CUDA", "-D_OMP", f"-I{ROOT / 'include'}", src, "-o", exe, f"-L{LIBDIR}", "-lhpnn",
                      f"-Wl,-rpath,{LIBDIR}", "-lm"], verbose)
            out[name] = exe
    for name in ("pmnist", "pdif"):
        src = ROOT / "tools" / f"{name}.cpp"
        if src.exists():
            exe = BINDIR / name
            if not exe.exists() or exe.stat().st_mtime < src.stat().st_mtime:
                _run([cxx, "-O2", "-std=c++17", f"-I{ROOT / 'include'}", f"-I{ROOT / 'tools'}", src, "-o", exe], verbose)
            out[name] = exe
    return out


if __name__ == "__main__":
    arts = build(verbose="-q" not in sys.argv, force="--force" in sys.argv, ptxas_v="--ptxas-v" in sys.argv)
    for k, v in arts.items():
        print(f"{k}: {v}")
