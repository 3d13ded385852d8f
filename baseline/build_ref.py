"""Stage and build the UNMODIFIED reference (ovhpa/hpnn) under ``baseline/_ref``.

``pip install /root/reference`` is not possible (plain C/autotools project: "Neither
'setup.py' nor 'pyproject.toml' found"), and autotools, MPI and CBLAS are not on the image,
so the sources are copied verbatim and compiled with the commands of SURVEY 6.2:

  * ``cuda``  : -D_OMP -D_CUDA -D_CUBLAS, nvcc for sm_100 (the reference passes no -arch at
                all, configure.ac:535)  -> libhpnn_ref_cuda.so, train_nn_cuda, run_nn_cuda
  * ``cpu``   : -D_OMP (no BLAS)        -> libhpnn_ref_cpu.so,  train_nn_cpu,  run_nn_cpu

``baseline/_ref`` is git-ignored (the reference is GPLv3 third-party code, not product
source) but travels to the GPU box with gpurun.  Nothing here is imported by hpnn_b200.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF_SRC = Path(os.environ.get("HPNN_REFERENCE", "/root/reference"))
DST = HERE / "_ref"
CUDA = Path(os.environ.get("CUDA_HOME", "/usr/local/cuda"))


def _run(cmd, cwd=None):
    r = subprocess.run([str(c) for c in cmd], cwd=cwd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("reference build step failed: " + " ".join(str(c) for c in cmd))


def stage() -> Path:
    src = DST / "src_tree"
    if not src.exists():
        if not REF_SRC.exists():
            raise FileNotFoundError(f"reference sources not found at {REF_SRC}")
        DST.mkdir(parents=True, exist_ok=True)
        shutil.copytree(REF_SRC, src, ignore=shutil.ignore_patterns(".git"))
    return src


def build(flavours=("cuda", "cpu")) -> dict:
    src = stage()
    out = DST / "build"
    out.mkdir(parents=True, exist_ok=True)
    inc = src / "include"
    arts = {}
    for fl in flavours:
        lib = out / f"libhpnn_ref_{fl}.so"
        if lib.exists() and (out / f"train_nn_{fl}").exists():
            arts[fl] = lib
            continue
        defs = ["-D_OMP"] + (["-D_CUDA", "-D_CUBLAS"] if fl == "cuda" else [])
        objs = []
        for f in ("libhpnn", "ann", "snn"):
            o = out / f"{f}_{fl}.o"
            cmd = ["gcc", "-O2", "-fPIC", "-fopenmp", *defs, f"-I{inc}", "-w", "-c", src / "src" / f"{f}.c", "-o", o]
            if fl == "cuda":
                cmd.insert(-4, f"-I{CUDA / 'include'}")
            _run(cmd)
            objs.append(o)
        link = ["-lm", "-lgomp"]
        if fl == "cuda":
            for f in ("cuda_ann", "cuda_snn"):
                o = out / f"{f}.o"
                # the reference's own NVCCFLAGS (configure.ac:535) + an explicit Blackwell target
                _run([CUDA / "bin" / "nvcc", "-O2", "-w", "--default-stream", "per-thread", "--use_fast_math",
                      "-lineinfo", "-D_CUDA", "-D_CUBLAS", "-D_OMP", "-Xcompiler", "-fPIC,-fopenmp",
                      "-gencode", "arch=compute_100,code=sm_100", f"-I{inc}", "-c", src / "src" / f"{f}.cu", "-o", o])
                objs.append(o)
            link += [f"-L{CUDA / 'lib64'}", "-lcudart", "-lcublas", f"-Wl,-rpath,{CUDA / 'lib64'}"]
        _run(["g++", "-shared", "-fopenmp", "-o", lib, *objs, *link])
        for drv in ("train_nn", "run_nn"):
            _run(["gcc", "-O2", "-fopenmp", *defs, f"-I{inc}", f"-I{CUDA / 'include'}", "-w",
                  src / "tests" / f"{drv}.c", "-o", out / f"{drv}_{fl}", f"-L{out}", f"-lhpnn_ref_{fl}",
                  f"-Wl,-rpath,{out}", *link])
        arts[fl] = lib
    return arts


if __name__ == "__main__":
    for k, v in build().items():
        print(f"{k}: {v}")
