"""Why does the unmodified reference abort on more than one GPU?  

Diagnostic only, never part of a benchmark arm: the staged reference tree is COPIED to a scratch
directory, one header line is changed there so that the reference's own ``CHK_ERR(label)`` checks
(compiled out unless DEBUG, include/libhpnn/common.h:324-335) stay active without the rest of the
DEBUG build (which fakes three GPUs on device 0), and ``train_nn`` is run on a small net with
``CUDA_LAUNCH_BLOCKING=1`` so that the first failing call names itself.

    python baseline/diag_ref_multigpu.py [n_gpus] > gpurun_out/ref_diag.log
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import tempfile
from pathlib import Path

HERE = Path(__file__).resolve().parent
SRC = HERE / "_ref" / "src_tree"
CUDA = Path(os.environ.get("CUDA_HOME", "/usr/local/cuda"))


def sh(cmd, **kw):
    r = subprocess.run([str(c) for c in cmd], capture_output=True, text=True, **kw)
    return r.returncode, r.stdout, r.stderr


def build(tree: Path, out: Path, cublas: bool):
    inc = tree / "include"
    defs = ["-D_OMP", "-D_CUDA"] + (["-D_CUBLAS"] if cublas else [])
    objs = []
    for f in ("libhpnn", "ann", "snn"):
        o = out / f"{f}.o"
        rc, so, se = sh(["gcc", "-O1", "-g", "-fPIC", "-fopenmp", *defs, f"-I{inc}", f"-I{CUDA / 'include'}", "-w", "-c",
                         tree / "src" / f"{f}.c", "-o", o])
        assert rc == 0, se
        objs.append(o)
    for f in ("cuda_ann", "cuda_snn"):
        o = out / f"{f}.o"
        rc, so, se = sh([CUDA / "bin" / "nvcc", "-O1", "-w", "--default-stream", "per-thread", "--use_fast_math", "-lineinfo",
                         *defs, "-Xcompiler", "-fPIC,-fopenmp", "-gencode", "arch=compute_100,code=sm_100", f"-I{inc}",
                         "-c", tree / "src" / f"{f}.cu", "-o", o])
        assert rc == 0, se
        objs.append(o)
    exe = out / "train_nn"
    link = ["-lm", "-lgomp", f"-L{CUDA / 'lib64'}", "-lcudart"] + (["-lcublas"] if cublas else []) + ["-lstdc++"]
    rc, so, se = sh(["g++", "-fopenmp", *defs, f"-I{inc}", f"-I{CUDA / 'include'}", "-w", "-x", "c",
                     tree / "tests" / "train_nn.c", "-x", "none", *objs, "-o", exe, *link])
    assert rc == 0, se
    return exe


def main():
    n_gpus = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    work = Path(tempfile.mkdtemp(prefix="hpnn_ref_diag_"))
    tree = work / "tree"
    shutil.copytree(SRC, tree)
    hdr = tree / "include" / "libhpnn" / "common.h"
    text = hdr.read_text()
    marker = "#ifdef   DEBUG\n#define CHK_ERR(func)"
    assert marker in text, "CHK_ERR block not found"
    hdr.write_text(text.replace(marker, "#if 1 /* diagnostic copy: keep the checks */\n#define CHK_ERR(func)"))
    # data: a small ANN, two samples
    import random

    random.seed(1)
    d = work / "run"
    (d / "samples").mkdir(parents=True)
    for i in range(2):
        x = " ".join("%7.5f" % random.random() for _ in range(64))
        t = " ".join("1.0" if j == i else "-1.0" for j in range(10))
        (d / "samples" / f"s{i}.txt").write_text(f"[input] 64\n{x}\n[output] 10\n{t}\n")
    (d / "nn.conf").write_text("# diag\n[name] diag\n[type] ANN\n[init] generate\n[seed] 1\n[input] 64\n[hidden] 48\n"
                               "[output] 10\n[train] BP\n[sample_dir] ./samples\n[test_dir] ./samples\n")
    for cublas in (True, False):
        out = work / ("cublas" if cublas else "nocublas")
        out.mkdir()
        exe = build(tree, out, cublas)
        for streams in (1, 2):
            for blocking in ("1", "0"):
                env = dict(os.environ, CUDA_VISIBLE_DEVICES=",".join(str(i) for i in range(n_gpus)),
                           CUDA_LAUNCH_BLOCKING=blocking, OMP_NUM_THREADS="2")
                try:
                    r = subprocess.run([str(exe), "-v", "-v", "-v", "-S", str(streams), "nn.conf"], cwd=d, env=env,
                                       capture_output=True, text=True, timeout=40)
                    rc, so, se = r.returncode, r.stdout, r.stderr
                except subprocess.TimeoutExpired:
                    rc, so, se = -999, "", "timeout"
                print(f"=== cublas={int(cublas)} n_gpus={n_gpus} -S {streams} CUDA_LAUNCH_BLOCKING={blocking}: exit {rc}")
                keep = [l for l in so.splitlines() if "GPU" in l or "multi" in l or "TRAINING" in l or "ERROR" in l][:12]
                print("\n".join(keep))
                print("--- stderr (first lines)")
                print("\n".join(se.splitlines()[:8]))
                sys.stdout.flush()
    shutil.rmtree(work, ignore_errors=True)


if __name__ == "__main__":
    main()
