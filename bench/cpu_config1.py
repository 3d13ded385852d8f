"""BASELINE.json config 1: ANN 784-300-10 BP via train_nn on CPU/OpenMP, single rank.
Reference CPU build (no BLAS, OpenMP) vs this repo's C++/OpenMP engine, same files."""
import json
import os
import re
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def run(cmd, cwd, env=None):
    t0 = time.perf_counter()
    r = subprocess.run(cmd, cwd=cwd, capture_output=True, text=True, env=env)
    dt = time.perf_counter() - t0
    iters = [int(m) for m in re.findall(r"N_ITER=\s*(\d+)", r.stdout)]
    return dt, iters, r


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    threads = sys.argv[2] if len(sys.argv) > 2 else "4"
    config = sys.argv[3] if len(sys.argv) > 3 else "ann_mnist"  # any bench.py config; config 1 is ann_mnist
    tmp = tempfile.mkdtemp(prefix="hpnn_cfg1_")
    conf = bench.make_dataset(bench.CONFIGS[config], n, 1, tmp)
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS=threads)
    out = []
    ref = os.path.join(ROOT, "baseline", "_ref", "build", "train_nn_cpu")
    mine = os.path.join(ROOT, "build", "bin", "train_nn")
    for name, exe in (("reference CPU/OpenMP (no BLAS)", ref), ("hpnn_b200 C++/OpenMP engine", mine)):
        if not os.path.exists(exe):
            continue
        dt, iters, r = run([exe, "-v", "-v", "-O" + threads, conf], tmp, env)
        out.append({"impl": name, "samples": len(iters), "iterations": sum(iters), "seconds": round(dt, 3),
                    "samples_per_s": round(len(iters) / dt, 3), "iterations_per_s": round(sum(iters) / dt, 1),
                    "threads": int(threads)})
        print(json.dumps(out[-1]), flush=True)


if __name__ == "__main__":
    main()
