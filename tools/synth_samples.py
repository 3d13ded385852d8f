#!/usr/bin/env python
"""Synthetic sample directories of a given shape (there is no network on the build boxes).

    synth_samples.py --out DIR --n 64 --inputs 784 --outputs 10 [--snn] [--seed 1] [--clusters]

``--clusters`` draws every class from its own Gaussian blob so that a trained network can
actually generalise to a held-out test directory (used by the tutorials)."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hpnn_b200.samples import write_sample  # noqa: E402


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", required=True)
    ap.add_argument("--n", type=int, default=64)
    ap.add_argument("--inputs", type=int, default=784)
    ap.add_argument("--outputs", type=int, default=10)
    ap.add_argument("--snn", action="store_true", help="0/1 targets instead of -1/+1")
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--clusters", action="store_true")
    ap.add_argument("--start", type=int, default=1, help="first sample number")
    a = ap.parse_args(argv)
    os.makedirs(a.out, exist_ok=True)
    rng = np.random.default_rng(a.seed)
    centres = np.random.default_rng(12345).random((a.outputs, a.inputs))
    lo = 0.0 if a.snn else -1.0
    for i in range(a.n):
        lab = int(rng.integers(a.outputs))
        x = np.clip(centres[lab] + 0.15 * rng.standard_normal(a.inputs), 0, 1) if a.clusters else rng.random(a.inputs)
        t = np.full(a.outputs, lo)
        t[lab] = 1.0
        write_sample(os.path.join(a.out, f"s{a.start + i:05d}.txt"), x, t, str(lab))
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
