#!/usr/bin/env python
"""gen_ann: write a random kernel file to stdout (scripts/gen_ann.bash of the reference).

    gen_ann.py [--seed S] [--fan-in] [--name NAME] n_in n_hid1 ... n_hidN n_out

The reference script scales the uniform draws by 1/sqrt(neurons of the layer) -- not by the
fan-in as ann_generate does -- and prints "%7.5f" values from /dev/urandom; that
scaling and precision are the default here, ``--fan-in`` switches to ann_generate's scaling."""
import argparse
import math
import random
import sys


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--seed", type=int, default=None)
    ap.add_argument("--fan-in", action="store_true")
    ap.add_argument("--name", default="auto")
    ap.add_argument("dims", type=int, nargs="+")
    a = ap.parse_args(argv)
    if len(a.dims) < 3 or min(a.dims) < 1:
        ap.error("need n_in, at least one hidden layer, n_out (all >= 1)")
    rng = random.Random(a.seed)
    out = sys.stdout
    out.write(f"[name] {a.name}\n[param] {' '.join(map(str, a.dims))}\n[input] {a.dims[0]}\n")
    for l in range(1, len(a.dims)):
        n, m = a.dims[l], a.dims[l - 1]
        out.write(f"[output] {n}\n" if l == len(a.dims) - 1 else f"[hidden {l}] {n}\n")
        scale = math.sqrt(m if a.fan_in else n)
        for j in range(n):
            out.write(f"[neuron {j + 1}] {m}\n")
            out.write(" ".join("%7.5f" % (2.0 * (rng.random() - 0.5) / scale) for _ in range(m)) + "\n")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
