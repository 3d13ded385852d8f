"""Same measurement as bench/ref_run_nn.py through hpnn_b200.run_kernel (files on disk, batch 1
and the batched extension)."""
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    import torch

    import bench
    import hpnn_b200 as hp

    cfg = bench.CONFIGS["snn_mnist"]
    tmp = tempfile.mkdtemp(prefix="hpnn_oursrun_")
    conf = hp.load_conf(bench.make_dataset(cfg, 3, n, tmp))
    from hpnn_b200.utils import log

    for batch, verbose in ((0, 0), (256, 0), (-1, 0), (-1, 2)):  # -1 = the default (whole directory batched); 2 = "-vv"
        log.set_verbose(verbose)
        sink = open(os.devnull, "w")
        conf.tests = os.path.join(tmp, "warm")
        hp.run_kernel(conf, backend="cuda", batch=batch, stream=sink)
        conf.tests = os.path.join(tmp, "timed")
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        st = hp.run_kernel(conf, backend="cuda", batch=batch, stream=sink)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print(json.dumps({"impl": "ours", "path": "hpnn_b200.run_kernel(batch=%d, verbosity=%d)" % (batch, verbose), "samples": st["total"],
                          "samples_per_s": st["total"] / dt, "ms_per_sample": dt * 1e3 / max(1, st["total"])}), flush=True)


if __name__ == "__main__":
    main()
