"""nn_run_kernel through THIS repo's libhpnn.so, measured exactly like bench/ref_run_nn.py measures the
reference's (same call sequence, same files).  HPNN_RUN_BATCH=0 gives the one-launch-per-file path."""
import ctypes as C
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    import bench

    cfg = bench.CONFIGS["snn_mnist"]
    tmp = tempfile.mkdtemp(prefix="hpnn_shimrun_")
    conf_path = bench.make_dataset(cfg, 3, n, tmp)
    lib = C.CDLL(os.path.join(ROOT, "hpnn_b200", "lib", "libhpnn.so"), mode=C.RTLD_GLOBAL)
    lib.nn_load_conf.restype = C.c_void_p
    lib.nn_load_conf.argtypes = [C.c_char_p]
    lib.nn_run_kernel.argtypes = [C.c_void_p]
    lib.nn_set_tests_directory.argtypes = [C.c_void_p, C.c_char_p]
    lib.nn_init_all(C.c_uint(0))
    lib.nn_set_cuda_streams(C.c_uint(1))
    conf = lib.nn_load_conf(conf_path.encode())
    lib.nn_set_tests_directory(conf, os.path.join(tmp, "warm").encode())
    lib.nn_run_kernel(conf)
    lib.nn_set_tests_directory(conf, os.path.join(tmp, "timed").encode())
    t0 = time.perf_counter()
    lib.nn_run_kernel(conf)
    dt = time.perf_counter() - t0
    print(json.dumps({"impl": "ours (libhpnn.so shim)", "path": "nn_run_kernel, HPNN_RUN_BATCH=%s" % os.environ.get("HPNN_RUN_BATCH", "1024"),
                      "samples": n, "samples_per_s": n / dt, "ms_per_sample": dt * 1e3 / n}), flush=True)


if __name__ == "__main__":
    main()
