"""Time the UNMODIFIED reference's nn_run_kernel (CUDA+cuBLAS build) on N synthetic test files.

Run as a child process (the reference calls exit() on errors).  Prints one JSON line."""
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
    tmp = tempfile.mkdtemp(prefix="hpnn_refrun_")
    conf_path = bench.make_dataset(cfg, 3, n, tmp)
    ref = C.CDLL(os.path.join(ROOT, "baseline", "_ref", "build", "libhpnn_ref_cuda.so"), mode=C.RTLD_GLOBAL)
    rt = C.CDLL("libcudart.so.12", mode=C.RTLD_GLOBAL)
    ref.nn_load_conf.restype = C.c_void_p
    ref.nn_load_conf.argtypes = [C.c_char_p]
    ref.nn_run_kernel.argtypes = [C.c_void_p]
    ref.nn_set_tests_directory.argtypes = [C.c_void_p, C.c_char_p]
    ref.nn_set_cuda_streams.argtypes = [C.c_uint]
    ref.nn_init_all(C.c_uint(0))
    ref.nn_set_cuda_streams(C.c_uint(1))
    conf = ref.nn_load_conf(conf_path.encode())
    ref.nn_set_tests_directory(conf, os.path.join(tmp, "warm").encode())
    ref.nn_run_kernel(conf)
    ref.nn_set_tests_directory(conf, os.path.join(tmp, "timed").encode())
    rt.cudaDeviceSynchronize()
    t0 = time.perf_counter()
    ref.nn_run_kernel(conf)
    rt.cudaDeviceSynchronize()
    dt = time.perf_counter() - t0
    print(json.dumps({"impl": "reference", "path": "nn_run_kernel (batch 1, file per sample)", "samples": n,
                      "samples_per_s": n / dt, "ms_per_sample": dt * 1e3 / n}), flush=True)


if __name__ == "__main__":
    main()
