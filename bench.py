#!/usr/bin/env python
"""Headline benchmark: ``train_nn`` samples/s on a BASELINE.json config, ours vs the reference.

    python bench.py --gpus N --steps K --warmup W [--config NAME] [--impl ours|reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

A *step* is one training sample taken to convergence by the reference's own stopping rule
(dE <= 1e-6 and argmax == target, 31/15..102399 iterations): exactly what ``nn_train_kernel``
does per file.  Both arms train the SAME synthetic samples (files on disk, same seed, same
pseudo-random visiting order) from the SAME glibc-random initial weights, in FP64.

  value        device-timed samples/s (CUDA events around each step, max over ranks); inputs are
               already on the device, one cooperative kernel launch per step.
  e2e.value    samples/s through the public API ``hpnn_b200.make_engine(...).train(x, t)``:
               every step copies that sample from pinned host memory, launches, and reads the
               step's result struct back (host-synchronised per step).
  reference    ``nn_train_kernel`` of the unmodified reference (CUDA+cuBLAS build, all visible
               GPUs in one process, as it is designed), same directories, wall clock bracketed
               by device synchronisation + CUDA events.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# name -> (type, train, n_in, hiddens, n_out, BASELINE.json config it implements)
CONFIGS = {
    "wide_ann": ("ANN", "BP", 784, [8192, 8192, 8192, 8192], 10,
                 "Wide ANN 8192x4 hidden layers BP, row split (BASELINE.json config 4)"),
    "wide_ann_bpm": ("ANN", "BPM", 784, [8192, 8192, 8192, 8192], 10,
                     "Wide ANN 8192x4 hidden layers, momentum (BPM) variant of BASELINE.json config 4"),
    "snn_mnist": ("SNN", "BPM", 784, [1000, 1000], 10,
                  "SNN 784-1000-1000-10 BPM online training (BASELINE.json config 2)"),
    "xrd_snn": ("SNN", "BPM", 851, [230], 230,
                "RRUFF-XRD tutorial-shaped SNN 851-230-230 BPM (BASELINE.json config 3)"),
    "ann_mnist": ("ANN", "BP", 784, [300], 10,
                  "ANN 784-300-10 MNIST tutorial shape BP (BASELINE.json config 1)"),
}
SEED = 10958


def rank_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


# ----------------------------------------------------------------------------- data
def write_sample_text(path, x, t, comment):
    """Sample file in the reference's text format (src/libhpnn.c:1088-1136; the writers are
    tutorials/mnist/prepare_mnist.c:47-60).  Plain Python on purpose: both arms of the benchmark
    create their data with this function, and the reference arm must not map any of this
    repository's native code (byte-identical to hpnn_sample_write, tests/test_bench_contract.py)."""
    with open(path, "w") as f:
        f.write("[input] %i\n" % len(x))
        f.write(" ".join("%7.5f" % v for v in x))
        f.write("\n[output] %i  #%s\n" % (len(t), comment))
        f.write(" ".join("%.1f" % v for v in t))
        f.write("\n")


def make_dataset(cfg, n_warm, n_timed, root):
    """Synthetic samples of the config's shape in two directories (+ nn.conf)."""
    import numpy as np

    write_sample = write_sample_text
    typ, train, n_in, hid, n_out, _ = cfg
    rng = np.random.default_rng(2024)
    lo = 0.0 if typ == "SNN" else -1.0
    for sub, n in (("warm", n_warm), ("timed", n_timed)):
        d = os.path.join(root, sub)
        os.makedirs(d, exist_ok=True)
        for i in range(n):
            x = rng.random(n_in)
            lab = int(rng.integers(n_out))
            t = np.full(n_out, lo)
            t[lab] = 1.0
            write_sample(os.path.join(d, f"s{i:05d}.txt"), x, t, str(lab))
    conf = os.path.join(root, "nn.conf")
    with open(conf, "w") as f:
        f.write(f"# bench\n[name] bench\n[type] {typ}\n[init] generate\n[seed] {SEED}\n[input] {n_in}\n"
                f"[hidden] {' '.join(str(h) for h in hid)}\n[output] {n_out}\n[train] {train}\n"
                f"[sample_dir] {root}/warm\n[test_dir] {root}/timed\n")
    return conf


def config_block(args, cfg, n_gpus):
    """The ``config`` object of the JSON line: key for key and value for value the same in both
    arms (everything arm-specific or measured goes to ``detail``)."""
    typ, train, n_in, hid, n_out, desc = cfg
    wmb = sum(a * b for a, b in zip([n_in, *hid], [*hid, n_out])) * 8 / 1e6
    return {"model": "%s %s" % (typ, "-".join(str(d) for d in [n_in, *hid, n_out])), "name": args.config,
            "baseline_config": desc, "global_batch": 1, "seq_len": None,
            "parallelism": "tp%d (row split of every layer over %d GPU(s))" % (n_gpus, n_gpus),
            "l2": "weights %.0f MB per pass (L2 is 126 MB); the ours arm also writes 256 MiB between timed steps, "
                  "the reference's directory loop cannot be interrupted" % wmb}


def metric_name(cfg):
    return "train_nn samples/s (per-sample convergent %s training, fp64)" % cfg[1]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.samples, self.proc, self.idx = [], None, gpu_index

    def start(self):
        exe = shutil.which("nvidia-smi")
        if not exe:
            return
        try:
            self.proc = subprocess.Popen([exe, "-i", str(self.idx), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            p = [x.strip() for x in s.split(",")]
            if len(p) < 6:
                continue
            try:
                sm.append(float(p[0]))
                mx = max(mx, float(p[1]))
            except ValueError:
                continue
            for nm, v in zip(names, p[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------- ours
def run_ours(args, cfg):
    import numpy as np
    import torch
    import torch.distributed as dist

    import hpnn_b200 as hp
    from hpnn_b200 import _native as N
    from hpnn_b200.ops import cuda as ops
    from hpnn_b200.samples import SampleSet, walk
    from hpnn_b200.trainer import CudaEngine

    rank, world, local = rank_world()
    torch.cuda.set_device(local)
    group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        group = dist.group.WORLD
    typ, train, n_in, hid, n_out, desc = cfg
    nn_type = hp.TYPE_SNN if typ == "SNN" else hp.TYPE_ANN
    nn_train = hp.TRAIN_BPM if train == "BPM" else hp.TRAIN_BP

    tmp = tempfile.mkdtemp(prefix="hpnn_bench_") if rank == 0 else None
    if world > 1:
        box = [tmp]
        dist.broadcast_object_list(box, src=0)
        tmp = box[0]
    if rank == 0:
        make_dataset(cfg, args.warmup, args.steps, tmp)
    if world > 1:
        dist.barrier()
    kernel, _ = hp.generate_kernel(SEED, n_in, hid, n_out)  # identical on every rank (glibc random)
    warm = SampleSet(walk(os.path.join(tmp, "warm"), SEED))
    timed = SampleSet(walk(os.path.join(tmp, "timed"), SEED))

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    # ---------------- device-timed arm: inputs resident, one launch per step
    eng = CudaEngine(kernel, nn_type, nn_train, group=group)
    Xw, Tw = warm.inputs.cuda(), warm.targets.cuda()
    Xt, Tt = timed.inputs.cuda(), timed.targets.cuda()
    for i in range(len(warm)):
        eng.net.load_sample(Xw[i], Tw[i])
        eng.net.train_sample_async(eng.opts, i)
    sync_all()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(len(timed))]
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    launches0 = ops.launch_count()
    sync_all()
    for i in range(len(timed)):
        flush.zero_()  # L2 flush between timed steps (outside the per-step event pair)
        ev[i][0].record()
        eng.net.load_sample(Xt[i], Tt[i])
        eng.net.train_sample_async(eng.opts, i)
        ev[i][1].record()
    sync_all()
    launches = ops.launch_count() - launches0
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    res_dev = eng.net.fetch_results(len(timed))
    iters_dev = sum(r.n_iter for r in res_dev)
    split_world = eng.net.world  # 1 when the net is too small to be worth splitting (every rank holds all of it)
    clk = clocks.stop() if rank == 0 else None
    eng.close()

    # ---------------- end-to-end arm: public per-sample API, pinned H2D + result D2H each step
    eng = CudaEngine(kernel, nn_type, nn_train, group=group)
    for i in range(len(warm)):
        eng.train(warm.inputs[i].numpy(), warm.targets[i].numpy())
    sync_all()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    res_e2e = []
    for i in range(len(timed)):
        eng.net.load_sample(timed.inputs[i], timed.targets[i])   # pinned host -> device
        eng.net.train_sample_async(eng.opts, 0)
        res_e2e.append(eng.net.fetch_results(1)[0])              # device -> host, host-synchronised
    e1.record()
    sync_all()
    e2e_ms = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3 if world == 1 else 0.0)
    iters_e2e = sum(r.n_iter for r in res_e2e)
    n_ok = sum(int(r.success) for r in res_e2e)
    eng.close()

    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = t.tolist()
    K = len(timed)
    if rank == 0:
        out = {
            "metric": metric_name(cfg),
            "value": K / (dev_ms * 1e-3), "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": len(warm),
            "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "fp64", "data": "synthetic (uniform [0,1) inputs, one-hot targets), random-init weights "
                                      "(glibc random(), seed %d)" % SEED,
            "impl": "ours",
            "config": config_block(args, cfg, world),
            "detail": {"iterations_timed": iters_dev, "iterations_per_s": iters_dev / (dev_ms * 1e-3),
                       "samples_converged": n_ok, "us_per_iteration": dev_ms * 1e3 / max(1, iters_dev),
                       "engine": ("one process per GPU, one cooperative kernel launch per sample; rows split over %d ranks, "
                                  "all-gather / reduce-scatter by peer or NVLS multicast stores from inside the kernel" % split_world)
                       if split_world > 1 else
                       ("one cooperative kernel launch per sample on one GPU" + ("" if world == 1 else
                        "; the net is below HPNN_TP_MIN_MB, so each of the %d ranks holds all of it and trains the same "
                        "samples (row-splitting it over NVLink is slower)" % world))},
            "clocks": clk,
            "e2e": {"value": K / (e2e_ms * 1e-3), "unit": "samples/s",
                    "h2d_bytes_per_step": 8 * (n_in + n_out), "d2h_bytes_per_step": C.sizeof(N.TrainResult),
                    "iterations_timed": iters_e2e, "api": "HpnnNet.load_sample + train_sample_async + fetch_results"},
            "gpu_launches": launches,
        }
        print(json.dumps(out), flush=True)
        shutil.rmtree(tmp, ignore_errors=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ----------------------------------------------------------------------------- reference
def run_reference(args, cfg):
    """Rank 0 runs the reference in a child process (the reference calls exit(-1) on any CUDA
    or allocation error, which must not take the reporter down); other ranks idle."""
    rank, world, local = rank_world()
    if rank != 0:
        # the reference drives every GPU from ONE process (src/libhpnn.c:201-305): ranks > 0 idle
        _wait_for_flag(args)
        return
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE",
                                                              "GROUP_RANK", "ROLE_RANK", "TORCHELASTIC_RUN_ID")}
    env["CUDA_VISIBLE_DEVICES"] = ",".join(str(i) for i in range(max(1, args.gpus)))
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", args.impl, "--child", "--gpus", str(args.gpus),
           "--steps", str(args.steps), "--warmup", str(args.warmup), "--config", args.config,
           "--ref-streams", str(args.ref_streams)]
    try:
        r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=args.ref_timeout)
        lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
        if r.returncode == 0 and lines:
            print(lines[-1], flush=True)
        else:
            tail = (r.stderr.strip().splitlines() or r.stdout.strip().splitlines() or ["no output"])[-1]
            print(json.dumps({"impl": args.impl, "n_gpus": args.gpus, "unavailable":
                              ("reference process exited with code %d: %s" % (r.returncode, tail))[:300]}), flush=True)
    except subprocess.TimeoutExpired:
        print(json.dumps({"impl": args.impl, "n_gpus": args.gpus,
                          "unavailable": "reference run exceeded %d s" % args.ref_timeout}), flush=True)
    _set_flag(args)


def run_reference_child(args, cfg):
    n_gpus = max(1, args.gpus)
    out = {"impl": args.impl}
    shim = args.impl == "shim"  # same host code, but linked against OUR drop-in libhpnn.so (not a reference number)
    try:
        lib_path = os.path.join(ROOT, "baseline", "_ref", "build", "libhpnn_ref_cuda.so")
        if shim:
            lib_path = os.path.join(ROOT, "hpnn_b200", "lib", "libhpnn.so")
        elif not os.path.exists(lib_path):
            sys.path.insert(0, os.path.join(ROOT, "baseline"))
            import build_ref
            build_ref.build(("cuda",))
        tmp = tempfile.mkdtemp(prefix="hpnn_bench_ref_")
        conf_path = make_dataset(cfg, args.warmup, args.steps, tmp)
        ref = C.CDLL(lib_path, mode=C.RTLD_GLOBAL)
        rt = C.CDLL("libcudart.so.12", mode=C.RTLD_GLOBAL)
        ref.nn_load_conf.restype = C.c_void_p
        ref.nn_load_conf.argtypes = [C.c_char_p]
        ref.nn_train_kernel.argtypes = [C.c_void_p]
        ref.nn_set_samples_directory.argtypes = [C.c_void_p, C.c_char_p]
        ref.nn_set_verbose.argtypes = [C.c_short]
        ref.nn_set_cuda_streams.argtypes = [C.c_uint]
        ref.nn_init_all(C.c_uint(0))
        ref.nn_set_cuda_streams(C.c_uint(args.ref_streams))
        conf = ref.nn_load_conf(conf_path.encode())
        if not conf:
            raise RuntimeError("reference nn_load_conf failed")
        typ, train, n_in, hid, n_out, desc = cfg
        ref.nn_set_verbose(C.c_short(2))
        log_path = os.path.join(tmp, "ref.log")

        def train_dir(d):
            ref.nn_set_samples_directory(conf, d.encode())
            sys.stdout.flush()
            saved = os.dup(1)
            fd = os.open(log_path, os.O_WRONLY | os.O_CREAT | os.O_TRUNC)
            os.dup2(fd, 1)
            try:
                rt.cudaDeviceSynchronize()
                ev0, ev1 = C.c_void_p(), C.c_void_p()
                rt.cudaEventCreate(C.byref(ev0))
                rt.cudaEventCreate(C.byref(ev1))
                t0 = time.perf_counter()
                rt.cudaEventRecord(ev0, None)
                ok = ref.nn_train_kernel(conf)
                rt.cudaEventRecord(ev1, None)
                rt.cudaDeviceSynchronize()
                wall = (time.perf_counter() - t0) * 1e3
                ms = C.c_float(0)
                rt.cudaEventElapsedTime(C.byref(ms), ev0, ev1)
                C.CDLL(None).fflush(None)
            finally:
                os.dup2(saved, 1)
                os.close(fd)
                os.close(saved)
            text = open(log_path).read()
            iters = [int(m) for m in re.findall(r"N_ITER=\s*(\d+)", text)]
            return ok, max(wall, float(ms.value)), iters, text

        train_dir(os.path.join(tmp, "warm"))
        # the reference arm must be the reference alone: none of this repository's native code
        # may be mapped into the process that is being timed
        mapped = sorted({l.split()[-1] for l in open("/proc/self/maps") if "hpnn_b200/lib" in l})
        if mapped and not shim:
            raise RuntimeError("reference arm has this repository's library mapped: %s" % mapped)
        clocks = ClockSampler(0)
        clocks.start()
        ok, ms, iters, text = train_dir(os.path.join(tmp, "timed"))
        clk = clocks.stop()
        K = args.steps
        out.update({
            "metric": metric_name(cfg),
            "value": K / (ms * 1e-3), "unit": "samples/s", "n_gpus": n_gpus, "steps": K, "warmup": args.warmup,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "fp64",
            "data": "synthetic (uniform [0,1) inputs, one-hot targets), random-init weights "
                    "(glibc random(), seed %d)" % SEED,
            "config": config_block(args, cfg, n_gpus),
            "detail": {"iterations_timed": sum(iters), "iterations_per_s": sum(iters) / (ms * 1e-3),
                       "samples_converged": len(re.findall(r"SUCCESS", text)),
                       "us_per_iteration": ms * 1e3 / max(1, sum(iters)),
                       "engine": ("hpnn-b200 through the libhpnn C API, row split over %d GPU(s), one process" % n_gpus) if shim else
                       "reference row split over %d GPU(s) x %d stream(s), one process" % (n_gpus, args.ref_streams),
                       "build": "hpnn_b200/lib/libhpnn.so (drop-in shim over the sm_100a engine)" if shim else
                       "unmodified sources, -D_OMP -D_CUDA -D_CUBLAS, nvcc sm_100 (baseline/build_ref.py)",
                       "native_libraries_of_this_repo_mapped": mapped},
            "clocks": clk,
            "e2e": {"value": K / (ms * 1e-3), "unit": "samples/s", "h2d_bytes_per_step": 8 * (n_in + n_out),
                    "d2h_bytes_per_step": 8 * n_out, "api": "nn_train_kernel(conf) (src/libhpnn.c:1149)"},
            "gpu_launches": None,
        })
        shutil.rmtree(tmp, ignore_errors=True)
    except Exception as e:  # the contract: report and exit 0
        out = {"impl": args.impl, "unavailable": ("%s: %s" % (type(e).__name__, e)).replace("\n", " ")[:300]}
    print(json.dumps(out), flush=True)


def _flag_path(args):
    return os.path.join(tempfile.gettempdir(), "hpnn_bench_ref_done_%s" % os.environ.get("MASTER_PORT", "0"))


def _set_flag(args):
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        open(_flag_path(args), "w").write("done")


def _wait_for_flag(args):
    p = _flag_path(args)
    t0 = time.time()
    while not os.path.exists(p) and time.time() - t0 < 3 * 3600:
        time.sleep(0.5)
    time.sleep(1.0)
    if int(os.environ.get("RANK", "0")) == 1:
        try:
            os.remove(p)
        except OSError:
            pass


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "shim"],
                    help="ours: the framework's own API; reference: unmodified reference build; shim: the reference "
                         "arm's host code (nn_load_conf + nn_train_kernel) linked against our drop-in libhpnn.so")
    ap.add_argument("--config", default=os.environ.get("HPNN_BENCH_CONFIG", "wide_ann"), choices=sorted(CONFIGS))
    ap.add_argument("--ref-streams", type=int, default=1)
    ap.add_argument("--ref-timeout", type=int, default=3000)
    ap.add_argument("--child", action="store_true", help=argparse.SUPPRESS)
    args = ap.parse_args()
    args.warmup = max(3, args.warmup)
    cfg = CONFIGS[args.config]
    if args.impl in ("reference", "shim"):
        (run_reference_child if args.child else run_reference)(args, cfg)
    else:
        run_ours(args, cfg)


if __name__ == "__main__":
    main()
