"""C-ABI shim (libhpnn.so), C drivers, dataset converters and the Python CLI."""
import os
import shutil
import struct
import subprocess
import sys

import numpy as np
import pytest

import hpnn_b200 as hp
from hpnn_b200 import _native as N
from hpnn_b200 import build as B

from conftest import ROOT, write_conf
from test_cpu_engine import REF_INIT, REF_NITER, REF_ORDER

REFERENCE = "/root/reference"
LIBDIR = os.path.join(ROOT, "hpnn_b200", "lib")
BIN = os.path.join(ROOT, "build", "bin")


@pytest.fixture(scope="module")
def artefacts():
    return B.build(verbose=False)


def _run(cmd, cwd, env_extra=None):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")  # CPU engine: golden numbers are CPU-rate numbers
    env.update(env_extra or {})
    return subprocess.run(cmd, cwd=cwd, capture_output=True, text=True, env=env, timeout=300)


def _golden_dir(tmp_path):
    """The toy problem whose reference-binary output is recorded in test_cpu_engine.py."""
    from conftest import make_problem

    X, T = make_problem(N.TYPE_ANN)
    (tmp_path / "samples").mkdir()
    (tmp_path / "tests").mkdir()
    # one file per directory entry position: write in REF_ORDER so that any readdir order works
    for i, (x, t) in enumerate(zip(X, T)):
        hp.write_sample(str(tmp_path / "samples" / f"s{i:03d}.txt"), x, t, str(i % 3))
        hp.write_sample(str(tmp_path / "tests" / f"s{i:03d}.txt"), x, t, str(i % 3))
    write_conf(tmp_path / "nn.conf")
    return tmp_path


def _parse_training(stdout):
    rows = {}
    for line in stdout.splitlines():
        if "TRAINING FILE:" in line:
            name = line.split("TRAINING FILE:")[1].split()[0]
            rows[name] = (float(line.split("init=")[1].split()[0]), int(line.split("N_ITER=")[1].split()[0]),
                          line.rstrip().endswith("SUCCESS!"))
    return rows


def test_c_driver_trains_and_runs(artefacts, tmp_path):
    d = _golden_dir(tmp_path)
    r = _run([str(artefacts["train_nn"]), "-v", "-v", "nn.conf"], d)
    assert r.returncode == 0, r.stderr
    rows = _parse_training(r.stdout)
    assert len(rows) == 6 and all(ok for _, _, ok in rows.values())
    assert (d / "kernel.tmp").exists() and (d / "kernel.opt").exists()
    # the order depends on readdir, the first visited sample's numbers do not depend on history
    first = r.stdout.split("TRAINING FILE:")[1].split()[0]
    i = int(first[1:4])
    assert abs(rows[first][0] - {k: v for k, v in zip(REF_ORDER, REF_INIT)}[i]) < 1e-9
    assert "[name] (null)" in (d / "kernel.tmp").read_text()
    # resume from kernel.opt and classify
    write_conf(d / "nn.conf", init="kernel.opt")
    r2 = _run([str(artefacts["run_nn"]), "-vv", "nn.conf"], d)
    assert r2.returncode == 0 and r2.stdout.count("TESTING FILE:") == 6
    assert r2.stdout.count("[PASS]") + r2.stdout.count("[FAIL idx=") == 6
    assert r2.stdout.count("[PASS]") >= 3


def test_python_cli_matches_c_driver(artefacts, tmp_path):
    d = _golden_dir(tmp_path)
    rc = _run([str(artefacts["train_nn"]), "-vv", "nn.conf"], d)
    shutil.copy(d / "kernel.opt", d / "kernel.c.opt")
    rp = _run([sys.executable, "-m", "hpnn_b200.cli.train_nn", "-v", "-v", "--backend", "cpu", "nn.conf"], d,
              {"PYTHONPATH": ROOT})
    assert rp.returncode == 0, rp.stderr
    lines_c = [l for l in rc.stdout.splitlines() if "TRAINING FILE" in l]
    lines_p = [l for l in rp.stdout.splitlines() if "TRAINING FILE" in l]
    assert lines_c == lines_p and len(lines_c) == 6  # byte-identical log lines
    assert (d / "kernel.opt").read_text() == (d / "kernel.c.opt").read_text()
    write_conf(d / "nn.conf", init="kernel.opt")
    rr = _run([sys.executable, "-m", "hpnn_b200.cli.run_nn", "-vv", "--backend", "cpu", "nn.conf"], d, {"PYTHONPATH": ROOT})
    rcr = _run([str(artefacts["run_nn"]), "-vv", "nn.conf"], d)
    assert [l for l in rr.stdout.splitlines() if "TESTING" in l] == [l for l in rcr.stdout.splitlines() if "TESTING" in l]


def test_snn_log_format(artefacts, tmp_path):
    d = _golden_dir(tmp_path)
    for f in (d / "samples").iterdir():  # SNN targets are 0/1
        f.write_text(f.read_text().replace("-1.0", "0.0"))
    write_conf(d / "nn.conf", nn_type="SNN", train="BP")
    r = _run([str(artefacts["train_nn"]), "-vv", "nn.conf"], d)
    lines = [l for l in r.stdout.splitlines() if "TRAINING FILE" in l]
    assert len(lines) == 6 and not any("SUCCESS" in l or "FAIL" in l for l in lines)  # src/snn.c:1496
    write_conf(d / "nn.conf", nn_type="SNN", train="BPM")
    r = _run([str(artefacts["train_nn"]), "-vv", "nn.conf"], d)
    assert sum("SUCCESS!" in l for l in r.stdout.splitlines()) == 6
    write_conf(d / "nn.conf", nn_type="SNN", train="BPM", init="kernel.opt")
    r = _run([str(artefacts["run_nn"]), "-vv", "nn.conf"], d)
    assert r.stdout.count(" BEST CLASS idx=") == 6


def test_verbosity_gates_output(artefacts, tmp_path):
    d = _golden_dir(tmp_path)
    quiet = _run([str(artefacts["train_nn"]), "nn.conf"], d)
    assert quiet.returncode == 0 and "TRAINING FILE" not in quiet.stdout
    h = _run([str(artefacts["train_nn"]), "-h"], d)
    assert h.returncode == 0 and "usage" in h.stdout
    bad = _run([str(artefacts["train_nn"]), "-O", "zero"], d)
    assert bad.returncode != 0 and "bad -O parameter" in bad.stderr


@pytest.mark.skipif(not os.path.exists(os.path.join(REFERENCE, "tests", "train_nn.c")), reason="reference tree not mounted")
def test_reference_drivers_compile_unmodified_against_the_shim(artefacts, tmp_path):
    """the reference's own tests/*.c must build against include/libhpnn.h as they are."""
    d = _golden_dir(tmp_path)
    for drv in ("train_nn", "run_nn"):
        exe = tmp_path / f"ref_{drv}"
        r = subprocess.run(["gcc", "-O2", "-D_OMP", "-D_CUDA", f"-I{ROOT}/include", f"{REFERENCE}/tests/{drv}.c", "-o",
                            str(exe), f"-L{LIBDIR}", "-lhpnn", f"-Wl,-rpath,{LIBDIR}", "-lm"], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    r = _run([str(tmp_path / "ref_train_nn"), "-v", "-v", "-O2", "-S", "2", "nn.conf"], d)
    assert r.returncode == 0 and len(_parse_training(r.stdout)) == 6
    mine = _run([str(artefacts["train_nn"]), "-vv", "nn.conf"], d)
    assert [l for l in r.stdout.splitlines() if "TRAINING" in l] == [l for l in mine.stdout.splitlines() if "TRAINING" in l]


def test_shim_api_surface():
    """Every nn_* symbol of the reference header is exported (include/libhpnn.h:126-215)."""
    import ctypes

    lib = ctypes.CDLL(os.path.join(LIBDIR, "libhpnn.so"))
    names = """inc_verbose dec_verbose set_verbose get_verbose return_verbose toggle_dry get_capabilities
    unset_capability return_capabilities init_OMP init_MPI init_CUDA init_BLAS init_all deinit_OMP deinit_MPI
    deinit_CUDA deinit_BLAS deinit_all set_omp_threads get_omp_threads return_omp_threads set_mpi_tasks get_mpi_tasks
    get_curr_mpi_task set_n_gpu get_n_gpu set_cuda_streams get_cuda_streams set_omp_blas get_omp_blas return_cudas
    init_conf deinit_conf set_name get_name return_name set_type get_type return_type set_need_init get_need_init
    return_need_init set_seed get_seed return_seed set_kernel_filename get_kernel_filename return_kernel_filename
    set_train get_train return_train set_samples_directory get_samples_directory return_samples_directory
    set_tests_directory get_tests_directory return_tests_directory load_conf dump_conf free_kernel generate_kernel
    load_kernel dump_kernel get_n_inputs get_n_hiddens get_n_outputs get_h_neurons read_sample train_kernel
    run_kernel""".split()
    for n in names:
        assert hasattr(lib, "nn_" + n), n
    for n in ("ann_load", "ann_generate", "ann_dump", "ann_kernel_run", "ann_kernel_train", "ann_train_BP", "ann_train_BPM",
              "ann_momentum_init", "ann_raz_momentum", "ann_momentum_free", "ann_kernel_train_momentum", "ann_act",
              "ann_dact", "ann_validate_kernel", "ann_kernel_allocate", "ann_kernel_free", "snn_kernel_run",
              "snn_kernel_train", "snn_kernel_train_momentum", "snn_train_BP", "snn_train_BPM", "lib_runtime"):
        assert hasattr(lib, n), n
    lib.nn_init_all(0)
    cap = lib.nn_return_capabilities()
    assert cap & 1 and not cap & 8 and not cap & 32 and not cap & 64  # OMP yes; never CUBLAS/PBLAS/SBLAS
    lib.nn_set_verbose(2)
    assert lib.nn_return_verbose() == 2
    lib.nn_inc_verbose(); lib.nn_inc_verbose(); lib.nn_inc_verbose()
    assert lib.nn_return_verbose() == 3
    lib.nn_set_verbose(0)
    lib.ann_act.restype = ctypes.c_double
    lib.ann_act.argtypes = [ctypes.c_double]
    assert abs(lib.ann_act(0.5) - (2 / (1 + np.exp(-0.5)) - 1)) < 1e-15


def _fake_idx(dirpath, n_train=9, n_test=5):
    rng = np.random.default_rng(3)
    labels = {}
    for stem, n in (("train", n_train), ("test", n_test)):
        lab = rng.integers(0, 10, n, dtype=np.uint8)
        img = rng.integers(0, 256, (n, 28, 28), dtype=np.uint8)
        (dirpath / f"{stem}_labels").write_bytes(struct.pack(">II", 2049, n) + lab.tobytes())
        (dirpath / f"{stem}_images").write_bytes(struct.pack(">IIII", 2051, n, 28, 28) + img.tobytes())
        labels[stem] = (lab, img)
    return labels


def test_pmnist(artefacts, tmp_path):
    data = _fake_idx(tmp_path)
    for sub in ("s", "t", "s2", "t2"):
        (tmp_path / sub).mkdir()
    r = subprocess.run([str(artefacts["pmnist"]), "s", "t"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert sorted(os.listdir(tmp_path / "s")) == [f"s{i:05d}.txt" for i in range(1, 10)]
    assert sorted(os.listdir(tmp_path / "t")) == [f"s{i:05d}.txt" for i in range(10, 15)]  # index keeps counting
    lab, img = data["train"]
    x, t = hp.read_sample(str(tmp_path / "s" / "s00003.txt"))
    np.testing.assert_array_equal(x, img[2].reshape(-1).astype(float))  # raw 0..255 like the reference
    assert t[lab[2]] == 1.0 and (t == -1.0).sum() == 9
    assert f"[output] 10  #{lab[2]}" in (tmp_path / "s" / "s00003.txt").read_text()
    labt, _ = data["test"]
    _, t = hp.read_sample(str(tmp_path / "t" / "s00010.txt"))
    assert t[labt[0]] == 1.0  # fixed: first test image carries the first test label
    subprocess.run([str(artefacts["pmnist"]), "--ref-quirks", "--scale", "s2", "t2"], cwd=tmp_path, check=True,
                   capture_output=True)
    x2, t2 = hp.read_sample(str(tmp_path / "t2" / "s00010.txt"))
    assert t2[labt[1]] == 1.0 and x2.max() <= 1.0  # reference quirk: test labels shifted by one
    assert len(os.listdir(tmp_path / "t2")) == 4


def test_pdif(artefacts, tmp_path):
    rng = np.random.default_rng(5)
    (tmp_path / "rruff" / "dif").mkdir(parents=True)
    (tmp_path / "rruff" / "raw").mkdir()
    (tmp_path / "out").mkdir()
    specs = {"R1.txt": ("P2_1/c", "T = 25 C", 1.541838), "R2.txt": ("Fm-3m", "T = 300 K", 1.541838),
             "R3.txt": ("Pnma", "T = 25 C", 0.710730), "R4.txt": ("NOPE", "nothing", 1.541838)}
    ts = np.arange(3.0, 92.0, 0.5)
    raws = {}
    for name, (sg, temp, lam) in specs.items():
        (tmp_path / "rruff" / "dif" / name).write_text(
            f"      Mineral\n      Sample: {temp}\n      CELL PARAMETERS:  5 5 5 90 90 90\n      SPACE GROUP: {sg}\n"
            f"            X-RAY WAVELENGTH:     {lam:.6f}\n               2-THETA      INTENSITY\n"
            "                10.00        50.00\n                20.00        60.00\n\n")
        inten = rng.random(ts.size) * 100
        raws[name] = inten
        (tmp_path / "rruff" / "raw" / name).write_text("##HEADER\n" + "".join(f"{t:.4f}, {v:.4f}\n" for t, v in zip(ts, inten)))
    r = subprocess.run([str(artefacts["pdif"]), "rruff", "-i", "17", "-o230", "-s", "out"], cwd=tmp_path,
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert sorted(os.listdir(tmp_path / "out")) == ["R1.txt", "R2.txt", "R4.txt"]  # 0.71073 A pattern skipped
    x, t = hp.read_sample(str(tmp_path / "out" / "R1.txt"))
    assert x.size == 18 and t.size == 230 and abs(x[0] - round((25 + 273.15) / 273.15, 5)) < 1e-9
    assert t[14 - 1] == 1.0 and (t == 1.0).sum() == 1  # P2_1/c is space group 14
    assert abs(x[1:].max() - 1.0) < 1e-9  # normalised by the largest bin
    step = 85.0 / 17
    bins = np.array([raws["R1.txt"][(np.round(ts, 4) >= 5 + b * step - 1e-9) & (np.round(ts, 4) < 5 + (b + 1) * step - 1e-9)].round(4).sum() for b in range(17)])
    np.testing.assert_allclose(x[1:], bins / bins.max(), atol=2e-5)
    x, t = hp.read_sample(str(tmp_path / "out" / "R2.txt"))
    assert abs(x[0] - round(300 / 273.15, 5)) < 1e-9 and t[225 - 1] == 1.0
    _, t = hp.read_sample(str(tmp_path / "out" / "R4.txt"))
    assert (t == -1.0).all()  # unknown space group: no class


def test_gen_ann(tmp_path):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_ann.py"), "--seed", "4", "6", "5", "4", "3"],
                         capture_output=True, text=True, check=True).stdout
    p = tmp_path / "k"
    p.write_text(out)
    k = hp.load_kernel(str(p))
    assert k.dims == [6, 5, 4, 3] and k.name == "auto"
    assert np.abs(k.weights[0]).max() <= 1 / np.sqrt(5) + 1e-5  # scaled by the neuron count (reference quirk)


def test_checkpoint_and_metrics(tmp_path):
    import json

    from hpnn_b200.kernel_io import load_kernel_binary

    d = _golden_dir(tmp_path)
    conf = hp.load_conf(str(d / "nn.conf"))
    conf.samples = str(d / "samples")
    ck = str(d / "kernel.ckpt")
    assert hp.train_kernel(conf, backend="cpu", checkpoint_every=2, checkpoint_path=ck, metrics_path=str(d / "m.jsonl"),
                           max_iter=200)
    recs = [json.loads(l) for l in open(d / "m.jsonl")]
    assert len(recs) == 6 and all(r["n_iter"] == 201 for r in recs)
    k_txt, k_bin = hp.load_kernel(ck), load_kernel_binary(ck + ".bin")
    assert k_txt.dims == k_bin.dims == [8, 5, 4, 3]
    for a, b in zip(k_bin.weights, conf.kernel.weights):  # last checkpoint == end of pass (6 % 2 == 0)
        assert (a == b).all()


MODEL_LAYER_C = r"""
#include <stdio.h>
#include <string.h>
#include <libhpnn.h>
#include <libhpnn/ann.h>
#include <libhpnn/snn.h>
/* the reference's contract for ann_kernel_allocate: the CALLER owns the struct (include/libhpnn/ann.h:59,
 * src/ann.c:113-202); and no entry point may read behind a kernel_ann it has never seen */
int main(void){
    kernel_ann mine, stranger;
    UINT hid[2]={5,4};
    DOUBLE in[8]={.1,.2,.3,.4,.5,.6,.7,.8}, out[3]={1.,-1.,-1.};
    memset(&mine,0,sizeof(mine)); memset(&stranger,0,sizeof(stranger));
    nn_init_all(0);
    /* unknown pointers: every call must refuse or do nothing */
    if(ann_kernel_free(&stranger)!=FALSE) return 1;
    ann_kernel_run(&stranger); ann_raz_momentum(&stranger); ann_momentum_init(&stranger); ann_momentum_free(&stranger);
    if(ann_kernel_train(&stranger,out)!=0.) return 2;
    if(ann_train_BP(&stranger,in,out,-1.)!=0.) return 3;
    ann_dump(&stranger,stdout);
    /* caller-owned struct */
    if(ann_kernel_allocate(&mine,8,2,hid,3)!=TRUE) return 4;
    if(mine.n_inputs!=8||mine.n_hiddens!=2||mine.n_outputs!=3||!mine.in||!mine.output.weights) return 5;
    if(mine.hiddens[0].n_neurons!=5||mine.hiddens[1].n_inputs!=5||mine.output.n_inputs!=4) return 6;
    if(ann_validate_kernel(&mine)!=TRUE) return 7;
    for(UINT i=0;i<5*8;i++) mine.hiddens[0].weights[i]=0.01*(double)(i%7)-0.02;
    for(UINT i=0;i<4*5;i++) mine.hiddens[1].weights[i]=0.02*(double)(i%5)-0.03;
    for(UINT i=0;i<3*4;i++) mine.output.weights[i]=0.03*(double)(i%3)-0.02;
    memcpy(mine.in,in,sizeof(in));
    ann_kernel_run(&mine);
    printf("o0=%.12f\n",mine.output.vec[0]);
    DOUBLE d=ann_kernel_train(&mine,out);
    printf("dE=%.6e\n",d);
    if(!(d>0.)) return 8;
    if(ann_kernel_free(&mine)!=TRUE) return 9;
    if(mine.in!=NULL||mine.n_inputs!=0) return 10;
    if(ann_kernel_free(&mine)!=FALSE) return 11; /* second free: unknown again */
    nn_deinit_all();
    printf("MODEL_LAYER_OK\n");
    return 0;
}
"""


def test_model_layer_caller_allocated_kernel(artefacts, tmp_path):
    """ann_kernel_allocate fills a struct the caller allocated, and pointers the library
    has never seen are refused instead of being reinterpreted."""
    src = tmp_path / "model_layer.c"
    src.write_text(MODEL_LAYER_C)
    exe = tmp_path / "model_layer"
    r = subprocess.run(["gcc", "-O1", "-g", f"-I{ROOT}/include", str(src), "-o", str(exe), f"-L{LIBDIR}", "-lhpnn",
                        f"-Wl,-rpath,{LIBDIR}", "-lm"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = _run([str(exe)], tmp_path)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    assert "MODEL_LAYER_OK" in r.stdout


def test_c_driver_under_a_multi_process_launcher(artefacts, tmp_path):
    """Several processes as torchrun / mpirun start them: only rank 0 prints and writes
    kernel files, every rank ends with the same kernel (here: a seed of 0 is agreed on through rank 0),
    and nn_get_mpi_tasks / nn_get_curr_mpi_task report the launcher's numbers."""
    import threading

    d = _golden_dir(tmp_path)
    write_conf(d / "nn.conf", seed=0)  # "pick a seed": the ranks must not pick different ones
    dirs, outs = [], {}
    for r in range(3):
        rd = tmp_path / f"rank{r}"
        rd.mkdir()
        for sub in ("samples", "tests"):
            os.symlink(d / sub, rd / sub)
        (rd / "nn.conf").write_text((d / "nn.conf").read_text())
        dirs.append(rd)

    def run(r):
        env = {"WORLD_SIZE": "3", "RANK": str(r), "LOCAL_RANK": str(r)}
        outs[r] = _run([str(artefacts["train_nn"]), "-v", "-v", "nn.conf"], dirs[r], env)

    ts = [threading.Thread(target=run, args=(r,)) for r in range(3)]  # same parent pid, like launcher workers
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert all(outs[r].returncode == 0 for r in range(3)), [outs[r].stderr for r in range(3)]
    assert outs[0].stdout.count("TRAINING FILE") == 6
    seed_file = "/tmp/.hpnn_seed_%d_0" % os.getpid()  # rank 0 published its choice for the others
    assert os.path.exists(seed_file) and int(open(seed_file).read()) > 0
    os.remove(seed_file)
    assert outs[1].stdout.strip() == "" and outs[2].stdout.strip() == ""  # rank-0-only output
    # the driver opens kernel.opt on every rank (so does the reference's), only rank 0 writes into it
    assert (dirs[0] / "kernel.opt").stat().st_size > 0
    assert all(not (dirs[r] / "kernel.opt").exists() or (dirs[r] / "kernel.opt").stat().st_size == 0 for r in (1, 2))


DEVICE_LAYER_C = r"""
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <libhpnn.h>
#include <libhpnn/cuda_snn.h>
/* the reference's device-layer entry points (include/libhpnn/cuda_ann.h:40-52, cuda_snn.h:33-39) */
int main(void){
    UINT hid[2]={5,4}, seed=77;
    DOUBLE in[8]={.1,.2,.3,.4,.5,.6,.7,.8}, t[3]={1.,-1.,-1.};
    nn_init_all(0);
    kernel_ann *k=ann_generate(&seed,8,2,3,hid);
    if(!k) return 1;
    scuda_ann_allocate(k,NULL);
    DOUBLE *w=(DOUBLE*)malloc(sizeof(DOUBLE)*5*8), *back=(DOUBLE*)malloc(sizeof(DOUBLE)*5*8);
    for(int i=0;i<40;i++) w[i]=0.01*(i%9)-0.03;
    scuda_ann_weight_transfer_C2G(k,0,w,NULL);
    scuda_ann_weight_transfer_G2C(k,0,&back,NULL);
    if(memcmp(w,back,sizeof(DOUBLE)*40)) return 2;
    memcpy(k->in,in,sizeof(in));
    scuda_ann_forward(k,NULL);
    double e0=scuda_ann_error(k,t,NULL);
    double d=scuda_ann_train(k,t,NULL);            /* one BP iteration: returns Ep - Epr */
    double e1=scuda_ann_error(k,t,NULL);
    printf("E0=%.12f dE=%.6e E1=%.12f\n",e0,d,e1);
    if(!(e0>0.) || !(d>0.) || fabs((e0-e1)-d)>1e-12) return 3;
    if(scuda_ann_allocate_momentum(k,NULL)<=0) return 4;
    scuda_ann_raz_momentum(k,NULL);
    if(!(scuda_ann_train_momentum(k,t,0.2,NULL)>0.)) return 5;
    if(scuda_ann_free_momentum(k,NULL)!=TRUE) return 6;
    scuda_snn_forward(k,NULL);
    double s=0.; for(UINT i=0;i<k->n_outputs;i++) s+=k->output.vec[i];
    if(fabs(s-1.)>1e-9) return 7;                   /* a soft-max came out */
    if(!(scuda_snn_error(k,t,NULL)==scuda_snn_error(k,t,NULL))) return 8;
    scuda_ann_deallocate(k,NULL);
    printf("DEVICE_LAYER_OK\n");
    return 0;
}
"""


def test_device_layer_entry_points(artefacts, tmp_path):
    """scuda_ann_* / scuda_snn_* exist with the reference's signatures and host-mirror semantics."""
    src = tmp_path / "device_layer.c"
    src.write_text(DEVICE_LAYER_C)
    exe = tmp_path / "device_layer"
    r = subprocess.run(["gcc", "-O1", f"-I{ROOT}/include", str(src), "-o", str(exe), f"-L{LIBDIR}", "-lhpnn",
                        f"-Wl,-rpath,{LIBDIR}", "-lm"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = _run([str(exe)], tmp_path)
    assert r.returncode == 0 and "DEVICE_LAYER_OK" in r.stdout, (r.returncode, r.stdout, r.stderr)
