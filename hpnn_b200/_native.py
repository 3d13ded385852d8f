"""ctypes binding of ``libhpnn_core.so`` (C ABI in ``include/hpnn_core.h``).

The library is built in-tree by :mod:`hpnn_b200.build`.  On a machine with a GPU a
missing library is a hard error: the CUDA path must never silently fall back.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

MAX_LAYERS = 16
MAX_WORLD = 8
TYPE_ANN, TYPE_LNN, TYPE_SNN, TYPE_UKN = 0, 1, 2, -1
TRAIN_BP, TRAIN_BPM, TRAIN_CG, TRAIN_SPLX, TRAIN_UKN = 0, 1, 2, 3, -1

_LIB_PATH = Path(__file__).resolve().parent / "lib" / "libhpnn_core.so"


class TrainResult(C.Structure):
    _fields_ = [("init_err", C.c_double), ("final_de", C.c_double), ("last_err", C.c_double),
                ("n_iter", C.c_int), ("first_ok", C.c_int), ("success", C.c_int), ("status", C.c_int)]


class TrainOpts(C.Structure):
    _fields_ = [("type", C.c_int), ("train", C.c_int), ("lr", C.c_double), ("alpha", C.c_double),
                ("delta", C.c_double), ("min_iter", C.c_int), ("max_iter", C.c_int),
                ("stable_softmax", C.c_int)]


class Conf(C.Structure):
    _fields_ = [("name", C.c_char_p), ("type", C.c_int), ("need_init", C.c_int), ("seed", C.c_uint),
                ("f_kernel", C.c_char_p), ("train", C.c_int), ("samples", C.c_char_p), ("tests", C.c_char_p),
                ("n_in", C.c_int), ("n_out", C.c_int), ("n_hid", C.c_int), ("hid", C.POINTER(C.c_int))]


class DevNet(C.Structure):
    _fields_ = [("n_layers", C.c_int), ("dims", C.c_int * (MAX_LAYERS + 1)), ("ld", C.c_int * MAX_LAYERS),
                ("row0", C.c_int * MAX_LAYERS), ("rows", C.c_int * MAX_LAYERS),
                ("W", C.c_void_p * MAX_LAYERS), ("dW", C.c_void_p * MAX_LAYERS),
                ("rank", C.c_int), ("world", C.c_int), ("arena", C.c_void_p * MAX_WORLD),
                ("arena_mc", C.c_void_p), ("local_sync", C.c_void_p),
                ("off_act", C.c_longlong * (MAX_LAYERS + 1)), ("off_part", C.c_longlong * MAX_LAYERS),
                ("off_target", C.c_longlong), ("off_out", C.c_longlong), ("off_flags", C.c_longlong),
                ("arena_doubles", C.c_longlong), ("part_stride", C.c_int * MAX_LAYERS),
                ("off_cta_part", C.c_longlong * MAX_LAYERS), ("cta_slots", C.c_int),
                ("off_hist_u", C.c_longlong * MAX_LAYERS), ("off_hist_v", C.c_longlong * MAX_LAYERS),
                ("off_hist_dot", C.c_longlong * MAX_LAYERS),
                ("hist_depth", C.c_int), ("off_panel", C.c_longlong), ("panel_stride", C.c_int),
                ("off_wlast", C.c_longlong), ("off_mlast", C.c_longlong), ("off_zpart", C.c_longlong)]


_lib = None


def lib_path() -> Path:
    return _LIB_PATH


def _declare(lib):
    vp, i, d, ll = C.c_void_p, C.c_int, C.c_double, C.c_longlong
    dp = C.POINTER(C.c_double)
    sig = {
        "hpnn_free": (None, [vp]),
        "hpnn_kernel_create": (vp, [i, C.POINTER(i)]),
        "hpnn_kernel_destroy": (None, [vp]),
        "hpnn_kernel_n_layers": (i, [vp]),
        "hpnn_kernel_dim": (i, [vp, i]),
        "hpnn_kernel_weights": (dp, [vp, i]),
        "hpnn_kernel_name": (C.c_char_p, [vp]),
        "hpnn_kernel_set_name": (None, [vp, C.c_char_p]),
        "hpnn_kernel_generate": (vp, [C.POINTER(C.c_uint), i, C.POINTER(i)]),
        "hpnn_kernel_load": (vp, [C.c_char_p]),
        "hpnn_kernel_dump_path": (i, [vp, C.c_char_p]),
        "hpnn_kernel_dump_binary": (i, [vp, C.c_char_p]),
        "hpnn_kernel_load_binary": (vp, [C.c_char_p]),
        "hpnn_conf_load": (i, [C.c_char_p, C.POINTER(Conf)]),
        "hpnn_conf_free": (None, [C.POINTER(Conf)]),
        "hpnn_sample_read": (i, [C.c_char_p, C.POINTER(dp), C.POINTER(i), C.POINTER(dp), C.POINTER(i)]),
        "hpnn_sample_write": (i, [C.c_char_p, dp, i, dp, i, C.c_char_p]),
        "hpnn_sample_read_many": (i, [C.POINTER(C.c_char_p), i, i, i, vp, vp, vp]),
        "hpnn_dir_list": (i, [C.c_char_p, C.POINTER(C.POINTER(C.c_char_p)), C.POINTER(i)]),
        "hpnn_dir_list_free": (None, [C.POINTER(C.c_char_p), i]),
        "hpnn_shuffle_order": (None, [C.c_uint, i, C.POINTER(i)]),
        "hpnn_train_opts_default": (None, [C.POINTER(TrainOpts), i, i, i]),
        "hpnn_set_omp_threads": (None, [i]),
        "hpnn_cpu_forward": (i, [vp, i, i, dp, dp]),
        "hpnn_cpu_act": (dp, [vp, i]),
        "hpnn_cpu_momentum_init": (None, [vp]),
        "hpnn_cpu_momentum_zero": (None, [vp]),
        "hpnn_cpu_forward_inplace": (i, [vp, i, i]),
        "hpnn_cpu_error": (C.c_double, [vp, i, dp]),
        "hpnn_cpu_train_iteration": (C.c_double, [vp, C.POINTER(TrainOpts), dp]),
        "hpnn_cpu_train_sample": (i, [vp, C.POINTER(TrainOpts), dp, dp, C.POINTER(TrainResult), dp]),
        "hpnn_dev_arena_layout": (ll, [C.POINTER(DevNet)]),
        "hpnn_cuda_device_count": (i, []),
        "hpnn_cuda_set_device": (i, [i]),
        "hpnn_dev_alloc": (vp, [C.c_size_t]),
        "hpnn_dev_free": (None, [vp]),
        "hpnn_dev_memset": (i, [vp, i, C.c_size_t, vp]),
        "hpnn_dev_h2d": (i, [vp, vp, C.c_size_t, vp]),
        "hpnn_dev_d2h": (i, [vp, vp, C.c_size_t, vp]),
        "hpnn_dev_sync": (i, [vp]),
        "hpnn_host_alloc_pinned": (vp, [C.c_size_t]),
        "hpnn_host_free_pinned": (None, [vp]),
        "hpnn_ipc_export": (i, [vp, C.c_char_p]),
        "hpnn_ipc_open": (vp, [C.c_char_p]),
        "hpnn_ipc_close": (i, [vp]),
        "hpnn_enable_peer_access": (i, [i, i]),
        "hpnn_last_error": (C.c_char_p, []),
        "hpnn_dev_train_sample": (i, [C.POINTER(DevNet), C.POINTER(TrainOpts), i, vp, vp]),
        "hpnn_dev_train_sample_vranks": (i, [C.POINTER(DevNet), i, C.POINTER(TrainOpts), i, C.POINTER(vp), vp]),
        "hpnn_dev_set_spin_timeout_ms": (None, [ll]),
        "hpnn_dev_set_fault": (None, [i, i]),
        "hpnn_dev_set_jitter_ns": (None, [i]),
        "hpnn_dev_persistent_grid": (i, []),
        "hpnn_dev_set_persistent_grid": (None, [i]),
        "hpnn_dev_set_lazy": (None, [i, C.c_longlong]),
        "hpnn_dev_set_lazy_panel": (None, [i]),
        "hpnn_dev_set_slicewise": (None, [i]),
        "hpnn_abi_sizeof": (C.c_longlong, [i]),
        "hpnn_dev_set_phase_trace": (None, [vp, i]),
        "hpnn_dev_forward_layer": (i, [vp, i, i, i, vp, vp, i, vp]),
        "hpnn_dev_output_stats": (i, [i, i, vp, vp, i, vp, vp, vp, vp]),
        "hpnn_dev_backward_update": (i, [vp, vp, i, i, i, vp, vp, vp, i, d, d, vp]),
        "hpnn_dev_delta_finalize": (i, [vp, i, i, vp, i, vp, vp]),
        "hpnn_dev_update_forward": (i, [vp, vp, i, i, i, vp, vp, vp, i, d, d, i, vp]),
        "hpnn_dev_forward_batch_f64": (i, [vp, i, i, i, vp, i, i, vp, i, i, vp]),
        "hpnn_dev_softmax_rows": (i, [vp, i, i, i, i, vp]),
        "hpnn_dev_forward_batch_f64_tp": (i, [vp, i, i, i, vp, i, i, C.POINTER(vp), i, i, i, i, vp]),
        "hpnn_dev_forward_batch_f64_tp_mc": (i, [vp, i, i, i, vp, i, i, C.POINTER(vp), vp, i, i, i, i, vp]),
        "hpnn_dev_world_barrier": (i, [C.POINTER(vp), i, i, C.c_ulonglong, vp]),
        "hpnn_dev_grad_merge": (i, [vp, ll, C.POINTER(vp), i, d, vp]),
        "hpnn_dev_launch_count": (ll, [i]),
        "hpnn_dev_dp_publish": (i, [vp, ll, C.POINTER(vp), vp, C.POINTER(vp), vp, i, C.c_ulonglong, vp]),
        "hpnn_dev_dp_apply": (i, [vp, i, i, i, vp, ll, vp, ll, i, vp, C.c_ulonglong, ll, vp, vp]),
        "hpnn_dev_grad_merge_nvls": (i, [vp, vp, vp, ll, i, i, d, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    # optional entry points (present once their translation units are built)
    opt = {
        "hpnn_dev_cast_bf16": (i, [vp, i, i, i, vp, i, vp]),
        "hpnn_dev_gemm_bf16": (i, [vp, i, i, i, vp, i, i, vp, i, i, vp]),
        "hpnn_dev_gemm_bf16_bf16": (i, [vp, i, i, i, vp, i, i, vp, i, i, vp]),
        "hpnn_dev_set_kt_cluster": (None, [i]),
        "hpnn_dev_set_kt_trace": (None, [vp]),
        "hpnn_net_forward_batch": (i, [vp, vp, i, vp, i]),
        "hpnn_dev_forward_noalloc": (i, [i, i, i, C.POINTER(i), C.POINTER(vp), vp, vp, vp]),
    }
    for name, (res, args) in opt.items():
        if hasattr(lib, name):
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args


def load(build_if_missing: bool = True):
    """Return the loaded core library (building it first if necessary)."""
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists() and build_if_missing and os.environ.get("HPNN_NO_AUTOBUILD") != "1":
        from . import build as _build
        _build.build(verbose=False)
    if not _LIB_PATH.exists():
        raise RuntimeError(
            f"native core {_LIB_PATH} is missing: run `python -m hpnn_b200.build` (nvcc, sm_100a)")
    lib = C.CDLL(str(_LIB_PATH), mode=C.RTLD_GLOBAL)
    _declare(lib)
    # a stale library (older struct layout) must fail here, not corrupt memory later
    for which, cls in enumerate((TrainResult, TrainOpts, Conf, DevNet)):
        got = lib.hpnn_abi_sizeof(which)
        if got != C.sizeof(cls):
            raise RuntimeError(f"{_LIB_PATH} was built from other sources: sizeof({cls.__name__}) is {got} there, "
                               f"{C.sizeof(cls)} here; rebuild with `python -m hpnn_b200.build`")
    _lib = lib
    return lib


def last_error() -> str:
    return (load().hpnn_last_error() or b"").decode()


def check(rc: int, what: str):
    if rc != 0:
        raise RuntimeError(f"{what} failed (rc={rc}): {last_error()}")
