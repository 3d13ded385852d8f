"""hpnn-b200: a Blackwell-native framework with the capabilities of libhpnn
(on-the-fly training of small fully connected ANN / SNN networks, BP and BPM).

Layers:  ``config`` / ``kernel_io`` / ``samples`` (file formats) -> ``models.net.HpnnNet``
(device-resident, optionally row-split network) -> ``trainer`` (``train_kernel`` /
``run_kernel`` drivers) -> ``cli`` (``train_nn`` / ``run_nn``).  Compute lives in the
native core (``csrc/``): hand-written sm_100a kernels + a C++/OpenMP CPU engine."""
from . import _native
from ._native import TRAIN_BP, TRAIN_BPM, TYPE_ANN, TYPE_SNN
from .config import NNConf, dump_conf, load_conf, parse_conf
from .kernel_io import Kernel, dump_kernel, generate_kernel, load_kernel
from .samples import read_sample, shuffle_order, walk, write_sample
from .trainer import make_engine, run_kernel, train_kernel

__version__ = "0.2.0"

__all__ = ["NNConf", "load_conf", "parse_conf", "dump_conf", "Kernel", "generate_kernel", "load_kernel",
           "dump_kernel", "read_sample", "write_sample", "walk", "shuffle_order", "train_kernel", "run_kernel",
           "make_engine", "TYPE_ANN", "TYPE_SNN", "TRAIN_BP", "TRAIN_BPM"]
