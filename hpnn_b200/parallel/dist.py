"""World bootstrap: one process per GPU, ``torch.distributed`` (NCCL) for the plumbing.

The reference's MPI layer (src/libhpnn.c:182-200, 431-457) maps as: ``MPI_Init`` ->
``init_world``; ``nn_get_mpi_tasks`` / ``nn_get_curr_mpi_task`` -> ``world_size`` / ``rank``;
rank-0-only I/O and ``MPI_Bcast`` of kernels (src/ann.c:549-614) -> ``broadcast_kernel``."""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.distributed as dist

from ..kernel_io import Kernel


def init_world(backend: Optional[str] = None):
    """Initialise the default process group from the torchrun environment (idempotent)."""
    if dist.is_initialized():
        return dist.group.WORLD
    if "RANK" not in os.environ:
        return None
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        dist.init_process_group(backend)
    return dist.group.WORLD


def rank() -> int:
    return dist.get_rank() if dist.is_initialized() else 0


def world_size() -> int:
    return dist.get_world_size() if dist.is_initialized() else 1


def broadcast_kernel(kernel: Optional[Kernel], src: int = 0, group=None) -> Kernel:
    """Rank ``src`` reads / generates the kernel, everybody else receives it."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return kernel
    # the reference sends a "bailout" int so that the other ranks do not hang when rank 0 cannot
    # read the kernel (src/ann.c:242-245,551-555); here a None broadcast plays that role
    box = [kernel if dist.get_rank(group) == src else None]
    dist.broadcast_object_list(box, src=src, group=group)
    if box[0] is None:
        raise IOError("rank %d could not provide the kernel (bailout)" % src)
    return box[0]


def shutdown():
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
