"""Fused row-split trainer on >= 2 GPUs (skipped on a single-GPU box)."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world", [2, 4, 8])
def test_fused_row_split_over_nvlink(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    here = os.path.dirname(os.path.abspath(__file__))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(here, "mp_tp_check.py")],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "TP_CHECK_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
