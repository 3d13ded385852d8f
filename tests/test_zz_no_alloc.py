"""The allocation-free forward entry point, called through the C ABI exactly as a C host would.

Some programs that embed this library cannot let it allocate device memory: they reserved the
whole GPU for their own pool at start-up, or they run other streams that a cudaMalloc would
stall.  ``hpnn_dev_forward_noalloc`` serves them from a fixed scratch array that lives inside
the library.  Because that array is shared by every call and keeps the previous call's
contents, the test makes two calls in a row -- a larger network first, then a smaller one whose
vectors land on top of the first one's leftovers -- and checks both against the regular entry
point (``HpnnNet.run`` -> ``hpnn_dev_train_sample`` with a caller-allocated arena).

The file name sorts last on purpose: a run that stops at the first failure has then already
executed every other test."""
import ctypes as C

import numpy as np
import pytest
import torch

import hpnn_b200 as hp
from hpnn_b200 import _native as N
from hpnn_b200.models.net import HpnnNet

pytestmark = pytest.mark.gpu


def _bind():
    lib = C.CDLL(str(N.lib_path()))  # plain dlopen of the in-tree library, as a C program would
    fn = lib.hpnn_dev_forward_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p,
                   C.c_void_p]
    lib.hpnn_last_error.restype = C.c_char_p
    return lib, fn


def _call(fn, nn_type, net: HpnnNet, x_dev: torch.Tensor) -> torch.Tensor:
    dims = (C.c_int * len(net.dims))(*net.dims)
    ws = (C.c_void_p * net.n_layers)(*[w.data_ptr() for w in net.W_full])
    out = torch.full((net.dims[-1],), float("nan"), dtype=torch.float64, device=x_dev.device)
    rc = fn(nn_type, 0, net.n_layers, dims, ws, x_dev.data_ptr(), out.data_ptr(),
            torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("nn_type", [hp.TYPE_ANN, hp.TYPE_SNN])
def test_two_calls_in_a_row_match_the_regular_entry_point(nn_type):
    lib, fn = _bind()
    rng = np.random.default_rng(11)
    # the second network is smaller: its buffers overlap what the first call left behind
    for n_in, hid, n_out in ((784, [300, 200], 10), (37, [29], 5)):
        k, _ = hp.generate_kernel(99, n_in, hid, n_out)
        net = HpnnNet(k, nn_type, momentum=False)
        x = torch.from_numpy(rng.random(n_in)).cuda()
        want = net.run(x).clone()          # existing entry point, caller-allocated arena
        got = _call(fn, nn_type, net, x)   # new entry point, built-in scratch
        torch.testing.assert_close(got, want, atol=1e-13, rtol=0)
        got2 = _call(fn, nn_type, net, x)  # and once more on top of its own leftovers
        torch.testing.assert_close(got2, want, atol=1e-13, rtol=0)
        net.close()


def test_networks_beyond_the_built_in_scratch_are_refused():
    lib, fn = _bind()
    dims_list = [20000, 30000, 30000, 10]  # needs far more than HPNN_NOALLOC_MAX_DOUBLES doubles
    dims = (C.c_int * 4)(*dims_list)
    dummy = torch.zeros(8, dtype=torch.float64, device="cuda")
    ws = (C.c_void_p * 3)(dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr())
    rc = fn(hp.TYPE_ANN, 0, 3, dims, ws, dummy.data_ptr(), dummy.data_ptr(), None)
    assert rc == -2
    msg = lib.hpnn_last_error().decode()
    assert "HPNN_NOALLOC_MAX_DOUBLES" in msg and "doubles" in msg
    assert fn(hp.TYPE_ANN, 0, 1, dims, ws, dummy.data_ptr(), dummy.data_ptr(), None) == -1  # no hidden layer
