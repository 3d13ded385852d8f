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


def _need(dims):
    """Work-space doubles of a forward pass (the formula documented in include/hpnn_core.h)."""
    pad8 = lambda v: (v + 7) & ~7
    return pad8(dims[0]) + 12 * sum(pad8(d) for d in dims[1:]) + 2 * pad8(dims[-1]) + 16


def test_documented_limit_is_the_real_one():
    """Largest accepted and smallest refused hidden width for an 8-h-8 network, straight from the
    documented formula; the accepted one must also compute the right answer."""
    lib, fn = _bind()
    limit = 128 * 1024
    h_ok = max(h for h in range(10000, 11000, 8) if _need([8, h, 8]) <= limit)
    h_bad = h_ok + 8
    assert _need([8, h_ok, 8]) <= limit < _need([8, h_bad, 8])
    k, _ = hp.generate_kernel(3, 8, [h_ok], 8)
    net = HpnnNet(k, hp.TYPE_ANN, momentum=False)
    x = torch.linspace(0.05, 0.95, 8, dtype=torch.float64, device="cuda")
    want = net.run(x).clone()
    torch.testing.assert_close(_call(fn, hp.TYPE_ANN, net, x), want, atol=1e-13, rtol=0)
    net.close()
    dims = (C.c_int * 3)(8, h_bad - 7, 8)  # pads up to h_bad
    dummy = torch.zeros(8, dtype=torch.float64, device="cuda")
    ws = (C.c_void_p * 2)(dummy.data_ptr(), dummy.data_ptr())
    assert fn(hp.TYPE_ANN, 0, 2, dims, ws, dummy.data_ptr(), dummy.data_ptr(), None) == -2
    # the examples quoted in the documentation
    assert _need([784, 5000, 5000, 10]) <= limit < _need([784, 20000, 10])


def test_repeated_calls_leave_device_memory_untouched():
    """The point of the entry point: once the first call on a device is behind us (module load,
    address look-up), further calls must not change the amount of free device memory."""
    lib, fn = _bind()
    k, _ = hp.generate_kernel(5, 64, [48, 40], 10)
    net = HpnnNet(k, hp.TYPE_SNN, momentum=False)
    x = torch.rand(64, dtype=torch.float64, device="cuda")
    dims = (C.c_int * len(net.dims))(*net.dims)
    ws = (C.c_void_p * net.n_layers)(*[w.data_ptr() for w in net.W_full])
    out = torch.empty(10, dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    assert fn(hp.TYPE_SNN, 0, net.n_layers, dims, ws, x.data_ptr(), out.data_ptr(), stream) == 0  # first call
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(20):
        assert fn(hp.TYPE_SNN, 0, net.n_layers, dims, ws, x.data_ptr(), out.data_ptr(), stream) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
    assert abs(float(out.sum()) - 1.0) < 1e-12  # a soft-max came out
    net.close()
