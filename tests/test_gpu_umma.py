"""K-T: the tcgen05 / TMEM / TMA batched forward (opt-in bf16) against PyTorch on the same
bf16-rounded operands (the only difference left is fp32 vs fp64 accumulation)."""
import numpy as np
import pytest
import torch

from hpnn_b200.ops import reference as R

pytestmark = pytest.mark.gpu


def _rand(*shape, seed=0):
    g = torch.Generator().manual_seed(seed)
    return (torch.rand(*shape, generator=g, dtype=torch.float64) * 2 - 1).cuda()


@pytest.fixture(params=[2, 1, 0], ids=["pair-mma", "pair-multicast", "single"])
def kt_cluster(request):
    """The tile mappings of K-T: CTA pairs on one 256 x 256 tile (tcgen05.mma.cta_group::2, the default for
    two or more batch tiles), CTA pairs sharing each block of W through TMA multicast, one CTA per tile."""
    from hpnn_b200 import _native as N

    lib = N.load()
    lib.hpnn_dev_set_kt_cluster(request.param)
    yield request.param
    lib.hpnn_dev_set_kt_cluster(-1)


@pytest.mark.parametrize("batch,rows,cols", [(128, 128, 64), (128, 128, 256), (1, 10, 1000), (333, 230, 851),
                                             (4096, 1000, 784), (257, 1000, 1000)])
@pytest.mark.parametrize("act_kind", [0, 1])
def test_gemm_bf16_umma(batch, rows, cols, act_kind, kt_cluster):
    from hpnn_b200.ops import cuda as ops

    W = _rand(rows, cols, seed=1) / np.sqrt(cols)
    X = _rand(batch, cols, seed=2)
    Wb, Xb = ops.cast_bf16(ops.padded_weights(W), cols), ops.cast_bf16(X, cols)
    torch.testing.assert_close(Wb[:, :cols].double(), W.to(torch.bfloat16).double(), atol=0, rtol=0)
    assert (Wb[:, cols:] == 0).all()
    Y = ops.gemm_bf16(Wb, Xb, cols, act_kind=act_kind)
    ref = Xb[:, :cols].double() @ Wb[:, :cols].double().T
    if act_kind == 0:
        ref = R.act(ref)
    torch.testing.assert_close(Y, ref, atol=2e-5, rtol=1e-5)


def test_forward_batch_bf16_close_to_fp64():
    import hpnn_b200 as hp
    from hpnn_b200.models.net import HpnnNet

    k, _ = hp.generate_kernel(5, 784, [300, 200], 10)
    net = HpnnNet(k, hp.TYPE_SNN, momentum=False)
    X = (_rand(64, 784, seed=3) + 1) / 2
    Y64 = net.forward_batch(X)
    Y16 = net.forward_batch(X, precision="bf16")
    assert (Y64.argmax(1) == Y16.argmax(1)).float().mean() > 0.9
    torch.testing.assert_close(Y16, Y64, atol=5e-3, rtol=0)
    torch.testing.assert_close(Y64.sum(1), torch.ones(64, dtype=torch.float64, device="cuda"), atol=1e-9, rtol=0)
    net.close()


@pytest.mark.parametrize("batch,rows,cols", [(128, 128, 64), (100, 70, 40), (333, 230, 851), (4096, 1000, 784),
                                             (8192, 2048, 512), (8192 + 128, 520, 192)])
@pytest.mark.parametrize("act_kind", [0, 1])
def test_gemm_bf16_chain_output(batch, rows, cols, act_kind, kt_cluster):
    """bf16-in / bf16-out variant (hidden layers of the tensor-core chain): persistent tile loop, two TMEM
    accumulators, swizzled staging + TMA store.  The largest shape gives every CTA about seven tiles."""
    from hpnn_b200.ops import cuda as ops

    W = _rand(rows, cols, seed=1) / np.sqrt(cols)
    X = _rand(batch, cols, seed=2)
    Wb, Xb = ops.cast_bf16(ops.padded_weights(W), cols), ops.cast_bf16(X, cols)
    Yb = ops.gemm_bf16_chain(Wb, Xb, cols, act_kind)
    assert Yb.dtype == torch.bfloat16 and Yb.shape == (batch, (rows + 63) & ~63)
    ref = Xb[:, :cols].double() @ Wb[:, :cols].double().T
    if act_kind == 0:
        ref = R.act(ref)
    torch.testing.assert_close(Yb[:, :rows].double(), ref, atol=1e-2, rtol=8e-3)  # bf16 output: 8 mantissa bits
    assert (Yb[:, rows:] == 0).all()
    # and twice in a row on the same buffers (cached tensor maps)
    Yb2 = ops.gemm_bf16_chain(Wb, Xb, cols, act_kind)
    assert torch.equal(Yb, Yb2) and Yb2.data_ptr() != Yb.data_ptr()
