"""b200dp_conv_bwd_noalloc, called through the C ABI, against b200dp_conv_bwd.

Some hosts cannot allocate device memory at the point where they run the convolution
backward: all of the GPU's memory may already belong to one pool allocator, or the call may
sit in a loop that cannot afford an allocation's locking and synchronisation.  For them the
library carries the partial-gradient rows in a built-in array and returns finished gradients.
This test drives that entry point with ctypes exactly as a C host would (raw pointers, raw
stream), twice in a row — the second time with a smaller batch, so rows left over from the
first call sit in the shared array — and compares each result with the gradients obtained
from the ordinary entry point and a caller-owned scratch buffer, to the tolerance the other
kernel tests use.  It is collected last so that a run stopping at the first failure has
already exercised everything else; it builds all of its inputs itself, so it also passes
alone in a fresh process.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

RANGES = (("ConvNet.1.weight", 150), ("ConvNet.1.bias", 6), ("ConvNet.4.weight", 2400), ("ConvNet.4.bias", 16))
MAX_STRIDE = 2688   # B200DP_NOALLOC_MAX_STRIDE in include/b200dp.h


def _bind(native):
    lib = native.load()
    vp, i, f = C.c_void_p, C.c_int, C.c_float
    fn = lib.b200dp_conv_bwd_noalloc
    fn.restype = C.c_int
    fn.argtypes = [vp, i, f, f, vp, vp, vp, vp, vp, vp, vp, i, i, i, i, i, native.Dims, vp]
    return lib, fn


def _reference_grads(net, x, y):
    """Existing path: forward, head backward, b200dp_conv_bwd into the caller-owned scratch rows;
    the rows are then summed here."""
    net.forward(x, y, train=True)
    net.backward(x)
    torch.cuda.synchronize()
    rows = net.partials[: net.cur_partials].double().sum(0)   # only the rows this batch's CTAs wrote
    return {name: rows[off: off + n].float() for (name, n), off in zip(RANGES, net.conv_offsets)}


def _call_noalloc(fn, native, net, x, out):
    d = net.dims(x.shape[0])
    ow1, ob1, ow2, ob2 = net.conv_offsets
    stream = torch.cuda.current_stream().cuda_stream
    return fn(x.data_ptr(), int(x.dtype == torch.uint8), net.spec.mean, net.spec.std, net.p1.data_ptr(),
              net.idx1.data_ptr(), net.p2.data_ptr(), net.idx2.data_ptr(), net.dp2.data_ptr(),
              net.w["ConvNet.4.weight"].data_ptr(), out.data_ptr(), out.numel(), ow1, ob1, ow2, ob2, d, stream)


def test_noalloc_matches_existing_entry_twice_in_a_row():
    from b200dp.models.fused import NativeLeNet
    from b200dp.models.lenet import MNIST_LENET
    from b200dp.ops import native

    lib, fn = _bind(native)
    dev = torch.device("cuda")
    torch.manual_seed(11)
    net = NativeLeNet(MNIST_LENET, 256, dev, loss_mode="xent")
    stride = net.partial_stride
    assert stride <= MAX_STRIDE
    out = torch.full((stride,), float("nan"), device=dev)

    for b in (256, 37):            # second call: fewer rows than the first left behind
        x = torch.randint(0, 256, (b, 1, 28, 28), dtype=torch.uint8, device=dev)
        y = torch.randint(0, 10, (b,), device=dev)
        want = _reference_grads(net, x, y)           # also leaves p1/idx1/p2/idx2/dp2 for this batch
        out.fill_(float("nan"))
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        rc = _call_noalloc(fn, native, net, x, out)
        assert rc == 0, lib.b200dp_last_error().decode()
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_before          # nothing was allocated on the device
        for (name, n), off in zip(RANGES, net.conv_offsets):
            got = out[off: off + n]
            scale = want[name].abs().max().clamp_min(1e-12)
            err = float((got - want[name]).abs().max() / scale)
            assert err < 2e-4, (b, name, err)
        owned = torch.zeros(stride, dtype=torch.bool, device=dev)
        for (_, n), off in zip(RANGES, net.conv_offsets):
            owned[off: off + n] = True
        assert torch.isnan(out[~owned]).all()                        # cells outside the four ranges untouched


def test_noalloc_rejects_oversize():
    from b200dp.models.fused import NativeLeNet
    from b200dp.models.lenet import MNIST_LENET
    from b200dp.ops import native

    lib, fn = _bind(native)
    dev = torch.device("cuda")
    net = NativeLeNet(MNIST_LENET, 8, dev)
    x = torch.zeros(8, 1, 28, 28, dtype=torch.uint8, device=dev)
    y = torch.zeros(8, dtype=torch.int64, device=dev)
    net.forward(x, y, train=True)
    net.backward(x)
    too_long = torch.zeros(MAX_STRIDE + 4, device=dev)
    rc = _call_noalloc(fn, native, net, x, too_long)
    assert rc != 0
    msg = lib.b200dp_last_error().decode()
    assert "grad_len" in msg and str(MAX_STRIDE) in msg
    torch.cuda.synchronize()
