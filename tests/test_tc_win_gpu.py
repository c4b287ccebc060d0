"""Window-tile tcgen05 convolutions (csrc/tc_win.cu: fprop+pool for both layers, dgrad) vs fp64 torch autograd."""
import pytest
import torch
import torch.nn.functional as F

from b200dp.models.lenet import CIFAR_WIDE_LENET, MNIST_LENET, LeNet

pytestmark = pytest.mark.gpu
CFGS = [MNIST_LENET, CIFAR_WIDE_LENET]
MEAN, STD = 0.1307, 0.3081


def _native():
    from b200dp.ops import native
    native.load()
    return native


def _dims(native, cfg, b, ld_p2=0):
    return native.Dims(b, cfg.in_ch, cfg.in_hw, cfg.c1, cfg.c2, cfg.c3, cfg.f1, cfg.num_classes, ld_p2)


def _pool(z):
    a = F.leaky_relu(z, 0.2)
    p, flat = F.max_pool2d(a, 2, 2, return_indices=True)
    W = a.shape[-1]
    return p, (((flat // W) % 2) * 2 + (flat % W) % 2).to(torch.uint8), flat


def _close(got, want, rel):
    scale = want.abs().max().clamp_min(1e-12)
    return float((got.double() - want).abs().max() / scale) < rel


def _pack(native, md, cfg, b, ld_p2=0):
    d = _dims(native, cfg, b, ld_p2)
    wpack = torch.empty(native.tcw_pack_floats(d), device="cuda")
    native.tcw_pack_weights(md.ConvNet[1].weight.float().contiguous(), md.ConvNet[4].weight.float().contiguous(), wpack, d)
    return d, wpack


@pytest.mark.parametrize("cfg", CFGS)
@pytest.mark.parametrize("b,u8", [(1, True), (5, True), (64, True), (7, False)])
def test_tcw_conv1_fwd(cfg, b, u8):
    native = _native()
    torch.manual_seed(0)
    dev = "cuda"
    md = LeNet(cfg).to(dev).double()
    if u8:
        x = torch.randint(0, 256, (b, cfg.in_ch, cfg.in_hw, cfg.in_hw), dtype=torch.uint8, device=dev)
        xf = (x.double().div(255) - MEAN) / STD
    else:
        x = torch.randn(b, cfg.in_ch, cfg.in_hw, cfg.in_hw, device=dev)
        xf = x.double()
    with torch.no_grad():
        p_ref, i_ref, _ = _pool(md.ConvNet[1](md.ConvNet[0](xf)))
    d, wpack = _pack(native, md, cfg, b)
    p1 = torch.full((b, cfg.c1, 14, 14), float("nan"), device=dev)
    idx = torch.full((b, cfg.c1, 14, 14), 9, dtype=torch.uint8, device=dev)
    native.tcw_conv1_fwd(x, wpack, md.ConvNet[1].bias.float(), p1, idx, d, MEAN, STD)
    torch.cuda.synchronize()
    assert _close(p1, p_ref, 1e-2)
    assert (idx != i_ref).float().mean() < 0.02        # tf32 rounding may flip near-ties
    assert int(idx.max()) <= 3


@pytest.mark.parametrize("cfg", CFGS)
@pytest.mark.parametrize("b,pad", [(1, 0), (3, 0), (64, 4), (17, 4)])
def test_tcw_conv2_fwd(cfg, b, pad):
    native = _native()
    torch.manual_seed(1)
    dev = "cuda"
    md = LeNet(cfg).to(dev).double()
    p1 = torch.randn(b, cfg.c1, 14, 14, device=dev)
    with torch.no_grad():
        p_ref, i_ref, _ = _pool(md.ConvNet[4](p1.double()))
    ld = cfg.c2 * 25 + pad
    d, wpack = _pack(native, md, cfg, b, ld if pad else 0)
    p2 = torch.full((b, ld), -7.0, device=dev)
    idx = torch.full((b, cfg.c2, 5, 5), 9, dtype=torch.uint8, device=dev)
    native.tcw_conv2_fwd(p1, wpack, md.ConvNet[4].bias.float(), p2, idx, d)
    torch.cuda.synchronize()
    assert _close(p2[:, : cfg.c2 * 25].reshape(b, cfg.c2, 5, 5), p_ref, 1e-2)
    assert (p2[:, cfg.c2 * 25:] == -7.0).all()
    assert (idx != i_ref).float().mean() < 0.02
    assert int(idx.max()) <= 3


@pytest.mark.parametrize("cfg", CFGS)
@pytest.mark.parametrize("b", [1, 2, 37])
def test_tcw_conv2_dgrad(cfg, b):
    native = _native()
    torch.manual_seed(2)
    dev = "cuda"
    md = LeNet(cfg).to(dev).double()
    conv = md.ConvNet[4]
    z1 = torch.randn(b, cfg.c1, 14, 14, device=dev, dtype=torch.float64, requires_grad=True)
    P1 = F.leaky_relu(z1, 0.2)
    p2, idx2, _ = _pool(conv(P1))
    dp2 = torch.randn_like(p2)
    p2.backward(dp2)
    d, wpack = _pack(native, md, cfg, b)
    g1 = torch.full((b, cfg.c1, 14, 14), float("nan"), device=dev)
    native.tcw_conv2_dgrad(dp2.float().contiguous(), p2.detach().float().contiguous(), idx2, wpack,
                           P1.detach().float().contiguous(), g1, d)
    torch.cuda.synchronize()
    assert _close(g1, z1.grad, 2e-2)


@pytest.mark.parametrize("cfg", CFGS)
@pytest.mark.parametrize("b", [1, 2, 37, 256])
def test_tcw_conv2_wgrad(cfg, b):
    """wgrad/bgrad of conv2 against autograd through conv2 -> LeakyReLU -> max-pool."""
    native = _native()
    torch.manual_seed(4)
    dev = "cuda"
    md = LeNet(cfg).to(dev).double()
    conv = md.ConvNet[4]
    P1 = F.leaky_relu(torch.randn(b, cfg.c1, 14, 14, device=dev, dtype=torch.float64), 0.2)
    p2, idx2, _ = _pool(conv(P1))
    dp2 = torch.randn_like(p2)
    conv.zero_grad()
    p2.backward(dp2)
    d = _dims(native, cfg, b)
    n_rows = native.tcw_conv_wgrad_num_partials(d)
    wn = cfg.c2 * cfg.c1 * 25
    partials = torch.zeros(n_rows, wn + cfg.c2 + 8, device=dev)
    native.tcw_conv_wgrad(2, P1.detach().float().contiguous(), 0.0, 1.0, dp2.float().contiguous(),
                          p2.detach().float().contiguous(), cfg.c2 * 25, idx2, partials, partials.shape[1], 0, wn, d)
    torch.cuda.synchronize()
    rows = partials.double().sum(0)
    assert _close(rows[:wn].view_as(conv.weight), conv.weight.grad, 2e-2)
    assert _close(rows[wn: wn + cfg.c2], conv.bias.grad, 2e-2)
    assert (partials[:, wn + cfg.c2:] == 0).all()


@pytest.mark.parametrize("cfg", CFGS)
@pytest.mark.parametrize("b,u8", [(1, True), (2, True), (37, True), (256, True), (5, False)])
def test_tcw_conv1_wgrad(cfg, b, u8):
    native = _native()
    torch.manual_seed(3)
    dev = "cuda"
    md = LeNet(cfg).to(dev).double()
    conv = md.ConvNet[1]
    if u8:
        x = torch.randint(0, 256, (b, cfg.in_ch, cfg.in_hw, cfg.in_hw), dtype=torch.uint8, device=dev)
        xf = (x.double().div(255) - MEAN) / STD
    else:
        x = torch.randn(b, cfg.in_ch, cfg.in_hw, cfg.in_hw, device=dev)
        xf = x.double()
    z = conv(md.ConvNet[0](xf))
    _, idx1, flat = _pool(z)
    g1 = torch.randn(b, cfg.c1, 14, 14, device=dev, dtype=torch.float64)      # gradient at the pooled cells
    dz = torch.zeros_like(z).view(b, cfg.c1, -1).scatter_(2, flat.view(b, cfg.c1, -1), g1.view(b, cfg.c1, -1)).view_as(z)
    gw, gb = torch.autograd.grad(z, [conv.weight, conv.bias], grad_outputs=dz)
    d = _dims(native, cfg, b)
    n_rows = native.tcw_conv_wgrad_num_partials(d)
    wn = cfg.c1 * cfg.in_ch * 25
    partials = torch.zeros(n_rows, wn + cfg.c1 + 3, device=dev)
    native.tcw_conv_wgrad(1, x, MEAN, STD, g1.float().contiguous(), None, 0, idx1, partials, partials.shape[1], 0, wn, d)
    torch.cuda.synchronize()
    rows = partials.double().sum(0)
    assert _close(rows[:wn].view_as(gw), gw, 2e-2)
    assert _close(rows[wn: wn + cfg.c1], gb, 2e-2)
