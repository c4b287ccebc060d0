"""tcgen05 implicit-GEMM convolutions (fprop+pool, dgrad, wgrad) vs fp64 torch autograd."""
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


@pytest.mark.parametrize("cfg", CFGS)
@pytest.mark.parametrize("b", [3, 64])
def test_tc_conv1_fwd(cfg, b):
    native = _native()
    torch.manual_seed(0)
    dev = "cuda"
    md = LeNet(cfg).to(dev).double()
    x = torch.randint(0, 256, (b, cfg.in_ch, cfg.in_hw, cfg.in_hw), dtype=torch.uint8, device=dev)
    xf = (x.double().div(255) - MEAN) / STD
    with torch.no_grad():
        p_ref, i_ref, _ = _pool(md.ConvNet[1](md.ConvNet[0](xf)))
    p1 = torch.empty(b, cfg.c1, 14, 14, device=dev)
    idx = torch.empty(b, cfg.c1, 14, 14, dtype=torch.uint8, device=dev)
    native.tc_conv1_fwd(x, md.ConvNet[1].weight.float().contiguous(), md.ConvNet[1].bias.float(), p1, idx,
                        _dims(native, cfg, b), MEAN, STD)
    torch.cuda.synchronize()
    assert _close(p1, p_ref, 1e-2)
    assert (idx != i_ref).float().mean() < 0.02        # tf32 rounding may flip near-ties


@pytest.mark.parametrize("cfg", CFGS)
@pytest.mark.parametrize("b,pad", [(3, 0), (64, 4)])
def test_tc_conv2_fwd(cfg, b, pad):
    native = _native()
    torch.manual_seed(1)
    dev = "cuda"
    md = LeNet(cfg).to(dev).double()
    p1 = torch.randn(b, cfg.c1, 14, 14, device=dev)
    with torch.no_grad():
        p_ref, i_ref, _ = _pool(md.ConvNet[4](p1.double()))
    ld = cfg.c2 * 25 + pad
    p2 = torch.full((b, ld), -7.0, device=dev)
    idx = torch.empty(b, cfg.c2, 5, 5, dtype=torch.uint8, device=dev)
    native.tc_conv2_fwd(p1, md.ConvNet[4].weight.float().contiguous(), md.ConvNet[4].bias.float(), p2, idx,
                        _dims(native, cfg, b, ld if pad else 0))
    torch.cuda.synchronize()
    assert _close(p2[:, : cfg.c2 * 25].reshape(b, cfg.c2, 5, 5), p_ref, 1e-2)
    assert (p2[:, cfg.c2 * 25:] == -7.0).all()
    assert (idx != i_ref).float().mean() < 0.02


@pytest.mark.parametrize("cfg", CFGS)
@pytest.mark.parametrize("b", [2, 37])
def test_tc_conv2_backward(cfg, b):
    """dgrad (-> g1) and wgrad/bgrad of conv2 against autograd through conv2 -> LeakyReLU -> max-pool."""
    native = _native()
    torch.manual_seed(2)
    dev = "cuda"
    md = LeNet(cfg).to(dev).double()
    conv = md.ConvNet[4]
    z1 = torch.randn(b, cfg.c1, 14, 14, device=dev, dtype=torch.float64, requires_grad=True)
    P1 = F.leaky_relu(z1, 0.2)
    p2, idx2, _ = _pool(conv(P1))
    dp2 = torch.randn_like(p2)
    conv.zero_grad()
    p2.backward(dp2)
    d = _dims(native, cfg, b)
    g1 = torch.empty(b, cfg.c1, 14, 14, device=dev)
    native.tc_conv2_dgrad(dp2.float().contiguous(), p2.detach().float().contiguous(), idx2, conv.weight.detach().float().contiguous(),
                          P1.detach().float().contiguous(), g1, d)
    n_rows = native.tc_conv_wgrad_num_partials(d)
    wn = cfg.c2 * cfg.c1 * 25
    partials = torch.zeros(n_rows, wn + cfg.c2 + 8, device=dev)
    native.tc_conv_wgrad(2, P1.detach().float().contiguous(), 0.0, 1.0, dp2.float().contiguous(),
                         p2.detach().float().contiguous(), cfg.c2 * 25, idx2, partials, partials.shape[1], 0, wn, d)
    torch.cuda.synchronize()
    assert _close(g1, z1.grad, 2e-2)
    rows = partials.double().sum(0)
    assert _close(rows[:wn].view_as(conv.weight), conv.weight.grad, 2e-2)
    assert _close(rows[wn: wn + cfg.c2], conv.bias.grad, 2e-2)
    assert (partials[:, wn + cfg.c2:] == 0).all()


@pytest.mark.parametrize("cfg", CFGS)
@pytest.mark.parametrize("b", [2, 37])
def test_tc_conv1_wgrad(cfg, b):
    native = _native()
    torch.manual_seed(3)
    dev = "cuda"
    md = LeNet(cfg).to(dev).double()
    conv = md.ConvNet[1]
    x = torch.randint(0, 256, (b, cfg.in_ch, cfg.in_hw, cfg.in_hw), dtype=torch.uint8, device=dev)
    x0 = md.ConvNet[0]((x.double().div(255) - MEAN) / STD)
    z = conv(x0)
    _, idx1, flat = _pool(z)
    g1 = torch.randn(b, cfg.c1, 14, 14, device=dev, dtype=torch.float64)      # gradient at the pooled cells
    dz = torch.zeros_like(z).view(b, cfg.c1, -1).scatter_(2, flat.view(b, cfg.c1, -1), g1.view(b, cfg.c1, -1)).view_as(z)
    gw, gb = torch.autograd.grad(z, [conv.weight, conv.bias], grad_outputs=dz)
    d = _dims(native, cfg, b)
    n_rows = native.tc_conv_wgrad_num_partials(d)
    wn = cfg.c1 * cfg.in_ch * 25
    partials = torch.zeros(n_rows, wn + cfg.c1 + 3, device=dev)
    native.tc_conv_wgrad(1, x, MEAN, STD, g1.float().contiguous(), None, 0, idx1, partials, partials.shape[1], 0, wn, d)
    torch.cuda.synchronize()
    rows = partials.double().sum(0)
    assert _close(rows[:wn].view_as(gw), gw, 2e-2)
    assert _close(rows[wn: wn + cfg.c1], gb, 2e-2)
