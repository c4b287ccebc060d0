"""Model / checkpoint contract vs the reference's cnn.LeNet (cnn.py:4-30, train.py:107-108)."""
import torch

from b200dp import ckpt
from b200dp.models.lenet import CIFAR_WIDE_LENET, LeNet, LeNetConfig, loss_fn


def test_state_dict_contract(ref_modules):
    ref, ours = ref_modules["cnn"].LeNet(), LeNet()
    rs, os_ = ref.state_dict(), ours.state_dict()
    assert list(rs.keys()) == list(os_.keys())
    for k in rs:
        assert rs[k].shape == os_[k].shape and rs[k].dtype == os_[k].dtype == torch.float32
    assert sum(p.numel() for p in ours.parameters()) == 61706 == LeNetConfig().num_params()
    assert [n for n, _ in ours.named_parameters()] == [n for n, _ in LeNetConfig().param_shapes()]
    assert LeNetConfig().flops_per_image_fwd() == 833040


def test_forward_matches_reference(ref_modules):
    torch.manual_seed(0)
    ref, ours = ref_modules["cnn"].LeNet(), LeNet()
    ours.load_state_dict(ref.state_dict(), strict=True)
    x = torch.randn(5, 1, 28, 28)
    assert torch.equal(ref(x), ours(x))
    y = torch.randint(0, 10, (5,))
    assert torch.allclose(loss_fn(ours(x), y, "ref"), -ours(x)[torch.arange(5), y].mean())


def test_checkpoint_roundtrip_both_ways(ref_modules, tmp_path):
    torch.manual_seed(1)
    ours = LeNet()
    path = str(tmp_path / "mnist_cnn.pt")
    ckpt.save_model(ours, path)
    ref = ref_modules["cnn"].LeNet()
    ref.load_state_dict(torch.load(path), strict=True)          # reference loads our file
    x = torch.randn(3, 1, 28, 28)
    assert torch.equal(ref(x), ours(x))
    ref2 = ref_modules["cnn"].LeNet()
    path2 = str(tmp_path / "ref.pt")
    torch.save(ref2.state_dict(), path2)                         # what train.py:108 writes
    ours2 = ckpt.load_model(LeNet(), path2)
    assert torch.equal(ref2(x), ours2(x))


def test_wide_variant_shapes():
    m = LeNet(CIFAR_WIDE_LENET)
    out = m(torch.randn(2, 3, 32, 32))
    assert out.shape == (2, 10)
    assert list(m.state_dict().keys())[0] == "ConvNet.1.weight"
    assert m.state_dict()["ConvNet.4.weight"].shape == (64, 24, 5, 5)


def test_full_state_resume(tmp_path):
    from b200dp.optim.flat import FlatOptimizer, OptHyper, StepLR
    from b200dp.parallel.arena import FlatArena
    torch.manual_seed(2)
    m = LeNet()
    arena = FlatArena(m)
    opt = FlatOptimizer(arena.param, arena.grad, OptHyper("adadelta", lr=0.1))
    sch = StepLR(opt, 1, 0.9)
    for q in m.parameters():
        q.grad.normal_()          # views into arena.grad; padding stays zero
    opt.step_torch()
    sch.step()
    p = str(tmp_path / "full.pt")
    ckpt.save_full(p, model=m, optimizer=opt, scheduler=sch, epoch=3, step=42)
    m2 = LeNet()
    a2 = FlatArena(m2)
    o2 = FlatOptimizer(a2.param, a2.grad, OptHyper("adadelta", lr=0.1))
    s2 = StepLR(o2, 1, 0.9)
    meta = ckpt.load_full(p, model=m2, optimizer=o2, scheduler=s2)
    assert meta["epoch"] == 3 and meta["step"] == 42
    assert torch.equal(a2.param, arena.param) and torch.equal(o2.state0, opt.state0)
    assert abs(o2.lr - 0.09) < 1e-12 and o2.step_count == 1
