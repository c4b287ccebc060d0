"""End-to-end GPU parity tier (last row): the native step against the reference's OWN training loop
(train.py:42-57, run unmodified from baseline/_ref or /root/reference) from the same initial weights, and the generic
DataParallel(comm="b200") path against torch.optim at world size 1."""
import contextlib
import io
import os
import re
import sys
import types

import pytest
import torch

pytestmark = pytest.mark.gpu


class _Loader:
    """What the reference's train() needs from a DataLoader: iteration, len(), .dataset with a len()."""

    def __init__(self, xs, ys):
        self.xs, self.ys = xs, ys
        self.dataset = range(sum(len(y) for y in ys))

    def __len__(self):
        return len(self.xs)

    def __iter__(self):
        return iter(zip(self.xs, self.ys))


@pytest.fixture(scope="module")
def single_rank_group():
    """The reference's average_gradients() calls dist.get_world_size(): give it a one-rank group."""
    import torch.distributed as dist
    from b200dp.parallel.launch import free_port
    created = False
    if not dist.is_initialized():
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(free_port())
        dist.init_process_group("gloo", rank=0, world_size=1)
        created = True
    yield
    if created:
        dist.destroy_process_group()


@pytest.mark.parametrize("graph", [True, False])
def test_native_step_matches_reference_training_loop(ref_modules, single_rank_group, graph):
    """50 steps, reference loss (nll on probabilities), Adadelta lr 0.1: per-step losses and final parameters of the
    fp32 CUDA-core path agree with the reference loop (cuDNN/cuBLAS/ATen kernels) to 1e-5."""
    from b200dp.data.synthetic import MNIST_SPEC, normalise_reference
    from b200dp.models.fused import NativeStepper
    from b200dp.models.lenet import MNIST_LENET, LeNet
    from b200dp.optim.flat import OptHyper
    from b200dp.parallel.launch import RankContext

    sys.path.insert(0, ref_modules["dir"])
    try:
        import train as ref_train            # the reference's module, unmodified
    finally:
        sys.path.remove(ref_modules["dir"])
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = torch.device("cuda", 0)
    steps, batch = 50, 64
    g = torch.Generator().manual_seed(11)
    xs = [torch.randint(0, 256, (batch, 1, 28, 28), dtype=torch.uint8, generator=g).to(dev) for _ in range(steps)]
    ys = [torch.randint(0, 10, (batch,), generator=g).to(dev) for _ in range(steps)]

    torch.manual_seed(5)
    ref_model = ref_modules["cnn"].LeNet().cuda(0)                         # train.py:98
    ours = LeNet(MNIST_LENET)
    ours.load_state_dict({k: v.detach().cpu().clone() for k, v in ref_model.state_dict().items()})
    ref_opt = torch.optim.Adadelta(ref_model.parameters(), lr=0.1)          # train.py:99
    args = types.SimpleNamespace(log_interval=1, dry_run=False)
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        ref_train.train(args, ref_model, 0, _Loader([normalise_reference(x, MNIST_SPEC) for x in xs], ys), ref_opt, 1)
    ref_losses = [float(m) for m in re.findall(r"Loss: (-?[0-9.]+)", buf.getvalue())]
    assert len(ref_losses) == steps

    st = NativeStepper(RankContext(0, 1, 0, dev, "nccl"), MNIST_LENET, OptHyper("adadelta", lr=0.1), batch=batch,
                       loss_mode="ref", graph=graph, dtype="fp32", module=ours)
    got = [float(st.train_step(x, y)) for x, y in zip(xs, ys)]
    torch.cuda.synchronize()
    for i, (a, b) in enumerate(zip(got, ref_losses)):
        assert abs(a - b) < 1e-5, (i, a, b)
    for k, v in ref_model.state_dict().items():
        assert torch.allclose(st.model.state_dict()[k], v, atol=1e-5, rtol=0), (k, float((st.model.state_dict()[k] - v).abs().max()))
    st.close()


@pytest.mark.parametrize("opt,lr", [("adadelta", 0.5), ("sgd", 0.05), ("adam", 0.01)])
@pytest.mark.parametrize("bucket_bytes", [-1, 4096])
def test_generic_dataparallel_world1_matches_torch_optim(opt, lr, bucket_bytes):
    """DataParallel(comm="b200") on a non-LeNet module at world size 1: the fused update kernel against torch.optim."""
    import copy
    from torch import nn
    from b200dp.optim.flat import OptHyper
    from b200dp.parallel.ddp import DataParallel
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = torch.device("cuda", 0)
    torch.manual_seed(3)
    model = nn.Sequential(nn.Linear(64, 96), nn.Tanh(), nn.Linear(96, 48), nn.Tanh(), nn.Linear(48, 10)).to(dev)
    ref = copy.deepcopy(model)
    ddp = DataParallel(model, OptHyper(opt, lr=lr), bucket_bytes=bucket_bytes, comm="b200")
    ropt = {"adadelta": lambda p: torch.optim.Adadelta(p, lr=lr), "sgd": lambda p: torch.optim.SGD(p, lr=lr, momentum=0.9),
            "adam": lambda p: torch.optim.Adam(p, lr=lr)}[opt](ref.parameters())
    lossf = nn.CrossEntropyLoss()
    g = torch.Generator().manual_seed(4)
    for _ in range(25):
        x, y = torch.randn(32, 64, generator=g).to(dev), torch.randint(0, 10, (32,), generator=g).to(dev)
        ddp.zero_grad(); lossf(ddp(x), y).backward(); ddp.step()
        ropt.zero_grad(); lossf(ref(x), y).backward(); ropt.step()
    torch.cuda.synchronize()
    ddp.check_error()
    assert int(ddp.step_counter[0]) == 25
    for p, q in zip(model.parameters(), ref.parameters()):
        assert torch.allclose(p, q, atol=2e-6, rtol=1e-5), float((p - q).abs().max())
    assert list(model.state_dict().keys()) == list(ref.state_dict().keys())


@pytest.mark.parametrize("dtype,conv,tol", [("auto", "auto", 2e-2), ("bf16", "auto", 2e-2), ("tf32", "simt", 5e-3)])
def test_reduced_precision_paths_track_fp32_over_200_steps(dtype, conv, tol):
    """The default kernel mix (bf16 tcgen05 convolutions, tf32 tcgen05 backward GEMMs) against the exact fp32 CUDA-core
    path from the same initial weights on the same data: the loss trajectories must stay together over 200 Adadelta
    steps (cross-entropy loss, learnable synthetic classes), and both must learn."""
    from b200dp.data.synthetic import MNIST_SPEC, SyntheticImages
    from b200dp.models.fused import NativeStepper
    from b200dp.models.lenet import MNIST_LENET, LeNet
    from b200dp.optim.flat import OptHyper
    from b200dp.parallel.launch import RankContext
    dev = torch.device("cuda", 0)
    batch, steps = 128, 200
    data = SyntheticImages(batch * 20, MNIST_SPEC, seed=9)
    xs = [data.images[i * batch:(i + 1) * batch].to(dev) for i in range(20)]
    ys = [data.targets[i * batch:(i + 1) * batch].to(dev) for i in range(20)]
    curves = {}
    for name, kw in (("fp32", dict(dtype="fp32", conv="simt")), ("test", dict(dtype=dtype, conv=conv))):
        torch.manual_seed(5)
        st = NativeStepper(RankContext(0, 1, 0, dev, "nccl"), MNIST_LENET, OptHyper("adadelta", lr=0.1), batch=batch,
                           loss_mode="xent", graph=True, module=LeNet(MNIST_LENET), **kw)
        curves[name] = [float(st.train_step(xs[i % 20], ys[i % 20])) for i in range(steps)]
        if name == "test" and (dtype != "tf32"):
            assert st.net.conv2_tc and st.net.convbwd_tc and not st.net.conv1_tc      # the hybrid really ran
        st.close()
    a, b = torch.tensor(curves["fp32"]), torch.tensor(curves["test"])
    assert float(a[-10:].mean()) < 0.5 * float(a[:10].mean())          # learning, not noise
    assert float(b[-10:].mean()) < 0.5 * float(b[:10].mean())
    drift = float((a - b).abs().max())
    assert drift < tol * max(1.0, float(a.abs().max())), drift
