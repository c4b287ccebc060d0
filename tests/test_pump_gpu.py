"""Native step pump (csrc/pump.cu, b200dp/data/pump.py): an epoch driven by the C++ loop must be the
same epoch as the Python loop — same sample order, same log lines, bit-identical parameters."""
import io

import pytest
import torch

from b200dp.models.lenet import MNIST_LENET

pytestmark = pytest.mark.gpu


def _run_epochs(no_pump: bool, n: int, batch: int, epochs: int, log_interval: int = 4):
    from b200dp.data.synthetic import MNIST_SPEC, PinnedRingLoader, SyntheticImages
    from b200dp.engine import Trainer
    from b200dp.models.fused import NativeStepper
    from b200dp.optim.flat import OptHyper
    from b200dp.parallel.launch import RankContext
    from b200dp.utils.logging import RankLogger
    torch.manual_seed(3)
    dev = torch.device("cuda", 0)
    ctx = RankContext(0, 1, 0, dev, "nccl")
    base = SyntheticImages(n, MNIST_SPEC, seed=11)
    loader = PinnedRingLoader(base.images, base.targets, 0, n, batch, MNIST_SPEC, shuffle=True, fmt="u8",
                              device=dev, seed=4)
    st = NativeStepper(ctx, MNIST_LENET, OptHyper("adadelta", lr=0.1), batch=batch, loss_mode="ref", graph=True,
                       dtype="fp32")
    out = io.StringIO()
    tr = Trainer(ctx, st, RankLogger(True, None, stream=out), log_interval=log_interval, no_pump=no_pump)
    for e in range(1, epochs + 1):
        tr.train_epoch(loader, e)
    torch.cuda.synchronize()
    used_pump = tr._pump_for(loader) is not None
    params = st.arena.param.clone()
    steps = int(st.step_counter[0])
    st.close()
    return params, out.getvalue(), steps, used_pump, tr.global_step


@pytest.mark.parametrize("n,batch", [(64 * 13, 64), (64 * 9 + 17, 64)])   # exact epochs, and a ragged last batch
def test_pumped_epoch_equals_python_epoch(n, batch):
    p_py, log_py, steps_py, pump_py, gs_py = _run_epochs(True, n, batch, epochs=2)
    p_pp, log_pp, steps_pp, pump_pp, gs_pp = _run_epochs(False, n, batch, epochs=2)
    assert not pump_py and pump_pp
    assert steps_py == steps_pp == gs_py == gs_pp == 2 * -(-n // batch)
    assert torch.equal(p_py, p_pp), (p_py - p_pp).abs().max()
    assert log_py == log_pp            # same lines, same losses to the printed precision
    assert log_pp.count("Train Epoch") >= 2


def test_pump_direct_api_losses():
    """StepPump.run: losses land in the pinned array, one per step, equal to the step-by-step path."""
    from b200dp.data.synthetic import MNIST_SPEC, SyntheticImages
    from b200dp.models.fused import NativeStepper
    from b200dp.optim.flat import OptHyper
    from b200dp.parallel.launch import RankContext
    dev = torch.device("cuda", 0)
    ctx = RankContext(0, 1, 0, dev, "nccl")
    base = SyntheticImages(32 * 10, MNIST_SPEC, seed=5)
    order = torch.randperm(32 * 10, generator=torch.Generator().manual_seed(1))

    def make():
        torch.manual_seed(6)
        return NativeStepper(ctx, MNIST_LENET, OptHyper("sgd", lr=0.05), batch=32, loss_mode="xent", graph=True,
                             dtype="fp32")
    a = make()
    ref = []
    for k in range(10):
        idx = order[k * 32:(k + 1) * 32]
        ref.append(float(a.train_step(base.images[idx].to(dev), base.targets[idx].to(dev))))
    b = make()
    pump = b.make_pump(base.images, base.targets, 32)
    losses = torch.full((10,), float("nan")).pin_memory()
    pump.run(order, 4, losses)            # includes the stepper's two eager warm-up steps
    pump.run(order[4 * 32:], 6, losses[4:])
    pump.drain()
    assert torch.equal(a.arena.param, b.arena.param)
    assert [float(v) for v in losses] == ref
    assert b.optimizer.step_count == 10 and int(b.step_counter[0]) == 10
    a.close(); b.close()


def test_pump_pregather_hit_and_miss():
    """StepPump.pregather puts the next run's first batches into their ring slots ahead of time.  A run whose order starts
    with the same indices uses them; a run with another order must notice and gather its own — in both cases the
    parameters and losses equal the step-by-step path."""
    from b200dp.data.synthetic import MNIST_SPEC, SyntheticImages
    from b200dp.models.fused import NativeStepper
    from b200dp.optim.flat import OptHyper
    from b200dp.parallel.launch import RankContext
    dev = torch.device("cuda", 0)
    ctx = RankContext(0, 1, 0, dev, "nccl")
    B, n = 32, 16
    base = SyntheticImages(B * n, MNIST_SPEC, seed=9)
    g = torch.Generator().manual_seed(2)
    orders = [torch.randperm(B * n, generator=g) for _ in range(3)]       # three "epochs" of 16, 5 and 5 steps
    steps = [n, 5, 5]

    def make():
        torch.manual_seed(7)
        return NativeStepper(ctx, MNIST_LENET, OptHyper("sgd", lr=0.05), batch=B, loss_mode="xent", graph=True,
                             dtype="fp32")
    a = make()
    ref = []
    for order, k_n in zip(orders, steps):
        for k in range(k_n):
            idx = order[k * B:(k + 1) * B]
            ref.append(float(a.train_step(base.images[idx].to(dev), base.targets[idx].to(dev))))
    b = make()
    pump = b.make_pump(base.images, base.targets, B)
    losses = torch.full((sum(steps),), float("nan")).pin_memory()
    pump.run(orders[0], steps[0], losses)
    pump.pregather(orders[1], steps[1])               # hit: the next run starts with these batches
    pump.run(orders[1], steps[1], losses[steps[0]:])
    pump.pregather(orders[0], steps[2])               # miss: prepared for another order than the one that runs
    pump.run(orders[2], steps[2], losses[steps[0] + steps[1]:])
    pump.drain()
    assert [float(v) for v in losses] == ref
    assert torch.equal(a.arena.param, b.arena.param)
    a.close(); b.close()
