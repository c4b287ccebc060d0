"""Collective tier (>=2 B200): the fused P2P all-reduce + update against a single-GPU run on the
global batch (mean of per-rank mean gradients == global-batch mean gradient), determinism under
rank skew, and replica consistency."""
import os

import pytest
import torch

from _p2p_worker import train_native

pytestmark = pytest.mark.gpu


def _need(n):
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs")


def _single_gpu_reference(cfg, world):
    from b200dp.engine import EagerStepper
    from b200dp.models.lenet import MNIST_LENET
    from b200dp.optim.flat import OptHyper
    from b200dp.parallel.launch import RankContext
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(cfg["seed"])
    dev = torch.device("cuda", 0)
    st = EagerStepper(RankContext(0, 1, 0, dev, "nccl"), MNIST_LENET, OptHyper(cfg["opt"], lr=cfg["lr"]),
                      loss_mode="xent")
    g = torch.Generator().manual_seed(77)
    b = cfg["per_rank_batch"]
    gx = torch.randint(0, 256, (cfg["n_batches"], world * b, 1, 28, 28), dtype=torch.uint8, generator=g)
    gy = torch.randint(0, 10, (cfg["n_batches"], world * b), generator=g)
    losses = []
    for i in range(cfg["steps"]):
        j = i % cfg["n_batches"]
        loss = st.train_step(gx[j].to(dev), gy[j].to(dev))
        if i < 8 or i == cfg["steps"] - 1:
            losses.append(float(loss))
    return losses, {k: v.detach().cpu().clone() for k, v in st.model.state_dict().items()}


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("opt,lr,algo", [("adadelta", 0.1, "oneshot"), ("sgd", 0.05, "oneshot"),
                                         ("adadelta", 0.1, "twoshot"), ("sgd", 0.05, "nvls")])
def test_p2p_matches_global_batch(tmp_path, world, opt, lr, algo):
    """``nvls`` falls back to two-shot (with a warning) where the box has no multicast support."""
    _need(world)
    from b200dp.parallel import launch
    cfg = {"seed": 5, "per_rank_batch": 64, "n_batches": 3, "steps": 12, "opt": opt, "lr": lr, "graph": True,
           "overlap": True, "out": str(tmp_path), "tag": "a", "algo": algo}
    launch.spawn(train_native, cfg, world, backend="nccl", use_cuda=True, port=0)
    got = torch.load(os.path.join(tmp_path, "native_a.pt"))
    ref_losses, ref_params = _single_gpu_reference(cfg, world)
    assert got["div"] == 0.0 and got["gen"] == cfg["steps"]
    for a, b in zip(got["losses"], ref_losses):
        assert abs(a - b) < 1e-3 * max(1.0, abs(b)), (got["losses"], ref_losses)
    for k in ref_params:
        assert torch.allclose(got["params"][k], ref_params[k], atol=1e-3, rtol=1e-2), k


@pytest.mark.parametrize("world", [2, 8])
def test_p2p_deterministic_under_skew(tmp_path, world):
    """Same data, with and without random per-rank delays and with/without graphs: bit-identical weights."""
    _need(world)
    from b200dp.parallel import launch
    base = {"seed": 5, "per_rank_batch": 32, "n_batches": 4, "steps": 150, "opt": "adadelta", "lr": 0.1,
            "out": str(tmp_path), "dtype": "tf32"}
    runs = {"plain": dict(graph=True, overlap=True, jitter=False), "skew": dict(graph=True, overlap=True, jitter=True),
            "nograph": dict(graph=False, overlap=False, jitter=True)}
    outs = {}
    for tag, extra in runs.items():
        launch.spawn(train_native, dict(base, tag=tag, **extra), world, backend="nccl", use_cuda=True, port=0)
        outs[tag] = torch.load(os.path.join(tmp_path, f"native_{tag}.pt"))
        assert outs[tag]["div"] == 0.0
    for k in outs["plain"]["params"]:
        assert torch.equal(outs["plain"]["params"][k], outs["skew"]["params"][k]), k
        assert torch.equal(outs["plain"]["params"][k], outs["nograph"]["params"][k]), k


@pytest.mark.parametrize("world,wide", [(2, False), (2, True), (8, False)])
def test_pumped_epochs_match_python_loop_multi_gpu(tmp_path, world, wide):
    """The native step pump drives every rank's epoch (graphs with the fused P2P reduce+update inside): same
    parameters, bit for bit, as the Python loop, and identical replicas."""
    _need(world)
    from _p2p_worker import train_epochs
    from b200dp.parallel import launch
    base = {"seed": 5, "per_rank_batch": 32, "n": 32 * 9 * world, "out": str(tmp_path), "wide": wide}
    outs = {}
    for tag, no_pump in (("pump", False), ("py", True)):
        launch.spawn(train_epochs, dict(base, tag=tag, no_pump=no_pump), world, backend="nccl", use_cuda=True, port=0)
        outs[tag] = torch.load(os.path.join(tmp_path, f"epochs_{tag}.pt"))
        assert outs[tag]["div"] == 0.0
        assert outs[tag]["gen"] == outs[tag]["steps"] == 18
    assert outs["pump"]["pumped"] and not outs["py"]["pumped"]
    assert torch.equal(outs["pump"]["param"], outs["py"]["param"])
