"""Collective tier (>= 2 B200): the fused P2P all-reduce + update against a single-GPU run on the global batch
(mean of per-rank mean gradients == global-batch mean gradient), determinism under rank skew, replica consistency,
the push protocol's bf16 wire, a dead rank, pumped epochs, and a non-LeNet module under DataParallel(comm="b200").

Process start-up and NCCL bootstrap dominate a multi-GPU test's run time, so all jobs of one world size run in ONE
spawn (``_p2p_worker.run_batch``); the tests below only read the per-job result files."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

MATCH_CASES = [("adadelta", 0.1, "push"), ("adam", 0.002, "push"), ("adadelta", 0.1, "oneshot"), ("sgd", 0.05, "oneshot"),
               ("adadelta", 0.1, "twoshot"), ("sgd", 0.05, "nvls")]
GENERIC_CASES = [("mlp", "adadelta", 0.5, -1, "auto"), ("mlp", "sgd", 0.05, 8192, "auto"), ("mlp", "adam", 0.01, 8192, "oneshot"),
                 ("resblock", "adadelta", 0.5, 1024, "auto"), ("resblock", "sgd", 0.05, 1024, "twoshot")]
GENERIC_AT_8 = (0, 4)            # the 8-GPU run keeps two of them
DEAD_CASES = [("push", True), ("push", False), ("oneshot", True)]


def _match_cfg(opt, lr, algo):
    return {"seed": 5, "per_rank_batch": 64, "n_batches": 3, "steps": 12, "opt": opt, "lr": lr, "graph": True,
            "overlap": True, "algo": algo, "tag": f"match_{opt}_{algo}"}


SKEW_BASE = {"seed": 5, "per_rank_batch": 32, "n_batches": 4, "steps": 150, "opt": "adadelta", "lr": 0.1, "dtype": "tf32"}
SKEW_RUNS = {"plain": dict(graph=True, overlap=True, jitter=False), "skew": dict(graph=True, overlap=True, jitter=True),
             "nograph": dict(graph=False, overlap=False, jitter=True)}
WIRE_BASE = {"seed": 5, "per_rank_batch": 32, "n_batches": 4, "steps": 60, "opt": "sgd", "lr": 0.05, "graph": True,
             "overlap": True, "algo": "push"}


def _generic_cfg(i):
    model, opt, lr, bucket_bytes, algo = GENERIC_CASES[i]
    return {"seed": 3, "model": model, "opt": opt, "lr": lr, "bucket_bytes": bucket_bytes, "algo": algo,
            "per_rank_batch": 16, "steps": 20, "tag": f"generic_{i}"}


def _jobs(world):
    jobs = [("train_native", _match_cfg(*c)) for c in MATCH_CASES]
    if world in (2, 8):
        jobs += [("train_native", dict(SKEW_BASE, tag=f"skew_{t}", **extra)) for t, extra in SKEW_RUNS.items()]
        jobs += [("train_native", dict(WIRE_BASE, tag=f"wire_{w}", wire=w)) for w in ("fp32", "bf16")]
        for wide in ((False, True) if world == 2 else (False,)):
            for tag, no_pump in (("pump", False), ("py", True)):
                jobs.append(("train_epochs", {"seed": 5, "per_rank_batch": 32, "n": 32 * 9 * world, "wide": wide,
                                              "tag": f"epochs_{int(wide)}_{tag}", "no_pump": no_pump}))
        for i in range(len(GENERIC_CASES)):
            if world == 2 or i in GENERIC_AT_8:
                jobs.append(("train_generic", _generic_cfg(i)))
    if world == 2:
        for algo, graph in DEAD_CASES:
            jobs.append(("train_with_dead_rank", {"seed": 5, "per_rank_batch": 32, "steps": 12, "alive": 5, "timeout_ms": 300,
                                                  "algo": algo, "graph": graph, "tag": f"dead_{algo}{int(graph)}"}))
    return jobs


_DONE = {}


@pytest.fixture(scope="session")
def batch(tmp_path_factory):
    """batch(world) -> directory with one ``<tag>.pt`` per job of that world size (spawned once per session)."""
    def get(world):
        if torch.cuda.device_count() < world:
            pytest.skip(f"needs {world} GPUs")
        if world not in _DONE:
            from _p2p_worker import run_batch
            from b200dp.parallel import launch
            out = str(tmp_path_factory.mktemp(f"w{world}"))
            launch.spawn(run_batch, {"out": out, "jobs": _jobs(world)}, world, backend="nccl", use_cuda=True, port=0)
            _DONE[world] = out
        return _DONE[world]
    return get


def _load(d, tag):
    err = os.path.join(d, f"{tag}.err")
    if os.path.exists(err):
        pytest.fail(f"job {tag} failed in the worker:\n{open(err).read()}")
    return torch.load(os.path.join(d, f"{tag}.pt"))


_REF = {}


def _single_gpu_reference(cfg, world):
    key = (cfg["opt"], cfg["lr"], world)
    if key in _REF:
        return _REF[key]
    from b200dp.engine import EagerStepper
    from b200dp.models.lenet import MNIST_LENET
    from b200dp.optim.flat import OptHyper
    from b200dp.parallel.launch import RankContext
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
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
    _REF[key] = (losses, {k: v.detach().cpu().clone() for k, v in st.model.state_dict().items()})
    return _REF[key]


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("opt,lr,algo", MATCH_CASES)
def test_p2p_matches_global_batch(batch, world, opt, lr, algo):
    """``nvls`` falls back to two-shot (with a warning) where the box has no multicast support."""
    cfg = _match_cfg(opt, lr, algo)
    got = _load(batch(world), cfg["tag"])
    ref_losses, ref_params = _single_gpu_reference(cfg, world)
    assert got["div"] == 0.0 and got["gen"] == cfg["steps"]
    for a, b in zip(got["losses"], ref_losses):
        assert abs(a - b) < 1e-3 * max(1.0, abs(b)), (got["losses"], ref_losses)
    for k in ref_params:
        if opt == "adam":
            # Adam divides by sqrt(v): where a gradient is ~0 every update is +-lr whatever its size, so two summation
            # orders (8 per-rank means vs one global mean) can walk single elements apart by up to lr per step; the bulk of
            # the parameters must still agree
            d = (got["params"][k] - ref_params[k]).abs()
            assert float(d.mean()) < 5e-4 and float(d.max()) <= 2 * lr * cfg["steps"], (k, float(d.mean()), float(d.max()))
        else:
            assert torch.allclose(got["params"][k], ref_params[k], atol=1e-3, rtol=1e-2), k


@pytest.mark.parametrize("world", [2, 8])
def test_p2p_deterministic_under_skew(batch, world):
    """Same data, with and without random per-rank delays and with/without graphs: bit-identical weights
    (default algorithm = the push protocol; default kernels = tcgen05 convolutions + tf32 GEMMs)."""
    d = batch(world)
    outs = {t: _load(d, f"skew_{t}") for t in SKEW_RUNS}
    for t in outs:
        assert outs[t]["div"] == 0.0
    for k in outs["plain"]["params"]:
        assert torch.equal(outs["plain"]["params"][k], outs["skew"]["params"][k]), k
        assert torch.equal(outs["plain"]["params"][k], outs["nograph"]["params"][k]), k


@pytest.mark.parametrize("world", [2, 8])
def test_push_bf16_wire_tracks_fp32_wire(batch, world):
    """bf16 gradient lines (half the NVLink bytes): replicas stay bit-identical, and the parameters stay close to
    the fp32-wire run (gradients rounded to 8 mantissa bits before the fp32 rank-order sum)."""
    d = batch(world)
    outs = {w: _load(d, f"wire_{w}") for w in ("fp32", "bf16")}
    for w in outs:
        assert outs[w]["div"] == 0.0 and outs[w]["gen"] == WIRE_BASE["steps"]
    worst = max(float((outs["fp32"]["params"][k] - outs["bf16"]["params"][k]).abs().max()) for k in outs["fp32"]["params"])
    assert 0.0 < worst < 5e-2, worst      # different bits (the wire really is bf16) ...
    for a, b in zip(outs["fp32"]["losses"], outs["bf16"]["losses"]):
        assert abs(a - b) < 2e-2 * max(1.0, abs(b)), (outs["fp32"]["losses"], outs["bf16"]["losses"])   # ... same trajectory


@pytest.mark.parametrize("world,wide", [(2, False), (2, True), (8, False)])
def test_pumped_epochs_match_python_loop_multi_gpu(batch, world, wide):
    """The native step pump drives every rank's epoch (graphs with the fused P2P reduce+update inside): same
    parameters, bit for bit, as the Python loop, and identical replicas."""
    d = batch(world)
    outs = {t: _load(d, f"epochs_{int(wide)}_{t}") for t in ("pump", "py")}
    for t in outs:
        assert outs[t]["div"] == 0.0
        assert outs[t]["gen"] == outs[t]["steps"] == 18
    assert outs["pump"]["pumped"] and not outs["py"]["pumped"]
    assert torch.equal(outs["pump"]["param"], outs["py"]["param"])


@pytest.mark.parametrize("algo,graph", DEAD_CASES)
def test_dead_rank_raises_instead_of_hanging(batch, algo, graph):
    """rank 1 stops stepping; rank 0 must raise from Trainer.check_peers within the time bound, with its
    generation counter still advancing (defined state), instead of hanging (reference train.py:38) or training on."""
    d = batch(2)
    tag = f"dead_{algo}{int(graph)}"
    r0, r1 = _load(d, f"{tag}_r0"), _load(d, f"{tag}_r1")
    assert r1["steps_done"] == 5 and not r1["raised"]
    assert r0["raised"] and r0["flag"] == 1
    assert r0["steps_done"] == 6                         # the first step without the peer is the one that times out
    assert r0["gen"] == r0["steps_done"]                 # bookkeeping went on: the ticket was still taken
    assert r0["t_detect"] < 30.0


@pytest.mark.parametrize("world", [2, 8])
@pytest.mark.parametrize("case", range(len(GENERIC_CASES)))
def test_generic_module_through_fused_kernels(batch, world, case):
    """DataParallel(comm="b200") serves any nn.Module (reference: average_gradients(model) + optimizer.step(),
    train.py:34-39, 51): parameters match NCCL all_reduce + torch.optim, replicas identical, --bucket-bytes changes
    the launch count."""
    if world == 8 and case not in GENERIC_AT_8:
        pytest.skip("two generic cases are kept for the 8-GPU run")
    cfg = _generic_cfg(case)
    out = _load(batch(world), cfg["tag"])
    assert out["div"] == 0.0 and out["gen"] == cfg["steps"]
    assert out["err"] < (2e-5 if cfg["opt"] == "adam" else 1e-5), out
    if cfg["bucket_bytes"] > 0:
        assert out["buckets"] > 1
        # push: one push per bucket + one final launch; handshake: one launch per bucket
        assert out["launches"] == (out["buckets"] + 1 if out["push"] else out["buckets"]), out
    else:
        assert out["buckets"] == 1
