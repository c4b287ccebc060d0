"""CPU unit tier: small components against the reference's own code or torch (tier "unit")."""
import argparse
import math
import os
import sys

import pytest
import torch

from b200dp.cli import build_parser
from b200dp.models.lenet import CIFAR_WIDE_LENET, MNIST_LENET, LeNet
from b200dp.optim.flat import FlatOptimizer, OptHyper, StepLR
from b200dp.parallel.arena import FlatArena, lenet_groups
from b200dp.parallel.p2p import oneshot_max_bytes
from b200dp.utils.logging import test_line as fmt_test_line
from b200dp.utils.logging import train_line as fmt_train_line


def _reference_parser(ref_dir):
    """Rebuild the reference's argparse exactly as train.py:113-133 does, by executing that block."""
    src = open(os.path.join(ref_dir, "train.py")).read()
    start = src.index("parser = argparse.ArgumentParser")
    end = src.rfind("\n", 0, src.index("parser.parse_args()")) + 1   # up to the line that parses
    import textwrap
    ns = {"argparse": argparse}
    line0 = src.rfind("\n", 0, start) + 1                     # keep the block's own indentation, then strip it
    exec(textwrap.dedent(src[line0:end]), ns)                 # only the parser construction: no torch, no side effects
    return ns["parser"]


def test_cli_has_the_reference_flags_with_the_same_defaults(ref_modules):
    ref = _reference_parser(ref_modules["dir"])
    ours = build_parser()
    ref_defaults = vars(ref.parse_args([]))
    our_defaults = vars(ours.parse_args([]))
    assert set(ref_defaults) <= set(our_defaults)
    for k, v in ref_defaults.items():
        assert our_defaults[k] == v, (k, our_defaults[k], v)
    # every reference option string exists here too
    ref_opts = {o for a in ref._actions for o in a.option_strings}
    our_opts = {o for a in ours._actions for o in a.option_strings}
    assert ref_opts <= our_opts, ref_opts - our_opts


def test_log_lines_are_the_reference_format(ref_modules):
    src = open(os.path.join(ref_modules["dir"], "train.py")).read()
    assert "'Train Epoch: {} [{}/{} ({:.0f}%)]\\tLoss: {:.6f}'" in src                      # train.py:53
    assert "'\\nTest set on {}: Average loss: {:.4f}, Accuracy: {}/{} ({:.0f}%)\\n'" in src  # train.py:74
    assert fmt_train_line(3, 15, 128, 7500, 59, -0.123456789) == \
        'Train Epoch: {} [{}/{} ({:.0f}%)]\tLoss: {:.6f}'.format(3, 15 * 128, 7500, 100. * 15 / 59, -0.123456789)
    assert fmt_test_line(0, -0.98765, 1200, 1250) == \
        '\nTest set on {}: Average loss: {:.4f}, Accuracy: {}/{} ({:.0f}%)\n'.format(0, -0.98765, 1200, 1250, 100. * 1200 / 1250)


@pytest.mark.parametrize("kind,kw,torch_cls,tkw", [
    ("adadelta", dict(lr=0.1), torch.optim.Adadelta, dict(lr=0.1)),                         # train.py:99
    ("sgd", dict(lr=0.05, momentum=0.9), torch.optim.SGD, dict(lr=0.05, momentum=0.9)),
    ("adam", dict(lr=1e-3), torch.optim.Adam, dict(lr=1e-3)),
    ("sgd", dict(lr=0.05, momentum=0.9, weight_decay=1e-2), torch.optim.SGD, dict(lr=0.05, momentum=0.9, weight_decay=1e-2)),
])
def test_flat_optimizer_matches_torch_optim(kind, kw, torch_cls, tkw):
    torch.manual_seed(0)
    m = LeNet(MNIST_LENET)
    ref = LeNet(MNIST_LENET)
    ref.load_state_dict(m.state_dict())
    arena = FlatArena(m)
    opt = FlatOptimizer(arena.param, arena.grad, OptHyper(kind, **kw))
    topt = torch_cls(ref.parameters(), **tkw)
    sched, tsched = StepLR(opt, 1, 0.9), torch.optim.lr_scheduler.StepLR(topt, 1, 0.9)     # train.py:101
    for step in range(6):
        x, y = torch.randn(8, 1, 28, 28), torch.randint(0, 10, (8,))
        for mod, o in ((m, None), (ref, topt)):
            for p in mod.parameters():
                if p.grad is not None:
                    p.grad.zero_()
            torch.nn.functional.nll_loss(mod(x), y).backward()
        opt.step_torch()
        topt.step()
        if step % 2 == 1:
            sched.step()
            tsched.step()
    for (n, a), (_, b) in zip(m.state_dict().items(), ref.state_dict().items()):
        assert torch.allclose(a, b, atol=1e-6, rtol=1e-5), n
    assert math.isclose(opt.lr, topt.param_groups[0]["lr"], rel_tol=1e-12)


@pytest.mark.parametrize("cfg", [MNIST_LENET, CIFAR_WIDE_LENET])
def test_arena_layout_contract(cfg):
    m = LeNet(cfg)
    names = [n for n, _ in m.named_parameters()]
    arena = FlatArena(m, lenet_groups(names))
    L = arena.layout
    assert [b.names for b in L.buckets] == [("FC.0.weight", "FC.0.bias", "FC.2.weight", "FC.2.bias"),
                                            ("ConvNet.7.weight", "ConvNet.7.bias"),
                                            ("ConvNet.1.weight", "ConvNet.1.bias", "ConvNet.4.weight", "ConvNet.4.bias")]
    covered = 0
    for b in L.buckets:                              # buckets: 128-float aligned, disjoint, cover every slot
        assert b.start % 128 == 0 and b.numel % 128 == 0
        for n in b.names:
            s = L.slots[n]
            assert b.start <= s.offset and s.offset + s.numel <= b.stop and s.offset % 4 == 0
            covered += s.numel
    assert covered == sum(p.numel() for p in m.parameters())
    # parameters are views of the arena: state_dict keeps the reference's keys and shapes (cnn.py:9-23)
    assert list(m.state_dict().keys()) == [n for n, _ in cfg.param_shapes()]
    for n, p in m.named_parameters():
        assert p.data_ptr() == arena.param.data_ptr() + 4 * L.slots[n].offset
        assert p.grad.data_ptr() == arena.grad.data_ptr() + 4 * L.slots[n].offset


def test_allreduce_algorithm_thresholds():
    assert oneshot_max_bytes(2) > (1 << 40)                      # two ranks: one-shot always (same bytes, one handshake)
    assert oneshot_max_bytes(8) == 2 << 20 and oneshot_max_bytes(4) == 4 << 20
    lenet_grad_bytes = 4 * sum(p.numel() for p in LeNet(MNIST_LENET).parameters())
    assert lenet_grad_bytes == 246824 and lenet_grad_bytes < oneshot_max_bytes(8)   # the whole LeNet gradient: one-shot


def test_pinned_ring_loader_cpu_semantics():
    """Per-rank batch streamer on the CPU path: every sample of the shard exactly once per epoch, ragged last batch,
    a different order every epoch, never a sample from another rank's shard (train.py:82-96)."""
    from b200dp.data.shard import shard_for_rank
    from b200dp.data.synthetic import MNIST_SPEC, PinnedRingLoader, SyntheticImages
    base = SyntheticImages(1000, MNIST_SPEC, seed=1)
    sh = shard_for_rank(base, 1, 3)
    lo, hi = sh.bounds()
    loader = PinnedRingLoader(base.images, base.targets, lo, hi, 64, MNIST_SPEC, shuffle=True, fmt="u8", device=None, seed=9)
    assert len(loader) == math.ceil((hi - lo) / 64) and len(loader.dataset) == hi - lo
    orders = []
    for _ in range(2):
        seen, sizes = [], []
        for x, y in loader:
            assert x.dtype == torch.uint8 and x.shape[1:] == (1, 28, 28) and y.dtype == torch.int64
            sizes.append(len(x))
            for img, t in zip(x, y):                       # identify each sample by content
                seen.append((int(t), int(img.sum())))
        assert sizes[:-1] == [64] * (len(sizes) - 1) and sum(sizes) == hi - lo
        orders.append(seen)
    want = sorted((int(base.targets[i]), int(base.images[i].sum())) for i in range(lo, hi))
    assert sorted(orders[0]) == want and sorted(orders[1]) == want and orders[0] != orders[1]


def test_bench_contract_without_gpu():
    """bench.py on a box without CUDA: both arms print one JSON line saying so and exit 0 (the driver's contract)."""
    import json
    import subprocess
    if torch.cuda.is_available():
        pytest.skip("only meaningful without a GPU")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for impl in ("b200dp", "reference"):
        r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", impl, "--steps", "2", "--warmup", "1"],
                           capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-500:]
        line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1]
        d = json.loads(line)
        assert d["impl"] == impl and "unavailable" in d


def test_native_library_builds_for_sm100a():
    """The in-tree CUDA library cross-compiles (nvcc, sm_100a) and exports the C ABI of include/b200dp.h."""
    import ctypes
    import shutil
    if shutil.which("nvcc") is None and not os.path.exists("/usr/local/cuda/bin/nvcc"):
        pytest.skip("no nvcc on this box")
    from b200dp import build as b
    lib = b.build()
    assert os.path.exists(lib) and not b.needs_build()
    h = ctypes.CDLL(lib)
    header = open(os.path.join(os.path.dirname(os.path.dirname(lib)), "..", "include", "b200dp.h")).read()
    import re
    names = set(re.findall(r"\b(b200dp_[a-z0-9_]+)\s*\(", header))
    assert len(names) > 30
    missing = [n for n in sorted(names) if not hasattr(h, n)]
    assert not missing, missing
    # and the other way round: every entry point the sources define, and every one the ctypes layer binds, is declared
    root = os.path.join(os.path.dirname(os.path.dirname(lib)), "..")
    import glob
    defined = set()
    for f in glob.glob(os.path.join(root, "csrc", "*.cu")):
        defined |= set(re.findall(r'extern "C"[^;{(]*?\b(b200dp_[a-z0-9_]+)\s*\(', open(f).read()))
    assert len(defined) > 30 and not sorted(defined - names), sorted(defined - names)
    bound = set(re.findall(r"\blib\.(b200dp_[a-z0-9_]+)\.", open(os.path.join(root, "b200dp", "ops", "native.py")).read()))
    assert len(bound) > 20 and not sorted(bound - names), sorted(bound - names)


def test_push_threshold_and_new_cli_flags():
    """The push protocol's auto threshold shrinks with the world size (measured cross-overs, profiles/allreduce_sweep.md)
    and always covers the reference LeNet; the round-2 flags parse and default to the measured choices."""
    from b200dp.cli import parse_args
    from b200dp.parallel.p2p import ll_max_bytes
    lenet_bytes = 61952 * 4                                   # the reference model's padded arena
    assert ll_max_bytes(2) == 1 << 20 and ll_max_bytes(8) == 256 << 10
    assert all(ll_max_bytes(w) >= lenet_bytes for w in range(1, 9))
    a = parse_args([])
    assert (a.wire, a.conv, a.peer_timeout_ms, a.allreduce_algo, a.data_root) == ("auto", "auto", 0, "auto", None)
    a = parse_args(["--wire", "bf16", "--conv", "hybrid", "--peer-timeout-ms", "500", "--allreduce-algo", "push",
                    "--data-root", "/tmp/x", "--dtype", "bf16"])
    assert (a.wire, a.conv, a.peer_timeout_ms, a.allreduce_algo, a.data_root, a.dtype) == ("bf16", "hybrid", 500, "push", "/tmp/x", "bf16")


def test_loader_prepares_next_epoch_order():
    """Epoch orders are the same permutations whether they were prepared ahead (thread for large shards, inline
    ``prepare_next`` for small ones) or computed on demand."""
    import torch
    from b200dp.data.synthetic import MNIST_SPEC, PinnedRingLoader, SyntheticImages
    ds = SyntheticImages(300, MNIST_SPEC, seed=0)
    mk = lambda: PinnedRingLoader(ds.images, ds.targets, 10, 290, 32, MNIST_SPEC, fmt="f32", seed=4)   # noqa: E731
    a, b = mk(), mk()
    for _ in range(3):
        assert b.peek_next() is None                           # nothing prepared on the on-demand loader
        oa = a._order(); a.epoch += 1; a.prepare_next()        # what the pumped training loop does
        ob = b._order_for(b.epoch); b.epoch += 1               # on demand
        assert torch.equal(oa, ob) and sorted(oa.tolist()) == list(range(10, 290))
        assert torch.equal(a.peek_next(), a._order_for(a.epoch))    # the coming epoch's order, not consumed by the peek


def test_core_shares_are_disjoint_and_cover_whole_slices():
    """--pin-cores: every rank of a node gets its own contiguous slice of the allowed cores; too few cores: no pinning."""
    from b200dp.cli import parse_args
    from b200dp.parallel.launch import core_share
    allowed = set(range(4, 36))                         # 32 cores, not starting at 0
    shares = [core_share(r, 8, allowed) for r in range(8)]
    assert all(len(s) == 4 for s in shares)
    assert set().union(*shares) == allowed and sum(len(s) for s in shares) == 32
    assert shares[0] == {4, 5, 6, 7} and shares[7] == {32, 33, 34, 35}
    assert core_share(0, 8, set(range(8))) is None      # one core per rank
    assert core_share(1, 3, set(range(8))) == {2, 3}    # the remainder stays unused
    assert parse_args(["--pin-cores"]).pin_cores and not parse_args([]).pin_cores
