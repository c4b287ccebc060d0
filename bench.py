#!/usr/bin/env python
"""Headline benchmark: cnn.py (LeNet) data-parallel training throughput, images/sec.

Metric / config are BASELINE.json's: "cnn.py images/sec (device-timed, max over ranks) at
1/2/4/8 B200", configs[1] = per-GPU batch 256 on synthetic data of train.py's dataset shape
(1x28x28, 10 classes), weak scaling.  Reference optimizer (Adadelta lr 0.1) and loss
(nll_loss on softmax probabilities) are kept, and every step runs forward, backward, the
gradient reduction and the optimizer update.

    python bench.py --gpus 1 --steps 200 --warmup 20
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
           --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...   # the unmodified reference from baseline/_ref

Two numbers per run:
  value  device-timed kernel path: inputs are already on the device (a pool larger than the
         126 MB L2, cycled), K steps between CUDA events, barrier + synchronize on both sides,
         max over ranks.
  e2e    through the public API (NativeStepper.train_step fed by PinnedRingLoader): every step
         copies its batch from pinned host memory to the device and reads the loss back.
"""
from __future__ import annotations

import argparse
import contextlib
import io
import json
import math
import os
import sys
import time
import types

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner to fd 1 on the first
# collective), so the real stdout is kept aside for the result line and fd 1 is pointed at stderr for everything else.
_RESULT_OUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def _emit(obj):
    print(json.dumps(obj), file=_RESULT_OUT, flush=True)


PER_GPU_BATCH = 256
L2_BYTES = 126 * 1024 * 1024


def _parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=2000)
    p.add_argument("--warmup", type=int, default=200)
    p.add_argument("--impl", choices=["b200dp", "reference"], default="b200dp")
    p.add_argument("--config", choices=["lenet_b256", "strong4096", "wide_cifar"], default="lenet_b256")
    p.add_argument("--optimizer", default="adadelta")
    p.add_argument("--no-graph", action="store_true")
    p.add_argument("--no-overlap", action="store_true")
    p.add_argument("--no-pump", action="store_true", help="e2e: drive every step from Python")
    p.add_argument("--pin-cores", action="store_true", help="pin each rank's host threads to its own slice of the cores")
    p.add_argument("--conv", choices=["auto", "simt", "tc", "hybrid"], default="auto",
                   help="convolution kernels: CUDA-core fp32, window-tile tcgen05 (bf16 operands), or hybrid = first layer "
                        "on CUDA cores + everything after it on tcgen05 (auto: measured-fastest for the model)")
    p.add_argument("--algo", choices=["auto", "push", "oneshot", "twoshot", "nvls"], default="auto",
                   help="gradient all-reduce protocol (auto: push for whole-model gradients up to 1 MB)")
    p.add_argument("--wire", choices=["auto", "fp32", "bf16"], default="auto", help="gradient wire format of the push protocol")
    p.add_argument("--input", choices=["u8", "f32"], default="u8", help="host->device input format (e2e)")
    p.add_argument("--dtype", choices=["auto", "fp32", "tf32", "bf16"], default="auto",
                   help="fp32 = exact CUDA-core kernels everywhere; tf32 = tcgen05 GEMMs on fp32 tensors; bf16 = bf16 "
                        "tcgen05 convolutions + bf16 gradient wire; auto = measured-fastest mix")
    return p.parse_args()


def _init_dist(n):
    """Join the group when launched by torchrun; single process otherwise."""
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", rank))
    if world != n:
        if world == 1 and n > 1:
            raise SystemExit(f"--gpus {n} needs torchrun with --nproc-per-node {n}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1 or "RANK" in os.environ:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    return rank, world, local, dev


def _sync_all():
    if dist.is_initialized():
        dist.barrier()
    torch.cuda.synchronize()


def _max_ranks(ms, dev):
    if dist.is_initialized() and dist.get_world_size() > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return ms


def _model_cfg(name):
    from b200dp.data.synthetic import CIFAR_SPEC, MNIST_SPEC
    from b200dp.models.lenet import CIFAR_WIDE_LENET, MNIST_LENET
    if name == "wide_cifar":
        return CIFAR_WIDE_LENET, CIFAR_SPEC
    return MNIST_LENET, MNIST_SPEC


def _dtype_label(stepper):
    """What actually computes: bf16 wherever a tcgen05 convolution runs, tf32 for tcgen05 GEMMs, fp32 elsewhere."""
    net = stepper.net
    parts = []
    if net.conv_tc:
        which = [n for n, on in (("conv1 fprop", net.conv1_tc), ("conv2 fprop", net.conv2_tc),
                                 ("conv dgrad+wgrad", net.convbwd_tc)) if on]
        parts.append("bf16 tcgen05 (" + ", ".join(which) + ")")
    if stepper.dtype != "fp32":
        parts.append("tf32 tcgen05 GEMMs")
    if not parts:
        return "fp32 CUDA cores everywhere"
    rest = "fp32 CUDA cores elsewhere" if not (net.conv1_tc and stepper.dtype != "fp32") else "fp32 head"
    return " + ".join(parts) + "; " + rest + "; fp32 accumulation, storage and master weights"


def run_own(args):
    from b200dp.data.synthetic import PinnedRingLoader
    from b200dp.models.fused import NativeStepper
    from b200dp.ops import native
    from b200dp.optim.flat import OptHyper
    from b200dp.parallel.launch import RankContext
    from b200dp.utils.clocks import ClockSampler

    rank, world, local, dev = _init_dist(args.gpus)
    if args.pin_cores and world > 1:
        from b200dp.parallel.launch import pin_to_core_share
        pin_to_core_share(local, int(os.environ.get("LOCAL_WORLD_SIZE", world)))
    cfg, spec = _model_cfg(args.config)
    batch = PER_GPU_BATCH if args.config != "strong4096" else math.ceil(4096 / world)
    ctx = RankContext(rank, world, local, dev, "nccl")
    torch.manual_seed(5)
    stepper = NativeStepper(ctx, cfg, OptHyper(args.optimizer, lr=0.1 if args.optimizer == "adadelta" else 0.01),
                            batch=batch, loss_mode="ref", spec=spec, graph=not args.no_graph,
                            overlap=not args.no_overlap, dtype=args.dtype, conv=args.conv, algo=args.algo,
                            wire=args.wire)
    K, W = args.steps, args.warmup

    # ---- device-resident input pool larger than L2, cycled (timing rule) ----
    img_bytes = spec.channels * spec.height * spec.width
    n_pool = L2_BYTES // (batch * img_bytes) + 8
    g = torch.Generator(device="cpu").manual_seed(1234 + rank)
    pool_x = torch.randint(0, 256, (n_pool, batch, spec.channels, spec.height, spec.width), dtype=torch.uint8,
                           generator=g)
    pool_y = torch.randint(0, spec.classes, (n_pool, batch), generator=g)
    # images and labels of a batch sit in ONE pool row, so a step's inputs arrive with a single device copy
    x_bytes = batch * img_bytes
    y_off = (x_bytes + 15) // 16 * 16
    pool = torch.zeros(n_pool, y_off + 8 * batch, dtype=torch.uint8)
    pool[:, :x_bytes] = pool_x.view(n_pool, -1)
    pool[:, y_off:] = pool_y.contiguous().view(torch.uint8).view(n_pool, -1)
    pool = pool.to(dev)
    slot = torch.empty(y_off + 8 * batch, dtype=torch.uint8, device=dev)
    sx = slot[:x_bytes].view(batch, spec.channels, spec.height, spec.width)
    sy = slot[y_off:].view(torch.int64)

    def dev_step(i):
        slot.copy_(pool[i % n_pool], non_blocking=True)
        return stepper.train_step(sx, sy)

    # clocks are sampled from before the warm-up until after the timed region (the timed region alone
    # can be shorter than one nvidia-smi sampling period)
    sampler = ClockSampler(local, period_ms=20).start() if rank == 0 else None
    for i in range(W):
        dev_step(i)
    _sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        dev_step(W + i)
    e1.record()
    _sync_all()
    ms = _max_ranks(e0.elapsed_time(e1), dev)
    clocks = sampler.stop() if sampler else None
    stepper.check_error()
    loss_dev = float(stepper.net.loss)

    # ---- end to end through the public API: pinned host -> device every step, loss read back ----
    # Trainer.train_epoch is the call a user makes (b200dp.engine.run makes it once per epoch); with the
    # native stepper it drives the C++ step pump: per step, gather into a pinned ring slot, one graph launch
    # (H2D copies of the slot + the whole step) and a 4-byte D2H copy of the loss into a pinned array.
    from b200dp.engine import Trainer
    from b200dp.utils.logging import RankLogger
    n_host = max(K, W) * batch                              # one "epoch" of the timed loader = exactly K steps
    hx = torch.randint(0, 256, (n_host, spec.channels, spec.height, spec.width), dtype=torch.uint8, generator=g)
    hy = torch.randint(0, spec.classes, (n_host,), generator=g)
    mk = lambda n: PinnedRingLoader(hx, hy, 0, n * batch, batch, spec, shuffle=True, fmt=args.input, device=dev,  # noqa: E731
                                    seed=rank, drop_last=True)
    warm_loader, loader = mk(W), mk(K)
    trainer = Trainer(ctx, stepper, RankLogger(False, None), log_interval=15, no_pump=args.no_pump)
    trainer.train_epoch(warm_loader, 0)
    trainer.train_epoch(loader, 0)                          # also warms the timed loader's own pump/graphs
    _sync_all()
    e0.record()
    t0 = time.perf_counter()
    trainer.train_epoch(loader, 1)                          # K steps; returns after the last step finished
    wall_ms = (time.perf_counter() - t0) * 1e3              # this rank's K steps, losses on the host; max over ranks below
    e1.record()
    _sync_all()
    e2e_ms = _max_ranks(max(e0.elapsed_time(e1), wall_ms), dev)
    stepper.check_error()
    pumped = trainer._pump_for(loader) is not None
    loss_e2e = float(trainer._loss_ring[K - 1]) if pumped else float(stepper.net.loss)
    div = stepper.replica_divergence()

    # ---- exposed communication per step: same device-timed loop with the cross-rank exchange elided ----
    exposed_us = None
    if world > 1:
        stepper.set_comm_enabled(False)
        for i in range(W):
            dev_step(i)
        _sync_all()
        e0.record()
        for i in range(K):
            dev_step(W + i)
        e1.record()
        _sync_all()
        ms_nocomm = _max_ranks(e0.elapsed_time(e1), dev)
        exposed_us = (ms - ms_nocomm) / K * 1e3

    if rank == 0:
        imgs = batch * world * K
        out = {
            "metric": "cnn.py images/sec (device-timed, max over ranks)",
            "value": imgs / (ms / 1e3), "unit": "images/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True,
            "scaling": "strong" if args.config == "strong4096" else "weak",
            "vs_baseline": None,
            "dtype": "bf16" if (stepper.net.conv_tc or stepper.dtype != "fp32") else "fp32",
            "dtype_detail": _dtype_label(stepper),
            "data": "synthetic", "impl": "b200dp",
            "config": {"model": "cnn.py LeNet" if args.config != "wide_cifar" else "cnn.py LeNet 4x wide, CIFAR-shaped",
                       "global_batch": batch * world, "per_gpu_batch": batch, "seq_len": None,
                       "input": f"{spec.channels}x{spec.height}x{spec.width}",
                       "parallelism": f"dp{world}", "optimizer": args.optimizer, "loss": "ref (nll on softmax probs)",
                       "l2": f"inputs larger than L2: they cycle through a {n_pool * batch * img_bytes >> 20} MiB device pool (> 126 MiB L2)"},
            # descriptive extras live outside `config` so that both arms print the same config keys
            "details": {"cuda_graph": not args.no_graph, "overlap": not args.no_overlap,
                        "allreduce": (("push protocol: P2P pushes + fused wait/sum/average/update kernels, "
                                       + stepper.wire + " wire") if stepper.fused_final else
                                      "fused per-bucket P2P reduce+average+update kernel") if world > 1 else "n/a (1 GPU)"},
            "clocks": clocks,
            "e2e": {"value": imgs / (e2e_ms / 1e3), "unit": "images/s", "ms_per_step": e2e_ms / K,
                    "h2d_bytes_per_step": loader.h2d_bytes_per_batch, "d2h_bytes_per_step": 4,
                    "input_format": args.input,
                    "api": "b200dp.engine.Trainer.train_epoch" + (" (native C++ step pump)" if pumped else " (Python loop)")},
            "gpu_launches": stepper.launches_per_step * K,
            "final_loss": {"device_path": loss_dev, "e2e_path": loss_e2e},
            "replica_divergence": div,
            "exposed_comm_us_per_step": exposed_us,
            "native_lib": os.path.relpath(native.lib_path(), ROOT),
        }
        _emit(out)
    stepper.close()
    if dist.is_initialized():
        dist.destroy_process_group()


class _RefLoader:
    """Feeds the reference's train() (train.py:42-57) W+K pinned fp32 batches of its dataset shape and
    brackets the timed region with CUDA events from inside the iteration."""

    def __init__(self, xs, ys, batch, warmup, steps):
        self.xs, self.ys, self.batch, self.W, self.K = xs, ys, batch, warmup, steps
        self.dataset = list(range((warmup + steps) * batch))
        self.e0, self.e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.t0 = 0.0

    def __len__(self):
        return self.W + self.K

    def __iter__(self):
        n = self.xs.shape[0]
        for i in range(self.W + self.K):
            if i == self.W:
                _sync_all()
                self.e0.record()
                self.t0 = time.perf_counter()
            j = i % n
            yield self.xs[j], self.ys[j]


def run_reference(args):
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isfile(os.path.join(ref_dir, "train.py")):
        _emit({"impl": "reference", "unavailable": "baseline/_ref not installed (see DESIGN.md)"})
        return
    if args.config != "lenet_b256":
        _emit({"impl": "reference", "unavailable": "the reference has a single fixed model/config"})
        return
    rank, world, local, dev = _init_dist(args.gpus)
    if not dist.is_initialized():   # the reference's average_gradients() needs a group even at size 1
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", str(29500 + os.getpid() % 2000))
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=dev)
    sys.path.insert(0, ref_dir)
    import train as ref_train          # unmodified reference modules
    from cnn import LeNet as RefLeNet
    from b200dp.utils.clocks import ClockSampler

    torch.manual_seed(5)
    torch.cuda.manual_seed(5)
    batch, K, W = PER_GPU_BATCH, args.steps, args.warmup
    n_pool = 64
    g = torch.Generator().manual_seed(1234 + rank)
    xs = torch.randn(n_pool, batch, 1, 28, 28, generator=g).pin_memory()      # what ToTensor+Normalize yields
    ys = torch.randint(0, 10, (n_pool, batch), generator=g).pin_memory()
    model = RefLeNet().cuda(local)                                             # train.py:98
    optimizer = torch.optim.Adadelta(model.parameters(), lr=0.1)               # train.py:99
    ref_args = types.SimpleNamespace(log_interval=15, dry_run=False)           # train.py defaults
    loader = _RefLoader(xs, ys, batch, W, K)
    sampler = ClockSampler(local, period_ms=20).start() if rank == 0 else None
    with contextlib.redirect_stdout(io.StringIO()):                            # keep its rank-0 log lines off stdout
        ref_train.train(ref_args, model, local, loader, optimizer, 1)          # the reference's own loop
    loader.e1.record()
    _sync_all()
    wall_ms = (time.perf_counter() - loader.t0) * 1e3
    ms = _max_ranks(loader.e0.elapsed_time(loader.e1), dev)
    clocks = sampler.stop() if sampler else None
    if rank == 0:
        imgs = batch * world * K
        h2d = batch * 784 * 4 + batch * 8
        _emit({
            "metric": "cnn.py images/sec (device-timed, max over ranks)",
            "value": imgs / (ms / 1e3), "unit": "images/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "fp32", "data": "synthetic", "impl": "reference",
            "config": {"model": "cnn.py LeNet", "global_batch": batch * world, "per_gpu_batch": batch,
                       "seq_len": None, "input": "1x28x28", "parallelism": f"dp{world}",
                       "optimizer": "adadelta", "loss": "ref (nll on softmax probs)",
                       "l2": "inputs larger than L2: they come from pinned host memory every step (never L2-resident)"},
            "details": {"path": "baseline/_ref train.train() + average_gradients(), NCCL, stock code path"},
            "clocks": clocks,
            "e2e": {"value": imgs / (max(ms, wall_ms) / 1e3), "unit": "images/s",
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4 / 15,
                    "note": "the reference loop is end to end by construction: synchronous H2D each step, "
                            "loss.item() every 15 steps on rank 0"},
            "gpu_launches": 0,
        })
    dist.destroy_process_group()


if __name__ == "__main__":
    a = _parse()
    if not torch.cuda.is_available():
        _emit({"impl": a.impl, "unavailable": "no CUDA device visible"})
        sys.exit(0)
    if a.impl == "reference":
        run_reference(a)
    else:
        run_own(a)
