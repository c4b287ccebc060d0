"""Training driver: per-rank run, train / eval loops, logging, checkpointing.

Maps onto the reference as follows:

=====================  ====================================================
reference              here
=====================  ====================================================
``run``  train.py:79   :func:`run` — seeds, per-rank batch ``ceil(B/size)``,
                       shards, loaders, model, optimizer, StepLR, epoch loop,
                       final save
``train`` train.py:42  :meth:`Trainer.train_epoch` — same step order; the only
                       host sync is ``loss.item()`` on rank 0 every
                       ``log_interval`` batches
``test``  train.py:60  :meth:`Trainer.test_epoch` — loss-sum / correct-count
                       accumulate on the device; one D2H per epoch instead of
                       two per batch; normalised by the local shard length
=====================  ====================================================

Two steppers implement one interface (``train_step``, ``eval_batch``,
``eval_result``): :class:`EagerStepper` (torch ops + autograd + bucketed
``torch.distributed`` reduction — CPU/gloo plumbing path and NCCL baseline) and
:class:`b200dp.models.fused.NativeStepper` (hand-written sm_100a kernels, fused
P2P reduce+update, CUDA-graph step).
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
import torch.distributed as dist

from . import ckpt
from .data.shard import common_batches, shard_for_rank
from .data.synthetic import (CIFAR_SPEC, MNIST_SPEC, PinnedRingLoader, SyntheticImages,
                             normalise_reference)
from .models.lenet import LeNet, LeNetConfig, loss_fn
from .optim.flat import OptHyper, StepLR
from .parallel.ddp import DataParallel
from .parallel.launch import RankContext
from .utils.logging import RankLogger, test_line, train_line


def _nvtx(name: str, push: bool) -> None:
    """NVTX range around a host-side phase (only once the native library is loaded: the CPU path has no use for it)."""
    from .ops import native
    if native.is_loaded():
        native.nvtx_push(name) if push else native.nvtx_pop()


class EagerStepper:
    """torch-eager forward/backward under :class:`DataParallel`."""

    name = "eager"

    def __init__(self, ctx: RankContext, cfg: LeNetConfig, hyper: OptHyper, *, loss_mode: str = "ref",
                 spec=MNIST_SPEC, bucket_bytes: int = -1, overlap: bool = True, comm: str = "torch",
                 algo: str = "auto"):
        """``comm="torch"``: gradients through ``torch.distributed`` (gloo / NCCL) + the flat torch optimizer;
        ``comm="b200"``: torch-eager forward/backward, gradients and update through the fused sm_100a P2P kernels."""
        self.ctx, self.cfg, self.loss_mode, self.spec = ctx, cfg, loss_mode, spec
        model = LeNet(cfg).to(ctx.device)
        self.ddp = DataParallel(model, hyper, bucket_bytes=bucket_bytes, overlap=overlap, device=ctx.device,
                                comm=comm, algo=algo)
        self.model = self.ddp.module
        self.optimizer = self.ddp.optimizer
        self._eval = torch.zeros(2, dtype=torch.float64, device=ctx.device)

    def _prep(self, x: torch.Tensor) -> torch.Tensor:
        return normalise_reference(x, self.spec) if x.dtype == torch.uint8 else x

    def _out(self, x):
        return self.model(x) if self.loss_mode == "ref" else self.model.logits(x)

    def train_step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        self.model.train()
        self.ddp.zero_grad()
        loss = loss_fn(self._out(self._prep(x)), y, self.loss_mode)
        loss.backward()
        self.ddp.step()
        return loss.detach()

    @torch.no_grad()
    def eval_batch(self, x: torch.Tensor, y: torch.Tensor) -> None:
        self.model.eval()
        out = self._out(self._prep(x))
        self._eval[0] += loss_fn(out, y, self.loss_mode, reduction="sum").double()
        self._eval[1] += out.argmax(dim=1).eq(y).sum().double()

    def eval_result(self, reset: bool = True) -> Tuple[float, int]:
        vals = self._eval.tolist()  # the one D2H of the epoch
        if reset:
            self._eval.zero_()
        return float(vals[0]), int(vals[1])

    def replica_divergence(self) -> float:
        return self.ddp.replica_divergence()

    def error_flag(self) -> int:
        return int(self.ddp.step_counter[2].item()) if self.ddp.native else 0

    def close(self) -> None:
        self.ddp.close()


class Trainer:
    """Epoch loops around a stepper (reference ``train`` / ``test``)."""

    def __init__(self, ctx: RankContext, stepper, logger: RankLogger, *, log_interval: int = 15,
                 dry_run: bool = False, global_eval: bool = False, no_pump: bool = False):
        self.ctx, self.stepper, self.log = ctx, stepper, logger
        self.log_interval, self.dry_run, self.global_eval = log_interval, dry_run, global_eval
        self.global_step = 0
        self.no_pump = no_pump
        self._t_log = None
        self._steps_at_log = 0
        self._pumps = {}
        self._loss_ring = None
        self._ev_log = None

    def _device_mark(self):
        """A CUDA event on the current stream (device time per step for the JSONL)."""
        if self.ctx.device.type != "cuda":
            return None
        ev = torch.cuda.Event(enable_timing=True)
        ev.record(torch.cuda.current_stream(self.ctx.device))
        return ev

    @staticmethod
    def _step_us(ev_a, ev_b, n_steps: int):
        if ev_a is None or ev_b is None or n_steps <= 0:
            return None
        ev_b.synchronize()
        return ev_a.elapsed_time(ev_b) * 1e3 / n_steps

    def _peer_wait_us(self):
        """Exposed part of the gradient exchange in the most recent step: how long the last bucket kept the final
        kernel's first CTA waiting for the peers (device-measured, csrc/ll_reduce.cu; None off the push protocol).
        Reads a device word, i.e. waits for the queued steps: only called when a JSONL log asked for it."""
        st = self.stepper
        if not (getattr(st, "fused_final", False) and getattr(st, "world", 1) > 1):
            return None
        return float(int(st.step_counter[3].item()) & 0xffffffff) / 1e3

    def _rate(self, n_steps: int, batch: int):
        """images/s of this rank since the previous log point (host wall clock; the log points are the only host
        syncs of the loop, so the interval covers exactly the steps between them)."""
        import time
        now = time.perf_counter()
        prev, self._t_log = self._t_log, now
        if prev is None or n_steps <= 0 or now <= prev:
            return None
        return n_steps * batch * self.ctx.world_size / (now - prev)

    def check_peers(self) -> None:
        """Raise if the device's sticky error word is set: a peer did not deliver its gradients within the time bound
        (csrc/ll_reduce.cu, csrc/reduce_update.cu) and this rank stopped updating its parameters.  Called where the
        loop synchronises anyway — rank 0's log points and every rank's epoch end — so it costs one 4-byte read.  No
        collective is needed for every rank to stop: the ranks that are still alive time out on the NEXT step's
        gradients of whichever rank raised first.  The reference hangs in ``all_reduce`` here until NCCL's watchdog
        fires (train.py:38)."""
        st = self.stepper
        if hasattr(st, "error_flag") and st.error_flag():
            raise RuntimeError(
                f"rank {self.ctx.rank}: a peer rank did not deliver its gradients within the time bound (dead or "
                f"stalled rank?); training stopped at global step <= {self.global_step}")

    def _pump_for(self, loader):
        """The native step pump for (stepper, loader), when the combination supports it: native stepper with
        CUDA graphs, raw uint8 samples streamed to a CUDA device."""
        st = self.stepper
        if not (hasattr(st, "make_pump") and getattr(st, "use_graph", False) and not self.no_pump):
            return None
        if not (isinstance(loader, PinnedRingLoader) and loader.fmt in ("u8", "pre") and loader.device is not None
                and loader.device.type == "cuda" and loader.batch_size == st.net.B):
            return None
        key = id(loader)
        if key not in self._pumps:
            self._pumps[key] = st.make_pump(loader.images, loader.targets, loader.batch_size)
        return self._pumps[key]

    def _train_epoch_pumped(self, pump, loader, epoch: int) -> Optional[float]:
        """Same epoch as the Python loop below — same sample order, same log lines — but the per-step host
        work runs in the C++ pump, one ctypes call per ``log_interval`` steps.  The log line of a chunk is
        printed once its first step has finished, while the rest of the chunk is still queued."""
        order = loader._order().contiguous()
        loader.epoch += 1
        B, n_batches = loader.batch_size, len(loader)
        n_full = min((loader.stop - loader.start) // B, n_batches)
        run_ragged = n_full < n_batches
        if self.dry_run and n_full >= 1:                      # train.py:56-57: one batch only
            n_full, run_ragged = 1, False
        if self._loss_ring is None or self._loss_ring.numel() < n_batches:
            self._loss_ring = torch.zeros(max(n_batches, 64), dtype=torch.float32).pin_memory()
        last = None
        timed = self.ctx.is_main and self.log._jsonl is not None      # device step times only when somebody records them
        marks = [self._device_mark()] if timed else []                 # marks[i] = end of chunk i-1 on the stream
        chunk_n = []
        for c0 in range(0, n_full, self.log_interval):
            n = min(self.log_interval, n_full - c0)
            pump.run(order[c0 * B:], n, self._loss_ring[c0:])
            self.global_step += n
            if timed:
                marks.append(self._device_mark())
                chunk_n.append(n)
            if self.ctx.is_main:
                pump.wait_first()
                self.check_peers()
                last = float(self._loss_ring[c0])
                self.log.print(train_line(epoch, c0, B, len(loader.dataset), n_batches, last))
                first = self.global_step - n + 1
                # device time per step of the PREVIOUS chunk (finished: this chunk's first step has); the current chunk is
                # still in flight
                ci = len(chunk_n) - 1
                step_us = self._step_us(marks[ci - 1], marks[ci], chunk_n[ci - 1]) if (timed and ci >= 1) else None
                self.log.event(kind="train", epoch=epoch, step=first, batch_idx=c0, loss=last,
                               images_per_s=self._rate(first - self._steps_at_log, B), step_us=step_us,
                               peer_wait_us=self._peer_wait_us() if timed else None)
                self._steps_at_log = first
        if run_ragged:                                        # ragged last batch: Python path
            idx = order[n_full * B:]
            x = loader.images.index_select(0, idx).to(self.ctx.device)
            y = loader.targets.index_select(0, idx).to(self.ctx.device)
            loss = self.stepper.train_step(x, y)
            self.global_step += 1
            if n_full % self.log_interval == 0 and self.ctx.is_main:
                last = float(loss.item())
                self.log.print(train_line(epoch, n_full, len(x), len(loader.dataset), n_batches, last))
        # next epoch's sample order, and its first batches into their pinned ring slots (host-side prefetch, the uploads
        # stay with their steps), while the GPU works through the queued steps
        loader.prepare_next()
        nxt = loader.peek_next()
        if nxt is not None and not self.dry_run:
            pump.pregather(nxt, n_full)
        pump.drain()
        self.check_peers()
        return last

    def train_epoch(self, loader, epoch: int) -> Optional[float]:
        _nvtx("b200dp.train_epoch", True)
        try:
            return self._train_epoch(loader, epoch)
        finally:
            _nvtx("b200dp.train_epoch", False)

    def _train_epoch(self, loader, epoch: int) -> Optional[float]:
        pump = self._pump_for(loader)
        if pump is not None:
            return self._train_epoch_pumped(pump, loader, epoch)
        last = None
        n_batches = len(loader)
        for batch_idx, (data, target) in enumerate(loader):
            loss = self.stepper.train_step(data, target)
            self.global_step += 1
            if batch_idx % self.log_interval == 0 and self.ctx.is_main:
                last = float(loss.item())  # the only host sync on the step path (train.py:55)
                self.check_peers()
                self.log.print(train_line(epoch, batch_idx, len(data), len(loader.dataset), n_batches, last))
                ev = self._device_mark() if self.log._jsonl is not None else None
                step_us = self._step_us(self._ev_log, ev, self.global_step - self._steps_at_log)
                self._ev_log = ev
                self.log.event(kind="train", epoch=epoch, step=self.global_step, batch_idx=batch_idx, loss=last,
                               images_per_s=self._rate(self.global_step - self._steps_at_log, len(data)),
                               step_us=step_us, peer_wait_us=self._peer_wait_us() if ev is not None else None)
                self._steps_at_log = self.global_step
            if self.dry_run:
                break
        self.check_peers()
        return last

    def test_epoch(self, loader) -> Tuple[float, int, int]:
        _nvtx("b200dp.test_epoch", True)
        for data, target in loader:
            self.stepper.eval_batch(data, target)
        loss_sum, correct = self.stepper.eval_result()
        _nvtx("b200dp.test_epoch", False)
        n = len(loader.dataset)
        if self.global_eval and self.ctx.world_size > 1:
            t = torch.tensor([loss_sum, correct, n], dtype=torch.float64, device=self.ctx.device)
            dist.all_reduce(t)
            loss_sum, correct, n = float(t[0]), int(t[1]), int(t[2])
        test_loss = loss_sum / n
        if self.ctx.is_main:
            self.log.print(test_line(self.ctx.local_rank if self.ctx.device.type == "cuda" else 0,
                                     test_loss, correct, n))
            self.log.event(kind="test", loss=test_loss, correct=correct, n=n)
        return test_loss, correct, n


def build_stepper(ctx: RankContext, args, cfg: LeNetConfig, hyper: OptHyper, spec, per_rank_batch: int):
    backend = getattr(args, "backend", None)
    use_native = ctx.device.type == "cuda" and not getattr(args, "eager", False) and backend in (None, "b200")
    if use_native:
        from .models.fused import NativeStepper
        return NativeStepper(ctx, cfg, hyper, batch=per_rank_batch, loss_mode=args.loss, spec=spec,
                             algo=getattr(args, "allreduce_algo", "auto"),
                             graph=not getattr(args, "no_graph", False),
                             overlap=not getattr(args, "no_overlap", False),
                             dtype=getattr(args, "dtype", "auto"), conv=getattr(args, "conv", "auto"),
                             wire=getattr(args, "wire", "auto"), timeout_ms=getattr(args, "peer_timeout_ms", 0))
    # --eager: torch kernels for the model; the gradient path is still ours (--backend b200, default on GPU) unless
    # --backend nccl/gloo asks for torch.distributed
    comm = "b200" if (ctx.device.type == "cuda" and backend in (None, "b200")) else "torch"
    return EagerStepper(ctx, cfg, hyper, loss_mode=args.loss, spec=spec,
                        bucket_bytes=getattr(args, "bucket_bytes", -1),
                        overlap=not getattr(args, "no_overlap", False), comm=comm,
                        algo=getattr(args, "allreduce_algo", "auto"))


def make_hyper(args) -> OptHyper:
    return OptHyper(kind=args.optimizer, lr=args.lr, momentum=getattr(args, "momentum", 0.9),
                    weight_decay=getattr(args, "weight_decay", 0.0))


def load_datasets(args, spec):
    """The train / test base datasets: real files under ``--data-root`` when they are there (the reference reads
    torchvision MNIST from ``./data``, train.py:88-93; no download here), else synthetic data of the same shape."""
    import os
    from .data.files import MNISTFiles, find_mnist, load_arrays
    root = getattr(args, "data_root", None)
    if root:
        npz = [os.path.join(root, f"{s}.npz") for s in ("train", "test")]
        if all(os.path.isfile(f) for f in npz):
            return load_arrays(npz[0], spec), load_arrays(npz[1], spec), f"{npz[0]}, {npz[1]}"
        if args.dataset == "mnist" and find_mnist(root, True) and find_mnist(root, False):
            tr, te = MNISTFiles(root, True, spec), MNISTFiles(root, False, spec)
            return tr, te, f"MNIST IDX files {tr.paths[0]} ({len(tr)} train / {len(te)} test samples)"
    note = f"synthetic {args.dataset}-shaped data" + (f" (no dataset files under {root!r})" if root else "")
    return (SyntheticImages(args.train_size, spec, seed=args.seed * 7919 + 1),
            SyntheticImages(args.test_size, spec, seed=args.seed * 7919 + 2), note)


def run(ctx: RankContext, args) -> dict:
    """Per-rank entry (reference ``run(rank, size, args)``, train.py:79-108)."""
    if getattr(args, "pin_cores", False) and ctx.world_size > 1:
        from .parallel.launch import pin_to_core_share
        n_local = int(__import__("os").environ.get("LOCAL_WORLD_SIZE", ctx.world_size))
        share = pin_to_core_share(ctx.local_rank % max(n_local, 1), n_local)
        if ctx.is_main:
            print(f"host threads pinned: rank 0 runs on cores {sorted(share)}" if share else
                  "host threads not pinned (fewer than two cores per rank)")
    torch.manual_seed(args.seed)                      # train.py:80 — same seed on every rank
    if ctx.device.type == "cuda":
        torch.cuda.manual_seed(args.seed)             # train.py:81
    per_rank_batch = math.ceil(args.batch_size / ctx.world_size)   # train.py:82

    spec = MNIST_SPEC if args.dataset == "mnist" else CIFAR_SPEC
    cfg = LeNetConfig(in_ch=spec.channels, width_mult=args.width_mult, in_hw=spec.height)
    base_train, base_test, source = load_datasets(args, spec)
    if ctx.is_main and getattr(args, "data_root", None):
        print(f"data: {source}")
    compat = getattr(args, "compat_shard", False)
    tr = shard_for_rank(base_train, ctx.rank, ctx.world_size, compat=compat and ctx.world_size > 1)
    te = shard_for_rank(base_test, ctx.rank, ctx.world_size, compat=compat and ctx.world_size > 1)

    on_gpu = ctx.device.type == "cuda"
    fmt = "u8" if on_gpu else "f32"
    mk = lambda base, sh, cap: PinnedRingLoader(  # noqa: E731  (train.py:95-96: both loaders shuffle)
        base.images, base.targets, *sh.bounds(), per_rank_batch, spec, shuffle=True, fmt=fmt,
        device=ctx.device if on_gpu else None, seed=args.seed + 17 * ctx.rank, max_batches=cap)
    # every training step is a collective: all ranks run the batch count the smallest shard allows
    train_cap = common_batches(base_train, ctx.world_size, per_rank_batch, compat=compat)
    train_loader, test_loader = mk(base_train, tr, train_cap), mk(base_test, te, None)

    stepper = build_stepper(ctx, args, cfg, make_hyper(args), spec, per_rank_batch)
    scheduler = StepLR(stepper.optimizer, step_size=1, gamma=args.gamma)     # train.py:101
    logger = RankLogger(ctx.is_main, getattr(args, "jsonl", None))
    trainer = Trainer(ctx, stepper, logger, log_interval=args.log_interval, dry_run=args.dry_run,
                      global_eval=getattr(args, "global_eval", False), no_pump=getattr(args, "no_pump", False))

    start_epoch = 1
    if getattr(args, "resume", None):
        meta = ckpt.load_full(args.resume, model=stepper.model, optimizer=stepper.optimizer,
                              scheduler=scheduler)
        if hasattr(stepper, "params_changed"):
            stepper.params_changed()
        start_epoch = meta["epoch"] + 1
        trainer.global_step = meta["step"]

    result = {}
    for epoch in range(start_epoch, args.epochs + 1):                         # train.py:102-105
        last = trainer.train_epoch(train_loader, epoch)
        test_loss, correct, n = trainer.test_epoch(test_loader)
        scheduler.step()
        result = {"epoch": epoch, "train_loss": last, "test_loss": test_loss, "correct": correct, "n": n}
        if getattr(args, "save_full", None):
            ckpt.save_full(args.save_full, model=stepper.model, optimizer=stepper.optimizer,
                           scheduler=scheduler, epoch=epoch, step=trainer.global_step,
                           is_main=ctx.is_main)

    if args.save_model:                                                       # train.py:107-108
        ckpt.save_model(stepper.model, getattr(args, "checkpoint", ckpt.DEFAULT_NAME), is_main=ctx.is_main)
    result["divergence"] = stepper.replica_divergence() if ctx.world_size > 1 else 0.0
    logger.close()
    return result
