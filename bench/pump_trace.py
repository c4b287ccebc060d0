"""Where a short epoch's fixed cost goes: host and device timeline of one pumped ``Trainer.train_epoch``.

Same set-up as the end-to-end part of ``bench.py`` (LeNet, batch 256, uint8 samples in pinned memory), with
``B200DP_PUMP_TRACE=1`` (csrc/pump.cu prints, per run, when each step was gathered / launched / finished on the device)
and wall-clock stamps around the Python-side calls of the epoch.  One GPU; run under ``gpurun``:

    B200DP_PUMP_TRACE=1 python bench/pump_trace.py --steps 20 --epochs 3
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--epochs", type=int, default=3)
    ap.add_argument("--log-interval", type=int, default=15)
    args = ap.parse_args()
    os.environ.setdefault("B200DP_PUMP_TRACE", "1")

    from b200dp.data.synthetic import MNIST_SPEC, PinnedRingLoader
    from b200dp.engine import Trainer
    from b200dp.models.fused import NativeStepper
    from b200dp.models.lenet import LeNetConfig
    from b200dp.optim.flat import OptHyper
    from b200dp.parallel.launch import RankContext
    from b200dp.utils.logging import RankLogger

    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    ctx = RankContext(0, 1, 0, dev, "nccl")
    spec, batch, K = MNIST_SPEC, 256, args.steps
    stepper = NativeStepper(ctx, LeNetConfig(), OptHyper("adadelta", lr=0.1), batch=batch, loss_mode="ref", spec=spec)
    g = torch.Generator().manual_seed(3)
    hx = torch.randint(0, 256, (K * batch, 1, 28, 28), dtype=torch.uint8, generator=g)
    hy = torch.randint(0, 10, (K * batch,), generator=g)
    loader = PinnedRingLoader(hx, hy, 0, K * batch, batch, spec, shuffle=True, fmt="u8", device=dev, seed=0,
                              drop_last=True)
    trainer = Trainer(ctx, stepper, RankLogger(False, None), log_interval=args.log_interval)
    trainer.train_epoch(loader, 0)
    trainer.train_epoch(loader, 0)
    pump = trainer._pump_for(loader)
    stamps = []

    def wrap(obj, name, label):
        fn = getattr(obj, name)

        def timed(*a, **k):
            t = time.perf_counter_ns()
            r = fn(*a, **k)
            stamps.append((label, t, time.perf_counter_ns()))
            return r
        setattr(obj, name, timed)

    wrap(pump, "run", "pump.run")
    wrap(pump, "wait_first", "pump.wait_first")
    wrap(pump, "drain", "pump.drain")
    wrap(loader, "_order", "loader._order")
    wrap(loader, "prepare_next", "loader.prepare_next")
    wrap(trainer, "check_peers", "check_peers")
    for ep in range(args.epochs):
        torch.cuda.synchronize()
        stamps.clear()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        t0 = time.perf_counter_ns()
        trainer.train_epoch(loader, ep + 1)
        t1 = time.perf_counter_ns()
        e1.record()
        torch.cuda.synchronize()
        print(f"[epoch {ep}] t0 {t0} ns; train_epoch returned +{(t1 - t0) / 1e3:.1f} us "
              f"({(t1 - t0) / 1e3 / K:.2f} us/step); events {e0.elapsed_time(e1) * 1e3:.1f} us", file=sys.stderr)
        for label, a, b in stamps:
            print(f"[epoch {ep}]   {label:20s} +{(a - t0) / 1e3:8.1f} .. +{(b - t0) / 1e3:8.1f} us", file=sys.stderr)
    stepper.close()


if __name__ == "__main__":
    main()
