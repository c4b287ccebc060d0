#!/usr/bin/env python
"""Warm-cache per-kernel times of one training step: every launch of the step is captured 20x back to back in a
CUDA graph and replayed (ncu's per-kernel durations are cold-cache and serialised, which overstates the small ones).

    python bench/step_breakdown.py [--config wide_cifar|lenet_b256] [--batch 256] [--dtype auto] [--conv auto]
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200dp.data.synthetic import CIFAR_SPEC, MNIST_SPEC  # noqa: E402
from b200dp.models.fused import NativeStepper  # noqa: E402
from b200dp.models.lenet import CIFAR_WIDE_LENET, MNIST_LENET  # noqa: E402
from b200dp.ops import native  # noqa: E402
from b200dp.optim.flat import OptHyper  # noqa: E402
from b200dp.parallel.launch import RankContext  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="wide_cifar")
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--dtype", default="auto")
ap.add_argument("--conv", default="auto")
a = ap.parse_args()
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
cfg, spec = (CIFAR_WIDE_LENET, CIFAR_SPEC) if a.config == "wide_cifar" else (MNIST_LENET, MNIST_SPEC)
st = NativeStepper(RankContext(0, 1, 0, dev, "nccl"), cfg, OptHyper("adadelta", lr=0.1), batch=a.batch, loss_mode="ref",
                   spec=spec, graph=False, overlap=False, dtype=a.dtype, conv=a.conv)
x = torch.randint(0, 256, (a.batch, spec.channels, spec.height, spec.width), dtype=torch.uint8, device=dev)
y = torch.randint(0, 10, (a.batch,), device=dev)
for _ in range(3):
    st.train_step(x, y)
torch.cuda.synchronize()

# record the launch sequence of one step by wrapping the native entry points
calls = []
names = [n for n in dir(native) if callable(getattr(native, n)) and not n.startswith("_") and n not in
         ("load", "check", "lib_path", "is_loaded", "tc_problem", "bwd_problem", "nvtx_push", "nvtx_pop")
         and not n.endswith("num_partials") and not n.endswith("floats") and not n[0].isupper()
         and not n.startswith(("symm_", "pump_"))]
orig = {n: getattr(native, n) for n in names}
for n in names:
    def mk(n):
        def w(*args, **kw):
            calls.append((n, args, kw))
            return orig[n](*args, **kw)
        return w
    setattr(native, n, mk(n))
st.train_step(x, y)
torch.cuda.synchronize()
for n in names:
    setattr(native, n, orig[n])

REP = 20
total = 0.0
print(f"{a.config} batch {a.batch} dtype {st.dtype} conv_tc {st.net.conv_tc}: {len(calls)} launches")
for n, args, kw in calls:
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(REP):
            orig[n](*args, **kw)
    for _ in range(2):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (5 * REP)
    total += us
    extra = ""
    if n in ("tc_gemm_group", "linear_bwd_group"):
        extra = f" ({len(args[0])} problems)"
    if n == "reduce_update":
        extra = f" (count {args[6]})"
    print(f"  {n:22s}{extra:18s} {us:8.2f} us")
print(f"  sum {total:.1f} us")
