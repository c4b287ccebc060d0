#!/usr/bin/env python
"""Smallest program that launches every kernel of the library once or twice (for compute-sanitizer):
two LeNet steps on the CUDA-core path and two steps of the wide model on the tcgen05 path, batch 8, no graphs."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200dp.data.synthetic import CIFAR_SPEC, MNIST_SPEC  # noqa: E402
from b200dp.models.fused import NativeStepper  # noqa: E402
from b200dp.models.lenet import CIFAR_WIDE_LENET, MNIST_LENET  # noqa: E402
from b200dp.optim.flat import OptHyper  # noqa: E402
from b200dp.parallel.launch import RankContext  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
ctx = RankContext(0, 1, 0, dev, "nccl")
for cfg, spec, kw in ((MNIST_LENET, MNIST_SPEC, {}), (CIFAR_WIDE_LENET, CIFAR_SPEC, {}),
                      (MNIST_LENET, MNIST_SPEC, dict(dtype="tf32", conv="tc"))):
    torch.manual_seed(0)
    st = NativeStepper(ctx, cfg, OptHyper("adadelta", lr=0.1), batch=8, loss_mode="ref", spec=spec, graph=False, **kw)
    x = torch.randint(0, 256, (8, spec.channels, spec.height, spec.width), dtype=torch.uint8, device=dev)
    y = torch.randint(0, 10, (8,), device=dev)
    losses = [float(st.train_step(x, y)) for _ in range(2)]
    st.eval_batch(x, y)
    torch.cuda.synchronize()
    print(cfg, kw, "losses", losses, "eval", st.eval_result(), flush=True)
    st.close()
print("done")
