#!/usr/bin/env python
"""Phase stamps (clock64 deltas, CTA 0) of the tcgen05 weight-gradient kernels, wide model, batch 256."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200dp.models.lenet import CIFAR_WIDE_LENET, MNIST_LENET  # noqa: E402
from b200dp.ops import native  # noqa: E402

lib = native.load()
lib.b200dp_tcw_wgrad_set_debug.argtypes = [C.c_void_p]
lib.b200dp_tcw_wgrad_set_debug.restype = None
dev = "cuda"
names = ["setup", "planes0", "dZ0", "chunk0", "chunk1", "all_staged", "acc_ready", "epi_done", "mma_first", "mma_all"]
for name, cfg in (("wide", CIFAR_WIDE_LENET), ("lenet", MNIST_LENET)):
    b = 256
    d = native.Dims(b, cfg.in_ch, cfg.in_hw, cfg.c1, cfg.c2, cfg.c3, cfg.f1, cfg.num_classes, 0)
    rows = native.tcw_conv_wgrad_num_partials(d)
    x = torch.randint(0, 256, (b, cfg.in_ch, cfg.in_hw, cfg.in_hw), dtype=torch.uint8, device=dev)
    p1 = torch.randn(b, cfg.c1, 14, 14, device=dev)
    g1 = torch.randn(b, cfg.c1, 14, 14, device=dev)
    idx1 = torch.randint(0, 4, (b, cfg.c1, 14, 14), dtype=torch.uint8, device=dev)
    dp2 = torch.randn(b, cfg.c2, 5, 5, device=dev)
    p2 = torch.randn(b, cfg.c2 * 25, device=dev)
    idx2 = torch.randint(0, 4, (b, cfg.c2, 5, 5), dtype=torch.uint8, device=dev)
    wn2, wn1 = cfg.c2 * cfg.c1 * 25, cfg.c1 * cfg.in_ch * 25
    part = torch.zeros(rows, wn2 + cfg.c2 + 64, device=dev)
    dbg = torch.zeros(16, dtype=torch.int64, device=dev)
    for layer in (2, 1):
        def f():
            if layer == 2:
                native.tcw_conv_wgrad(2, p1, 0.0, 1.0, dp2, p2, cfg.c2 * 25, idx2, part, part.shape[1], 0, wn2, d)
            else:
                native.tcw_conv_wgrad(1, x, 0.1307, 0.3081, g1, None, 0, idx1, part, part.shape[1], 0, wn1, d)
        for _ in range(3):
            f()
        torch.cuda.synchronize()
        lib.b200dp_tcw_wgrad_set_debug(dbg.data_ptr())
        f()
        torch.cuda.synchronize()
        lib.b200dp_tcw_wgrad_set_debug(None)
        t = dbg.tolist()
        print(f"{name} layer {layer}: " + " ".join(f"{n}={t[i + 1] - t[0]}" for i, n in enumerate(names)), flush=True)
