#!/usr/bin/env python
"""Probe for csrc/head_fused.cu: CUDA-event time of the fused head at the reference geometry (batch 256), back to back
in a CUDA graph (warm) and as isolated launches; the target of `ncu -k regex:head_fused`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200dp.ops import native  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
native.load()
B, K3, C3, F1, NC = int(os.environ.get("B", 256)), 400, 120, 84, 10
torch.manual_seed(0)
f = dict(device=dev, dtype=torch.float32)
p2 = torch.randn(B, K3, **f)
w3, b3 = torch.randn(C3, K3, **f) / 20, torch.zeros(C3, **f)
w1, b1 = torch.randn(F1, C3, **f) / 11, torch.zeros(F1, **f)
w2, b2 = torch.randn(NC, F1, **f) / 9, torch.zeros(NC, **f)
t = torch.randint(0, NC, (B,), device=dev)
a3, h1, probs = torch.empty(B, C3, **f), torch.empty(B, F1, **f), torch.empty(B, NC, **f)
rows, loss = torch.empty(B, **f), torch.zeros(1, **f)
dl, dz1, dz3 = torch.empty(B, NC, **f), torch.empty(B, F1, **f), torch.empty(B, C3, **f)
ticket = torch.zeros(1, dtype=torch.int32, device=dev)


def run(train=True):
    native.head_fused(p2, 0, w3, b3, w1, b1, w2, b2, t, 0, a3, 0, h1, 0, probs, rows, loss, dl if train else None, 0,
                      dz1 if train else None, dz3 if train else None, None, ticket, B)


for _ in range(5):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for tma, train in ((1, True), (1, False), (0, True)):
    native.load().b200dp_head_fused_set_tma(tma)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(20):
            run(train)
    g.replay()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    print(f"head_fused tma={tma} train={train} B={B}: {e0.elapsed_time(e1) * 10:.2f} us per launch (graph, back to back)")
native.load().b200dp_head_fused_set_tma(1)

# phase stamps of CTA 0 (cycles since its start), one isolated launch after warm-up
dbg = torch.zeros(16, dtype=torch.int64, device=dev)
native.load().b200dp_head_fused_set_debug(dbg.data_ptr())
run()
torch.cuda.synchronize()
native.load().b200dp_head_fused_set_debug(None)
d = dbg.tolist()
names = ["start", "ring primed", "pdl wait", "inputs staged", "conv3 slabs", "conv3 epilogue", "FC.0 slabs", "FC.0 epilogue",
         "row tail", "dz3 slabs", "dz3 written"]
print("phase stamps (cycles from start):", ", ".join(f"{nm} {d[i] - d[0]}" for i, nm in enumerate(names)), f", end {d[15] - d[0]}")
