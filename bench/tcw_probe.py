#!/usr/bin/env python
"""Bring-up probe for csrc/tc_win.cu: relative error of every window-tile kernel against fp64 torch, plus CUDA-event timings at batch 256.  Prints; asserts nothing."""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200dp.models.lenet import CIFAR_WIDE_LENET, MNIST_LENET, LeNet  # noqa: E402
from b200dp.ops import native  # noqa: E402

MEAN, STD = 0.1307, 0.3081
native.load()
lib = native.load()
dev = "cuda"


def pool(z):
    a = F.leaky_relu(z, 0.2)
    p, flat = F.max_pool2d(a, 2, 2, return_indices=True)
    W = a.shape[-1]
    return p, (((flat // W) % 2) * 2 + (flat % W) % 2).to(torch.uint8)


def rel(got, want):
    return float((got.double() - want).abs().max() / want.abs().max().clamp_min(1e-12))


dbg = torch.zeros(16, dtype=torch.int64, device=dev)


def phases(fn):
    """clock64 stamps of CTA 0: cycles spent in set-up, staging, barrier, MMA issue, MMA drain, epilogue, tear-down."""
    lib.b200dp_tcw_set_debug(dbg.data_ptr())
    fn(); torch.cuda.synchronize()
    lib.b200dp_tcw_set_debug(None)
    t = dbg.tolist()
    return "[" + " ".join(str(t[i + 1] - t[i]) for i in range(7)) + "]"


def timed(fn, n=20):
    if os.environ.get("TCW_ONLY"):
        return 0.0
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


ONLY = os.environ.get("TCW_ONLY")            # "wide" / "lenet": one configuration, batch 256, no timing loops (for ncu)
for _once in (0,):
    for name, cfg in (("lenet", MNIST_LENET), ("wide", CIFAR_WIDE_LENET)):
        if ONLY and name != ONLY:
            continue
        for b in ((256,) if ONLY else (3, 256)):
            torch.manual_seed(0)
            md = LeNet(cfg).to(dev).double()
            d = native.Dims(b, cfg.in_ch, cfg.in_hw, cfg.c1, cfg.c2, cfg.c3, cfg.f1, cfg.num_classes, 0)
            wpack = torch.empty(native.tcw_pack_floats(d), device=dev)
            w1, w2 = md.ConvNet[1].weight.float().contiguous(), md.ConvNet[4].weight.float().contiguous()
            native.tcw_pack_weights(w1, w2, wpack, d)
            x = torch.randint(0, 256, (b, cfg.in_ch, cfg.in_hw, cfg.in_hw), dtype=torch.uint8, device=dev)
            xf = (x.double().div(255) - MEAN) / STD
            out = {}
            try:
                with torch.no_grad():
                    p_ref, i_ref = pool(md.ConvNet[1](md.ConvNet[0](xf)))
                p1 = torch.zeros(b, cfg.c1, 14, 14, device=dev)
                idx1 = torch.zeros(b, cfg.c1, 14, 14, dtype=torch.uint8, device=dev)
                f1 = lambda: native.tcw_conv1_fwd(x, wpack, md.ConvNet[1].bias.float(), p1, idx1, d, MEAN, STD)  # noqa: E731
                f1(); torch.cuda.synchronize()
                out["fwd1"] = (rel(p1, p_ref), float((idx1 != i_ref).float().mean()), timed(f1) if b == 256 else 0, phases(f1))
                p1r = torch.randn(b, cfg.c1, 14, 14, device=dev)
                with torch.no_grad():
                    p2_ref, i2_ref = pool(md.ConvNet[4](p1r.double()))
                p2 = torch.zeros(b, cfg.c2 * 25, device=dev)
                idx2 = torch.zeros(b, cfg.c2, 5, 5, dtype=torch.uint8, device=dev)
                f2 = lambda: native.tcw_conv2_fwd(p1r, wpack, md.ConvNet[4].bias.float(), p2, idx2, d)  # noqa: E731
                f2(); torch.cuda.synchronize()
                out["fwd2"] = (rel(p2.view(b, cfg.c2, 5, 5), p2_ref), float((idx2 != i2_ref).float().mean()),
                               timed(f2) if b == 256 else 0, phases(f2))
                z1 = torch.randn(b, cfg.c1, 14, 14, device=dev, dtype=torch.float64, requires_grad=True)
                P1 = F.leaky_relu(z1, 0.2)
                pp2, ii2 = pool(md.ConvNet[4](P1))
                dp2 = torch.randn_like(pp2)
                pp2.backward(dp2)
                g1 = torch.zeros(b, cfg.c1, 14, 14, device=dev)
                a_dp2, a_p2, a_P1 = dp2.float().contiguous(), pp2.detach().float().contiguous(), P1.detach().float().contiguous()
                f3 = lambda: native.tcw_conv2_dgrad(a_dp2, a_p2, ii2, wpack, a_P1, g1, d)  # noqa: E731
                f3(); torch.cuda.synchronize()
                out["dgrad"] = (rel(g1, z1.grad), 0.0, timed(f3) if b == 256 else 0, phases(f3))
                tp = timed(lambda: native.tcw_pack_weights(w1, w2, wpack, d)) if b == 256 else 0
                out["pack"] = (0.0, 0.0, tp, "")
            except Exception as e:  # noqa: BLE001
                out["error"] = repr(e)[:300]
            print(f"{name} b={b}: " + "  ".join(
                f"{k}: rel={v[0]:.2e} idxdiff={v[1]:.3f} us={v[2]:.1f} cyc={v[3]}" if isinstance(v, tuple) else f"{k}: {v}"
                for k, v in out.items()), flush=True)
            if "error" in out:
                sys.exit(1)
