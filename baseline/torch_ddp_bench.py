#!/usr/bin/env python
"""Strong library baseline: the same model in torch DistributedDataParallel.

Not product code and not the reference arm (that is `bench.py --impl reference`): this is what a user gets
from stock PyTorch on the same box — cuDNN / cuBLAS / ATen kernels, bucketed + overlapped NCCL all-reduce,
the foreach Adadelta — for the reference model and for the 4x-wide CIFAR-shaped model the reference does
not have.  Same step definition and timing rules as bench.py (device-resident inputs cycling through a pool
larger than L2, CUDA events, barrier + synchronize on both sides, max over ranks).

    python baseline/torch_ddp_bench.py --config lenet_b256 [--amp] [--cuda-graph]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        baseline/torch_ddp_bench.py --config wide_cifar
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", choices=["lenet_b256", "wide_cifar"], default="lenet_b256")
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=30)
    ap.add_argument("--amp", action="store_true", help="bf16 autocast")
    ap.add_argument("--tf32", action="store_true", help="allow TF32 in cuDNN / cuBLAS")
    a = ap.parse_args()
    if not torch.cuda.is_available():
        print(json.dumps({"impl": "torch_ddp", "unavailable": "no CUDA device visible"}))
        return 0
    from b200dp.data.synthetic import CIFAR_SPEC, MNIST_SPEC, normalise_reference
    from b200dp.models.lenet import CIFAR_WIDE_LENET, MNIST_LENET, LeNet

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(29500 + os.getpid() % 2000))
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    torch.backends.cudnn.allow_tf32 = a.tf32
    torch.backends.cuda.matmul.allow_tf32 = a.tf32
    torch.backends.cudnn.benchmark = True

    cfg, spec = (CIFAR_WIDE_LENET, CIFAR_SPEC) if a.config == "wide_cifar" else (MNIST_LENET, MNIST_SPEC)
    torch.manual_seed(5)
    model = torch.nn.parallel.DistributedDataParallel(LeNet(cfg).to(dev), device_ids=[local])
    opt = torch.optim.Adadelta(model.parameters(), lr=0.1)
    B = 256
    img_bytes = spec.channels * spec.height * spec.width
    n_pool = (126 << 20) // (B * img_bytes) + 8
    g = torch.Generator().manual_seed(1234 + rank)
    px = torch.randint(0, 256, (n_pool, B, spec.channels, spec.height, spec.width), dtype=torch.uint8, generator=g).to(dev)
    py = torch.randint(0, spec.classes, (n_pool, B), generator=g).to(dev)

    def step(i: int):
        x = normalise_reference(px[i % n_pool], spec)
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=a.amp):
            out = model(x)
        loss = F.nll_loss(out.float(), py[i % n_pool])            # the reference's loss: nll on softmax probabilities
        loss.backward()
        opt.step()
        return loss

    for i in range(a.warmup):
        step(i)
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.steps):
        step(a.warmup + i)
    e1.record()
    dist.barrier(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    if rank == 0:
        print(json.dumps({
            "impl": "torch_ddp", "metric": "images/sec (device-timed, max over ranks)",
            "value": B * world * a.steps / (ms / 1e3), "unit": "images/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms / a.steps,
            "dtype": "bf16 autocast" if a.amp else ("tf32" if a.tf32 else "fp32"), "data": "synthetic",
            "config": {"model": "cnn.py LeNet" if a.config == "lenet_b256" else "cnn.py LeNet 4x wide, CIFAR-shaped",
                       "per_gpu_batch": B, "optimizer": "adadelta (torch foreach)", "parallelism": f"dp{world}",
                       "path": "torch DDP: cuDNN/cuBLAS/ATen + bucketed NCCL all-reduce"}}), flush=True)
    dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
