"""Command line.

The nine reference flags keep their names, types and defaults (reference
train.py:113-133): ``--batch-size 1024`` (global; per-rank = ceil(/size),
train.py:82), ``--epochs 20``, ``--lr 0.1``, ``--gamma 0.9``, ``--no-cuda``,
``--dry-run``, ``--seed 5``, ``--log-interval 15``, ``--save-model``.
Everything the reference hard-codes becomes a flag with the
reference's value as default.

Run:  ``python -m b200dp.cli [flags]``   (spawns one process per GPU, like
``python train.py``), or under ``torchrun`` (rank layout from the env).
"""
from __future__ import annotations

import argparse
import sys
from typing import Optional, Sequence


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="b200dp: Blackwell-native data-parallel CNN training")
    # ---- reference flags, verbatim (train.py:114-131) ----
    p.add_argument("--batch-size", type=int, default=1024, metavar="N",
                   help="input batch size for training, summed over ranks (default: 1024)")
    p.add_argument("--epochs", type=int, default=20, metavar="N",
                   help="number of epochs to train (default: 20)")
    p.add_argument("--lr", type=float, default=1e-1, metavar="LR", help="learning rate (default: 0.1)")
    p.add_argument("--gamma", type=float, default=0.9, metavar="M",
                   help="Learning rate step gamma (default: 0.9)")
    p.add_argument("--no-cuda", action="store_true", default=False, help="disables CUDA training")
    p.add_argument("--dry-run", action="store_true", default=False, help="quickly check a single pass")
    p.add_argument("--seed", type=int, default=5, metavar="S", help="random seed (default: 5)")
    p.add_argument("--log-interval", type=int, default=15, metavar="N",
                   help="how many batches to wait before logging training status")
    p.add_argument("--save-model", action="store_true", default=False, help="For Saving the current Model")
    # ---- extensions ----
    p.add_argument("--backend", choices=["b200", "nccl", "gloo"], default=None,
                   help="gradient path: b200 = fused P2P kernels (default on GPU), nccl/gloo = torch.distributed")
    p.add_argument("--world-size", type=int, default=0, help="ranks to spawn (0 = one per visible GPU)")
    p.add_argument("--cpu", action="store_true", help="run on CPU with gloo instead of exiting when no GPU")
    p.add_argument("--dtype", choices=["auto", "fp32", "tf32", "bf16"], default="auto",
                   help="fp32 = exact CUDA-core kernels; tf32 / bf16 = the tcgen05 tensor-core path (tf32 dense GEMMs, "
                        "bf16 convolutions when the model is wide; fp32 storage, accumulation and master weights); "
                        "auto = whichever is faster for the model size")
    p.add_argument("--optimizer", choices=["adadelta", "sgd", "adam"], default="adadelta")
    p.add_argument("--momentum", type=float, default=0.9)
    p.add_argument("--weight-decay", type=float, default=0.0)
    p.add_argument("--loss", choices=["ref", "xent"], default="ref",
                   help="ref = nll_loss on softmax probabilities like the reference; xent = cross-entropy")
    p.add_argument("--dataset", choices=["mnist", "cifar"], default="mnist", help="synthetic data shape")
    p.add_argument("--data-root", default=None,
                   help="directory with the MNIST IDX files (torchvision layout <root>/MNIST/raw or flat; .gz accepted) "
                        "or train.npz/test.npz dumps; the reference reads ./data (train.py:88).  Absent: synthetic data")
    p.add_argument("--train-size", type=int, default=60000)
    p.add_argument("--test-size", type=int, default=10000)
    p.add_argument("--width-mult", type=int, default=1)
    p.add_argument("--bucket-bytes", type=int, default=-1, help="<=0: layer-aligned LeNet buckets")
    p.add_argument("--allreduce-algo", choices=["auto", "push", "oneshot", "twoshot", "nvls"], default="auto",
                   help="push = per-bucket P2P pushes + one wait/sum/update launch per step (whole-model gradients up to "
                        "1 MB); oneshot/twoshot/nvls = one handshake launch per bucket; auto = by size")
    p.add_argument("--wire", choices=["auto", "fp32", "bf16"], default="auto",
                   help="gradient wire format of the push protocol (auto: bf16 only with --dtype bf16)")
    p.add_argument("--conv", choices=["auto", "simt", "tc", "hybrid"], default="auto",
                   help="convolution kernels: CUDA cores, tcgen05 window-tile kernels, or hybrid (first layer on CUDA cores, "
                        "the rest on tcgen05); auto = measured-fastest for the model")
    p.add_argument("--peer-timeout-ms", type=int, default=0,
                   help="bound on every wait for a peer rank's gradients (0 = 20 s); a rank that does not deliver raises")
    p.add_argument("--pin-cores", action="store_true",
                   help="pin every rank's host threads (Python, step pump) to a disjoint slice of the node's CPU cores")
    p.add_argument("--no-overlap", action="store_true", help="reduce after backward, not during it")
    p.add_argument("--no-graph", action="store_true", help="do not capture the step in a CUDA graph")
    p.add_argument("--no-pump", action="store_true",
                   help="drive every step from Python instead of the native C++ step pump")
    p.add_argument("--eager", action="store_true", help="torch eager kernels instead of the native ones")
    p.add_argument("--global-eval", action="store_true",
                   help="sum eval metrics over ranks (the reference prints rank 0's shard only)")
    p.add_argument("--compat-shard", action="store_true", help="reproduce SplitDataset quirks exactly")
    p.add_argument("--checkpoint", default="mnist_cnn.pt", help="compat checkpoint path (train.py:108)")
    p.add_argument("--resume", default=None, help="full-state checkpoint to resume from")
    p.add_argument("--save-full", default=None, help="write a full-state checkpoint here every epoch")
    p.add_argument("--jsonl", default=None, help="machine-readable per-step log (rank 0)")
    p.add_argument("--master-port", type=int, default=29500)
    return p


def parse_args(argv: Optional[Sequence[str]] = None) -> argparse.Namespace:
    return build_parser().parse_args(argv)


def main(argv: Optional[Sequence[str]] = None) -> int:
    import torch
    from .engine import run
    from .parallel import launch

    args = parse_args(argv)
    use_cuda = not args.no_cuda and torch.cuda.is_available()
    if not use_cuda and not args.cpu:
        # reference train.py:139-142
        print("No available GPU instance!")
        return 0
    if launch.launched_by_torchrun():
        launch.init_from_env(run, args, backend=None if use_cuda else "gloo", use_cuda=use_cuda)
        return 0
    world = args.world_size or (launch.discover_world_size(args.no_cuda) if use_cuda else 2)
    launch.spawn(run, args, world, backend=None if use_cuda else "gloo", use_cuda=use_cuda,
                 port=args.master_port)
    return 0


if __name__ == "__main__":
    sys.exit(main())
