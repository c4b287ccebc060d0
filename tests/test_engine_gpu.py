"""End-to-end tier on one B200: the CLI / engine on the native path (kernels, CUDA graph, step pump) — learns the
synthetic task, prints the reference's log lines, writes a checkpoint the reference model loads, resumes."""
import os

import pytest
import torch

from b200dp.cli import parse_args
from b200dp.parallel import launch

from _gloo_worker import full_run

pytestmark = pytest.mark.gpu


def test_cli_native_run_learns_and_checkpoints(tmp_path, capfd, ref_modules):
    ck = str(tmp_path / "mnist_cnn.pt")
    args = parse_args(["--epochs", "3", "--batch-size", "64", "--train-size", "2048", "--test-size", "512", "--loss", "xent",
                       "--lr", "1.0", "--log-interval", "8", "--save-model", "--checkpoint", ck,
                       "--save-full", str(tmp_path / "full.pt"), "--jsonl", str(tmp_path / "log.jsonl")])
    args.out_dir = str(tmp_path)
    launch.spawn(full_run, args, 1, backend="nccl", use_cuda=True, port=0)
    out = capfd.readouterr().out
    assert "Train Epoch: 1 [0/2048 (0%)]\tLoss: " in out and "Train Epoch: 3 [" in out      # train.py:53-55 format
    assert "Test set on 0: Average loss: " in out                                           # train.py:73-76 format
    res = torch.load(os.path.join(tmp_path, "res0.pt"))
    assert res["epoch"] == 3 and res["n"] == 512
    assert res["correct"] / res["n"] > 0.9, res                                             # the native kernels learn the task
    # the checkpoint is the reference's format: its own LeNet loads it strictly
    sd = torch.load(ck)
    ref = ref_modules["cnn"].LeNet()
    ref.load_state_dict(sd, strict=True)
    assert all(v.dtype == torch.float32 and not v.is_cuda for v in sd.values())
    # resume one more epoch from the full-state file
    args2 = parse_args(["--epochs", "4", "--batch-size", "64", "--train-size", "2048", "--test-size", "512", "--loss", "xent",
                        "--lr", "1.0", "--resume", str(tmp_path / "full.pt")])
    args2.out_dir = str(tmp_path)
    launch.spawn(full_run, args2, 1, backend="nccl", use_cuda=True, port=0)
    res2 = torch.load(os.path.join(tmp_path, "res0.pt"))
    assert res2["epoch"] == 4 and res2["correct"] / res2["n"] > 0.9


def test_cli_native_wide_cifar_runs(tmp_path):
    """The 4x-wide CIFAR-shaped model through the CLI: tensor-core path end to end."""
    args = parse_args(["--epochs", "2", "--batch-size", "64", "--train-size", "1024", "--test-size", "256", "--loss", "xent",
                       "--lr", "1.0", "--dataset", "cifar", "--width-mult", "4", "--log-interval", "100"])
    args.out_dir = str(tmp_path)
    launch.spawn(full_run, args, 1, backend="nccl", use_cuda=True, port=0)
    res = torch.load(os.path.join(tmp_path, "res0.pt"))
    assert res["epoch"] == 2 and res["correct"] / res["n"] > 0.8, res
