"""BASELINE.json configs[0]: the full loop on synthetic MNIST-shaped data, gloo, world_size=2, CPU."""
import os

import pytest
import torch

from b200dp.cli import parse_args
from b200dp.parallel import launch

from _gloo_worker import engine_vs_reference, full_run


@pytest.mark.parametrize("overlap", [True, False])
def test_ws2_matches_reference_loop(tmp_path, ref_modules, overlap):
    cfg = {"seed": 5, "n": 256, "batch": 32, "out": str(tmp_path), "overlap": overlap}
    launch.spawn(engine_vs_reference, cfg, 2, backend="gloo", use_cuda=False, port=0)
    r0, r1 = (torch.load(os.path.join(tmp_path, f"rank{r}.pt")) for r in (0, 1))
    assert len(r0["ours_losses"]) == 4 == len(r0["ref_losses"])
    assert r0["div"] == 0.0                                       # replicas bit-identical
    for r in (r0, r1):
        assert torch.allclose(torch.tensor(r["ours_losses"]), torch.tensor(r["ref_losses"]), atol=1e-6)
        for k in r["ours"]:
            assert torch.allclose(r["ours"][k], r["ref"][k], atol=2e-6, rtol=1e-5), k
    for k in r0["ours"]:
        assert torch.equal(r0["ours"][k], r1["ours"][k])


def test_cli_full_run_ws2(tmp_path, capfd):
    ck = str(tmp_path / "mnist_cnn.pt")
    args = parse_args(["--cpu", "--epochs", "2", "--batch-size", "64", "--train-size", "512", "--test-size", "128",
                       "--log-interval", "2", "--save-model", "--checkpoint", ck,
                       "--save-full", str(tmp_path / "full.pt"), "--jsonl", str(tmp_path / "log.jsonl")])
    args.out_dir = str(tmp_path)
    launch.spawn(full_run, args, 2, backend="gloo", use_cuda=False, port=0)
    out = capfd.readouterr().out
    assert "Train Epoch: 1 [0/256 (0%)]\tLoss: -0." in out          # shard size 256 = 512/2 (train.py:54)
    assert "Test set on 0: Average loss: -0." in out
    res = torch.load(os.path.join(tmp_path, "res0.pt"))
    assert res["epoch"] == 2 and res["divergence"] == 0.0 and res["n"] == 64
    sd = torch.load(ck)
    assert list(sd.keys())[0] == "ConvNet.1.weight" and sd["FC.2.bias"].shape == (10,)
    assert os.path.getsize(tmp_path / "log.jsonl") > 0
    # resume from the full-state file for one more epoch
    args2 = parse_args(["--cpu", "--epochs", "3", "--batch-size", "64", "--train-size", "512", "--test-size", "128",
                        "--resume", str(tmp_path / "full.pt")])
    args2.out_dir = str(tmp_path)
    launch.spawn(full_run, args2, 2, backend="gloo", use_cuda=False, port=0)
    res = torch.load(os.path.join(tmp_path, "res0.pt"))
    assert res["epoch"] == 3


def test_learns_single_process(tmp_path):
    args = parse_args(["--cpu", "--epochs", "3", "--batch-size", "64", "--train-size", "1024", "--test-size", "256",
                       "--loss", "xent", "--optimizer", "adadelta", "--lr", "1.0", "--log-interval", "100"])
    args.out_dir = str(tmp_path)
    launch.spawn(full_run, args, 1, backend="gloo", use_cuda=False, port=0)
    res = torch.load(os.path.join(tmp_path, "res0.pt"))
    assert res["correct"] / res["n"] > 0.9
