"""Worker bodies for the gloo world_size=2 plumbing tests (run in spawned processes)."""
import os
import sys
import types

import torch


def _ref_dir():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for d in (os.path.join(root, "baseline", "_ref"), "/root/reference"):
        if os.path.isfile(os.path.join(d, "train.py")):
            return d
    return None


def engine_vs_reference(ctx, cfg):
    """Run N steps with our eager DataParallel and with the reference's own
    train()/average_gradients() (train.py:34-57) on identical batches; dump both."""
    import torch.distributed as dist
    from b200dp.data.shard import shard_for_rank
    from b200dp.data.synthetic import MNIST_SPEC, PinnedRingLoader, SyntheticImages
    from b200dp.engine import EagerStepper
    from b200dp.models.lenet import LeNetConfig
    from b200dp.optim.flat import OptHyper

    torch.manual_seed(cfg["seed"])
    base = SyntheticImages(cfg["n"], MNIST_SPEC, seed=11)
    sh = shard_for_rank(base, ctx.rank, ctx.world_size)
    mk = lambda: PinnedRingLoader(base.images, base.targets, *sh.bounds(), cfg["batch"], MNIST_SPEC,  # noqa: E731
                                  shuffle=False, fmt="f32", device=None)
    stepper = EagerStepper(ctx, LeNetConfig(), OptHyper("adadelta", lr=0.1), loss_mode="ref",
                           overlap=cfg.get("overlap", True))
    init = {k: v.clone() for k, v in stepper.model.state_dict().items()}
    ours_losses = [float(stepper.train_step(x, y)) for x, y in mk()]
    div = stepper.replica_divergence()

    # ---- the reference, unmodified, with .cuda() patched to identity ----
    d = _ref_dir()
    sys.path.insert(0, d)
    import train as ref_train            # noqa: E402  (imports torchvision at module scope)
    from cnn import LeNet as RefLeNet    # noqa: E402
    torch.Tensor.cuda = lambda self, *a, **k: self
    model = RefLeNet()
    model.load_state_dict(init)
    opt = torch.optim.Adadelta(model.parameters(), lr=0.1)
    ref_losses = []
    real_nll = torch.nn.functional.nll_loss

    def spy(*a, **k):
        out = real_nll(*a, **k)
        ref_losses.append(float(out))
        return out
    ref_train.F.nll_loss = spy
    args = types.SimpleNamespace(log_interval=10 ** 9, dry_run=False)
    ref_train.train(args, model, 1, mk(), opt, 1)   # device=1 -> never prints
    ref_train.F.nll_loss = real_nll

    out = {"ours_losses": ours_losses, "ref_losses": ref_losses, "div": div,
           "ours": {k: v.clone() for k, v in stepper.model.state_dict().items()},
           "ref": {k: v.detach().clone() for k, v in model.state_dict().items()}}
    torch.save(out, os.path.join(cfg["out"], f"rank{ctx.rank}.pt"))
    dist.barrier()


def full_run(ctx, args):
    from b200dp.engine import run
    res = run(ctx, args)
    torch.save(res, os.path.join(args.out_dir, f"res{ctx.rank}.pt"))
