"""Worker bodies for the multi-GPU tests (spawned, one process per GPU)."""
import os

import torch


def train_native(ctx, cfg):
    """N steps of the native stepper with the fused P2P reduce+update; rank r trains on slice r of
    every global batch.  Optionally injects rank-dependent device delays to shake out races."""
    import torch.distributed as dist
    from b200dp.models.fused import NativeStepper
    from b200dp.models.lenet import MNIST_LENET
    from b200dp.optim.flat import OptHyper

    torch.manual_seed(cfg["seed"])
    b = cfg["per_rank_batch"]
    st = NativeStepper(ctx, MNIST_LENET, OptHyper(cfg["opt"], lr=cfg["lr"]), batch=b, loss_mode="xent",
                       graph=cfg["graph"], overlap=cfg["overlap"], dtype=cfg.get("dtype", "fp32"),
                       algo=cfg.get("algo", "auto"))
    g = torch.Generator().manual_seed(77)
    n_batches = cfg["n_batches"]
    gx = torch.randint(0, 256, (n_batches, ctx.world_size * b, 1, 28, 28), dtype=torch.uint8, generator=g)
    gy = torch.randint(0, 10, (n_batches, ctx.world_size * b), generator=g)
    xs = [gx[i, ctx.rank * b:(ctx.rank + 1) * b].to(ctx.device).contiguous() for i in range(n_batches)]
    ys = [gy[i, ctx.rank * b:(ctx.rank + 1) * b].to(ctx.device).contiguous() for i in range(n_batches)]
    losses = []
    rng = torch.Generator().manual_seed(1000 + ctx.rank)
    for i in range(cfg["steps"]):
        if cfg.get("jitter"):
            cycles = int(torch.randint(0, 400_000, (1,), generator=rng))
            torch.cuda._sleep(cycles)          # rank-dependent skew of up to ~0.2 ms
        loss = st.train_step(xs[i % n_batches], ys[i % n_batches])
        if i < 8 or i == cfg["steps"] - 1:
            losses.append(float(loss))
    torch.cuda.synchronize()
    st.check_error()
    div = st.replica_divergence()
    lt = torch.tensor(losses, device=ctx.device)
    dist.all_reduce(lt)                        # mean over ranks of per-rank mean loss == global-batch mean
    out = {"div": div, "transport": st.comm.transport, "multicast": bool(st.comm.multicast_ptr), "losses": (lt / ctx.world_size).tolist(),
           "params": {k: v.detach().cpu().clone() for k, v in st.model.state_dict().items()},
           "gen": int(st.step_counter[0])}
    if ctx.rank == 0:
        torch.save(out, os.path.join(cfg["out"], f"native_{cfg['tag']}.pt"))
    dist.barrier()
    st.close()


def train_epochs(ctx, cfg):
    """Two epochs through Trainer.train_epoch on this rank's shard, with the native step pump or the Python loop;
    optionally the 4x-wide model on the tensor-core path."""
    import torch.distributed as dist
    from b200dp.data.shard import shard_for_rank
    from b200dp.data.synthetic import CIFAR_SPEC, MNIST_SPEC, PinnedRingLoader, SyntheticImages
    from b200dp.engine import Trainer
    from b200dp.models.fused import NativeStepper
    from b200dp.models.lenet import CIFAR_WIDE_LENET, MNIST_LENET
    from b200dp.optim.flat import OptHyper
    from b200dp.utils.logging import RankLogger

    wide = cfg.get("wide", False)
    mcfg, spec = (CIFAR_WIDE_LENET, CIFAR_SPEC) if wide else (MNIST_LENET, MNIST_SPEC)
    torch.manual_seed(cfg["seed"])
    b = cfg["per_rank_batch"]
    base = SyntheticImages(cfg["n"], spec, seed=3)
    sh = shard_for_rank(base, ctx.rank, ctx.world_size)
    loader = PinnedRingLoader(base.images, base.targets, *sh.bounds(), b, spec, shuffle=True, fmt="u8",
                              device=ctx.device, seed=17 * ctx.rank, drop_last=True)
    st = NativeStepper(ctx, mcfg, OptHyper("adadelta", lr=0.1), batch=b, loss_mode="ref", spec=spec, graph=True)
    tr = Trainer(ctx, st, RankLogger(False, None), log_interval=4, no_pump=cfg["no_pump"])
    for e in (1, 2):
        tr.train_epoch(loader, e)
    torch.cuda.synchronize()
    st.check_error()
    out = {"div": st.replica_divergence(), "pumped": tr._pump_for(loader) is not None, "gen": int(st.step_counter[0]),
           "param": st.arena.param.detach().cpu().clone(), "steps": tr.global_step}
    if ctx.rank == 0:
        torch.save(out, os.path.join(cfg["out"], f"epochs_{cfg['tag']}.pt"))
    dist.barrier()
    st.close()
