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
