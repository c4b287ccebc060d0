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
                       algo=cfg.get("algo", "auto"), wire=cfg.get("wire", "auto"), conv=cfg.get("conv", "auto"))
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
        torch.save(out, os.path.join(cfg["out"], f"{cfg['tag']}.pt"))
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
        torch.save(out, os.path.join(cfg["out"], f"{cfg['tag']}.pt"))
    dist.barrier()
    st.close()


def _make_generic(kind):
    """Two models that are not LeNet: a 3-layer MLP and a small residual conv block with a linear head."""
    from torch import nn
    if kind == "mlp":
        return nn.Sequential(nn.Flatten(), nn.Linear(64, 96), nn.Tanh(), nn.Linear(96, 48), nn.Tanh(), nn.Linear(48, 10))

    class Block(nn.Module):
        def __init__(self):
            super().__init__()
            self.c1 = nn.Conv2d(4, 4, 3, padding=1)
            self.c2 = nn.Conv2d(4, 4, 3, padding=1)
            self.head = nn.Linear(4 * 4 * 4, 10)
            self.unused = nn.Parameter(torch.zeros(7))      # never touched by forward: its hook never fires

        def forward(self, x):
            x = x.view(-1, 4, 4, 4)
            h = torch.relu(self.c1(x))
            h = x + self.c2(h)
            return self.head(torch.relu(h).flatten(1))
    return Block()


def train_generic(ctx, cfg):
    """A non-LeNet module under DataParallel(comm="b200") against the same module trained with
    dist.all_reduce (NCCL) + torch.optim on the same data: parameters after N steps, replica divergence,
    native launches per step."""
    import copy
    import torch.distributed as dist
    from b200dp.optim.flat import OptHyper
    from b200dp.parallel.ddp import DataParallel

    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(cfg["seed"])
    model = _make_generic(cfg["model"]).to(ctx.device)
    ref = copy.deepcopy(model)
    opt_name, lr = cfg["opt"], cfg["lr"]
    ddp = DataParallel(model, OptHyper(opt_name, lr=lr), bucket_bytes=cfg["bucket_bytes"], comm="b200",
                       algo=cfg.get("algo", "auto"), overlap=cfg.get("overlap", True))
    ropt = {"adadelta": lambda p: torch.optim.Adadelta(p, lr=lr), "sgd": lambda p: torch.optim.SGD(p, lr=lr, momentum=0.9),
            "adam": lambda p: torch.optim.Adam(p, lr=lr)}[opt_name](ref.parameters())
    g = torch.Generator().manual_seed(99)
    b, W = cfg["per_rank_batch"], ctx.world_size
    gx = torch.randn(cfg["steps"], W * b, 64, generator=g)
    gy = torch.randint(0, 10, (cfg["steps"], W * b), generator=g)
    lossf = torch.nn.CrossEntropyLoss()
    for i in range(cfg["steps"]):
        x = gx[i, ctx.rank * b:(ctx.rank + 1) * b].to(ctx.device)
        y = gy[i, ctx.rank * b:(ctx.rank + 1) * b].to(ctx.device)
        ddp.zero_grad()
        lossf(ddp(x), y).backward()
        ddp.step()
        ropt.zero_grad()
        lossf(ref(x), y).backward()
        for p in ref.parameters():                       # the reference's average_gradients (train.py:34-39)
            if p.grad is None:
                p.grad = torch.zeros_like(p)
            if W > 1:
                dist.all_reduce(p.grad, op=dist.ReduceOp.SUM)
            p.grad /= W
        ropt.step()
    torch.cuda.synchronize()
    ddp.check_error()
    err = max(float((p.detach() - q.detach()).abs().max()) for p, q in zip(model.parameters(), ref.parameters()))
    out = {"err": err, "div": ddp.replica_divergence(), "launches": ddp._launches, "push": ddp.push,
           "buckets": len(ddp.arena.layout.buckets), "gen": int(ddp.step_counter[0])}
    if ctx.rank == 0:
        torch.save(out, os.path.join(cfg["out"], f"{cfg['tag']}.pt"))
    if W > 1:
        dist.barrier()
    ddp.close()


def train_with_dead_rank(ctx, cfg):
    """Rank 1 stops stepping after ``alive`` steps; the other ranks must notice within the time bound, stop
    updating, and raise from Trainer.check_peers instead of training on (or hanging)."""
    import time
    import torch.distributed as dist
    from b200dp.engine import Trainer
    from b200dp.models.fused import NativeStepper
    from b200dp.models.lenet import MNIST_LENET
    from b200dp.optim.flat import OptHyper
    from b200dp.utils.logging import RankLogger

    torch.manual_seed(cfg["seed"])
    b = cfg["per_rank_batch"]
    st = NativeStepper(ctx, MNIST_LENET, OptHyper("adadelta", lr=0.1), batch=b, loss_mode="ref", graph=cfg["graph"],
                       algo=cfg["algo"], timeout_ms=cfg["timeout_ms"], dtype="fp32")
    tr = Trainer(ctx, st, RankLogger(False, None))
    x = torch.randint(0, 256, (b, 1, 28, 28), dtype=torch.uint8, device=ctx.device)
    y = torch.randint(0, 10, (b,), device=ctx.device)
    raised, t_detect, steps_done = False, None, 0
    t0 = time.perf_counter()
    for i in range(cfg["steps"]):
        if ctx.rank == 1 and i >= cfg["alive"]:
            break                                        # the "dead" rank: no more steps, no more pushes
        st.train_step(x, y)
        steps_done += 1
        if i >= cfg["alive"] - 1:
            try:
                torch.cuda.synchronize()
                tr.check_peers()
            except RuntimeError:
                raised, t_detect = True, time.perf_counter() - t0
                break
    torch.cuda.synchronize()
    out = {"raised": raised, "t_detect": t_detect, "steps_done": steps_done, "flag": st.error_flag(),
           "gen": int(st.step_counter[0])}
    torch.save(out, os.path.join(cfg["out"], f"{cfg['tag']}_r{ctx.rank}.pt"))
    dist.barrier()
    st.close()


def run_batch(ctx, spec):
    """Run a list of (worker name, cfg) jobs in this one process group; a job that raises leaves ``<tag>.err``."""
    import traceback
    import torch.distributed as dist
    fns = {"train_native": train_native, "train_epochs": train_epochs, "train_generic": train_generic,
           "train_with_dead_rank": train_with_dead_rank}
    for name, cfg in spec["jobs"]:
        cfg = dict(cfg, out=spec["out"])
        try:
            fns[name](ctx, cfg)
        except Exception:
            with open(os.path.join(spec["out"], f"{cfg['tag']}.err"), "a") as f:
                f.write(f"rank {ctx.rank}:\n{traceback.format_exc()}\n")
        torch.cuda.synchronize()
        dist.barrier()
