"""DataParallel — the model wrap that keeps replicas in lock-step.

Reference semantics (train.py:34-39, 49-51): after ``loss.backward()`` every
``param.grad`` is replaced by its mean over ranks (ten blocking per-tensor
``all_reduce`` + ``/= size``), then every rank applies the identical optimizer
update to its full replica.  Replica consistency relies on same-seed init;
there is no broadcast and no check (train.py:80-81).

This wrap provides the same contract with a different mechanism:

* parameters / gradients live in flat arenas (:mod:`b200dp.parallel.arena`);
* gradients are reduced per *bucket*, as soon as backward has produced the
  bucket (autograd post-accumulate hooks), overlapping the rest of backward;
* ``comm="torch"`` issues one async ``dist.all_reduce`` per bucket on
  gloo/NCCL and then runs the flat torch optimizer — the CPU plumbing path and
  the measured NCCL baseline;
* ``comm="b200"`` hands each bucket to the fused sm_100a peer-to-peer kernel
  (reduce + average + optimizer update in one launch, no NCCL on the step
  path) — see :class:`b200dp.parallel.p2p.P2PReducer`;
* extensions the reference lacks: :meth:`broadcast_parameters` (rank-0 init
  wins) and :meth:`replica_divergence` (max |param - rank0 param|).
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist
from torch import nn

from ..optim.flat import FlatOptimizer, OptHyper
from .arena import FlatArena, lenet_groups, plan_buckets


def _world() -> int:
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def average_gradients(model: nn.Module) -> None:
    """Literal reference behaviour (train.py:34-39): per-tensor blocking all-reduce + divide.
    Kept as the ``ref-literal`` baseline; nothing in the framework's own step path calls it."""
    size = float(_world())
    for p in model.parameters():
        if size > 1:
            dist.all_reduce(p.grad.data, op=dist.ReduceOp.SUM)
        p.grad.data /= size


class DataParallel(nn.Module):
    """Eager-mode wrap: autograd computes gradients, buckets are reduced as they
    complete, the flat optimizer applies the update.

    Usage (the train.py:46-51 sequence)::

        ddp = DataParallel(LeNet(), OptHyper("adadelta", lr=0.1))
        out = ddp(x); loss = loss_fn(out, y); ddp.zero_grad()
        loss.backward()          # bucket all-reduces start inside backward
        ddp.step()               # wait, average, optimizer update
    """

    def __init__(self, module: nn.Module, hyper: OptHyper = None, *, bucket_bytes: int = -1,
                 layer_buckets: bool = True, overlap: bool = True,
                 device: Optional[torch.device] = None):
        super().__init__()
        self.module = module
        names = [n for n, _ in module.named_parameters()]
        if layer_buckets and any(n.startswith("ConvNet.7.") for n in names):
            groups = lenet_groups(names)
        else:
            groups = plan_buckets([(n, tuple(p.shape)) for n, p in module.named_parameters()], bucket_bytes)
        self.arena = FlatArena(module, groups, device)
        self.optimizer = FlatOptimizer(self.arena.param, self.arena.grad, hyper or OptHyper())
        self.world_size = _world()
        self.overlap = overlap and self.world_size > 1
        self._pending: List = []
        self._bucket_of = {}
        self._remaining: List[int] = []
        self._hooks = []
        for b in self.arena.layout.buckets:
            for n in b.names:
                self._bucket_of[n] = b.index
        if self.overlap:
            for name, p in module.named_parameters():
                self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(name)))
        self._reset_counts()

    # -- module plumbing -----------------------------------------------------
    def forward(self, *a, **kw):
        return self.module(*a, **kw)

    def state_dict(self, *a, **kw):  # checkpoint keys must be the bare model's (train.py:108)
        return self.module.state_dict(*a, **kw)

    def load_state_dict(self, sd, strict: bool = True):
        with torch.no_grad():
            for k, v in self.module.state_dict().items():
                if k in sd:
                    v.copy_(sd[k])
                elif strict:
                    raise KeyError(k)
        return self

    def zero_grad(self, set_to_none: bool = False) -> None:
        self.arena.grad.zero_()
        self._reset_counts()

    # -- gradient reduction ----------------------------------------------------
    def _reset_counts(self) -> None:
        self._remaining = [len(b.names) for b in self.arena.layout.buckets]
        self._pending = []

    def _make_hook(self, name: str):
        def hook(_p):
            bi = self._bucket_of[name]
            self._remaining[bi] -= 1
            if self._remaining[bi] == 0:
                self._launch_bucket(bi)
        return hook

    def _launch_bucket(self, bi: int) -> None:
        b = self.arena.layout.buckets[bi]
        work = dist.all_reduce(self.arena.grad[b.start:b.stop], op=dist.ReduceOp.SUM, async_op=True)
        self._pending.append(work)

    def reduce_gradients(self) -> None:
        """Make ``grad`` the SUM over ranks (the 1/size is folded into the update)."""
        if self.world_size == 1:
            return
        if self.overlap:
            for bi, left in enumerate(self._remaining):  # buckets whose hooks never all fired
                if left > 0:
                    self._launch_bucket(bi)
        else:
            for bi in range(len(self.arena.layout.buckets)):
                self._launch_bucket(bi)
        for w in self._pending:
            w.wait()
        self._pending = []

    def step(self) -> None:
        """``average_gradients(model); optimizer.step()`` (train.py:50-51)."""
        self.reduce_gradients()
        self.optimizer.step_torch(grad_scale=1.0 / self.world_size)
        self._reset_counts()

    # -- extensions ------------------------------------------------------------
    @torch.no_grad()
    def broadcast_parameters(self, src: int = 0) -> None:
        if self.world_size > 1:
            dist.broadcast(self.arena.param, src=src)

    @torch.no_grad()
    def replica_divergence(self) -> float:
        """max over ranks of max|param - rank0's param|; 0.0 means bit-identical replicas."""
        if self.world_size == 1:
            return 0.0
        ref = self.arena.param.clone()
        dist.broadcast(ref, src=0)
        d = (self.arena.param - ref).abs().max().reshape(1)
        dist.all_reduce(d, op=dist.ReduceOp.MAX)
        return float(d.item())
