"""DataParallel — the model wrap that keeps replicas in lock-step, for ANY ``nn.Module``.

Reference semantics (train.py:34-39, 49-51): after ``loss.backward()`` every
``param.grad`` is replaced by its mean over ranks (ten blocking per-tensor
``all_reduce`` + ``/= size``), then every rank applies the identical optimizer
update to its full replica.  Replica consistency relies on same-seed init;
there is no broadcast and no check (train.py:80-81).

This wrap provides the same contract with a different mechanism:

* parameters / gradients live in flat arenas (:mod:`b200dp.parallel.arena`);
  buckets are layer-aligned for LeNet or size-capped (``bucket_bytes``) for any
  other module;
* gradients leave per *bucket*, as soon as autograd has produced the bucket
  (post-accumulate hooks), overlapping the rest of backward on a side stream;
* ``comm="b200"`` (default on a CUDA device) runs the hand-written sm_100a
  kernels — no NCCL call and no elementwise optimizer kernel on the step path:
  whole-model gradients up to 1 MB use the push protocol
  (``b200dp_grad_push`` per bucket from the hooks, one
  ``b200dp_reduce_update_ll`` in :meth:`step`: wait + rank-order sum + average
  + Adadelta/SGD/Adam); larger models use one fused handshake launch per
  bucket (``b200dp_reduce_update``: one-shot / two-shot / NVLS by size).
  Symmetric memory comes from :class:`b200dp.parallel.p2p.P2PComm`;
* ``comm="torch"`` issues one async ``dist.all_reduce`` per bucket on
  gloo/NCCL and then runs the flat torch optimizer — the CPU plumbing path and
  the measured NCCL baseline;
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
    complete, the optimizer update is fused with the reduction (``comm="b200"``)
    or applied by the flat torch optimizer (``comm="torch"``).

    Usage (the train.py:46-51 sequence)::

        ddp = DataParallel(model, OptHyper("adadelta", lr=0.1))
        out = ddp(x); loss = loss_fn(out, y); ddp.zero_grad()
        loss.backward()          # buckets leave for the peers inside backward
        ddp.step()               # wait, average, optimizer update
    """

    def __init__(self, module: nn.Module, hyper: OptHyper = None, *, bucket_bytes: int = -1,
                 layer_buckets: bool = True, overlap: bool = True,
                 device: Optional[torch.device] = None, comm: str = "auto", algo: str = "auto",
                 wire: str = "fp32", timeout_ms: int = 0):
        super().__init__()
        self.module = module
        names = [n for n, _ in module.named_parameters()]
        if layer_buckets and bucket_bytes <= 0 and any(n.startswith("ConvNet.7.") for n in names):
            groups = lenet_groups(names)
        else:
            groups = plan_buckets([(n, tuple(p.shape)) for n, p in module.named_parameters()], bucket_bytes)
        self.arena = FlatArena(module, groups, device)
        self.optimizer = FlatOptimizer(self.arena.param, self.arena.grad, hyper or OptHyper())
        self.world_size = _world()
        dev = self.arena.param.device
        if comm not in ("auto", "b200", "torch"):
            raise ValueError(f"unknown comm {comm!r}")
        if comm == "auto":
            comm = "b200" if dev.type == "cuda" else "torch"
        if comm == "b200" and dev.type != "cuda":
            raise ValueError('comm="b200" needs the model on a CUDA device')
        self.comm_kind = comm
        self.native = comm == "b200"
        self.overlap = overlap and (self.world_size > 1 or self.native)
        self._pending: List = []
        self._bucket_of = {}
        self._remaining: List[int] = []
        self._hooks = []
        for b in self.arena.layout.buckets:
            for n in b.names:
                self._bucket_of[n] = b.index
        if self.native:
            self._init_native(algo, wire, timeout_ms)
        if self.overlap:
            for name, p in module.named_parameters():
                self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(name)))
        self._reset_counts()

    # -- native (sm_100a) state ------------------------------------------------
    def _init_native(self, algo: str, wire: str, timeout_ms: int) -> None:
        from ..ops import native
        from .p2p import P2PComm, ll_max_bytes
        native.load()
        self._nat = native
        dev = self.arena.param.device
        rank = dist.get_rank() if self.world_size > 1 else 0
        with torch.cuda.device(dev):
            self.p2p = (P2PComm(self.arena.layout.total, rank, self.world_size, algo, wire=wire, timeout_ms=timeout_ms)
                        if self.world_size > 1 else None)
        small = self.arena.layout.total * 4 <= ll_max_bytes(self.world_size)
        self.push = self.p2p.push if self.p2p is not None else (small and algo in ("auto", "push"))
        self.side = torch.cuda.Stream(device=dev)
        self.hyper_dev = torch.zeros(8, dtype=torch.float32, device=dev)
        self.hyper_host = torch.zeros(8, dtype=torch.float32).pin_memory()
        self._hyper_cache = None
        self.step_counter = torch.zeros(4, dtype=torch.int32, device=dev)
        self._sent: List[bool] = []
        self._launches = 0               # native launches of the last step (tests: bucket_bytes changes it)

    def _sync_hyper(self) -> None:
        vec = self.optimizer.hyper_vector()
        if vec != self._hyper_cache:
            self.hyper_host[:5] = torch.tensor(vec, dtype=torch.float32)
            self.hyper_dev.copy_(self.hyper_host, non_blocking=True)
            self._hyper_cache = vec

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
        if self.native:
            self._sent = [False] * len(self.arena.layout.buckets)
            self._step_launches = 0

    def _make_hook(self, name: str):
        def hook(_p):
            bi = self._bucket_of[name]
            self._remaining[bi] -= 1
            if self._remaining[bi] == 0:
                self._launch_bucket(bi)
        return hook

    def _launch_bucket(self, bi: int) -> None:
        b = self.arena.layout.buckets[bi]
        if not self.native:
            work = dist.all_reduce(self.arena.grad[b.start:b.stop], op=dist.ReduceOp.SUM, async_op=True)
            self._pending.append(work)
            return
        # native: a bucket leaves on the side stream as soon as autograd has written it; the last bucket of the
        # handshake protocol is held back for step(), where it is known to be the last
        self._sync_hyper()       # handshake launches UPDATE inside backward: the scalars must be current before the first
        if self.push:
            if self.p2p is None:
                return
            self._on_side(lambda: self._nat.grad_push(self.arena.grad, b.start, b.numel, self.step_counter,
                                                      self.p2p.struct_for(0)))
            self._sent[bi] = True
        elif sum(1 for s in self._sent if not s) > 1:
            self._on_side(lambda: self._handshake_bucket(bi, last=False, overlapped=True))
            self._sent[bi] = True

    def _on_side(self, fn) -> None:
        main = torch.cuda.current_stream(self.arena.param.device)
        self.side.wait_stream(main)
        with torch.cuda.stream(self.side):
            fn()
        self._step_launches += 1

    def _handshake_bucket(self, bi: int, last: bool, overlapped: bool) -> None:
        b, opt = self.arena.layout.buckets[bi], self.optimizer
        self._nat.reduce_update(self.arena.param, self.arena.grad, opt.state0, opt.state1, None, b.start, b.numel,
                                None, 0, 0, opt.hyper.code(), self.hyper_dev, self.step_counter, last,
                                self.p2p.struct_for(b.numel * 4, overlapped=overlapped) if self.p2p else None)

    def _step_native(self) -> None:
        opt = self.optimizer
        self._sync_hyper()
        opt.step_count += 1
        nb = len(self.arena.layout.buckets)
        main = torch.cuda.current_stream(self.arena.param.device)
        if self.push:
            rest = [bi for bi in range(nb) if not self._sent[bi]]
            all_here = len(rest) == nb             # nothing was sent from the hooks: the final launch pushes everything
            if not all_here and self.p2p is not None:
                for bi in rest:                    # buckets whose hooks never all fired (unused parameters: zero gradient)
                    b = self.arena.layout.buckets[bi]
                    self._on_side(lambda b=b: self._nat.grad_push(self.arena.grad, b.start, b.numel, self.step_counter,
                                                                  self.p2p.struct_for(0)))
            self._on_side(lambda: self._nat.reduce_update_ll(
                self.arena.param, self.arena.grad, opt.state0, opt.state1, None, 0, self.arena.layout.total, None, 0, 0,
                0, 0, all_here, opt.hyper.code(), self.hyper_dev, self.step_counter,
                self.p2p.struct_for(0) if self.p2p else None, last=True))
        else:
            rest = [bi for bi in range(nb) if not self._sent[bi]]
            for k, bi in enumerate(rest):
                self._on_side(lambda bi=bi, k=k: self._handshake_bucket(bi, last=k == len(rest) - 1, overlapped=False))
        main.wait_stream(self.side)
        self._launches = self._step_launches

    def reduce_gradients(self) -> None:
        """Make ``grad`` the SUM over ranks (the 1/size is folded into the update).  ``comm="torch"`` only: the
        native kernels reduce, average and update in one launch (:meth:`step`)."""
        if self.native:
            raise RuntimeError('comm="b200" fuses the reduction with the update: call step()')
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
        if self.native:
            self._step_native()
        else:
            self.reduce_gradients()
            self.optimizer.step_torch(grad_scale=1.0 / self.world_size)
        self._reset_counts()

    def check_error(self) -> None:
        """Raise if a peer did not deliver its gradients within the time bound (native path; one 4-byte D2H read)."""
        if self.native and int(self.step_counter[2].item()) != 0:
            raise RuntimeError("a peer rank did not deliver its gradients within the time bound (dead or stalled rank?)")

    def close(self) -> None:
        if self.native and getattr(self, "p2p", None) is not None:
            self.p2p.close()
            self.p2p = None

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
