"""Process launch and process-group bootstrap.

Reference behaviour (train.py:26-31, 134-147): the parent counts GPUs,
``torch.multiprocessing.spawn`` starts one child per GPU, each child sets
``MASTER_ADDR=127.0.0.1`` / ``MASTER_PORT=29500`` and calls
``dist.init_process_group(backend, rank, world_size)``; rank == device index;
there is no ``set_device`` and no teardown.

This module keeps that model (one OS process per GPU, single node, loopback
rendezvous) and fixes the gaps: the device is bound explicitly, the port is
configurable (and probed when 0), the group is destroyed on exit, a timeout
is passed so a dead peer is an error rather than a hang, and
the same entry works under ``torchrun`` (reads RANK / LOCAL_RANK / WORLD_SIZE /
MASTER_* from the environment).  torch.distributed is used for rendezvous,
handle exchange and the CPU/gloo data path only — on B200 the step path never
calls it (see :mod:`b200dp.parallel.ddp`).
"""
from __future__ import annotations

import datetime
import os
import socket
from dataclasses import dataclass
from typing import Any, Callable, Optional

import torch
import torch.distributed as dist

DEFAULT_ADDR = "127.0.0.1"
DEFAULT_PORT = 29500


@dataclass
class RankContext:
    rank: int
    world_size: int
    local_rank: int
    device: torch.device
    backend: str            # torch.distributed backend used for plumbing: "nccl" | "gloo"

    @property
    def is_main(self) -> bool:
        """The reference gates logging and checkpointing on rank 0 (train.py:52, 73, 107)."""
        return self.rank == 0


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind((DEFAULT_ADDR, 0))
        return s.getsockname()[1]


def pick_plumbing_backend(device: torch.device) -> str:
    return "nccl" if device.type == "cuda" else "gloo"


def init_process(rank: int, size: int, fn: Callable[[RankContext, Any], Any], pyargs: Any,
                 backend: Optional[str] = None, *, use_cuda: Optional[bool] = None,
                 addr: str = DEFAULT_ADDR, port: int = DEFAULT_PORT,
                 timeout_s: float = 600.0, local_rank: Optional[int] = None):
    """Child entry: bind device, join the group, run ``fn(ctx, pyargs)``, tear down.

    Signature mirrors the reference's ``init_process(rank, size, fn, pyargs,
    backend)`` (train.py:26); ``fn`` receives a :class:`RankContext` instead of
    the bare ``(rank, size)`` pair.
    """
    if use_cuda is None:
        use_cuda = torch.cuda.is_available() and backend != "gloo"
    local_rank = rank if local_rank is None else local_rank
    if use_cuda:
        device = torch.device("cuda", local_rank % max(torch.cuda.device_count(), 1))
        torch.cuda.set_device(device)
    else:
        device = torch.device("cpu")
    backend = backend or pick_plumbing_backend(device)
    os.environ.setdefault("MASTER_ADDR", addr)
    os.environ.setdefault("MASTER_PORT", str(port))
    kwargs = {}
    if backend == "nccl":
        kwargs["device_id"] = device
    dist.init_process_group(backend, rank=rank, world_size=size,
                            timeout=datetime.timedelta(seconds=timeout_s), **kwargs)
    ctx = RankContext(rank, size, local_rank, device, backend)
    try:
        return fn(ctx, pyargs)
    finally:
        if dist.is_initialized():
            try:
                dist.barrier()
            except Exception:  # a failed peer must not mask the original error
                pass
            dist.destroy_process_group()


def core_share(local_rank: int, ranks_on_node: int, allowed=None):
    """This rank's slice of the CPU cores the process may run on: the allowed set (``sched_getaffinity``) in ascending
    order, cut into ``ranks_on_node`` contiguous, disjoint slices.  ``None`` when there are fewer than two cores per
    rank (nothing to gain from pinning)."""
    if allowed is None:
        allowed = os.sched_getaffinity(0) if hasattr(os, "sched_getaffinity") else ()
    cores = sorted(allowed)
    per = len(cores) // max(ranks_on_node, 1)
    if per < 2:
        return None
    return set(cores[local_rank * per:(local_rank + 1) * per])


def pin_to_core_share(local_rank: int, ranks_on_node: int) -> Optional[set]:
    """Opt-in (``--pin-cores``): restrict this process — the Python thread and every thread it starts afterwards, i.e.
    the step pump's launcher, helper and gather workers — to its own slice of the node's cores, so that eight ranks'
    host threads never share a core.  Off by default: see DESIGN.md (the end-to-end step at 8 GPUs is within 1 us of the
    device-timed one at steady state without it)."""
    share = core_share(local_rank, ranks_on_node)
    if share and hasattr(os, "sched_setaffinity"):
        os.sched_setaffinity(0, share)
        return share
    return None


def _spawn_entry(rank, size, fn, pyargs, backend, use_cuda, addr, port, timeout_s):
    os.environ["MASTER_ADDR"] = addr
    os.environ["MASTER_PORT"] = str(port)
    init_process(rank, size, fn, pyargs, backend, use_cuda=use_cuda, addr=addr, port=port,
                 timeout_s=timeout_s)


def spawn(fn: Callable[[RankContext, Any], Any], pyargs: Any, world_size: int,
          backend: Optional[str] = None, *, use_cuda: Optional[bool] = None,
          addr: str = DEFAULT_ADDR, port: int = DEFAULT_PORT, timeout_s: float = 600.0) -> None:
    """Parent entry (reference train.py:147): one process per rank, join all.

    ``world_size == 1`` runs in-process (no fork) — the reference cannot run
    at size 1 at all (splitdataset.py:38-39).
    """
    if port == 0:
        port = free_port()
    if world_size == 1:
        os.environ["MASTER_ADDR"] = addr
        os.environ["MASTER_PORT"] = str(port)
        init_process(0, 1, fn, pyargs, backend, use_cuda=use_cuda, addr=addr, port=port,
                     timeout_s=timeout_s)
        return
    torch.multiprocessing.spawn(
        _spawn_entry, args=(world_size, fn, pyargs, backend, use_cuda, addr, port, timeout_s),
        nprocs=world_size, join=True)


def launched_by_torchrun() -> bool:
    return "RANK" in os.environ and "WORLD_SIZE" in os.environ


def init_from_env(fn: Callable[[RankContext, Any], Any], pyargs: Any, backend: Optional[str] = None,
                  *, use_cuda: Optional[bool] = None, timeout_s: float = 600.0):
    """Entry for ``python -m torch.distributed.run``: rank layout comes from the env."""
    rank = int(os.environ["RANK"])
    size = int(os.environ["WORLD_SIZE"])
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    return init_process(rank, size, fn, pyargs, backend, use_cuda=use_cuda,
                        addr=os.environ.get("MASTER_ADDR", DEFAULT_ADDR),
                        port=int(os.environ.get("MASTER_PORT", DEFAULT_PORT)),
                        timeout_s=timeout_s, local_rank=local_rank)


def discover_world_size(no_cuda: bool = False) -> int:
    """Reference train.py:134-142: world size = number of visible GPUs (0 if none / disabled)."""
    if no_cuda or not torch.cuda.is_available():
        return 0
    return torch.cuda.device_count()
