"""Device-side timing helpers.

The reference has no timers at all.  Every number this
framework reports is taken with CUDA events on the launching stream, after a
synchronize on both sides, and multi-GPU numbers are the MAX over ranks.
"""
from __future__ import annotations

import contextlib
import time
from typing import Callable, List

import torch
import torch.distributed as dist


def _dist_on() -> bool:
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def max_over_ranks(value: float, device=None) -> float:
    if not _dist_on():
        return value
    dev = device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu")
    t = torch.tensor([value], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier_sync(device=None) -> None:
    if _dist_on():
        dist.barrier()
    if torch.cuda.is_available() and (device is None or torch.device(device).type == "cuda"):
        torch.cuda.synchronize()


class DeviceTimer:
    """``with DeviceTimer() as t: ...`` -> ``t.ms`` (CUDA events; wall clock on CPU)."""

    def __init__(self, device=None):
        self.cuda = torch.cuda.is_available() and (device is None or torch.device(device).type == "cuda")
        self.ms = 0.0

    def __enter__(self):
        if self.cuda:
            self._a, self._b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            self._a.record()
        else:
            self._t0 = time.perf_counter()
        return self

    def __exit__(self, *exc):
        if self.cuda:
            self._b.record()
            torch.cuda.synchronize()
            self.ms = self._a.elapsed_time(self._b)
        else:
            self.ms = (time.perf_counter() - self._t0) * 1e3
        return False


def time_steps(step: Callable[[int], None], steps: int, warmup: int, device=None) -> float:
    """ms per step: ``warmup`` untimed calls, then exactly ``steps`` timed calls bracketed by
    barrier + synchronize, max over ranks."""
    for i in range(warmup):
        step(i)
    barrier_sync(device)
    with DeviceTimer(device) as t:
        for i in range(steps):
            step(warmup + i)
    barrier_sync(device)
    return max_over_ranks(t.ms, device) / max(steps, 1)


@contextlib.contextmanager
def nvtx_range(name: str):
    """NVTX range when CUDA is present (visible to ncu/nsys), no-op otherwise."""
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()


def flush_l2(buf: List[torch.Tensor]) -> None:
    """Write a buffer larger than the 126 MB L2 (allocated lazily into ``buf``)."""
    if not buf:
        buf.append(torch.empty(192 * 1024 * 1024, dtype=torch.uint8, device="cuda"))
    buf[0].fill_(1)
