"""Flat parameter / gradient arenas and the bucket plan.

The reference all-reduces ten separate tensors per step, five of them under
500 bytes (train.py:37-39).  Here all parameters live in ONE
contiguous fp32 arena, all gradients in a second arena with the same layout,
and the modules' ``Parameter`` objects are views into them.  A *bucket* is a
contiguous ``[start, stop)`` range of the arena; buckets are listed in the
order backward produces them (head first, conv1 last) so that a bucket can be
reduced while earlier layers are still running backward.

Every tensor's offset is aligned to 4 elements (16 B) and every bucket to 128
elements (512 B) so kernels can use 128-bit accesses and buckets never share a
cache line.  Padding elements are zero and stay zero (gradient 0 => update 0).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Sequence, Tuple

import torch
from torch import nn

TENSOR_ALIGN = 4
BUCKET_ALIGN = 128


def _round_up(x: int, a: int) -> int:
    return (x + a - 1) // a * a


@dataclass(frozen=True)
class Slot:
    name: str
    shape: Tuple[int, ...]
    offset: int
    numel: int


@dataclass(frozen=True)
class Bucket:
    index: int
    start: int
    stop: int
    names: Tuple[str, ...]

    @property
    def numel(self) -> int:
        return self.stop - self.start


class ArenaLayout:
    """Offsets of named tensors inside a flat arena, plus its buckets."""

    def __init__(self, named_shapes: Sequence[Tuple[str, Tuple[int, ...]]],
                 groups: Sequence[Sequence[str]]):
        """``groups`` lists parameter names per bucket in *arena order*; the
        backward order of the buckets is the reverse of that list."""
        shapes = dict(named_shapes)
        self.slots: Dict[str, Slot] = {}
        ranges = []
        off = 0
        for g in groups:
            off = _round_up(off, BUCKET_ALIGN)
            start = off
            for name in g:
                off = _round_up(off, TENSOR_ALIGN)
                n = 1
                for s in shapes[name]:
                    n *= s
                self.slots[name] = Slot(name, tuple(shapes[name]), off, n)
                off += n
            ranges.append((start, _round_up(off, BUCKET_ALIGN), tuple(g)))
            off = ranges[-1][1]
        self.total = off
        # bucket 0 is the one backward finishes first == last group in arena order
        self.buckets: List[Bucket] = [Bucket(i, s, e, n) for i, (s, e, n) in enumerate(reversed(ranges))]
        missing = set(shapes) - set(self.slots)
        if missing:
            raise ValueError(f"parameters not assigned to a bucket: {sorted(missing)}")

    def view(self, flat: torch.Tensor, name: str) -> torch.Tensor:
        s = self.slots[name]
        return flat[s.offset:s.offset + s.numel].view(s.shape)


def plan_buckets(named_shapes: Sequence[Tuple[str, Tuple[int, ...]]], bucket_bytes: int) -> List[List[str]]:
    """Greedy size-capped grouping in parameter order (what torch DDP does in
    reverse).  ``bucket_bytes <= 0`` puts everything in one bucket."""
    groups: List[List[str]] = [[]]
    size = 0
    for name, shp in named_shapes:
        n = 4
        for s in shp:
            n *= s
        if groups[-1] and bucket_bytes > 0 and size + n > bucket_bytes:
            groups.append([])
            size = 0
        groups[-1].append(name)
        size += n
    return groups


def lenet_groups(names: Sequence[str]) -> List[List[str]]:
    """Layer-aligned buckets for LeNet: {conv1, conv2}, {conv3}, {fc1, fc2}.

    Backward order is the reverse: the head's 11 K parameters first, conv3's
    48 K (78 % of the bytes) second, the convs last."""
    pick = lambda pre: [n for n in names if any(n.startswith(p) for p in pre)]  # noqa: E731
    return [pick(["ConvNet.1.", "ConvNet.4."]), pick(["ConvNet.7."]), pick(["FC."])]


class FlatArena:
    """Owns the flat param and grad tensors and re-points a module into them."""

    def __init__(self, module: nn.Module, groups: Sequence[Sequence[str]] = None,
                 device: torch.device = None):
        named = [(n, tuple(p.shape)) for n, p in module.named_parameters()]
        if groups is None:
            groups = [[n for n, _ in named]]
        self.layout = ArenaLayout(named, groups)
        device = device or next(module.parameters()).device
        self.param = torch.zeros(self.layout.total, dtype=torch.float32, device=device)
        self.grad = torch.zeros_like(self.param)
        self.module = module
        with torch.no_grad():
            for name, p in module.named_parameters():
                v = self.layout.view(self.param, name)
                v.copy_(p.detach().to(device=device, dtype=torch.float32))
                p.data = v
                p.grad = self.layout.view(self.grad, name)
