"""LeNet — the reference's only model, parameterised for B200 work.

Architecture (reference cnn.py:9-23): bilinear upsample to 32x32
(align_corners=True) -> conv5x5(Cin->6w) -> LeakyReLU(0.2) -> maxpool2 ->
conv5x5(6w->16w) -> LeakyReLU(0.2) -> maxpool2 -> conv5x5(16w->120w) ->
LeakyReLU(0.2) -> flatten -> Linear(120w->84w) -> LeakyReLU(0.2) ->
Linear(84w->10) -> Softmax(-1).

``width_mult=1, in_ch=1, in_hw=28`` is the reference model; its
``state_dict`` has exactly the reference's ten keys / shapes / fp32 dtype
(``ConvNet.{1,4,7}.{weight,bias}``, ``FC.{0,2}.{weight,bias}``), which is the
checkpoint compatibility contract.  ``width_mult=4, in_ch=3,
in_hw=32`` is the widened CIFAR-shaped variant of BASELINE.json configs[3];
with a 32x32 input the upsample is the identity and is elided (slot 0 of
``ConvNet`` stays a parameter-less module so the key indices do not move).

This module is the *eager* definition: it is the oracle for every native
kernel, the CPU / gloo execution path, and the container whose parameters the
native runner (:mod:`b200dp.models.fused`) re-points into its flat arena.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Tuple

import torch
from torch import nn

LEAKY_SLOPE = 0.2


@dataclass(frozen=True)
class LeNetConfig:
    in_ch: int = 1
    width_mult: int = 1
    in_hw: int = 28
    num_classes: int = 10

    @property
    def c1(self) -> int:
        return 6 * self.width_mult

    @property
    def c2(self) -> int:
        return 16 * self.width_mult

    @property
    def c3(self) -> int:
        return 120 * self.width_mult

    @property
    def f1(self) -> int:
        return 84 * self.width_mult

    def param_shapes(self) -> List[Tuple[str, Tuple[int, ...]]]:
        """(state_dict key, shape) in ``parameters()`` order (reference train.py:37)."""
        return [
            ("ConvNet.1.weight", (self.c1, self.in_ch, 5, 5)),
            ("ConvNet.1.bias", (self.c1,)),
            ("ConvNet.4.weight", (self.c2, self.c1, 5, 5)),
            ("ConvNet.4.bias", (self.c2,)),
            ("ConvNet.7.weight", (self.c3, self.c2, 5, 5)),
            ("ConvNet.7.bias", (self.c3,)),
            ("FC.0.weight", (self.f1, self.c3)),
            ("FC.0.bias", (self.f1,)),
            ("FC.2.weight", (self.num_classes, self.f1)),
            ("FC.2.bias", (self.num_classes,)),
        ]

    def num_params(self) -> int:
        n = 0
        for _, shp in self.param_shapes():
            k = 1
            for s in shp:
                k *= s
            n += k
        return n

    def flops_per_image_fwd(self) -> int:
        """2*MACs of the five GEMM-shaped layers (833,040 for the reference model)."""
        c0, c1, c2, c3, f1, nc = self.in_ch, self.c1, self.c2, self.c3, self.f1, self.num_classes
        macs = 28 * 28 * c1 * c0 * 25 + 10 * 10 * c2 * c1 * 25 + c3 * c2 * 25 + f1 * c3 + nc * f1
        return 2 * macs


MNIST_LENET = LeNetConfig()
CIFAR_WIDE_LENET = LeNetConfig(in_ch=3, width_mult=4, in_hw=32)


class LeNet(nn.Module):
    """Reference-compatible LeNet (reference cnn.py:4-30)."""

    def __init__(self, config: LeNetConfig = MNIST_LENET):
        super().__init__()
        self.config = config
        c = config
        first = nn.UpsamplingBilinear2d(size=32) if c.in_hw != 32 else nn.Identity()
        self.ConvNet = nn.Sequential(
            first,
            nn.Conv2d(c.in_ch, c.c1, 5, padding=0, stride=1),
            nn.LeakyReLU(LEAKY_SLOPE),
            nn.MaxPool2d(2, stride=2, padding=0),
            nn.Conv2d(c.c1, c.c2, 5, padding=0, stride=1),
            nn.LeakyReLU(LEAKY_SLOPE),
            nn.MaxPool2d(2, stride=2, padding=0),
            nn.Conv2d(c.c2, c.c3, 5, padding=0, stride=1),
            nn.LeakyReLU(LEAKY_SLOPE),
        )
        self.FC = nn.Sequential(
            nn.Linear(c.c3, c.f1),
            nn.LeakyReLU(LEAKY_SLOPE),
            nn.Linear(c.f1, c.num_classes),
            nn.Softmax(-1),
        )

    def forward(self, img: torch.Tensor) -> torch.Tensor:
        out = self.ConvNet(img)
        out = out.view(out.shape[0], -1)
        return self.FC(out)

    def logits(self, img: torch.Tensor) -> torch.Tensor:
        """Pre-softmax activations (used by the ``xent`` loss mode)."""
        out = self.ConvNet(img)
        out = out.view(out.shape[0], -1)
        return self.FC[2](self.FC[1](self.FC[0](out)))


def loss_fn(probs_or_logits: torch.Tensor, target: torch.Tensor, mode: str = "ref",
            reduction: str = "mean") -> torch.Tensor:
    """``ref``: ``nll_loss`` on softmax *probabilities* = ``-mean p[target]``
    (reference cnn.py:23 + train.py:48); ``xent``: cross-entropy on logits."""
    if mode == "ref":
        return torch.nn.functional.nll_loss(probs_or_logits, target, reduction=reduction)
    if mode == "xent":
        return torch.nn.functional.cross_entropy(probs_or_logits, target, reduction=reduction)
    raise ValueError(f"unknown loss mode {mode!r}")
