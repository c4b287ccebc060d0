"""Flat-arena optimizers and the StepLR schedule.

The reference uses ``optim.Adadelta(model.parameters(), lr=args.lr)``
(rho=0.9, eps=1e-6, weight_decay=0; reference train.py:99) stepped once per
batch (train.py:51) and ``StepLR(optimizer, step_size=1, gamma)`` stepped once
per epoch (train.py:101, 105).  BASELINE.json additionally names SGD-momentum
and Adam.

Here an optimizer owns *one* flat fp32 parameter arena, one flat gradient
arena and up to two flat state arenas (Adadelta: square_avg, acc_delta;
SGD: momentum buffer; Adam: exp_avg, exp_avg_sq).  Two executors share this
state:

* :meth:`FlatOptimizer.step_torch` — a handful of fused torch ops over the
  whole arena; the CPU / gloo path and the oracle for the CUDA kernels;
* the native kernels (``csrc/opt_math.cuh``, used by ``csrc/reduce_update.cu``
  and ``csrc/ll_reduce.cu``), which apply the same arithmetic per element,
  fused with the cross-rank gradient reduction; their ``hyper[]`` vector
  ``{lr, h0, h1, h2, weight_decay}`` is built by :meth:`FlatOptimizer.hyper_vector`
  (Adam's bias correction is computed on the device from the step number).

The arithmetic follows torch.optim (single-tensor variants) exactly so that
per-parameter torch optimizers and the flat one agree to fp32 round-off.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict

import torch

OPT_ADADELTA, OPT_SGD, OPT_ADAM = 0, 1, 2
_KINDS = {"adadelta": OPT_ADADELTA, "sgd": OPT_SGD, "adam": OPT_ADAM}


@dataclass
class OptHyper:
    """Hyper-parameters, laid out the way the kernels take them (h0..h3)."""
    kind: str = "adadelta"
    lr: float = 0.1
    rho: float = 0.9          # adadelta
    eps: float = 1e-6         # adadelta (adam default is set in __post_init__)
    momentum: float = 0.9     # sgd
    nesterov: bool = False    # sgd
    beta1: float = 0.9        # adam
    beta2: float = 0.999      # adam
    weight_decay: float = 0.0
    _adam_eps_default: float = field(default=1e-8, repr=False)

    def code(self) -> int:
        return _KINDS[self.kind]

    def num_states(self) -> int:
        return {"adadelta": 2, "sgd": 1, "adam": 2}[self.kind]


class FlatOptimizer:
    """Optimizer over a flat (param, grad) arena pair."""

    def __init__(self, param: torch.Tensor, grad: torch.Tensor, hyper: OptHyper):
        assert param.dim() == 1 and grad.shape == param.shape
        assert param.dtype == torch.float32 and grad.dtype == torch.float32
        if hyper.kind not in _KINDS:
            raise ValueError(f"unknown optimizer {hyper.kind!r}")
        if hyper.kind == "adam" and hyper.eps == 1e-6:
            hyper.eps = hyper._adam_eps_default
        self.param, self.grad, self.hyper = param, grad, hyper
        self.state0 = torch.zeros_like(param)
        self.state1 = torch.zeros_like(param) if hyper.num_states() > 1 else param.new_zeros(1)
        self.step_count = 0
        self.initial_lr = hyper.lr
        # torch.optim look-alike so generic schedulers / user code can poke at lr
        self.param_groups = [{"lr": hyper.lr, "initial_lr": hyper.lr}]

    # lr lives in param_groups[0] so that a scheduler mutation is seen by both executors
    @property
    def lr(self) -> float:
        return self.param_groups[0]["lr"]

    @lr.setter
    def lr(self, v: float) -> None:
        self.param_groups[0]["lr"] = float(v)

    def zero_grad(self) -> None:
        self.grad.zero_()

    @torch.no_grad()
    def step_torch(self, grad_scale: float = 1.0) -> None:
        """Reference arithmetic in torch ops. ``grad_scale`` folds in 1/world_size."""
        h, p = self.hyper, self.param
        g = self.grad if grad_scale == 1.0 else self.grad * grad_scale
        if h.weight_decay != 0.0:
            g = g.add(p, alpha=h.weight_decay)
        self.step_count += 1
        lr = self.lr
        if h.kind == "adadelta":
            sq, acc = self.state0, self.state1
            sq.mul_(h.rho).addcmul_(g, g, value=1 - h.rho)
            std = sq.add(h.eps).sqrt_()
            delta = acc.add(h.eps).sqrt_().div_(std).mul_(g)
            acc.mul_(h.rho).addcmul_(delta, delta, value=1 - h.rho)
            p.add_(delta, alpha=-lr)
        elif h.kind == "sgd":
            buf = self.state0
            if h.momentum != 0.0:
                buf.mul_(h.momentum).add_(g)  # zero-initialised buffer => first step buf = g, as in torch
                d = g.add(buf, alpha=h.momentum) if h.nesterov else buf
            else:
                d = g
            p.add_(d, alpha=-lr)
        else:  # adam
            m, v = self.state0, self.state1
            m.mul_(h.beta1).add_(g, alpha=1 - h.beta1)
            v.mul_(h.beta2).addcmul_(g, g, value=1 - h.beta2)
            bc1 = 1 - h.beta1 ** self.step_count
            bc2 = 1 - h.beta2 ** self.step_count
            denom = (v.sqrt() / (bc2 ** 0.5)).add_(h.eps)
            p.addcdiv_(m, denom, value=-lr / bc1)

    def hyper_vector(self):
        """The kernels' ``hyper[]`` contract, ``(lr, h0, h1, h2, weight_decay)`` — include/b200dp.h."""
        h = self.hyper
        if h.kind == "adadelta":
            return (self.lr, h.rho, h.eps, 0.0, h.weight_decay)
        if h.kind == "sgd":
            return (self.lr, h.momentum, 1.0 if h.nesterov else 0.0, 0.0, h.weight_decay)
        return (self.lr, h.beta1, h.beta2, h.eps, h.weight_decay)

    # -- checkpointing -------------------------------------------------------
    def state_dict(self) -> Dict:
        return {"kind": self.hyper.kind, "lr": self.lr, "initial_lr": self.initial_lr,
                "step_count": self.step_count, "hyper": dict(self.hyper.__dict__),
                "state0": self.state0.detach().cpu().clone(),
                "state1": self.state1.detach().cpu().clone()}

    def load_state_dict(self, sd: Dict) -> None:
        if sd["kind"] != self.hyper.kind:
            raise ValueError(f"optimizer kind mismatch: {sd['kind']} vs {self.hyper.kind}")
        self.lr = sd["lr"]
        self.initial_lr = sd["initial_lr"]
        self.step_count = sd["step_count"]
        self.state0.copy_(sd["state0"])
        self.state1.copy_(sd["state1"])


class StepLR:
    """``lr = initial_lr * gamma ** (epoch // step_size)`` (reference train.py:101, 105)."""

    def __init__(self, optimizer: FlatOptimizer, step_size: int = 1, gamma: float = 0.9):
        self.optimizer, self.step_size, self.gamma = optimizer, step_size, gamma
        self.last_epoch = 0

    def step(self) -> None:
        self.last_epoch += 1
        self.optimizer.lr = self.optimizer.initial_lr * self.gamma ** (self.last_epoch // self.step_size)

    def get_last_lr(self):
        return [self.optimizer.lr]

    def state_dict(self) -> Dict:
        return {"last_epoch": self.last_epoch, "step_size": self.step_size, "gamma": self.gamma}

    def load_state_dict(self, sd: Dict) -> None:
        self.last_epoch, self.step_size, self.gamma = sd["last_epoch"], sd["step_size"], sd["gamma"]
        self.optimizer.lr = self.optimizer.initial_lr * self.gamma ** (self.last_epoch // self.step_size)
