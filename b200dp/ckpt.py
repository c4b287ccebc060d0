"""Checkpointing.

Reference (train.py:107-108): at the end of training rank 0 runs
``torch.save(model.state_dict(), "mnist_cnn.pt")`` — weights only, ten fp32
tensors with the ``nn.Sequential`` key names, no barrier, no resume path
(SURVEY.md §5.4).

Two files here:

* the **compat file** — exactly that format (an ``OrderedDict`` of the ten
  fp32 tensors, on CPU so no ``map_location`` is needed), loadable by the
  reference's ``LeNet().load_state_dict(strict=True)`` and accepting a file the
  reference wrote (CUDA tensors are mapped to CPU on load);
* the optional **full-state file** (``<name>.full``) — optimizer state arenas,
  scheduler, epoch / step counters, RNG state — for resume.  It is separate so
  the compat file stays pure.
"""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Any, Dict, Optional

import torch

DEFAULT_NAME = "mnist_cnn.pt"  # reference train.py:108


def model_state_dict_cpu(model) -> "OrderedDict[str, torch.Tensor]":
    sd = OrderedDict()
    for k, v in model.state_dict().items():
        sd[k] = v.detach().to("cpu", torch.float32).clone().contiguous()
    return sd


def save_model(model, path: str = DEFAULT_NAME, *, is_main: bool = True) -> Optional[str]:
    """Write the compat file (rank 0 only, like train.py:107)."""
    if not is_main:
        return None
    tmp = path + ".tmp"
    torch.save(model_state_dict_cpu(model), tmp)
    os.replace(tmp, path)  # atomic: a reader never sees a half-written file
    return path


def load_model(model, path: str = DEFAULT_NAME, *, strict: bool = True):
    """Load a compat file written by us or by the reference (maps CUDA tensors to CPU)."""
    sd = torch.load(path, map_location="cpu", weights_only=True)
    own = model.state_dict()
    if strict:
        missing, extra = set(own) - set(sd), set(sd) - set(own)
        if missing or extra:
            raise KeyError(f"state_dict mismatch: missing={sorted(missing)} unexpected={sorted(extra)}")
    with torch.no_grad():
        for k, v in own.items():
            if k in sd:
                if tuple(sd[k].shape) != tuple(v.shape):
                    raise ValueError(f"shape mismatch for {k}: {tuple(sd[k].shape)} vs {tuple(v.shape)}")
                v.copy_(sd[k].to(v.dtype))
    return model


def save_full(path: str, *, model, optimizer, scheduler=None, epoch: int = 0, step: int = 0,
              extra: Optional[Dict[str, Any]] = None, is_main: bool = True) -> Optional[str]:
    if not is_main:
        return None
    state = {
        "format": "b200dp.full.v1",
        "model": model_state_dict_cpu(model),
        "optimizer": optimizer.state_dict(),
        "scheduler": scheduler.state_dict() if scheduler is not None else None,
        "epoch": epoch, "step": step,
        "rng_cpu": torch.get_rng_state(),
        "rng_cuda": torch.cuda.get_rng_state_all() if torch.cuda.is_available() else None,
        "extra": extra or {},
    }
    tmp = path + ".tmp"
    torch.save(state, tmp)
    os.replace(tmp, path)
    return path


def load_full(path: str, *, model, optimizer, scheduler=None, restore_rng: bool = True) -> Dict[str, Any]:
    state = torch.load(path, map_location="cpu", weights_only=False)
    if state.get("format") != "b200dp.full.v1":
        raise ValueError(f"{path} is not a b200dp full-state checkpoint")
    own = model.state_dict()
    with torch.no_grad():
        for k, v in own.items():
            v.copy_(state["model"][k])
    optimizer.load_state_dict(state["optimizer"])
    if scheduler is not None and state["scheduler"] is not None:
        scheduler.load_state_dict(state["scheduler"])
    if restore_rng:
        torch.set_rng_state(state["rng_cpu"])
        if state["rng_cuda"] is not None and torch.cuda.is_available():
            try:
                torch.cuda.set_rng_state_all(state["rng_cuda"])
            except Exception:
                pass  # different GPU count than at save time
    return {"epoch": state["epoch"], "step": state["step"], "extra": state["extra"]}
