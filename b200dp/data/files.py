"""On-disk datasets and the bridge from any map-style ``Dataset`` to the loader's host arrays.

The reference feeds ``DataLoader(SplitDataset(datasets.MNIST('./data', ...)))`` (train.py:88-96): torchvision
parses the four MNIST IDX files under ``./data/MNIST/raw`` and every ``__getitem__`` decodes one sample through
PIL + ``ToTensor`` + ``Normalize`` (train.py:85).  Here the per-sample work is done ONCE, up front:

* :func:`read_idx` / :class:`MNISTFiles` parse the IDX files (plain or ``.gz``) straight into the
  ``(N,C,H,W) uint8`` + ``(N,) int64`` tensors that :class:`~b200dp.data.synthetic.PinnedRingLoader` and the
  native step pump stream from (normalisation is folded into the first device kernel);
* :func:`load_arrays` does the same for ``.npy`` / ``.npz`` / ``.pt`` dumps;
* :func:`materialise` turns ANY map-style dataset of ``(image, label)`` pairs into such arrays — zero-copy when the
  dataset exposes its storage (ours: ``.images``; torchvision's: ``.data``), otherwise by calling ``__getitem__``
  for every index on a small thread pool (what the reference's four DataLoader workers do every epoch).

There is no download path (no network); :func:`find_mnist` returns ``None`` when the files are absent and the
caller falls back to synthetic data of the same shape.
"""
from __future__ import annotations

import gzip
import os
import struct
from concurrent.futures import ThreadPoolExecutor
from typing import Optional, Tuple

import numpy as np
import torch
from torch.utils.data import Dataset

from .synthetic import MNIST_SPEC, ImageSpec

_IDX_DTYPES = {0x08: np.uint8, 0x09: np.int8, 0x0B: ">i2", 0x0C: ">i4", 0x0D: ">f4", 0x0E: ">f8"}
MNIST_FILES = {True: ("train-images-idx3-ubyte", "train-labels-idx1-ubyte"),
               False: ("t10k-images-idx3-ubyte", "t10k-labels-idx1-ubyte")}


def read_idx(path: str) -> np.ndarray:
    """Parse one IDX file (the MNIST container format: magic ``0x0000TTDD``, DD big-endian dimension sizes, data).
    ``path`` may be gzip-compressed (``.gz``)."""
    opener = gzip.open if path.endswith(".gz") else open
    with opener(path, "rb") as f:
        raw = f.read()
    if len(raw) < 4:
        raise ValueError(f"{path}: too short for an IDX header")
    zero, tcode, ndim = struct.unpack(">HBB", raw[:4])
    if zero != 0 or tcode not in _IDX_DTYPES:
        raise ValueError(f"{path}: not an IDX file (magic {raw[:4].hex()})")
    dims = struct.unpack(">" + "I" * ndim, raw[4:4 + 4 * ndim])
    dt = np.dtype(_IDX_DTYPES[tcode])
    n = int(np.prod(dims)) if dims else 1
    body = raw[4 + 4 * ndim:]
    if len(body) < n * dt.itemsize:
        raise ValueError(f"{path}: {len(body)} data bytes, header promises {n * dt.itemsize}")
    return np.frombuffer(body, dtype=dt, count=n).reshape(dims)


def _first_existing(cands):
    for c in cands:
        if os.path.isfile(c):
            return c
    return None


def find_mnist(root: str, train: bool) -> Optional[Tuple[str, str]]:
    """Paths of the (images, labels) IDX files of a split under ``root``: torchvision's layout
    (``root/MNIST/raw/``) or the files directly in ``root``; plain or ``.gz``."""
    out = []
    for name in MNIST_FILES[train]:
        cands = [os.path.join(d, name + ext) for d in (os.path.join(root, "MNIST", "raw"), root) for ext in ("", ".gz")]
        p = _first_existing(cands)
        if p is None:
            return None
        out.append(p)
    return out[0], out[1]


class TensorImages(Dataset):
    """Map-style dataset over in-memory ``(N,C,H,W) uint8`` images and ``(N,) int64`` labels; ``__getitem__`` applies
    the reference transform (train.py:85) and returns ``(float32 CHW, int)`` like the torchvision dataset."""

    def __init__(self, images: torch.Tensor, targets: torch.Tensor, spec: ImageSpec = MNIST_SPEC):
        if images.dtype != torch.uint8 or images.dim() != 4:
            raise TypeError("images must be (N,C,H,W) uint8")
        if targets.shape != (images.shape[0],):
            raise ValueError("one label per image")
        self.images = images.contiguous()
        self.targets = targets.to(torch.int64).contiguous()
        self.spec = spec

    def __len__(self) -> int:
        return self.images.shape[0]

    def __getitem__(self, idx: int):
        x = self.images[idx].to(torch.float32).div(255)
        return (x - self.spec.mean) / self.spec.std, int(self.targets[idx])


class MNISTFiles(TensorImages):
    """The MNIST IDX files of ``root`` (see :func:`find_mnist`) as a dataset: what ``datasets.MNIST(root, train)``
    yields in the reference (train.py:88, 91), without torchvision, PIL or a download."""

    def __init__(self, root: str, train: bool = True, spec: ImageSpec = MNIST_SPEC):
        found = find_mnist(root, train)
        if found is None:
            raise FileNotFoundError(f"no MNIST IDX files for the {'train' if train else 'test'} split under {root!r}")
        imgs, labels = read_idx(found[0]), read_idx(found[1])
        if imgs.ndim != 3 or imgs.shape[0] != labels.shape[0]:
            raise ValueError(f"{found[0]} / {found[1]}: inconsistent shapes {imgs.shape} / {labels.shape}")
        super().__init__(torch.from_numpy(np.array(imgs, copy=True)).unsqueeze(1),
                         torch.from_numpy(labels.astype(np.int64)), spec)
        self.paths = found


def load_arrays(path: str, spec: ImageSpec = MNIST_SPEC) -> TensorImages:
    """A dataset dump: ``.npz`` with ``images``/``targets`` (or ``x``/``y``), ``.pt`` holding a dict with the same keys
    or an ``(images, targets)`` pair.  Images ``(N,H,W)``, ``(N,C,H,W)`` or ``(N,H,W,C)`` uint8."""
    if path.endswith(".npz"):
        z = np.load(path)
        x = z["images"] if "images" in z else z["x"]
        y = z["targets"] if "targets" in z else z["y"]
        x, y = torch.from_numpy(x), torch.from_numpy(y)
    else:
        obj = torch.load(path, map_location="cpu")
        x, y = (obj["images"], obj["targets"]) if isinstance(obj, dict) else obj
    if x.dim() == 3:
        x = x.unsqueeze(1)
    elif x.dim() == 4 and x.shape[-1] == spec.channels and x.shape[1] != spec.channels:
        x = x.permute(0, 3, 1, 2)
    return TensorImages(x.contiguous(), y, spec)


def materialise(dataset, spec: ImageSpec = MNIST_SPEC, *, workers: int = 4) -> Tuple[torch.Tensor, torch.Tensor, str]:
    """``(images, targets, fmt)`` host arrays of a whole map-style dataset.

    ``fmt == "u8"``: raw ``(N,C,H,W) uint8`` (the device normalises) — when the dataset exposes uint8 storage.
    ``fmt == "pre"``: ``(N,C,H,W) float32`` exactly as ``__getitem__`` returns them (already transformed) — any other
    dataset, e.g. one with augmentation or a custom transform; every index is fetched once, on ``workers`` threads."""
    imgs = getattr(dataset, "images", None)
    if imgs is None:
        imgs = getattr(dataset, "data", None)               # torchvision MNIST / CIFAR10
    tg = getattr(dataset, "targets", None)
    plain = getattr(dataset, "transform", None) is None or isinstance(dataset, TensorImages)
    if imgs is not None and tg is not None and len(imgs) == len(dataset):
        x = torch.as_tensor(np.asarray(imgs)) if not torch.is_tensor(imgs) else imgs
        if x.dtype == torch.uint8 and (plain or _is_reference_transform(dataset, spec)):
            if x.dim() == 3:
                x = x.unsqueeze(1)
            elif x.dim() == 4 and x.shape[-1] == spec.channels and x.shape[1] != spec.channels:
                x = x.permute(0, 3, 1, 2)
            return x.contiguous(), torch.as_tensor(tg).to(torch.int64).contiguous(), "u8"
    n = len(dataset)
    x0, _ = dataset[0]
    x0 = torch.as_tensor(x0, dtype=torch.float32)
    out_x = torch.empty((n,) + tuple(x0.shape), dtype=torch.float32)
    out_y = torch.empty(n, dtype=torch.int64)

    def fetch(lo_hi):
        for i in range(*lo_hi):
            x, y = dataset[i]
            out_x[i] = torch.as_tensor(x, dtype=torch.float32)
            out_y[i] = int(y)

    step = max(1, -(-n // max(workers, 1)))
    with ThreadPoolExecutor(max_workers=max(workers, 1)) as ex:
        list(ex.map(fetch, [(lo, min(lo + step, n)) for lo in range(0, n, step)]))
    return out_x, out_y, "pre"


def _is_reference_transform(dataset, spec: ImageSpec) -> bool:
    """True when ``dataset.transform`` is torchvision's ``Compose([ToTensor(), Normalize((mean,), (std,))])`` with the
    spec's constants (train.py:85): then the raw bytes + on-device normalisation give the same numbers."""
    t = getattr(dataset, "transform", None)
    steps = getattr(t, "transforms", None)
    if not steps or len(steps) != 2 or type(steps[0]).__name__ != "ToTensor" or type(steps[1]).__name__ != "Normalize":
        return False
    mean, std = np.atleast_1d(steps[1].mean), np.atleast_1d(steps[1].std)
    return bool(np.allclose(mean, spec.mean) and np.allclose(std, spec.std))
