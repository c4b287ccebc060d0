"""Dataset front end (CPU): IDX reader, any map-style Dataset -> shard -> loader, against the reference's own
SplitDataset + DataLoader (train.py:86-96, splitdataset.py:30-87)."""
import gzip
import os
import struct

import numpy as np
import pytest
import torch
from torch.utils.data import DataLoader, Dataset

from b200dp.data.files import MNISTFiles, TensorImages, find_mnist, load_arrays, materialise, read_idx
from b200dp.data.shard import common_batches, shard_for_rank
from b200dp.data.synthetic import MNIST_SPEC, PinnedRingLoader, SyntheticImages


def _write_idx(path, arr, gz=False):
    codes = {np.dtype(np.uint8): 0x08, np.dtype(">i4"): 0x0C}
    head = struct.pack(">HBB", 0, codes[arr.dtype], arr.ndim) + struct.pack(">" + "I" * arr.ndim, *arr.shape)
    with (gzip.open if gz else open)(path, "wb") as f:
        f.write(head + arr.tobytes())


def _fake_mnist(root, n_train=96, n_test=40, gz=False, layout="torchvision"):
    d = os.path.join(root, "MNIST", "raw") if layout == "torchvision" else root
    os.makedirs(d, exist_ok=True)
    rng = np.random.RandomState(0)
    data = {}
    for split, n, names in (("train", n_train, ("train-images-idx3-ubyte", "train-labels-idx1-ubyte")),
                            ("test", n_test, ("t10k-images-idx3-ubyte", "t10k-labels-idx1-ubyte"))):
        x = rng.randint(0, 256, (n, 28, 28)).astype(np.uint8)
        y = rng.randint(0, 10, (n,)).astype(np.uint8)
        ext = ".gz" if gz else ""
        _write_idx(os.path.join(d, names[0] + ext), x, gz)
        _write_idx(os.path.join(d, names[1] + ext), y, gz)
        data[split] = (x, y)
    return data


@pytest.mark.parametrize("gz", [False, True])
@pytest.mark.parametrize("layout", ["torchvision", "flat"])
def test_mnist_idx_files_round_trip(tmp_path, gz, layout):
    data = _fake_mnist(str(tmp_path), gz=gz, layout=layout)
    assert find_mnist(str(tmp_path), True) is not None
    for train, split in ((True, "train"), (False, "test")):
        ds = MNISTFiles(str(tmp_path), train)
        x, y = data[split]
        assert ds.images.shape == (len(x), 1, 28, 28) and ds.images.dtype == torch.uint8
        assert torch.equal(ds.images[:, 0], torch.from_numpy(x)) and torch.equal(ds.targets, torch.from_numpy(y).long())
        xi, yi = ds[3]                                           # reference transform (train.py:85)
        assert torch.allclose(xi, (torch.from_numpy(x[3]).float().div(255).unsqueeze(0) - 0.1307) / 0.3081)
        assert yi == int(y[3])


def test_read_idx_rejects_garbage(tmp_path):
    p = tmp_path / "bad"
    p.write_bytes(b"\x12\x34\x56\x78rest")
    with pytest.raises(ValueError):
        read_idx(str(p))
    assert find_mnist(str(tmp_path), True) is None
    with pytest.raises(FileNotFoundError):
        MNISTFiles(str(tmp_path), True)


def test_load_arrays_npz_and_pt(tmp_path):
    x = np.random.RandomState(1).randint(0, 256, (10, 28, 28)).astype(np.uint8)
    y = np.arange(10) % 10
    np.savez(tmp_path / "train.npz", images=x, targets=y)
    torch.save({"images": torch.from_numpy(x), "targets": torch.from_numpy(y)}, tmp_path / "d.pt")
    for f in ("train.npz", "d.pt"):
        ds = load_arrays(str(tmp_path / f))
        assert ds.images.shape == (10, 1, 28, 28) and torch.equal(ds.targets, torch.from_numpy(y).long())


class _Opaque(Dataset):
    """A user dataset that exposes nothing but __len__/__getitem__ (e.g. with its own augmentation)."""

    def __init__(self, n):
        g = torch.Generator().manual_seed(7)
        self.x = torch.randn(n, 1, 28, 28, generator=g)
        self.y = torch.randint(0, 10, (n,), generator=g)

    def __len__(self):
        return len(self.y)

    def __getitem__(self, i):
        return self.x[i] * 2 + 1, int(self.y[i])


def test_materialise_paths():
    syn = SyntheticImages(50, MNIST_SPEC, seed=1)
    x, y, fmt = materialise(syn)
    assert fmt == "u8" and x.data_ptr() == syn.images.data_ptr()          # zero-copy for exposed uint8 storage
    op = _Opaque(37)
    x, y, fmt = materialise(op, workers=3)
    assert fmt == "pre" and x.dtype == torch.float32
    assert torch.equal(x, op.x * 2 + 1) and torch.equal(y, op.y)


@pytest.mark.parametrize("world", [2, 3, 4])
@pytest.mark.parametrize("kind", ["exposed", "opaque"])
def test_dataset_through_shard_and_loader_matches_reference_dataloader(ref_modules, world, kind):
    """Same batches, in the same order, as DataLoader(SplitDataset(dataset, {str(i): 1/size}), shuffle=False) of the
    reference (train.py:86-96) — for a dataset with exposed storage and for an opaque one."""
    n, bs = 120, 16
    syn = SyntheticImages(n, MNIST_SPEC, seed=2)
    ds = TensorImages(syn.images, syn.targets) if kind == "exposed" else _Opaque(n)
    for rank in range(world):
        ref = ref_modules["splitdataset"].SplitDataset(ds, {str(i): 1.0 / world for i in range(world)})
        ref.select(str(rank))
        ref_batches = list(DataLoader(ref, batch_size=bs, shuffle=False))
        ours = PinnedRingLoader.from_dataset(shard_for_rank(ds, rank, world), bs, MNIST_SPEC, shuffle=False,
                                             host_normalise=True)
        assert len(ref_batches) == len(ours)
        seen = 0
        for (gx, gy), (rx, ry) in zip(ours, ref_batches):       # the loader yields views of a pinned ring: compare in step
            assert torch.allclose(gx, rx, atol=1e-6) and torch.equal(gy, ry)
            seen += 1
        assert seen == len(ref_batches)


def test_common_batches_caps_uneven_shards():
    ds = SyntheticImages(1025, MNIST_SPEC, seed=0)
    per_rank = [len(shard_for_rank(ds, r, 2)) for r in range(2)]
    assert sorted(per_rank) == [512, 513]                        # ceil(513/512) = 2, ceil(512/512) = 1
    assert common_batches(ds, 2, 512) == 1
    sh = shard_for_rank(ds, per_rank.index(513), 2)
    ld = PinnedRingLoader(ds.images, ds.targets, *sh.bounds(), 512, MNIST_SPEC, fmt="f32", max_batches=1)
    assert len(ld) == 1 and sum(1 for _ in ld) == 1


def test_cli_trains_from_idx_files_on_cpu(tmp_path):
    """`python -m b200dp.cli --data-root DIR` reads real files when they are there (reference: ./data, train.py:88)."""
    from b200dp.cli import parse_args
    from b200dp.engine import load_datasets, run
    from b200dp.parallel import launch
    _fake_mnist(str(tmp_path), n_train=64, n_test=32)
    args = parse_args(["--data-root", str(tmp_path), "--cpu", "--epochs", "1", "--batch-size", "16", "--log-interval", "2",
                       "--world-size", "1"])
    tr, te, src = load_datasets(args, MNIST_SPEC)
    assert len(tr) == 64 and len(te) == 32 and "IDX" in src
    out = {}
    launch.spawn(lambda ctx, a: out.update(run(ctx, a)), args, 1, backend="gloo", use_cuda=False, port=0)
    assert out["n"] == 32 and out["epoch"] == 1
