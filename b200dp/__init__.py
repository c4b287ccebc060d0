"""b200dp — a Blackwell-native (sm_100a) data-parallel CNN training framework.

Capabilities mirror ownzonefeng/pytorch-distributed-training-example
(reference `train.py`, `cnn.py`, `splitdataset.py`), re-designed for B200:
flat parameter / gradient arenas, hand-written CUDA kernels for the LeNet
forward / backward, and a peer-to-peer all-reduce over NVLink that is fused
with the gradient average and the optimizer update.

Sub-packages
    data      deterministic contiguous sharder + synthetic pinned-ring loader
    models    LeNet (reference-compatible state_dict) and its fused native runner
    ops       ctypes bindings to the in-tree CUDA library (csrc/, include/)
    optim     flat fused optimizers (Adadelta / SGD-momentum / Adam) + StepLR
    parallel  process-group bootstrap, symmetric arena, DataParallel wrap
    utils     logging, clocks sampling
"""

__version__ = "0.1.0"

from .data.shard import ShardedDataset, make_rank_partition  # noqa: F401
from .models.lenet import LeNet, LeNetConfig  # noqa: F401
