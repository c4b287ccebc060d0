from .shard import ShardedDataset, make_rank_partition, rank_key, shard_for_rank  # noqa: F401
from .synthetic import (SyntheticImages, PinnedRingLoader, ImageSpec, MNIST_SPEC, CIFAR_SPEC,  # noqa: F401
                        normalise_reference)
from .files import MNISTFiles, TensorImages, find_mnist, load_arrays, materialise, read_idx  # noqa: F401
