from .shard import ShardedDataset, make_rank_partition, rank_key, shard_for_rank  # noqa: F401
from .synthetic import (SyntheticImages, PinnedRingLoader, ImageSpec, MNIST_SPEC, CIFAR_SPEC,  # noqa: F401
                        normalise_reference)
