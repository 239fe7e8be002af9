"""Input pipeline: CIFAR-10-shaped datasets (real files or synthetic), DistributedSampler-exact sharding, and a batch
loader with a host path (pinned, prefetched) and a device path (dataset resident on the GPU, one augmentation kernel
per batch).  Stands in for torchvision's CIFAR10 + transforms + DataLoader + DistributedSampler as the reference wires
them (reference main.py:71-98, main_gather.py:123-128)."""
from .datasets import CIFAR_MEAN, CIFAR_STD, ImageDataset, augment_draws, cifar10, synthetic_cifar10  # noqa: F401
from .loader import BatchLoader, ShardedSampler  # noqa: F401
