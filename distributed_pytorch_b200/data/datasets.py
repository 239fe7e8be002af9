"""CIFAR-10-shaped datasets and the reference's transforms.

The reference trains on ``torchvision.datasets.CIFAR10(root="./data", download=True)`` with
``RandomCrop(32, padding=4) -> RandomHorizontalFlip -> ToTensor -> Normalize(mean, std)`` for the training split and
``ToTensor -> Normalize`` for the test split (reference main.py:71-84, 91-92).  There is no network here, so
:func:`cifar10` reads the standard ``cifar-10-batches-py`` pickles when they are already under ``root`` and otherwise
falls back to :func:`synthetic_cifar10`: uint8 images of the same shape whose classes are separable (a per-class
colour/frequency pattern plus noise), so a short run shows a falling loss like the real data does.

Images are kept as uint8 ``(N, 3, 32, 32)`` -- 150 MB for the 50 000-image training split -- on the host and, for the
device path of :class:`~distributed_pytorch_b200.data.loader.BatchLoader`, on the GPU.

Augmentation draws are *counter based*: the crop offset and the flip bit of a sample are a hash of
``(seed, epoch, dataset index)`` (:func:`augment_draws`).  The host transform below and the sm_100a kernel
(``csrc/augment.cuh``) evaluate the same hash, so both paths produce bit-identical batches and a run is reproducible
whatever the batch composition or the number of loader threads.
"""
from __future__ import annotations

import os
import pickle
from typing import Optional, Tuple

import numpy as np
import torch

# reference main.py:72-73: mean/std given on the 0..255 scale, divided by 255
CIFAR_MEAN = tuple(x / 255.0 for x in (125.3, 123.0, 113.9))
CIFAR_STD = tuple(x / 255.0 for x in (63.0, 62.1, 66.7))
PAD = 4  # RandomCrop(32, padding=4)


def augment_draws(indices, seed: int, epoch: int) -> np.ndarray:
    """``(len(indices), 3)`` int32: crop offsets ``dx, dy`` in ``[0, 2*PAD]`` and the flip bit of every sample.

    32-bit mix (murmur3 finaliser) of the dataset index, the seed and the epoch; ``csrc/augment.cuh`` holds the same
    arithmetic for the device path."""
    idx = np.asarray(indices, dtype=np.uint32)
    with np.errstate(over="ignore"):
        h = (idx * np.uint32(0x9E3779B1) + np.uint32(seed & 0xFFFFFFFF) * np.uint32(0x85EBCA77)
             + np.uint32(epoch & 0xFFFFFFFF) * np.uint32(0xC2B2AE3D))
        h ^= h >> np.uint32(16)
        h *= np.uint32(0x85EBCA6B)
        h ^= h >> np.uint32(13)
        h *= np.uint32(0xC2B2AE35)
        h ^= h >> np.uint32(16)
    out = np.empty((idx.shape[0], 3), dtype=np.int32)
    out[:, 0] = (h & np.uint32(0xFF)) % np.uint32(2 * PAD + 1)
    out[:, 1] = ((h >> np.uint32(8)) & np.uint32(0xFF)) % np.uint32(2 * PAD + 1)
    out[:, 2] = (h >> np.uint32(16)) & np.uint32(1)
    return out


class ImageDataset:
    """uint8 images ``(N, 3, H, W)`` + int64 labels, with the reference's train / test transform."""

    def __init__(self, images: torch.Tensor, labels: torch.Tensor, train: bool, seed: int = 0, source: str = "array"):
        if images.dtype != torch.uint8 or images.dim() != 4 or images.shape[1] != 3:
            raise ValueError("images must be uint8 (N, 3, H, W)")
        self.images = images.contiguous()
        self.labels = labels.to(torch.int64).contiguous()
        self.train = bool(train)
        self.seed = int(seed)
        self.epoch = 0
        self.source = source
        self._mean = torch.tensor(CIFAR_MEAN, dtype=torch.float32).view(1, 3, 1, 1)
        self._inv_std = (1.0 / torch.tensor(CIFAR_STD, dtype=torch.float32)).view(1, 3, 1, 1)
        self._device_images: Optional[torch.Tensor] = None
        self._device_labels: Optional[torch.Tensor] = None

    def __len__(self) -> int:
        return self.images.shape[0]

    def set_epoch(self, epoch: int) -> None:
        self.epoch = int(epoch)

    # ------------------------------------------------------------------------------------------------ host path
    def get_batch(self, indices, draws: Optional[np.ndarray] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        """Transformed fp32 NCHW batch + labels for these dataset indices (the oracle of the device kernel)."""
        idx = torch.as_tensor(np.asarray(indices, dtype=np.int64))
        img = self.images[idx]  # (B, 3, H, W) uint8
        b, _, h, w = img.shape
        x = img.to(torch.float32).div_(255.0)
        if self.train:
            if draws is None:
                draws = augment_draws(np.asarray(indices), self.seed, self.epoch)
            d = torch.as_tensor(draws, dtype=torch.int64)
            xp = torch.zeros((b, 3, h + 2 * PAD, w + 2 * PAD), dtype=torch.float32)  # zero (black) padding, pre-normalise
            xp[:, :, PAD:PAD + h, PAD:PAD + w] = x
            rows = torch.arange(h).view(1, h) + d[:, 1:2]                                 # (B, H)
            cols = torch.arange(w).view(1, w)
            cols = torch.where(d[:, 2:3] > 0, (w - 1) - cols, cols) + d[:, 0:1]           # flip after the crop
            x = xp[torch.arange(b).view(b, 1, 1, 1), torch.arange(3).view(1, 3, 1, 1), rows.view(b, 1, h, 1),
                   cols.view(b, 1, 1, w)]
        x = (x - self._mean) * self._inv_std
        return x.contiguous(), self.labels[idx]

    # ------------------------------------------------------------------------------------------------ device path
    def to_device(self, device: torch.device):
        """uint8 images and labels resident on ``device`` (once)."""
        if self._device_images is None or self._device_images.device != device:
            self._device_images = self.images.to(device)
            self._device_labels = self.labels.to(device)
        return self._device_images, self._device_labels


def synthetic_cifar10(train: bool, size: Optional[int] = None, seed: int = 0) -> ImageDataset:
    """Deterministic CIFAR-10-shaped data: 50 000 / 10 000 uint8 images, 10 balanced classes.  Each class has its own
    mean colour and a low-frequency pattern; noise on top keeps the task from being trivial."""
    n = size if size is not None else (50000 if train else 10000)
    g = torch.Generator().manual_seed(1000 + seed + (0 if train else 1))
    labels = torch.arange(n, dtype=torch.int64) % 10
    labels = labels[torch.randperm(n, generator=g)]
    pg = torch.Generator().manual_seed(77)  # class prototypes: shared by the train and the test split
    colour = torch.rand(10, 3, 1, 1, generator=pg) * 0.6 + 0.2
    yy, xx = torch.meshgrid(torch.linspace(0, 1, 32), torch.linspace(0, 1, 32), indexing="ij")
    fx = torch.randint(1, 4, (10,), generator=pg).float().view(10, 1, 1)
    fy = torch.randint(1, 4, (10,), generator=pg).float().view(10, 1, 1)
    pattern = 0.15 * torch.sin(2 * np.pi * (fx * xx + fy * yy))          # (10, 32, 32)
    proto = (colour + pattern.unsqueeze(1)).clamp_(0, 1)                 # (10, 3, 32, 32)
    images = torch.empty((n, 3, 32, 32), dtype=torch.uint8)
    chunk = 8192
    for lo in range(0, n, chunk):
        hi = min(lo + chunk, n)
        noise = torch.randn(hi - lo, 3, 32, 32, generator=g) * 0.12
        images[lo:hi] = ((proto[labels[lo:hi]] + noise).clamp_(0, 1) * 255.0).round_().to(torch.uint8)
    return ImageDataset(images, labels, train, seed=seed, source="synthetic")


def _read_cifar_pickles(folder: str, train: bool) -> Tuple[torch.Tensor, torch.Tensor]:
    names = [f"data_batch_{i}" for i in range(1, 6)] if train else ["test_batch"]
    xs, ys = [], []
    for name in names:
        with open(os.path.join(folder, name), "rb") as fh:
            entry = pickle.load(fh, encoding="latin1")
        xs.append(np.asarray(entry["data"], dtype=np.uint8).reshape(-1, 3, 32, 32))
        ys.extend(entry["labels"] if "labels" in entry else entry["fine_labels"])
    return torch.from_numpy(np.concatenate(xs, 0)), torch.tensor(ys, dtype=torch.int64)


def cifar10(root: str = "./data", train: bool = True, synthetic: Optional[bool] = None, size: Optional[int] = None,
            seed: int = 0) -> ImageDataset:
    """The reference's ``datasets.CIFAR10(root, train=...)`` (main.py:83-84, 91-92) without the download: real files if
    ``root/cifar-10-batches-py`` exists (and ``synthetic`` is not forced), synthetic data of the same shape otherwise."""
    folder = os.path.join(root, "cifar-10-batches-py")
    if synthetic is not True and os.path.isdir(folder):
        images, labels = _read_cifar_pickles(folder, train)
        if size is not None:
            images, labels = images[:size], labels[:size]
        return ImageDataset(images, labels, train, seed=seed, source="cifar10")
    if synthetic is False:
        raise FileNotFoundError(f"{folder} not found and synthetic data was disabled (no network to download CIFAR-10)")
    return synthetic_cifar10(train, size, seed)
