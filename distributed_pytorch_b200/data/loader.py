"""Sampler and batch loader.

:class:`ShardedSampler` reproduces ``torch.utils.data.DistributedSampler`` index for index (the reference shards its
training set with ``DistributedSampler(training_set, num_replicas=N, rank=r)`` and keeps the defaults shuffle=True,
seed=0, drop_last=False; main_gather.py:123): one global ``randperm`` seeded with ``seed + epoch``, wrap-around padding
to a multiple of the world size, rank ``r`` takes ``indices[r::N]``.

:class:`BatchLoader` replaces ``DataLoader(num_workers=2, batch_size=256, pin_memory=True)`` (main.py:85-90) with two
paths:

* **device** (``device=cuda``): the uint8 dataset lives on the GPU; a batch is ONE launch of the augmentation kernel
  (gather by index + RandomCrop(pad 4) + flip + Normalize -> fp32 NCHW) on the current stream.  No per-batch host work
  besides slicing an index tensor that was uploaded once per epoch, so a ~1 ms training step is not loader-bound.
* **host**: the same transform with torch ops on a background thread, results in pinned memory, ``prefetch`` batches
  ahead.  This is the oracle of the device kernel and the path for CPU runs.
"""
from __future__ import annotations

import math
import queue
import threading
from typing import Iterator, List, Optional

import numpy as np
import torch

from .datasets import ImageDataset, augment_draws


class ShardedSampler:
    def __init__(self, num_examples: int, num_replicas: int, rank: int, shuffle: bool = True, seed: int = 0,
                 drop_last: bool = False):
        if rank < 0 or rank >= num_replicas:
            raise ValueError(f"rank {rank} outside [0, {num_replicas})")
        self.n, self.world, self.rank = int(num_examples), int(num_replicas), int(rank)
        self.shuffle, self.seed, self.drop_last = shuffle, int(seed), drop_last
        self.epoch = 0
        if drop_last and self.n % self.world != 0:
            self.num_samples = math.ceil((self.n - self.world) / self.world)
        else:
            self.num_samples = math.ceil(self.n / self.world)
        self.total_size = self.num_samples * self.world

    def set_epoch(self, epoch: int) -> None:
        self.epoch = int(epoch)

    def indices(self) -> List[int]:
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            idx = torch.randperm(self.n, generator=g).tolist()
        else:
            idx = list(range(self.n))
        if not self.drop_last:
            pad = self.total_size - len(idx)
            if pad <= len(idx):
                idx += idx[:pad]
            else:
                idx += (idx * math.ceil(pad / len(idx)))[:pad]
        else:
            idx = idx[:self.total_size]
        return idx[self.rank:self.total_size:self.world]

    def __iter__(self) -> Iterator[int]:
        return iter(self.indices())

    def __len__(self) -> int:
        return self.num_samples


class BatchLoader:
    def __init__(self, dataset: ImageDataset, batch_size: int, sampler: Optional[ShardedSampler] = None,
                 shuffle: bool = False, pin_memory: bool = True, prefetch: int = 2, max_batches: Optional[int] = None,
                 device: Optional[torch.device] = None, seed: int = 0):
        self.dataset, self.batch_size, self.sampler, self.shuffle = dataset, int(batch_size), sampler, shuffle
        self.pin_memory = bool(pin_memory) and torch.cuda.is_available()
        self.prefetch, self.max_batches = int(prefetch), max_batches
        self.device = device if (device is not None and device.type == "cuda") else None
        self._gen = torch.Generator().manual_seed(seed)
        self._epochs_served = 0

    # ------------------------------------------------------------------------------------------------ bookkeeping
    @property
    def num_examples(self) -> int:
        """Examples one pass yields (what the reference prints as ``len(test_loader.dataset)``, main.py:64-66)."""
        n = len(self.sampler) if self.sampler is not None else len(self.dataset)
        if self.max_batches is not None:
            n = min(n, self.max_batches * self.batch_size)
        return n

    def __len__(self) -> int:
        n = len(self.sampler) if self.sampler is not None else len(self.dataset)
        b = math.ceil(n / self.batch_size)
        return b if self.max_batches is None else min(b, self.max_batches)

    def _epoch_indices(self) -> np.ndarray:
        if self.sampler is not None:
            idx = self.sampler.indices()
        elif self.shuffle:
            idx = torch.randperm(len(self.dataset), generator=self._gen).tolist()
        else:
            idx = range(len(self.dataset))
        return np.asarray(idx, dtype=np.int64)

    # ------------------------------------------------------------------------------------------------ iteration
    def __iter__(self):
        idx = self._epoch_indices()
        epoch = self.sampler.epoch if self.sampler is not None else self._epochs_served
        self._epochs_served += 1
        self.dataset.set_epoch(epoch)
        nb = len(self)
        if self.device is not None:
            return self._iter_device(idx, nb)
        if self.prefetch <= 0:
            return self._iter_sync(idx, nb)
        return self._iter_prefetch(idx, nb)

    def _host_batch(self, idx: np.ndarray, i: int):
        sl = idx[i * self.batch_size:(i + 1) * self.batch_size]
        x, y = self.dataset.get_batch(sl)
        if self.pin_memory:
            x, y = x.pin_memory(), y.pin_memory()
        return x, y

    def _iter_sync(self, idx, nb):
        for i in range(nb):
            yield self._host_batch(idx, i)

    def _iter_prefetch(self, idx, nb):
        q: "queue.Queue" = queue.Queue(maxsize=self.prefetch)
        stop = threading.Event()

        def work():
            try:
                for i in range(nb):
                    if stop.is_set():
                        return
                    q.put(self._host_batch(idx, i))
                q.put(None)
            except BaseException as exc:  # noqa: BLE001 - handed to the consumer
                q.put(exc)

        t = threading.Thread(target=work, daemon=True)
        t.start()
        try:
            while True:
                item = q.get()
                if item is None:
                    return
                if isinstance(item, BaseException):
                    raise item
                yield item
        finally:
            stop.set()
            while t.is_alive():  # unblock a producer waiting on a full queue
                try:
                    q.get_nowait()
                except queue.Empty:
                    t.join(timeout=0.01)

    def _iter_device(self, idx, nb):
        from ..ops import kernels as K
        images, labels = self.dataset.to_device(self.device)
        idx_dev = torch.from_numpy(idx.astype(np.int32)).to(self.device)
        for i in range(nb):
            sl = idx_dev[i * self.batch_size:(i + 1) * self.batch_size]
            yield K.augment_batch(images, labels, sl, self.dataset.train, self.dataset.seed, self.dataset.epoch)


def host_draws(dataset: ImageDataset, indices) -> np.ndarray:
    """Crop / flip draws the loader would use for these indices in the dataset's current epoch."""
    return augment_draws(indices, dataset.seed, dataset.epoch)
