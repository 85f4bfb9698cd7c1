"""Data sharding and the tutorial's toy datasets.

``DistributedSampler`` gives every rank a disjoint, equally sized slice of the index space and
reproduces ``torch.utils.data.DistributedSampler`` index for index (same permutation generator,
same wrap-around padding), which is what the reference builds at
/root/reference/ddp_gpus.py:73-79 and reseeds at ddp_gpus.py:45.
"""
from __future__ import annotations

import math
from typing import Iterator, Optional, Sized

import torch
from torch.utils.data import Dataset, Sampler

from . import group as _group


class DistributedSampler(Sampler[int]):
    """Strided shard of a (optionally shuffled) index list, padded or truncated to equal length."""

    def __init__(self, dataset: Sized, num_replicas: Optional[int] = None, rank: Optional[int] = None,
                 shuffle: bool = True, seed: int = 0, drop_last: bool = False) -> None:
        if num_replicas is None:
            num_replicas = _group.world_size(require_init=True)
        if rank is None:
            rank = _group.rank(require_init=True)
        if not 0 <= rank < num_replicas:
            raise ValueError(f"Invalid rank {rank}, rank should be in the interval [0, {num_replicas - 1}]")
        self.dataset = dataset
        self.num_replicas = int(num_replicas)
        self.rank = int(rank)
        self.shuffle = shuffle
        self.seed = seed
        self.drop_last = drop_last
        self.epoch = 0
        n = len(dataset)
        if drop_last and n % self.num_replicas:
            # keep only whole rounds so that nobody sees duplicated samples
            self.num_samples = math.ceil((n - self.num_replicas) / self.num_replicas)
        else:
            self.num_samples = math.ceil(n / self.num_replicas)
        self.total_size = self.num_samples * self.num_replicas

    def _global_order(self) -> list:
        n = len(self.dataset)
        if not self.shuffle:
            return list(range(n))
        gen = torch.Generator()
        gen.manual_seed(self.seed + self.epoch)
        return torch.randperm(n, generator=gen).tolist()

    def __iter__(self) -> Iterator[int]:
        order = self._global_order()
        if self.drop_last:
            order = order[: self.total_size]
        else:
            missing = self.total_size - len(order)
            if missing > 0:
                reps = math.ceil(missing / max(len(order), 1))
                order = order + (order * reps)[:missing]
        assert len(order) == self.total_size
        mine = order[self.rank : self.total_size : self.num_replicas]
        assert len(mine) == self.num_samples
        return iter(mine)

    def __len__(self) -> int:
        return self.num_samples

    def set_epoch(self, epoch: int) -> None:
        """Reseed the permutation; call once per epoch before iterating (ddp_gpus.py:45)."""
        self.epoch = int(epoch)


class MyTrainDataset(Dataset):
    """``size`` pairs ``(rand(20), rand(1))`` held in memory (reference: ddp_gpus.py:56-67).

    Like the reference, the data is drawn from the process-global RNG at construction, so every
    rank holds different samples unless the caller seeds; pass ``seed`` for reproducible tests.
    """

    def __init__(self, size: int, in_features: int = 20, out_features: int = 1,
                 seed: Optional[int] = None) -> None:
        super().__init__()
        self.size = size
        gen = None
        if seed is not None:
            gen = torch.Generator().manual_seed(seed)
        self.data = [(torch.rand(in_features, generator=gen), torch.rand(out_features, generator=gen))
                     for _ in range(size)]

    def __len__(self) -> int:
        return self.size

    def __getitem__(self, index):
        return self.data[index]


class RandomDataset(Dataset):
    """``length`` rows of ``randn(size)`` (reference: 01.data_parallel.ipynb:118-127)."""

    def __init__(self, size: int, length: int, seed: Optional[int] = None) -> None:
        self.len = length
        gen = torch.Generator().manual_seed(seed) if seed is not None else None
        self.data = torch.randn(length, size, generator=gen)

    def __getitem__(self, index):
        return self.data[index]

    def __len__(self) -> int:
        return self.len


class SyntheticImageDataset(Dataset):
    """ImageNet-shaped synthetic samples ``([3,H,W] fp32, label)`` for the ResNet-50 configs."""

    def __init__(self, length: int, image_size: int = 224, num_classes: int = 1000, seed: int = 0) -> None:
        self.length = length
        self.image_size = image_size
        self.num_classes = num_classes
        self.seed = seed

    def __len__(self) -> int:
        return self.length

    def __getitem__(self, index):
        gen = torch.Generator().manual_seed(self.seed * 1_000_003 + index)
        x = torch.randn(3, self.image_size, self.image_size, generator=gen)
        y = int(torch.randint(0, self.num_classes, (1,), generator=gen))
        return x, y


class DevicePrefetcher:
    """Iterate host batches (tuples of tensors) as device tensors, uploading batch i+1 on a copy stream
    while the consumer's kernels for batch i run.

    Two fixed device slots per position: the upload into a slot waits (on the device, through an
    event) until the work the consumer launched on the previous batch of that slot has finished, and
    the consumer's stream waits for the upload of the slot it is handed.  The consumer must not block
    on the device between two batches, otherwise the next upload is issued too late to overlap
    (``Trainer.run_batches`` reads each loss one step later for that reason).  The reference uploads
    synchronously in the loop body (/root/reference/ddp_gpus.py:48-51); with 154 MB ImageNet-sized
    batches that is ~2.6 ms of idle GPU per step on PCIe Gen5.  On CPU devices it degrades to ``.to``.
    """

    def __init__(self, batches, device) -> None:
        self.batches = batches
        self.device = torch.device(device)
        self._stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None

    def __len__(self) -> int:
        return len(self.batches)

    def __iter__(self):
        dev = self.device
        if dev.type != "cuda":
            for batch in self.batches:
                yield tuple(t.to(dev) if torch.is_tensor(t) else t for t in batch)
            return
        cs = self._stream
        slots = [None, None]
        ready = [torch.cuda.Event(), torch.cuda.Event()]
        done = [None, None]

        def start(batch, s):
            batch = tuple(batch)
            if done[s] is not None:
                cs.wait_event(done[s])
            with torch.cuda.stream(cs):
                bufs = slots[s]
                if bufs is None or len(bufs) != len(batch) or any(
                        b.shape != t.shape or b.dtype != t.dtype for b, t in zip(bufs, batch)):
                    bufs = tuple(torch.empty(t.shape, dtype=t.dtype, device=dev) for t in batch)
                    slots[s] = bufs
                for b, t in zip(bufs, batch):
                    b.copy_(t, non_blocking=True)
                ready[s].record(cs)

        it = iter(self.batches)
        first = next(it, None)
        if first is None:
            return
        k = 0
        start(first, 0)
        while True:
            s = k % 2
            compute = torch.cuda.current_stream(dev)
            compute.wait_event(ready[s])
            for b in slots[s]:
                b.record_stream(compute)
            yield slots[s]
            # the consumer has queued its work for batch k: mark the slot, then overlap the next upload
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(dev))
            done[s] = ev
            nxt = next(it, None)
            if nxt is None:
                return
            k += 1
            start(nxt, k % 2)
