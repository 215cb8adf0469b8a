"""Resumable sampler + infinite loader (reference: core/data/dataloader.py:1-162).

The sampler is a pure function of (seed, consumed_samples, dp layout): epoch = consumed // usable, the dataset is
re-seeded with ``seed + epoch`` and each data-parallel rank walks ``arange(n/dp)*dp + dp_rank`` from the in-epoch
offset — restarting from a checkpoint reproduces the exact sample order.
"""
from __future__ import annotations

from typing import Any, Iterator, Optional

import torch

from ..topology import Topology
from .base import BaseDataset


class RandomSampler:
    def __init__(self, dataset: BaseDataset, seed: int, consumed_samples: int, topology: Topology, shuffle: bool = True):
        self.dataset = dataset
        self.seed = seed
        self.consumed_samples = consumed_samples
        self.topology = topology
        self.shuffle = shuffle
        c = topology.config
        self.dp = c.data_parallel_size
        self.dp_rank = topology.data_parallel_rank
        self.micro_batch_size = c.micro_batch_size
        self.total_samples = len(dataset)
        per_step = self.micro_batch_size * self.dp
        self.usable_total_samples = self.total_samples - self.total_samples % per_step
        assert self.usable_total_samples > 0, (
            f"not enough samples ({self.total_samples}) for micro_batch_size {self.micro_batch_size} x dp {self.dp}"
        )

    def __len__(self) -> int:
        return self.total_samples

    def seed_current_epoch(self) -> None:
        """Re-seed the dataset for the epoch `consumed_samples` lies in.  Called before a torch iterator is built:
        worker processes copy the dataset when the iterator is created, i.e. before the sampler's first index."""
        self.dataset.set_seed(seed=self.seed + self.consumed_samples // self.usable_total_samples, shuffle=self.shuffle)

    def __iter__(self) -> Iterator[int]:
        """ONE epoch (the rest of the current one after a resume).  The iterator ends at the epoch boundary so that
        the torch DataLoader — and its worker processes, which hold their own copy of the dataset — is rebuilt with
        the dataset re-seeded for the next epoch (`DataLoader.__next__` below)."""
        epoch = self.consumed_samples // self.usable_total_samples
        in_epoch = self.consumed_samples % self.usable_total_samples
        self.dataset.set_seed(seed=self.seed + epoch, shuffle=self.shuffle)
        start = in_epoch // self.dp
        n_per_rank = self.usable_total_samples // self.dp
        for i in range(start, n_per_rank):
            self.consumed_samples += self.dp
            yield i * self.dp + self.dp_rank


class DataLoader:
    """Infinite iterator of collated micro batches for this data-parallel rank."""

    def __init__(self, seed: int, consumed_samples: int, dataset: BaseDataset, topology: Topology, shuffle: bool = True,
                 num_workers: int = 0, pin_memory: bool = True, prefetch_factor: Optional[int] = None):
        self.dataset = dataset
        self.sampler = RandomSampler(dataset, seed, consumed_samples, topology, shuffle)
        kwargs: dict[str, Any] = {}
        if num_workers > 0 and prefetch_factor is not None:
            kwargs["prefetch_factor"] = prefetch_factor
        self.loader = torch.utils.data.DataLoader(
            dataset, batch_size=topology.config.micro_batch_size, sampler=self.sampler, num_workers=num_workers,
            pin_memory=pin_memory and torch.cuda.is_available(), collate_fn=dataset.collate, drop_last=True, **kwargs,
        )
        self._it = self._fresh_iterator()

    def _fresh_iterator(self) -> Iterator[Any]:
        """New torch iterator for the sampler's current epoch.  Torch draws the workers' base seed from the default CPU
        generator when an iterator is made; a resumed run makes its iterator at a different point of the random stream
        than the run it continues (after the checkpoint's RNG state is restored, or not at an epoch boundary), so the
        draw is taken back: everything that samples from the default generator afterwards (CPU dropout) stays
        bit-identical across a resume."""
        self.sampler.seed_current_epoch()
        state = torch.get_rng_state()
        it = iter(self.loader)
        torch.set_rng_state(state)
        return it

    def __iter__(self) -> "DataLoader":
        return self

    def __next__(self) -> Any:
        try:
            return next(self._it)
        except StopIteration:
            # epoch boundary: a fresh torch iterator forks fresh workers AFTER the sampler re-seeded the dataset
            self._it = self._fresh_iterator()
            return next(self._it)
