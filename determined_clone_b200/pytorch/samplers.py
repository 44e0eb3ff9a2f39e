"""Sampler building blocks for trials that assemble their own torch DataLoader and still want what
determined.pytorch.DataLoader gives them: the same shuffle on every rank, batches dealt to the ranks round-robin,
an endless stream for training, and a skip when training continues from a checkpoint (reference
harness/determined/pytorch/samplers.py).

Each class wraps another sampler (index sampler or batch sampler -- all of them only iterate it) and has the
`torch.utils.data` base class the reference gives it, so that isinstance checks in user code keep working.
Chain them in this order: shuffle -> repeat -> shard -> skip.  Shuffling after repeating never finishes;
skipping before sharding would skip a different number of global batches on every rank."""
import itertools
from typing import Any, Iterator

import numpy as np
from torch.utils import data as tud


def _dealt(source: Any, players: int, seat: int) -> Iterator[Any]:
    """Every players-th item of `source`, starting with item `seat`."""
    return iter(source) if players == 1 else itertools.islice(iter(source), seat, None, players)


def _share(total: int, players: int, seat: int) -> int:
    """How many of `total` items dealt round-robin land on `seat`."""
    return total // players + (1 if total % players > seat else 0)


def _after(source: Any, skip: int) -> Iterator[Any]:
    it = iter(source)
    for _ in itertools.islice(it, skip):  # stops quietly if the source is shorter than the skip
        pass
    return it


class RepeatSampler(tud.Sampler):
    """Endless: starts the wrapped sampler over whenever it runs out.  len() is one pass."""

    def __init__(self, sampler: Any) -> None:
        self._sampler = sampler

    def __len__(self) -> int:
        return len(self._sampler)

    def __iter__(self) -> Iterator[Any]:
        while True:
            yield from self._sampler


class RepeatBatchSampler(tud.BatchSampler):
    """RepeatSampler for a batch sampler."""

    def __init__(self, batch_sampler: Any) -> None:
        self.batch_sampler = batch_sampler

    def __len__(self) -> int:
        return len(self.batch_sampler)

    def __iter__(self) -> Iterator[Any]:
        while True:
            yield from self.batch_sampler


class DistributedSampler(tud.Sampler):
    """This rank's share of another sampler: items rank, rank + num_workers, ...  Unlike torch's
    DistributedSampler it neither shuffles nor pads, and the wrapped sampler may be endless -- the position keeps
    counting across its passes."""

    def __init__(self, sampler: Any, num_workers: int, rank: int) -> None:
        self._sampler, self._num_workers, self._rank = sampler, num_workers, rank

    def __len__(self) -> int:
        return _share(len(self._sampler), self._num_workers, self._rank)

    def __iter__(self) -> Iterator[Any]:
        return _dealt(self._sampler, self._num_workers, self._rank)


class DistributedBatchSampler(tud.BatchSampler):
    """This rank's share of a batch sampler: global batches rank, rank + num_workers, ...  Over a finite batch
    sampler whose length is not a multiple of num_workers the ranks get different numbers of batches: fine for
    validation, and training always repeats first."""

    def __init__(self, batch_sampler: Any, num_workers: int, rank: int) -> None:
        if rank < 0:
            raise ValueError("rank must be non-negative.")
        if num_workers < 1:
            raise ValueError("num_workers must be greater than zero.")
        if rank >= num_workers:
            raise ValueError("rank must be less than num_workers.")
        self.batch_sampler, self.num_workers, self.rank = batch_sampler, num_workers, rank

    def __len__(self) -> int:
        return _share(len(self.batch_sampler), self.num_workers, self.rank)

    def __iter__(self) -> Iterator[Any]:
        return _dealt(self.batch_sampler, self.num_workers, self.rank)


class SkipSampler(tud.BatchSampler):
    """Drops the first `skip` RECORDS.  Only reproduces a continued run if every batch has the same size (batches
    trained are counted, records are not): prefer SkipBatchSampler.  len() ignores the skip, since the wrapped
    sampler is expected to be endless."""

    def __init__(self, sampler: Any, skip: int) -> None:
        self._sampler, self._skip = sampler, skip

    def __len__(self) -> int:
        return len(self._sampler)

    def __iter__(self) -> Iterator[Any]:
        return _after(self._sampler, self._skip)


class SkipBatchSampler(tud.BatchSampler):
    """Drops the first `skip` batches: where a trial that already trained `skip` batches picks up.  len() ignores
    the skip."""

    def __init__(self, batch_sampler: Any, skip: int) -> None:
        self.batch_sampler, self.skip = batch_sampler, skip
        self.length = len(batch_sampler)

    def __len__(self) -> int:
        return len(self.batch_sampler)

    def __iter__(self) -> Iterator[Any]:
        return _after(self.batch_sampler, self.skip)


class ReproducibleShuffleSampler(tud.Sampler):
    """A new permutation of the wrapped sampler's output on every pass, drawn from numpy's RandomState(seed): the
    same sequence of permutations in every process that uses the same seed."""

    def __init__(self, sampler: Any, seed: int) -> None:
        self._sampler, self._rng = sampler, np.random.RandomState(seed)

    def __len__(self) -> int:
        return len(self._sampler)

    def __iter__(self) -> Iterator[Any]:
        items = list(self._sampler)
        self._rng.shuffle(items)
        return iter(items)


class ReproducibleShuffleBatchSampler(tud.Sampler):
    """The same for whole batches: their order changes from pass to pass, their contents do not -- shuffle records
    with ReproducibleShuffleSampler instead whenever you can."""

    def __init__(self, batch_sampler: Any, seed: int) -> None:
        self._batch_sampler, self._rng = batch_sampler, np.random.RandomState(seed)

    def __len__(self) -> int:
        return len(self._batch_sampler)

    def __iter__(self) -> Iterator[Any]:
        batches = list(self._batch_sampler)
        self._rng.shuffle(batches)
        return iter(batches)
