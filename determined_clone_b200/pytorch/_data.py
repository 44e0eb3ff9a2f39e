"""Training data loader with Determined's sharding rule: batches are dealt to the ranks round-robin by
batch index, so rank r sees global batches r, r+N, r+2N, ... of a loader that every rank builds with the
same seed (reference harness/determined/pytorch/_data.py:244-347 and samplers.py:43-120).

The input pipeline is outside the grad-sync path; this is the small part of it needed for a rank to see
the same batches it would see under the reference."""
from typing import Any, Iterator, Optional

import torch
from torch.utils import data as tud


class _ShardedBatchSampler(tud.Sampler):
    def __init__(self, batch_sampler: Any, rank: int, size: int, skip: int = 0, repeat: bool = True) -> None:
        self.batch_sampler, self.rank, self.size, self.skip, self.repeat = batch_sampler, rank, size, skip, repeat

    def __len__(self) -> int:
        n = len(self.batch_sampler)
        return (n - self.rank + self.size - 1) // self.size

    def __iter__(self) -> Iterator[Any]:
        skipped = 0
        while True:
            produced = False
            for i, batch in enumerate(self.batch_sampler):
                if i % self.size != self.rank:
                    continue
                produced = True
                if skipped < self.skip:
                    skipped += 1
                    continue
                yield batch
            if not self.repeat or not produced:
                return


class DataLoader:
    """Recipe for a torch DataLoader (same constructor arguments as determined.pytorch.DataLoader for the
    common cases); the controller turns it into a sharded, repeating loader."""

    def __init__(self, dataset: tud.Dataset, batch_size: Optional[int] = 1, shuffle: bool = False, sampler: Any = None,
                 batch_sampler: Any = None, num_workers: int = 0, collate_fn: Any = None, pin_memory: bool = False,
                 drop_last: bool = False, **kwargs: Any) -> None:
        if batch_sampler is None:
            if sampler is None:
                sampler = tud.RandomSampler(dataset) if shuffle else tud.SequentialSampler(dataset)  # type: ignore
            batch_sampler = tud.BatchSampler(sampler, batch_size, drop_last)  # type: ignore
        self.dataset, self.batch_sampler = dataset, batch_sampler
        self.num_workers, self.collate_fn, self.pin_memory, self.kwargs = num_workers, collate_fn, pin_memory, kwargs

    def __len__(self) -> int:
        return len(self.batch_sampler)

    def get_data_loader(self, repeat: bool = False, skip: int = 0, num_replicas: int = 1, rank: int = 0) -> tud.DataLoader:
        sampler = _ShardedBatchSampler(self.batch_sampler, rank, num_replicas, skip=skip, repeat=repeat)
        return tud.DataLoader(self.dataset, batch_sampler=sampler, num_workers=self.num_workers, collate_fn=self.collate_fn,
                              pin_memory=self.pin_memory, **self.kwargs)


def data_length(data: Any) -> int:
    """Number of samples in a (nested) batch: the length of its first tensor-like leaf."""
    if isinstance(data, torch.Tensor):
        return int(data.shape[0])
    if isinstance(data, dict):
        return data_length(next(iter(data.values())))
    if isinstance(data, (list, tuple)):
        return data_length(data[0])
    if hasattr(data, "__len__"):
        return len(data)
    raise TypeError(f"cannot infer the batch length of {type(data).__name__}")
