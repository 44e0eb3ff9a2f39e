"""Input side of the path: determined.pytorch.DataLoader and the helpers around a batch (reference
harness/determined/pytorch/_data.py: DataLoader :58-286, adapt_batch_sampler :288-313, data_length :316-347,
to_device :350-383).

Determined's sharding rule: every rank builds the same batch sequence (same seed), the sequence is repeated
endlessly, and global batch i goes to rank i % N -- counted along the endless sequence, not per epoch, so with an
epoch length that N does not divide the ranks swap batches from one epoch to the next.  A continued run skips the
batches this rank has already trained."""
import logging
from typing import Any, Iterator, Optional, Set, Type

import numpy as np
import torch
from torch.utils import data as tud

from determined_clone_b200.pytorch import samplers

logger = logging.getLogger("determined.pytorch")

_ADVANCED_GUIDE = "https://docs.determined.ai/latest/training-apis/api-pytorch-advanced.html"


def _dataset_repro_warning(fn: str, data_obj: Any, is_deepspeed_trial: bool = False) -> str:
    switch = "context.disable_dataset_reproducibility_checks()" if is_deepspeed_trial else "context.experimental.disable_dataset_reproducibility_checks()"
    return (
        f"{fn}() returned an instance of {type(data_obj).__name__}, which is not a "
        "subclass of det.pytorch.DataLoader.  For most non-Iterable DataSets, "
        "det.pytorch.DataLoader is a drop-in replacement for torch.utils.data.DataLoader "
        "but which offers easy and transparent reproducibility in Determined experiments. "
        "It is highly recommended that you use det.pytorch.DataLoader if possible.  If "
        f"not, you can disable this check by calling {switch} at some point "
        "in your trial's __init__() method."
    )


def adapt_batch_sampler(batch_sampler: Any, repeat: bool = False, skip: int = 0, num_replicas: int = 1, rank: int = 0) -> Any:
    """repeat -> shard -> skip around a batch sampler.  The skip comes last because it counts batches this rank
    trained, i.e. what the sharding stage emits."""
    if repeat:
        batch_sampler = samplers.RepeatBatchSampler(batch_sampler)
    if num_replicas > 1:
        batch_sampler = samplers.DistributedBatchSampler(batch_sampler, num_replicas, rank)
    if skip > 0:
        batch_sampler = samplers.SkipBatchSampler(batch_sampler, skip)
    return batch_sampler


class DataLoader:
    """What a trial returns from build_*_data_loader: the arguments of a torch.utils.data.DataLoader (same names,
    same defaults, same argument checks), kept until the training loop asks for the real loader with
    get_data_loader(repeat, skip, num_replicas, rank)."""

    def __init__(
        self,
        dataset: tud.Dataset,
        batch_size: Optional[int] = 1,
        shuffle: bool = False,
        sampler: Any = None,
        batch_sampler: Any = None,
        num_workers: int = 0,
        collate_fn: Any = None,
        pin_memory: bool = False,
        drop_last: bool = False,
        timeout: float = 0,
        worker_init_fn: Any = None,
        multiprocessing_context: Any = None,
        generator: Any = None,
        *,
        prefetch_factor: Optional[int] = None,
        persistent_workers: bool = False,
    ) -> None:
        if isinstance(dataset, tud.IterableDataset):
            raise ValueError(
                "IterableDatasets are not supported through det.pytorch.DataLoader().  Read about "
                "the difference between IterableDatasets and MapDatasets at: "
                "pytorch.org/docs/stable/data. "
                "You can use an IterableDataset in Determined, but you will be responsible for "
                "shuffling, sharding, repeating, and skipping to ensure reproducibility and "
                f"correctness of training.  See the in-depth guide at: {_ADVANCED_GUIDE}"
            )
        if batch_sampler is None and batch_size is None:
            raise ValueError(
                "The case of batch_sampler=None and batch_size=None is not supported through "
                "det.pytorch.DataLoader(). For customizing your data loader beyond what is "
                f"supported by det.pytorch.DataLoader, see the in-depth guide at: {_ADVANCED_GUIDE}"
            )
        # the argument checks of torch.utils.data.DataLoader.__init__, made now rather than when the loop starts
        if num_workers < 0:
            raise ValueError("num_workers option should be non-negative; use num_workers=0 to disable multiprocessing.")
        if timeout < 0:
            raise ValueError("timeout option should be non-negative")
        if num_workers == 0 and prefetch_factor is not None:
            raise ValueError(
                "prefetch_factor option could only be specified in multiprocessing. "
                "let num_workers > 0 to enable multiprocessing."
            )
        if prefetch_factor is not None and prefetch_factor < 0:
            raise ValueError("prefetch_factor option should be non-negative")
        if persistent_workers and num_workers == 0:
            raise ValueError("persistent_workers option needs num_workers > 0")
        if sampler is not None and shuffle:
            raise ValueError("sampler option is mutually exclusive with shuffle")
        if batch_sampler is not None:
            if batch_size != 1 or shuffle or sampler is not None or drop_last:
                raise ValueError("batch_sampler option is mutually exclusive with batch_size, shuffle, sampler, and drop_last")
            batch_size, drop_last = None, False
        else:
            if sampler is None:
                sampler = tud.RandomSampler(dataset, generator=generator) if shuffle else tud.SequentialSampler(dataset)  # type: ignore
            batch_sampler = tud.BatchSampler(sampler, batch_size, drop_last)  # type: ignore
        self.dataset, self.sampler, self.batch_sampler = dataset, sampler, batch_sampler
        self.batch_size, self.drop_last, self.generator = batch_size, drop_last, generator
        self.num_workers, self.collate_fn, self.pin_memory = num_workers, collate_fn, pin_memory
        self.timeout, self.worker_init_fn, self.multiprocessing_context = timeout, worker_init_fn, multiprocessing_context
        self.prefetch_factor, self.persistent_workers = prefetch_factor, persistent_workers

    def get_data_loader(self, repeat: bool = False, skip: int = 0, num_replicas: int = 1, rank: int = 0) -> tud.DataLoader:
        return tud.DataLoader(
            self.dataset,
            batch_sampler=adapt_batch_sampler(self.batch_sampler, repeat=repeat, skip=skip, num_replicas=num_replicas, rank=rank),
            num_workers=self.num_workers, collate_fn=self.collate_fn, pin_memory=self.pin_memory, timeout=self.timeout,
            worker_init_fn=self.worker_init_fn, multiprocessing_context=self.multiprocessing_context,
            generator=self.generator, prefetch_factor=self.prefetch_factor, persistent_workers=self.persistent_workers,
        )

    def __iter__(self) -> Iterator[Any]:
        """So that a trial's loaders also work in a plain torch loop outside this harness."""
        return iter(self.get_data_loader())

    def __len__(self) -> int:
        return len(self.batch_sampler)


def data_length(data: Any) -> int:
    """Number of records in a batch: the length of the first array found going down the first element of every
    dict / list / tuple (all leaves of a batch are assumed to have the same length)."""
    if isinstance(data, (np.ndarray, torch.Tensor)):
        return len(data)
    for kind, name in ((dict, "dict"), (list, "list"), (tuple, "tuple")):
        if isinstance(data, kind):
            if len(data) == 0:
                raise ValueError(f"`PyTorchTrial` must have at least one `np.ndarray` or `torch.Tensor` in it's {name} of inputs.")
            first = next(iter(data.values())) if kind is dict else data[0]
            return len(first)
    raise TypeError("Data of incorrect type: {}".format(type(data)))


def to_device(data: Any, device: Any, warned_types: Optional[Set[Type]] = None) -> Any:
    """Move a (nested) batch to `device`: dicts, lists and tuples are rebuilt, numeric numpy arrays become tensors,
    anything with a callable .to() is asked to move itself; other leaves are returned as they are, with one
    warning per type (per call unless the caller keeps `warned_types` between calls)."""
    if warned_types is None:
        warned_types = set()
    if isinstance(data, dict):
        return {k: to_device(v, device, warned_types) for k, v in data.items()}
    if isinstance(data, list):
        return [to_device(v, device, warned_types) for v in data]
    if isinstance(data, tuple):
        return tuple(to_device(v, device, warned_types) for v in data)
    if isinstance(data, np.ndarray):
        if data.dtype.kind in "fciub":  # float, complex, int, uint, bool: the kinds torch has tensors for
            return torch.from_numpy(data).to(device)
    elif callable(getattr(data, "to", None)):
        return data.to(device)
    if type(data) not in warned_types:
        warned_types.add(type(data))
        logger.warning(f"Was not able to move data item of type '{type(data).__name__}' to device.")
    return data
