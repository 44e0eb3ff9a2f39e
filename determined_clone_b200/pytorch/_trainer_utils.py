"""Lengths of training: Batch / Epoch / Record (reference harness/determined/pytorch/_trainer_utils.py)."""
from typing import Any, Optional


class TrainUnit:
    def __init__(self, value: int) -> None:
        self.value = int(value)

    def __repr__(self) -> str:
        return f"{type(self).__name__}({self.value})"

    @staticmethod
    def _from_searcher_unit(length: int, unit: str, global_batch_size: Optional[int] = None) -> "TrainUnit":
        if unit == "records":
            return Record(length, global_batch_size)
        return {"batches": Batch, "epochs": Epoch}[unit](length)

    @staticmethod
    def _from_values(batches: Optional[int] = None, records: Optional[int] = None, epochs: Optional[int] = None,
                     global_batch_size: Optional[int] = None) -> "TrainUnit":
        if sum(v is not None for v in (batches, records, epochs)) != 1:
            raise ValueError(f"invalid config: batches={batches} records={records} epochs={epochs}")
        if batches is not None:
            return Batch(batches)
        if records is not None:
            return Record(records, global_batch_size)
        return Epoch(epochs)  # type: ignore


class Batch(TrainUnit):
    pass


class Epoch(TrainUnit):
    pass


class Record(TrainUnit):
    def __init__(self, value: int, global_batch_size: Optional[int] = None) -> None:
        super().__init__(value)
        self.global_batch_size = global_batch_size


def to_batches(unit: Any, epoch_len: Optional[int], global_batch_size: Optional[int] = None) -> int:
    """A length in batches: ints pass through, epochs need the training loader's length, records the global batch
    size.  Lengths that would round down to nothing become one batch."""
    if isinstance(unit, int):
        return unit
    if isinstance(unit, Batch):
        return unit.value
    if isinstance(unit, Record):
        gbs = unit.global_batch_size or global_batch_size
        if not gbs:
            raise ValueError("global_batch_size not specified in the hyperparameters")
        return max(unit.value // gbs, 1)
    if isinstance(unit, Epoch):
        if epoch_len is None:
            raise ValueError("max_length in epochs needs a training data loader with a length")
        return unit.value * epoch_len
    raise TypeError(f"expected a number of batches or a TrainUnit, got {type(unit).__name__}")
