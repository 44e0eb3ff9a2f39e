"""Metric reducers: Reducer enum, MetricReducer API and the reducer context.

Drop-in for reference harness/determined/pytorch/_reducer.py (Reducer :9-25, _simple_reduce_metrics
:28-46, MetricReducer :49-166, _SimpleReducer :169-208, _PyTorchReducerContext :273-518): same names,
signatures, assertion texts and ordering guarantees.  User-defined reducers remain arbitrary Python;
the numeric cross-rank reductions of *tensors* go through libdgs (see _metric_utils.DeviceMetrics).
"""
import abc
import enum
import itertools
from typing import Any, Callable, Dict, List, Optional, Union

import numpy as np


class Reducer(enum.Enum):
    """How per-batch (or per-slot) values of a validation metric are combined."""

    AVG = 1
    SUM = 2
    MAX = 3
    MIN = 4


def _simple_reduce_metrics(reducer: Reducer, metrics: np.ndarray, num_batches: Optional[List[int]] = None) -> np.float64:
    if reducer == Reducer.AVG:
        if num_batches and len(metrics) != len(num_batches):
            raise RuntimeError(
                f"Lengths of metrics and num_batches are not equal: {len(metrics)} != {len(num_batches)}."
            )
        return np.average(metrics, weights=num_batches)  # type: ignore
    if reducer == Reducer.SUM:
        return np.sum(metrics)  # type: ignore
    if reducer == Reducer.MAX:
        return np.max(metrics)  # type: ignore
    if reducer == Reducer.MIN:
        return np.min(metrics)  # type: ignore
    raise NotImplementedError


class MetricReducer(metaclass=abc.ABCMeta):
    """Efficient custom metric: accumulate locally with your own update(), shrink with
    per_slot_reduce() on every slot, finish with cross_slot_reduce() on the gathered pieces."""

    @abc.abstractmethod
    def reset(self) -> None:
        """Called before every training or validation workload."""

    @abc.abstractmethod
    def per_slot_reduce(self) -> Any:
        """Return this slot's (picklable) partial result; it is allgathered to every slot."""

    @abc.abstractmethod
    def cross_slot_reduce(self, per_slot_metrics: List) -> Any:
        """Combine the per-slot partials (ordered by rank) into the metric: a dict of metrics when
        wrap_reducer() got name=None, a single non-dict value otherwise."""


class _SimpleReducer(MetricReducer):
    """MetricReducer around a plain function of the flat list of update() values."""

    class _NotMetric:
        pass

    NOT_METRIC = _NotMetric()

    def __init__(self, fn: Callable) -> None:
        self.fn = fn
        self.reset()

    def reset(self) -> None:
        self.values: List[Any] = []

    def update(self, value: Any) -> None:
        self.values.append(value)

    def per_slot_reduce(self) -> Any:
        return self.values

    def cross_slot_reduce(self, per_slot_metrics: List) -> Any:
        # Batches were dealt to the slots round-robin; dealing the values back in the same way
        # restores the original order (provided update() ran once per batch).
        rounds = itertools.zip_longest(*per_slot_metrics, fillvalue=self.NOT_METRIC)
        flat = [v for one_round in rounds for v in one_round if v is not self.NOT_METRIC]
        return self.fn(flat)


def default_allgather_fn(metrics: Any) -> List:
    return [metrics]


class _WrappedReducer:
    def __init__(self, reducer: MetricReducer, name: Optional[str], for_training: bool, for_validation: bool) -> None:
        self.reducer = reducer
        self.name = name
        self.for_training = for_training
        self.for_validation = for_validation

    def _guard(self, what: str, fn: Callable, *args: Any) -> Any:
        try:
            return fn(*args)
        except Exception as e:
            kind = "function" if isinstance(self.reducer, _SimpleReducer) else type(self.reducer).__name__
            raise RuntimeError(f"failure in {what}() of custom reducer {kind} (name={self.name})") from e

    def reset(self) -> None:
        self._guard("reset", self.reducer.reset)

    def per_slot_reduce(self) -> Any:
        return self._guard("per_slot_reduce", self.reducer.per_slot_reduce)

    def cross_slot_reduce(self, per_slot_metrics: List) -> Any:
        return self._guard("cross_slot_reduce", self.reducer.cross_slot_reduce, per_slot_metrics)


class _PyTorchReducerContext:
    """Registry of custom reducers and the one-allgather reduction over all of them."""

    def __init__(self, allgather_fn: Callable[[Any], List[Any]] = default_allgather_fn) -> None:
        self._wrapped_reducers: List[_WrappedReducer] = []
        self._allgather_fn = allgather_fn

    def reset_reducers(self) -> None:
        for wrapped in self._wrapped_reducers:
            wrapped.reset()

    def wrap_reducer(
        self,
        reducer: Union[Callable, MetricReducer],
        name: Optional[str] = None,
        for_training: bool = True,
        for_validation: bool = True,
    ) -> MetricReducer:
        """Register a custom reducer; a plain callable is wrapped in a _SimpleReducer whose
        update() the caller feeds once per batch."""
        if any(reducer is w.reducer for w in self._wrapped_reducers):
            raise AssertionError(
                f"Detected the same reducer of type {type(reducer).__name__} in "
                "context.wrap_reducer(), please avoid calling wrap_reducer() on the "
                "same reducer object twice."
            )
        if not isinstance(reducer, MetricReducer):
            if not callable(reducer):
                raise AssertionError(
                    f"Detected invalid reducer in wrap_reducer() of type {type(reducer).__name__}. "
                    "Reducers must either be a function or a subclass of pytorch.MetricReducer."
                )
            reducer = _SimpleReducer(reducer)
        self._wrapped_reducers.append(_WrappedReducer(reducer, name, for_training, for_validation))
        return reducer

    @staticmethod
    def _describe(wrapped: _WrappedReducer) -> str:
        if isinstance(wrapped.reducer, _SimpleReducer):
            return f"function {wrapped.reducer.fn.__name__}()"
        return f"MetricReducer {type(wrapped.reducer).__name__}"

    def run_cross_slot_reduction(self, reducables: List[_WrappedReducer], gathered: List[Any]) -> Dict[str, Any]:
        metrics: Dict[str, Any] = {}
        # gathered is per slot a list over reducers; regroup to per reducer a tuple over slots
        for wrapped, per_slot in zip(reducables, zip(*gathered)):
            reduced = wrapped.cross_slot_reduce(per_slot)
            if wrapped.name is None:
                if not isinstance(reduced, dict):
                    extra = "" if isinstance(wrapped.reducer, _SimpleReducer) else "return "
                    raise AssertionError(
                        f"The custom reduction {self._describe(wrapped)} was "
                        "wrapped by a call to wrap_reducer() with name=None but it did not "
                        f"{extra}return a dict (it returned a {type(reduced).__name__}).  Please set "
                        "name if you wish to return a single metric or return a dictionary "
                        "mapping names to metrics if you with to return multiple metrics from "
                        "a single reducer."
                    )
                metrics.update(reduced)
            else:
                if isinstance(reduced, dict):
                    raise AssertionError(
                        f"The custom reduction {self._describe(wrapped)} was "
                        "wrapped by a call to wrap_reducer() with name set but it returned a "
                        "dict anyway.  Please leave name=None (the default value) if you wish "
                        "to return a dict of multiple metrics or return a single metric (not "
                        "a dict) if you wish to return a single named metric."
                    )
                metrics[wrapped.name] = reduced
        return metrics

    def reduce_metrics(self, for_training: bool) -> Dict[str, Any]:
        reducables = [
            w for w in self._wrapped_reducers if (w.for_training if for_training else w.for_validation)
        ]
        if not reducables:
            return {}
        # one allgather carries every reducer's per-slot value
        gathered = self._allgather_fn([w.per_slot_reduce() for w in reducables])
        return self.run_cross_slot_reduction(reducables, gathered)
