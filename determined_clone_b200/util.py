"""Small helpers the per-step path touches (reference harness/determined/util.py:114-141 transpose
helpers, :171-196 make_metrics, :308-341 calculate_batch_sizes)."""
import logging
from typing import Any, Dict, List, Optional, Tuple

from determined_clone_b200 import errors

logger = logging.getLogger("determined")


def _list_to_dict(list_of_dicts: List[Dict[str, Any]]) -> Dict[str, List[Any]]:
    """[{a: 1, b: 3}, {a: 2, b: 4}] -> {a: [1, 2], b: [3, 4]}"""
    out: Dict[str, List[Any]] = {}
    for d in list_of_dicts:
        for k, v in d.items():
            out.setdefault(k, []).append(v)
    return out


def _dict_to_list(dict_of_lists: Dict[str, List]) -> List[Dict[str, Any]]:
    """{a: [1, 2], b: [3, 4]} -> [{a: 1, b: 3}, {a: 2, b: 4}]; every list must have the same length."""
    lengths = {len(v) for v in dict_of_lists.values()}
    if len(lengths) > 1:
        raise AssertionError(f"All lists in the dict must be the same length. Found lengths {sorted(lengths)}.")
    n = lengths.pop() if lengths else 0
    return [{k: v[i] for k, v in dict_of_lists.items()} for i in range(n)]


def validate_batch_metrics(batch_metrics: List[Dict[str, Any]]) -> None:
    names = _list_to_dict(batch_metrics).keys()
    for idx, d in enumerate(batch_metrics):
        if d.keys() != names:
            raise AssertionError(f"inconsistent training metrics: index: {idx}: {set(names)} != {set(d.keys())}")


def make_metrics(num_inputs: Optional[int], batch_metrics: List[Dict[str, Any]]) -> Dict[str, Any]:
    """Per-period summary: avg_metrics[name] is the mean over batches of the values that are not None,
    or None when the values are not scalars."""
    import numpy as np

    columns = _list_to_dict(batch_metrics)
    validate_batch_metrics(batch_metrics)
    avg: Dict[str, Optional[float]] = {}
    for name, values in columns.items():
        mean: Optional[float] = None
        try:
            arr = np.array(values)
            mean = np.mean(arr[arr != None]).item()  # noqa: E711
        except (TypeError, ValueError):
            pass  # non-scalar metric: keep the name, no average
        avg[name] = mean
    metrics: Dict[str, Any] = {"batch_metrics": batch_metrics, "avg_metrics": avg}
    if num_inputs is not None:
        metrics["num_inputs"] = num_inputs
    return metrics


def calculate_batch_sizes(hparams: Dict[str, Any], slots_per_trial: int, trialname: str) -> Tuple[int, int]:
    """(per-slot batch, effective global batch): global_batch_size floor-divided over the slots."""
    if "global_batch_size" not in hparams:
        raise errors.InvalidExperimentException(
            "Please specify an integer `global_batch_size` hyperparameter in your experiment config."
        )
    global_batch_size = hparams["global_batch_size"]
    if not isinstance(global_batch_size, int):
        raise errors.InvalidExperimentException("`global_batch_size` hyperparameter must be an integer value.")
    slots = max(slots_per_trial, 1)
    if global_batch_size < slots:
        raise errors.InvalidExperimentException(
            f"Please set the `global_batch_size` hyperparameter to be greater or equal to the number of slots. "
            f"Current batch_size: {global_batch_size}, slots_per_trial: {slots_per_trial}."
        )
    per_slot = global_batch_size // slots
    effective = per_slot * slots
    if effective != global_batch_size:
        logger.warning(
            f"`global_batch_size` changed from {global_batch_size} to {effective} to divide equally across "
            f"{slots_per_trial} slots."
        )
    return per_slot, effective


def is_overridden(method: Any, parent_class: Any) -> bool:
    """Whether `method` (looked up on an instance or subclass) is a redefinition of parent_class's method
    (reference util.py:84-92 compares qualified names; here: the underlying function objects).  Anything that is
    not a function -- e.g. a user attribute shadowing the name -- counts as not overridden."""
    fn = getattr(method, "__func__", method)
    name = getattr(fn, "__name__", None)
    if not callable(fn) or name is None:
        return False
    return getattr(parent_class, name, None) is not fn


def has_param(fn: Any, name: str, pos: Optional[int] = None) -> bool:
    """Whether fn accepts an argument called `name`, or -- failing that -- has a positional argument at index `pos`
    (counting `self` for unbound methods exactly as inspect.getfullargspec does; reference util.py:95-104)."""
    import inspect

    args = inspect.getfullargspec(fn)[0]
    return name in args or (pos is not None and pos < len(args))


def is_numerical_scalar(n: Any) -> bool:
    """int / float, numpy scalars and 0-d numeric numpy arrays: what can be written to TensorBoard as a scalar or
    compared as a searcher metric (reference util.py:462-485)."""
    import numpy as np

    if isinstance(n, (int, float, np.number)):
        return True
    return isinstance(n, np.ndarray) and n.ndim == 0 and bool(np.issubdtype(n.dtype, np.number))
