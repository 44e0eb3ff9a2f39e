"""Training / validation metric aggregation across ranks.

Two routes with the same results:

* DeviceMetricBuffer -- the B200 route for scalar *device tensors* returned by train_batch: every
  batch's values are packed into a device ring by one tiny kernel (no host sync), and once per
  reporting period the ring is allgathered over NCCL and averaged by the metric kernels of libdgs;
  one device->host copy returns the per-batch and per-period means.  This replaces the per-batch
  `metric.cpu()` of reference harness/determined/pytorch/_pytorch_trial.py:920-927 and the
  pickle/ZMQ gather + numpy loops of harness/determined/pytorch/_metric_utils.py:29-102.

* the object route -- metrics that are already host values (Python numbers, numpy arrays, None)
  keep the reference's exact semantics: gather to the chief through DistributedContext and average
  there.  Same function names and behaviour as the reference module (:13-176).
"""
import ctypes
import logging
from typing import Any, Dict, List, Optional, Tuple, Union, cast

import numpy as np
import torch

from determined_clone_b200 import _lib, core, errors, util
from determined_clone_b200.pytorch._reducer import Reducer, _simple_reduce_metrics

logger = logging.getLogger("determined.pytorch")


# ---- object route (reference semantics) -----------------------------------------------------
def _process_combined_metrics_and_batches(
    combined_metrics_and_batches: List[Any],
) -> Tuple[Dict[str, Any], List[int]]:
    # ranks that ran no batches (num_batches == 0) do not take part
    live = [entry for entry in combined_metrics_and_batches if entry[1]]
    all_metrics = [entry[0] for entry in live]
    names = all_metrics[0].keys()
    return {k: [m[k] for m in all_metrics] for k in names}, [entry[1] for entry in live]


def _average_training_metrics(
    combined_timeseries: Dict[str, Any], combined_num_batches: List[int]
) -> List[Dict[str, Any]]:
    """combined_timeseries[name][process][batch] -> one dict per batch with the mean over processes
    of the values that are not None.  A metric that arrives as 0-d arrays is returned as 0-d arrays."""
    n_batches = combined_num_batches[0]  # identical on every data-parallel rank
    n_procs = len(combined_num_batches)
    per_batch: List[Dict[str, Any]] = [{} for _ in range(n_batches)]
    for name, series in combined_timeseries.items():
        keep_array = isinstance(series[0][0], np.ndarray)
        for b in range(n_batches):
            column = np.array([series[p][b] for p in range(n_procs)])
            mean = np.mean(column[column != None])  # noqa: E711
            per_batch[b][name] = np.array(mean) if keep_array else mean
    return per_batch


def _combine_metrics_across_processes(
    context: core.DistributedContext, metrics: Dict[str, Any], num_batches: int
) -> Tuple[Optional[Dict[str, Any]], Optional[List[int]]]:
    assert (
        context.size > 1
    ), "_combine_metrics_across_processes should only be called if context.distributed > 1"
    gathered = context.gather((metrics, num_batches))
    if context.rank != 0:
        return None, None
    assert gathered is not None, "gathered metrics should not be None"
    return _process_combined_metrics_and_batches(gathered)


def _combine_and_average_training_metrics(
    context: core.DistributedContext, per_batch_metrics: List[Dict[str, Any]]
) -> List[Dict[str, Any]]:
    assert context.size > 1, "Can only average training metrics in multi-GPU training."
    timeseries, num_batches = _combine_metrics_across_processes(
        context, util._list_to_dict(per_batch_metrics), num_batches=len(per_batch_metrics)
    )
    if context.rank == 0:
        per_batch_metrics = _average_training_metrics(
            cast(Dict[str, List[List[Any]]], timeseries), cast(List[int], num_batches)
        )
    return per_batch_metrics


def _prepare_metrics_reducers(reducer: Union[Reducer, Dict[str, Any]], keys: Any) -> Dict[str, Reducer]:
    """One Reducer per validation metric, from what evaluation_reducer() returned: a single Reducer for all of
    them, or a dict naming every metric (reference _metric_utils.py:105-127, same exception type and texts)."""
    if isinstance(reducer, dict):
        table = dict(reducer)
        if table.keys() != set(keys):
            raise errors.InvalidExperimentException(
                "Please provide a single evaluation reducer or "
                "provide a reducer for every validation metric. "
                f"Expected keys: {keys}, provided keys: {table.keys()}.",
            )
    else:
        table = {k: reducer for k in keys}
    for r in table.values():
        if not isinstance(r, Reducer):
            raise errors.InvalidExperimentException(
                "Please select `determined.pytorch.Reducer` for reducing validation metrics.",
            )
    return table


def _convert_metrics_to_numpy(metrics: Dict[str, Any]) -> Dict[str, Any]:
    for name, value in metrics.items():
        if isinstance(value, torch.Tensor):
            metrics[name] = value.detach().cpu().numpy()
    return metrics


def _reduce_metrics(
    context: core.DistributedContext,
    batch_metrics: List,
    keys: Any,
    metrics_reducers: Dict[str, Reducer],
) -> Dict[str, Any]:
    """Validation: reduce over this rank's batches, then (chief only) over ranks, AVG weighted by the
    number of batches each rank saw."""
    metrics: Dict[str, Any] = {}
    if len(batch_metrics):
        metrics = {
            name: _simple_reduce_metrics(
                reducer=metrics_reducers[name],
                metrics=np.stack([b[name] for b in batch_metrics], axis=0),
                num_batches=None,
            )
            for name in keys or []
        }
    if context.size > 1:
        combined, batches_per_process = _combine_metrics_across_processes(context, metrics, len(batch_metrics))
        if context.rank != 0:
            return {}
        assert combined is not None
        combined = _convert_metrics_to_numpy(combined)
        metrics = {
            name: _simple_reduce_metrics(
                reducer=metrics_reducers[name], metrics=combined[name], num_batches=batches_per_process
            )
            for name in keys or []
        }
    return metrics


# ---- device route ---------------------------------------------------------------------------
def _stream() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    return ctypes.c_void_p(t.data_ptr() if t is not None else None)


def is_device_scalar(value: Any) -> bool:
    return isinstance(value, torch.Tensor) and value.is_cuda and value.numel() == 1


class DeviceMetricBuffer:
    """Per-period ring of per-batch metric scalars that never leaves the GPU until report time."""

    def __init__(self, device: torch.device, capacity: int = 128) -> None:
        self.device = device
        self.names: List[str] = []
        self._array_like: Dict[str, bool] = {}
        self._cap = capacity
        self._rows = 0
        self._ring: Optional[torch.Tensor] = None
        self._valid: Optional[torch.Tensor] = None

    def __len__(self) -> int:
        return self._rows

    def _ensure(self, m: int) -> None:
        if self._ring is None or self._ring.shape[1] != m or self._rows >= self._cap:
            cap = self._cap * 2 if self._ring is not None and self._rows >= self._cap else self._cap
            ring = torch.zeros(cap, m, dtype=torch.float64, device=self.device)
            valid = torch.zeros(cap, m, dtype=torch.uint8, device=self.device)
            if self._ring is not None and self._rows and self._ring.shape[1] == m:
                ring[: self._rows] = self._ring[: self._rows]
                valid[: self._rows] = self._valid[: self._rows]  # type: ignore
            self._ring, self._valid, self._cap = ring, valid, cap

    def accepts(self, metrics: Dict[str, Any]) -> bool:
        """True when every value is a scalar device tensor (or None) and names match earlier batches."""
        if not metrics or len(metrics) > _lib.MAX_METRICS:
            return False
        if self._rows and list(metrics.keys()) != self.names:
            return False
        return all(v is None or is_device_scalar(v) for v in metrics.values()) and any(
            v is not None for v in metrics.values()
        )

    def append(self, metrics: Dict[str, Any]) -> None:
        """Record one batch.  Launches one kernel on the current stream; no synchronisation."""
        if self._rows == 0:
            self.names = list(metrics.keys())
        m = len(self.names)
        self._ensure(m)
        refs = (_lib.ScalarRef * m)()
        keep = []  # the scalars must outlive the launch; same-stream allocator reuse keeps this safe
        for j, name in enumerate(self.names):
            v = metrics[name]
            if v is None:
                refs[j].ptr, refs[j].dtype = None, _lib.F32
                continue
            v = v.detach()
            if v.dtype not in (torch.float32, torch.bfloat16, torch.float16, torch.float64, torch.int64):
                v = v.to(torch.float64)
            keep.append(v)
            refs[j].ptr, refs[j].dtype = v.data_ptr(), _lib.dtype_code(v.dtype)
            self._array_like.setdefault(name, True)  # `.cpu().numpy()` of a tensor is a 0-d array
        assert self._ring is not None and self._valid is not None
        _lib.calls.dgs_metric_pack(refs, m, _p(self._ring[self._rows]), _p(self._valid[self._rows]), _stream())
        self._rows += 1

    def reset(self) -> None:
        self._rows = 0

    def reduce(self, distributed: core.DistributedContext) -> Tuple[List[Dict[str, Any]], Dict[str, Any], List[int]]:
        """Average over ranks per batch, then over batches.  Returns (per-batch dicts, avg dict,
        per-metric batch counts); identical on every rank.  One device->host copy."""
        return self.reduce_async(distributed).result()

    def reduce_async(self, distributed: core.DistributedContext) -> "PendingMetrics":
        """Enqueue the cross-rank reduction and the copy of its result into pinned host memory, and return
        without waiting: the caller keeps launching the next batches and calls .result() later."""
        b, m = self._rows, len(self.names)
        if b == 0:
            return PendingMetrics(None, None, 0, 0, [])
        assert self._ring is not None and self._valid is not None
        n = distributed.size
        # one result block -> one device->host copy: [per_batch b*m f64 | avg m f64 | count m i64 | valid b*m u8]
        n_f64 = b * m + m
        block = torch.empty(8 * n_f64 + 8 * m + b * m, dtype=torch.uint8, device=self.device)
        base = block.data_ptr()
        outs = (ctypes.c_void_p(base), ctypes.c_void_p(base + 8 * n_f64 + 8 * m), ctypes.c_void_p(base + 8 * b * m),
                ctypes.c_void_p(base + 8 * n_f64))
        if n > 1:
            comm = distributed.comm
            if comm is None:
                raise _lib.DgsError("DeviceMetricBuffer.reduce: no NCCL communicator attached to the DistributedContext")
            # one packed record per rank -> ONE allgather: [b*m f64 values | int64 num_batches | b*m validity bytes]
            rec = (8 * b * m + 8 + b * m + 7) // 8 * 8
            send = torch.empty(rec, dtype=torch.uint8, device=self.device)
            send[: 8 * b * m].view(torch.float64).copy_(self._ring[:b].reshape(-1))
            send[8 * b * m: 8 * b * m + 8].view(torch.int64).fill_(b)
            send[8 * b * m + 8: 9 * b * m + 8].copy_(self._valid[:b].reshape(-1))
            recv = torch.empty(n * rec, dtype=torch.uint8, device=self.device)
            comm.allgather(send, recv)
            comm.wait()
            _lib.calls.dgs_metric_average_packed(_p(recv), rec, n, b, m, *outs, _stream())
        else:
            ring, valid = self._ring[:b].contiguous(), self._valid[:b].contiguous()
            nb = torch.full((1,), b, dtype=torch.int64, device=self.device)
            _lib.calls.dgs_metric_average(_p(ring), _p(valid), _p(nb), 1, b, m, *outs, _stream())
        host = torch.empty(block.numel(), dtype=torch.uint8, pin_memory=True)
        host.copy_(block, non_blocking=True)
        done = torch.cuda.Event()
        done.record(torch.cuda.current_stream())
        return PendingMetrics(host, done, b, m, list(self.names))


def reduce_validation_on_device(buffer: "DeviceMetricBuffer", distributed: core.DistributedContext,
                                reducers: Dict[str, Reducer]) -> Dict[str, Any]:
    """Validation metrics kept on the device: reduce every metric over this rank's batches with its Reducer
    (dgs_metric_reduce), allgather the per-rank results and batch counts, reduce again over ranks -- AVG weighted
    by batches per rank, as _reduce_metrics above does on the chief (reference _metric_utils.py:137-176) -- and
    read the answer back once.  Every rank gets the same dict; non-chief callers drop it."""
    b, m = len(buffer), len(buffer.names)
    if b == 0 or m == 0:
        return {}
    assert buffer._ring is not None
    dev, n = buffer.device, distributed.size
    ring = buffer._ring[:b].contiguous()
    ops = sorted({reducers[name].value for name in buffer.names})
    local = torch.empty(len(ops), m, dtype=torch.float64, device=dev)  # one row per reducer kind in use
    for k, op in enumerate(ops):
        _lib.calls.dgs_metric_reduce(_p(ring), None, b, m, op, _p(local[k]), _stream())
    if n > 1:
        comm = distributed.comm
        if comm is None:
            raise _lib.DgsError("reduce_validation_on_device: no NCCL communicator attached to the DistributedContext")
        send = torch.cat([local.reshape(-1), torch.full((1,), float(b), dtype=torch.float64, device=dev)])
        recv = torch.empty(n, send.numel(), dtype=torch.float64, device=dev)
        comm.allgather(send, recv)
        comm.wait()
        weights = recv[:, -1].to(torch.int64).contiguous()
        out = torch.empty(len(ops), m, dtype=torch.float64, device=dev)
        for k, op in enumerate(ops):
            vals = recv[:, k * m:(k + 1) * m].contiguous()
            _lib.calls.dgs_metric_reduce(_p(vals), _p(weights) if op == Reducer.AVG.value else None, n, m, op, _p(out[k]), _stream())
    else:
        out = local
    host = out.cpu().numpy()
    return {name: np.float64(host[ops.index(reducers[name].value), j]) for j, name in enumerate(buffer.names)}


class PendingMetrics:
    """A metric reduction whose result is on its way to pinned host memory."""

    def __init__(self, host: Optional[torch.Tensor], done: Any, b: int, m: int, names: List[str]) -> None:
        self._host, self._done, self._b, self._m, self._names = host, done, b, m, names

    def result(self) -> Tuple[List[Dict[str, Any]], Dict[str, Any], List[int]]:
        if self._host is None:
            return [], {}, []
        self._done.synchronize()
        b, m = self._b, self._m
        n_f64 = b * m + m
        raw = self._host.numpy()
        host = raw[: 8 * n_f64].view(np.float64)
        counts = raw[8 * n_f64: 8 * n_f64 + 8 * m].view(np.int64).tolist()
        host_valid = raw[8 * n_f64 + 8 * m:].reshape(b, m).astype(bool)
        per_batch_vals = host[: b * m].reshape(b, m)
        per_batch: List[Dict[str, Any]] = []
        for i in range(b):
            row: Dict[str, Any] = {}
            for j, name in enumerate(self._names):
                row[name] = np.array(per_batch_vals[i, j]) if host_valid[i, j] else None
            per_batch.append(row)
        avg = {name: (float(host[b * m + j]) if counts[j] else None) for j, name in enumerate(self._names)}
        return per_batch, avg, counts


def _log_tb_metrics(writer: Any, metric_type: str, steps_completed: int, metrics: Dict[str, Any],
                    batch_metrics: Optional[List[Dict[str, Any]]] = None) -> None:
    """Scalars for TensorBoard under "Determined/<name>" (reference _metric_utils.py:179-211): every per-batch
    training value at the batch it belongs to; reduced values that have no per-batch series (custom reducers,
    validation metrics -- the latter prefixed "val_" unless they already are) at steps_completed."""
    if metric_type not in ("train", "val"):
        logger.warning("Unrecognized tensorboard metric type: " + metric_type)
    have_series = set()
    first = steps_completed - len(batch_metrics or [])
    for k, batch in enumerate(batch_metrics or []):
        for name, value in batch.items():
            have_series.add(name)
            if util.is_numerical_scalar(value):
                writer.add_scalar("Determined/" + name, value, first + k)
    for name, value in metrics.items():
        if name in have_series:
            continue
        if metric_type == "val" and not name.startswith("val"):
            name = "val_" + name
        if util.is_numerical_scalar(value):
            writer.add_scalar("Determined/" + name, value, steps_completed)
