"""PyTorchTrial and the per-batch training loop around it.

Mirrors the slice of the reference controller that sits on the grad-sync path
(reference harness/determined/pytorch/_pytorch_trial.py): pre_execute_hook / _set_random_seeds
:286-322, _aggregate_training_metrics :324-365, _train_with_boundaries / _train_for_op :695-820,
_auto_step_lr_scheduler_per_batch :845-870, _train_batch :877-939, _validate :941-1115 (metrics
part), and the PyTorchTrial interface :1416-1663.  Searcher operations, checkpoints, preemption and
the master connection are the reference's control plane and are not rebuilt: the loop runs a fixed
number of batches and hands every report to `on_report`.

What changes on the hot path: scalar device metrics are not pulled to the host every batch
(:920-927 does `.cpu()` per metric per batch); they are packed into a device ring and reduced across
ranks by kernels once per reporting period (_metric_utils.DeviceMetricBuffer).
"""
import abc
import contextlib
import inspect
import json
import logging
import os
import pathlib
import pickle
import random
import sys
import tempfile
import time
import uuid
import warnings
from typing import Any, Callable, Dict, List, Optional, Union

import numpy as np
import torch

from determined_clone_b200 import errors, util
from determined_clone_b200.pytorch import _metric_utils
from determined_clone_b200.pytorch._callback import PyTorchCallback
from determined_clone_b200.pytorch._data import DataLoader, data_length
from determined_clone_b200.pytorch._lr_scheduler import LRScheduler
from determined_clone_b200.pytorch._pytorch_context import PyTorchTrialContext
from determined_clone_b200.pytorch._reducer import Reducer
from determined_clone_b200.pytorch._trainer_utils import to_batches

logger = logging.getLogger("determined.pytorch")

TorchData = Union[Dict[str, torch.Tensor], List[torch.Tensor], torch.Tensor]


class PyTorchTrial(abc.ABC):
    """User-facing trial interface (same method names and contracts as the reference)."""

    @abc.abstractmethod
    def __init__(self, context: PyTorchTrialContext) -> None:
        pass

    @abc.abstractmethod
    def train_batch(self, batch: TorchData, epoch_idx: int, batch_idx: int) -> Union[torch.Tensor, Dict[str, Any]]:
        """Train on one batch: forward, context.backward(loss), context.step_optimizer(opt); return metrics."""

    @abc.abstractmethod
    def build_training_data_loader(self) -> DataLoader:
        pass

    def build_validation_data_loader(self) -> Optional[DataLoader]:
        return None

    def evaluate_batch(self, batch: TorchData, batch_idx: int = 0) -> Dict[str, Any]:
        raise NotImplementedError

    def evaluation_reducer(self) -> Union[Reducer, Dict[str, Reducer]]:
        return Reducer.AVG

    def get_batch_length(self, batch: Any) -> int:
        return data_length(batch)

    def evaluate_full_dataset(self, data_loader: Any) -> Dict[str, Any]:
        """Alternative to evaluate_batch: the chief alone computes the validation metrics over the whole loader
        (reference :1616-1633).  Exactly one of the two is overridden."""
        pass

    def build_callbacks(self) -> Dict[str, Any]:
        """name -> PyTorchCallback; see _callback.py for the hooks and when the loop calls them."""
        return {}


class _TrialState:
    """What trial_state.pkl holds (reference _pytorch_trial.py:164-186).  trial_id tells a pause/restart (same
    trial: counters continue) from a warm start off another trial's checkpoint (counters restart)."""

    def __init__(self, trial_id: int = 0, last_ckpt: int = 0, step_id: int = 0, last_val: int = 0,
                 batches_trained: int = 0, epochs_trained: int = 0) -> None:
        self.trial_id = trial_id
        self.last_ckpt = last_ckpt
        self.step_id = step_id
        self.last_val = last_val
        self.batches_trained = batches_trained
        self.epochs_trained = epochs_trained


def _clone_tensors(obj: Any) -> Any:
    if isinstance(obj, torch.Tensor):
        return obj.detach().clone()
    if isinstance(obj, dict):
        return {k: _clone_tensors(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_clone_tensors(v) for v in obj)
    return obj


class _BatchFeed:
    """next() of the training iterator, with the host-to-device copy of the FOLLOWING batch started on a side stream
    as soon as the current one is handed out, so that it overlaps the current batch's forward and backward instead of
    preceding the next one (the reference copies at the top of _train_batch, _pytorch_trial.py:883-884, on the
    compute stream).  Only on a CUDA device with automatic to_device; nothing is fetched beyond the last batch the
    loop will train."""

    def __init__(self, controller: "_PyTorchTrialController", it: Any) -> None:
        self.c, self.it = controller, it
        ctx = controller.context
        self.on = ctx.device.type == "cuda" and ctx.experimental._auto_to_device and torch.cuda.is_available()
        self.stream = torch.cuda.Stream(device=ctx.device) if self.on else None
        self.ahead: Any = None  # (host batch, device batch, copy-done event)

    def _pull(self) -> Any:
        try:
            return next(self.it)
        except StopIteration:
            self.it = iter(self.c.training_loader)
            return next(self.it)

    def _stage(self) -> None:
        host = self._pull()
        with torch.cuda.stream(self.stream):
            dev = self.c.context.to_device(host)
            done = torch.cuda.Event()
            done.record(self.stream)
        self.ahead = (host, dev, done)

    @staticmethod
    def _used_on(data: Any, stream: Any) -> None:
        if isinstance(data, torch.Tensor):
            if data.is_cuda:
                data.record_stream(stream)  # allocated on the copy stream, consumed on the compute stream
        elif isinstance(data, dict):
            for v in data.values():
                _BatchFeed._used_on(v, stream)
        elif isinstance(data, (list, tuple)):
            for v in data:
                _BatchFeed._used_on(v, stream)

    def next(self, more_after: bool) -> Any:
        if not self.on:
            return self._pull(), None
        if self.ahead is None:
            self._stage()
        host, dev, done = self.ahead
        self.ahead = None
        cur = torch.cuda.current_stream()
        cur.wait_event(done)
        self._used_on(dev, cur)
        if more_after:
            self._stage()
        return host, dev


class _PyTorchTrialController:
    def __init__(
        self,
        trial_inst: PyTorchTrial,
        context: PyTorchTrialContext,
        max_length: int,
        reporting_period: int = 100,
        on_report: Optional[Callable[[int, Dict[str, Any]], None]] = None,
        steps_completed: int = 0,
        device_metrics: bool = True,
        test_mode: bool = False,
        trial_id: int = 1,
        val_from_previous_run: int = 0,
        validation_period: Optional[int] = None,
        checkpoint_period: Optional[int] = None,
        checkpoint_policy: str = "none",
        searcher_metric_name: Optional[str] = None,
        smaller_is_better: bool = True,
        step_zero_validation: bool = False,
        checkpoint_dir: Optional[Any] = None,
        latest_checkpoint: Optional[str] = None,
        final_validation: bool = False,
        final_checkpoint: bool = False,
        on_validation: Optional[Callable[[int, Dict[str, Any]], None]] = None,
        local_training: bool = True,
        det_profiler: Any = None,
        global_batch_size: Optional[int] = None,
    ) -> None:
        """max_length / reporting_period / validation_period / checkpoint_period: a number of batches, or the
        reference's Batch / Epoch / Record units (resolved once the training loader's length is known).  The loop of
        the reference (_pytorch_trial.py:626-820) without a master: checkpoints go to directories named by a fresh
        uuid (under `checkpoint_dir`, or wherever the core context's checkpoint storage puts them), reports go to
        on_report / on_validation and the core context's train reporter, and are kept in .reports /
        .validation_reports.  With the defaults nothing but training and reporting happens.  local_training and
        det_profiler are accepted for signature compatibility: there is no other mode and no system profiler here."""
        self.trial = trial_inst
        self.context = context
        # test_mode trains exactly one batch (reference _pytorch_trial.py:742-743,755-756)
        self.max_length = steps_completed + 1 if test_mode else max_length
        self.reporting_period = reporting_period
        self.on_report = on_report
        self.batches_trained = steps_completed
        self.is_chief = context.distributed.rank == 0
        self.reports: List[Dict[str, Any]] = []
        self._device_metrics = device_metrics and context.device.type == "cuda"
        self._metric_buffer = _metric_utils.DeviceMetricBuffer(
            context.device, capacity=max(reporting_period, 1) if isinstance(reporting_period, int) else 128)
        self._global_batch_size = global_batch_size or context._global_batch_size
        self.training_loader: Any = None
        self.validation_loader: Any = None
        self.trial_id = trial_id
        self.state = _TrialState(trial_id=trial_id, batches_trained=steps_completed)
        self.val_from_previous_run = val_from_previous_run
        self.test_mode = test_mode
        self.validation_period = validation_period
        self.checkpoint_period = checkpoint_period
        if checkpoint_policy not in ("all", "best", "none"):
            raise ValueError(f"checkpoint_policy must be one of 'all', 'best', 'none'; got {checkpoint_policy!r}")
        self.ckpt_policy = checkpoint_policy
        self.searcher_metric_name = searcher_metric_name
        self.smaller_is_better = smaller_is_better
        self.step_zero_validation = step_zero_validation
        self.checkpoint_dir = checkpoint_dir
        self.latest_checkpoint = latest_checkpoint
        self.final_validation = final_validation
        self.final_checkpoint = final_checkpoint
        self.on_validation = on_validation
        self.validation_reports: List[Dict[str, Any]] = []
        self.checkpoints: List[str] = []  # uuids written in this run, oldest first
        self._best_validation: Optional[float] = None
        # duck-typed trials (anything with train_batch / build_training_data_loader) get PyTorchTrial's defaults
        self.callbacks: Dict[str, Any] = getattr(trial_inst, "build_callbacks", dict)() or {}
        for name, cb in self.callbacks.items():
            if not isinstance(cb, PyTorchCallback):
                raise TypeError(f"build_callbacks() must map names to PyTorchCallback objects; '{name}' is a {type(cb).__name__}")
            if util.is_overridden(cb.on_checkpoint_end, PyTorchCallback):
                warnings.warn(
                    "The on_checkpoint_end callback is deprecated, please use on_checkpoint_write_end instead.",
                    FutureWarning, stacklevel=2,
                )
        self._workload_end_hooks = any(util.is_overridden(cb.on_training_workload_end, PyTorchCallback) for cb in self.callbacks.values())

    # ---- set-up -------------------------------------------------------------------------------------
    @classmethod
    def pre_execute_hook(cls, trial_seed: int, distributed_backend: Any = None) -> None:
        # communicator bring-up happens lazily in the context (NCCL id over DistributedContext.broadcast)
        cls._set_random_seeds(trial_seed)

    @staticmethod
    def _set_random_seeds(seed: int) -> None:
        # every rank uses the same seed: identical initial weights, identical shuffles
        random.seed(seed)
        np.random.seed(seed)
        torch.random.manual_seed(seed)

    def _set_data_loaders(self) -> None:
        """Reference _pytorch_trial.py:446-506: a det DataLoader is adapted (repeat / shard / skip) here, anything
        else is taken as is once the trial has opted out of the reproducibility checks; the epoch length is the
        chief's len() of the loader it will iterate (per rank already), sys.maxsize when it has none."""
        from determined_clone_b200.pytorch._data import _dataset_repro_warning

        d = self.context.distributed
        checked = not self.context.experimental._data_repro_checks_disabled
        train = self.trial.build_training_data_loader()
        if isinstance(train, DataLoader):
            self.training_loader = train.get_data_loader(repeat=True, skip=self.batches_trained, num_replicas=d.size, rank=d.rank)
        else:
            if checked:
                raise RuntimeError(_dataset_repro_warning("build_training_data_loader", train))
            self.training_loader = train
        try:
            epoch_len = len(self.training_loader)
        except TypeError:
            epoch_len = sys.maxsize
        self.context._epoch_len = d.broadcast(epoch_len) if d.size > 1 else epoch_len  # ranks must agree on epoch boundaries
        val = self.trial.build_validation_data_loader()
        if isinstance(val, DataLoader):
            self.validation_loader = val.get_data_loader(repeat=False, skip=0, num_replicas=d.size, rank=d.rank)
        else:
            if val is not None and checked:
                raise RuntimeError(_dataset_repro_warning("build_validation_data_loader", val))
            self.validation_loader = val

    # ---- training loop ------------------------------------------------------------------------------
    def run(self) -> None:
        restore = self._checkpoint_path(self.latest_checkpoint) if self.latest_checkpoint is not None else None
        if restore is not None and self.training_loader is None:
            self.batches_trained = self._peek_batches_trained(restore)  # the loader skips what was trained
        with contextlib.ExitStack() as stack:
            for cb in self.callbacks.values():
                cb.on_trial_startup(self.batches_trained, self.latest_checkpoint)
                stack.callback(cb.on_trial_shutdown)  # also on an exception or a stop request
            if self.training_loader is None:
                self._set_data_loaders()
            self._resolve_units()
            it = iter(self.training_loader)
            if restore is not None:
                # after the training iterator exists, as in the reference (:583-612): creating it draws from torch's
                # global generator, and a restored run must continue the generator exactly where the checkpoint left it
                self._load(restore)
            for cb in self.callbacks.values():
                cb.on_training_start()
            if self.context.profiler:
                stack.enter_context(self.context.profiler)  # torch.profiler set through context.set_profiler()
            if self.step_zero_validation and not self.val_from_previous_run and self.batches_trained == 0:
                self._validation_step()
            self.context.reset_reducers()
            self._train_loop(it, [], None, 0, time.time())
            # end of training: the reference validates and checkpoints whatever is not current (:810-814)
            if self.final_validation and self.state.last_val != self.batches_trained:
                self._validation_step()
            if self.final_checkpoint and self.state.last_ckpt != self.batches_trained:
                self._checkpoint()

    def _resolve_units(self) -> None:
        """Batch / Epoch / Record lengths -> batches, now that the epoch length is known."""
        epoch_len, gbs = self.context._epoch_len, self._global_batch_size
        self.max_length = to_batches(self.max_length, epoch_len, gbs)
        self.reporting_period = max(to_batches(self.reporting_period, epoch_len, gbs), 1)
        if self.validation_period is not None:
            self.validation_period = to_batches(self.validation_period, epoch_len, gbs)
        if self.checkpoint_period is not None:
            self.checkpoint_period = to_batches(self.checkpoint_period, epoch_len, gbs)

    def _train_loop(self, it: Any, pending: List[Dict[str, Any]], previous: Any, samples: int, t0: float) -> None:
        epoch_len = self.context._epoch_len
        feed = _BatchFeed(self, it)
        while self.batches_trained < self.max_length:
            batch_idx = self.batches_trained
            # `staged`: the batch already on the device (its copy ran on a side stream during the previous batch)
            batch, staged = feed.next(more_after=self.batches_trained + 1 < self.max_length)
            self.context._current_batch_idx = batch_idx
            epoch_idx = batch_idx // epoch_len if epoch_len else 0
            if self.callbacks and epoch_len and batch_idx % epoch_len == 0:
                self._on_epoch_start(epoch_idx)
            metrics = self._train_batch(batch, epoch_idx, batch_idx, staged=staged)
            if self.context.profiler:
                self.context.profiler.step()
            samples += self.trial.get_batch_length(batch)
            if metrics is not None:
                pending.append(metrics)
            self.batches_trained += 1
            n = self.batches_trained
            if self.callbacks and epoch_len and n % epoch_len == 0:
                for cb in self.callbacks.values():
                    cb.on_training_epoch_end(epoch_idx)
            # context.set_stop_requested(True) from user code ends training after the current batch (reference
            # _pytorch_trial.py:521-525 raises ShouldExit there); what has been trained is still reported
            stopping = self._stop_requested()
            validate_due = bool(self.validation_period) and n % self.validation_period == 0
            checkpoint_due = bool(self.checkpoint_period) and n % self.checkpoint_period == 0
            if n % self.reporting_period == 0 or n == self.max_length or stopping or validate_due or checkpoint_due:
                finish = self._begin_report(pending)
                self._deliver(previous)
                previous = (finish, n, samples, time.time() - t0)
                if self._workload_end_hooks or validate_due or checkpoint_due:
                    # callbacks want the metrics now, or a validation / checkpoint (which drain the GPU anyway) comes next
                    self._deliver(previous)
                    previous = None
                # otherwise the reduction of this period is enqueued now and read one period later (or at the end),
                # so the host keeps launching batches instead of waiting for the GPU to drain
                pending, samples, t0 = [], 0, time.time()
                self.context.reset_reducers()
                if validate_due and self.state.last_val != n:
                    self._validation_step()
                if checkpoint_due and self.state.last_ckpt != n:
                    self._checkpoint()
                if validate_due or checkpoint_due:
                    t0 = time.time()
            if stopping:
                break
        self._deliver(previous)

    def _on_epoch_start(self, epoch_idx: int) -> None:
        for cb in self.callbacks.values():
            if inspect.signature(cb.on_training_epoch_start).parameters:
                cb.on_training_epoch_start(epoch_idx)
            else:  # the pre-0.17.8 signature, still accepted by the reference (:378-386)
                logger.warning("on_training_epoch_start() without parameters is deprecated since 0.17.8. Please add epoch_idx parameter.")
                cb.on_training_epoch_start()

    def _stop_requested(self) -> bool:
        """As in the reference, the flag is per rank: user code must make the same request on every rank (the next
        batch starts with collectives)."""
        preempt = getattr(self.context._core, "preempt", None)
        if preempt is not None and preempt.should_preempt():
            return True
        return bool(self.context.get_stop_requested())

    def _begin_report(self, training_metrics: List[Dict]) -> Callable[[], Dict]:
        """Start aggregating the period's metrics; returns the function that completes it."""
        if len(self._metric_buffer) and not self.context._wrapped_reducers:
            handle = self._metric_buffer.reduce_async(self.context.distributed)
            self._metric_buffer.reset()

            def finish() -> Dict:
                batch_metrics, avg, _ = handle.result()
                return {"batch_metrics": batch_metrics, "avg_metrics": avg} if self.is_chief else {}

            return finish
        report = self._aggregate_training_metrics(training_metrics)  # custom reducers / host values: synchronous
        return lambda: report

    def _deliver(self, previous: Any) -> None:
        if previous is None:
            return
        finish, steps, samples, dur = previous
        report = finish()
        if self._workload_end_hooks:
            d = self.context.distributed
            everyone = d.broadcast(report) if d.size > 1 else report  # non-chief ranks hold {} (:771)
            for cb in self.callbacks.values():
                cb.on_training_workload_end(avg_metrics=everyone["avg_metrics"], batch_metrics=everyone["batch_metrics"])
        if self.is_chief:
            report["samples_per_second"] = samples / dur * self.context.distributed.size if dur > 0 else 0.0
            report["steps_completed"] = steps
            self.reports.append(report)
            train = getattr(self.context._core, "train", None)
            if train is not None:  # a core.Context was given: report where the reference reports (:358-365)
                train.report_training_metrics(steps_completed=steps, metrics=report["avg_metrics"], batch_metrics=report["batch_metrics"])
            if self.context.get_enable_tensorboard_logging():
                _metric_utils._log_tb_metrics(self.context.get_tensorboard_writer(), "train", steps, report["avg_metrics"], report["batch_metrics"])
            if self.on_report:
                self.on_report(steps, report)

    def _auto_step_lr_scheduler_per_batch(self, batch_idx: int, lr_scheduler: LRScheduler) -> None:
        """Never step the LR when the optimizer did not step; STEP_EVERY_BATCH catches up A steps at the
        boundary, STEP_EVERY_EPOCH steps when the next optimizer step lands in a later epoch."""
        ctx = self.context
        if not ctx._should_communicate_and_update():
            return
        agg, freq, mode = ctx._aggregation_frequency, lr_scheduler._frequency, lr_scheduler._step_mode
        if mode == LRScheduler.StepMode.STEP_EVERY_BATCH:
            for i in range(batch_idx - agg + 1, batch_idx + 1):
                if (i + 1) % freq == 0:
                    lr_scheduler.step()
        elif mode == LRScheduler.StepMode.STEP_EVERY_OPTIMIZER_STEP:
            if (batch_idx + 1) % freq == 0:
                lr_scheduler.step()
        elif mode == LRScheduler.StepMode.STEP_EVERY_EPOCH:
            assert ctx._epoch_len, "Training dataloader uninitialized."
            for e in range(batch_idx // ctx._epoch_len, (batch_idx + agg) // ctx._epoch_len):
                if (e + 1) % freq == 0:
                    lr_scheduler.step()

    def _should_update_scaler(self) -> bool:
        if not self.context._scaler or not self.context.experimental._auto_amp:
            return False
        return self.context._should_communicate_and_update()

    def _train_batch(self, batch: TorchData, epoch_idx: int, batch_idx: int, staged: Any = None) -> Optional[Dict[str, Any]]:
        self.context._loss_ids = {}
        if staged is not None:
            batch = staged
        elif self.context.experimental._auto_to_device:
            batch = self.context.to_device(batch)
        training_metrics = self.trial.train_batch(batch=batch, epoch_idx=epoch_idx, batch_idx=batch_idx)
        if self._should_update_scaler():
            self.context._scaler.update()
            self.context._scaler_updated()
        if isinstance(training_metrics, torch.Tensor):
            training_metrics = {"loss": training_metrics}
        for lr_scheduler in self.context.lr_schedulers:
            self._auto_step_lr_scheduler_per_batch(batch_idx, lr_scheduler)
        if not isinstance(training_metrics, dict):
            raise TypeError(
                f"train_batch() must return a dictionary mapping string names to Tensor metrics, "
                f"got {type(training_metrics).__name__}"
            )
        if self._device_metrics and self._metric_buffer.accepts(training_metrics):
            self._metric_buffer.append(training_metrics)  # stays on the GPU: no host sync in the batch loop
            return None
        if len(self._metric_buffer):
            raise RuntimeError("train_batch() must return the same kind of metrics for every batch of a reporting period")
        for name, metric in training_metrics.items():
            if isinstance(metric, torch.Tensor):
                metric = metric.cpu().detach().numpy()
            training_metrics[name] = metric
        return training_metrics

    def _aggregate_training_metrics(self, training_metrics: List[Dict]) -> Dict:
        """Average per-batch metrics over ranks and over the period, add the custom reducers."""
        d = self.context.distributed
        if len(self._metric_buffer):
            batch_metrics, avg, _ = self._metric_buffer.reduce(d)
            self._metric_buffer.reset()
            metrics: Dict[str, Any] = {"batch_metrics": batch_metrics, "avg_metrics": avg}
        else:
            if d.size > 1:
                batch_metrics = _metric_utils._combine_and_average_training_metrics(d, training_metrics)
            else:
                batch_metrics = training_metrics
            metrics = util.make_metrics(None, batch_metrics)
        metrics["avg_metrics"].update(_metric_utils._convert_metrics_to_numpy(self.context.reduce_metrics(for_training=True)))
        if not self.is_chief:
            return {}
        return metrics

    # ---- checkpoints ----------------------------------------------------------------------------------
    def _storage_dir(self) -> pathlib.Path:
        """Where this run's checkpoints live, one directory per uuid: the checkpoint_dir argument, else the
        experiment config's shared_fs checkpoint_storage, else DET_CHECKPOINT_DIR, else the system temp dir."""
        if self.checkpoint_dir is not None:
            return pathlib.Path(self.checkpoint_dir)
        storage = (self.context._exp_conf or {}).get("checkpoint_storage") or {}
        if storage.get("type", "shared_fs") == "shared_fs" and storage.get("host_path"):
            return pathlib.Path(storage["host_path"]) / (storage.get("storage_path") or "")
        return pathlib.Path(os.environ.get("DET_CHECKPOINT_DIR") or os.path.join(tempfile.gettempdir(), "determined-checkpoints"))

    def _checkpoint_path(self, uuid_or_path: str) -> pathlib.Path:
        direct = pathlib.Path(uuid_or_path)
        if direct.is_dir():
            return direct
        storage = getattr(self.context._core, "checkpoint", None) if self.checkpoint_dir is None else None
        if storage is not None:
            with storage.restore_path(uuid_or_path) as path:  # local storage: the path stays valid after the block
                return pathlib.Path(path)
        return self._storage_dir() / uuid_or_path

    def _checkpoint(self) -> str:
        """Write a checkpoint now (reference :395-423): the chief saves into <storage>/<uuid>/, every rank learns
        the uuid and runs on_checkpoint_upload_end.  Returns the uuid."""
        self.state.last_ckpt = self.batches_trained
        d = self.context.distributed
        name = ""
        collective_state = any(eng.sharded for eng in self.context._engines.values())
        if self.is_chief:
            metadata = {"steps_completed": self.batches_trained, "framework": f"torch-{torch.__version__}", "format": "pickle"}
            storage = getattr(self.context._core, "checkpoint", None) if self.checkpoint_dir is None else None
            if storage is not None:  # a core.Context was given: its checkpoint storage decides where (:410-415)
                with storage.store_path(metadata) as (path, name):
                    self._save(path)
            else:
                name = str(uuid.uuid4())
                path = self._storage_dir() / name
                self._save(path)
                with (path / "metadata.json").open("w") as f:
                    json.dump(metadata, f)
        elif collective_state:
            # sharded step: optimizer.state_dict() gathers the other ranks' slices, so every rank has to ask for it
            for opt in self.context.optimizers:
                opt.state_dict()
        if d.size > 1:
            name = d.broadcast(name)
        self.checkpoints.append(name)
        for cb in self.callbacks.values():
            cb.on_checkpoint_upload_end(uuid=name)
        return name

    def _save(self, path: Any) -> None:
        """Write state_dict.pth + trial_state.pkl + load_data.json with the reference's keys
        (_pytorch_trial.py:1312-1385), so that checkpoints move between the two harnesses.  Parameters and
        optimizer state are views of the engine's flat buffers; they are cloned so that torch.save stores
        ordinary per-tensor entries instead of one copy of the whole flat storage per view."""
        path = pathlib.Path(path)
        path.mkdir(parents=True, exist_ok=True)
        rng_state = {
            "cpu_rng_state": torch.random.get_rng_state(),
            "np_rng_state": np.random.get_state(),
            "random_rng_state": random.getstate(),
        }
        if torch.cuda.is_available():
            rng_state["gpu_rng_state"] = torch.cuda.get_rng_state(self.context.device)
        checkpoint = {
            "models_state_dict": [{k: v.detach().clone() for k, v in m.state_dict().items()} for m in self.context.models],
            "optimizers_state_dict": [_clone_tensors(o.state_dict()) for o in self.context.optimizers],
            "lr_schedulers_state_dict": [s.state_dict() for s in self.context.lr_schedulers],
            "callbacks": {name: cb.state_dict() for name, cb in self.callbacks.items()},
            "rng_state": rng_state,
        }
        if self.context._scaler:
            checkpoint["scaler_state_dict"] = self.context._scaler.state_dict()
        for cb in self.callbacks.values():
            cb.on_checkpoint_save_start(checkpoint)
        torch.save(checkpoint, str(path / "state_dict.pth"))
        self.state.batches_trained = self.batches_trained
        self.state.epochs_trained = self._get_epoch_idx(self.batches_trained)
        self.state.last_ckpt = self.batches_trained
        with (path / "trial_state.pkl").open("wb") as f:
            pickle.dump(vars(self.state), f)
        cls = type(self.trial)
        with (path / "load_data.json").open("w") as f:
            json.dump({
                "trial_type": "PyTorchTrial",
                "experiment_config": self.context._exp_conf,
                "hparams": self.context._hparams,
                "trial_cls_spec": f"{cls.__module__}:{cls.__qualname__}",
                "is_trainer": True,
            }, f)
        for cb in self.callbacks.values():
            cb.on_checkpoint_end(str(path))
            cb.on_checkpoint_write_end(str(path))

    # where state_dict.pth has lived over the reference's releases, newest first (:1118-1125)
    _CHECKPOINT_FILES = (("state_dict.pth",), ("determined", "state_dict.pth"), ("pedl", "state_dict.pth"), ("checkpoint.pt",))
    # (current list-valued key, pre-multi-model singular key, what it restores)
    _CHECKPOINT_PARTS = (
        ("models_state_dict", "model_state_dict", "model"),
        ("optimizers_state_dict", "optimizer_state_dict", "optimizer"),
        ("lr_schedulers_state_dict", "lr_scheduler", "LR scheduler"),
    )

    def _load(self, path: Any) -> None:
        """Restore from a checkpoint directory written by _save() or by any reference release the reference's
        own _load accepts (:1117-1271): the older file locations, the singular pre-multi-model keys, model keys
        with or without DDP's "module." prefix, and workload_sequencer.pkl from before trial_state.pkl.  Loaded
        optimizer state is copied back into the engine's flat buffers on the next step
        (GradSyncEngine._adopt_state)."""
        path = pathlib.Path(path)
        checkpoint = None
        for parts in self._CHECKPOINT_FILES:
            if path.joinpath(*parts).exists():
                checkpoint = torch.load(str(path.joinpath(*parts)), map_location="cpu", weights_only=False)
                break
        if not isinstance(checkpoint, dict):
            return  # same as the reference: nothing to restore, not an error
        for cb in self.callbacks.values():
            cb.on_checkpoint_load_start(checkpoint)

        targets = {"model": self.context.models, "optimizer": self.context.optimizers, "LR scheduler": self.context.lr_schedulers}
        for plural, singular, what in self._CHECKPOINT_PARTS:
            if singular in checkpoint:
                if plural in checkpoint:
                    raise RuntimeError(f"Both {singular} and {plural} in checkpoint.")
                if len(targets[what]) > 1:
                    raise RuntimeError(f"Old-format checkpoint cannot be loaded into a context with more than one {what}.")
                saved = [checkpoint[singular]]
            else:
                saved = checkpoint[plural] if targets[what] else []  # a missing key only matters if something needs it
            for obj, sd in zip(targets[what], saved):
                if what == "model":
                    # this harness never wraps the model, so its keys are the bare ones: a checkpoint taken from a
                    # DDP-wrapped model (reference, slots > 1) loses the prefix, everything else loads unchanged
                    sd = dict(sd)
                    torch.nn.modules.utils.consume_prefix_in_state_dict_if_present(sd, "module.")
                obj.load_state_dict(sd)

        scaler = self.context._scaler
        if "scaler_state_dict" in checkpoint:
            if scaler:
                scaler.load_state_dict(checkpoint["scaler_state_dict"])
            else:
                logger.warning("There exists scaler_state_dict in checkpoint but the experiment is not using AMP.")
        elif scaler:
            logger.warning("The experiment is using AMP but scaler_state_dict does not exist in the checkpoint.")

        rng = checkpoint.get("rng_state")
        if rng:
            np.random.set_state(rng["np_rng_state"])
            random.setstate(rng["random_rng_state"])
            torch.random.set_rng_state(rng["cpu_rng_state"])
            if torch.cuda.is_available() and self.context.device.type == "cuda":
                if "gpu_rng_state" in rng:
                    torch.cuda.set_rng_state(rng["gpu_rng_state"], self.context.device)
                else:
                    logger.warning("The system has a gpu but no gpu_rng_state exists in the checkpoint.")
            elif "gpu_rng_state" in rng:
                logger.warning("There exists gpu_rng_state in checkpoint but the system has no gpu.")
        else:
            logger.warning("The checkpoint has no random state to restore.")

        saved_callbacks = checkpoint.get("callbacks", {})
        for name, cb in self.callbacks.items():
            if name in saved_callbacks:
                cb.load_state_dict(saved_callbacks[name])
            elif util.is_overridden(cb.load_state_dict, PyTorchCallback):
                logger.warning(
                    f"Callback '{name}' implements load_state_dict(), but no callback state "
                    "was found for that name when restoring from checkpoint. This "
                    "callback will be initialized from scratch."
                )

        if (path / "trial_state.pkl").exists():
            with (path / "trial_state.pkl").open("rb") as f:
                self._load_state(pickle.load(f))
        elif (path / "workload_sequencer.pkl").exists():
            with (path / "workload_sequencer.pkl").open("rb") as f:
                self._load_wlsq_state(pickle.load(f))

    def _peek_batches_trained(self, path: pathlib.Path) -> int:
        """batches_trained a checkpoint will restore, read ahead of _load() because the data loader needs it."""
        for name, key in (("trial_state.pkl", "batches_trained"), ("workload_sequencer.pkl", "steps_completed")):
            if (path / name).exists():
                with (path / name).open("rb") as f:
                    state = pickle.load(f)
                return int(state.get(key) or 0) if state.get("trial_id") == self.trial_id else 0
        return self.batches_trained

    def _get_epoch_idx(self, batch_id: int) -> int:
        return batch_id // self.context._epoch_len if self.context._epoch_len else 0

    def _adopt_trial_state(self, state: "_TrialState") -> None:
        self.state = state
        # a validation made against exactly this checkpoint is known to the master but not to the file (:1287-1293)
        if state.batches_trained == self.val_from_previous_run:
            state.last_val = state.batches_trained
        self.batches_trained = state.batches_trained

    def _load_state(self, state: Dict[str, Any]) -> None:
        """trial_state.pkl (:1273-1293).  A checkpoint of ANOTHER trial only warm-starts the weights: the counters
        start from zero."""
        if state.get("trial_id") != self.trial_id:
            self.state = _TrialState(trial_id=self.trial_id)
            return
        self._adopt_trial_state(_TrialState(**state))

    def _load_wlsq_state(self, state: Dict[str, Any]) -> None:
        """workload_sequencer.pkl, the pre-Trainer file (:1295-1313): `steps_completed` there is what
        trial_state.pkl calls batches_trained; epochs are recomputed from it."""
        if state.get("trial_id") != self.trial_id:
            self.state = _TrialState(trial_id=self.trial_id)
            return
        done = state.get("steps_completed")
        self._adopt_trial_state(_TrialState(
            trial_id=state.get("trial_id"), last_ckpt=state.get("last_ckpt"), last_val=state.get("last_val"),
            step_id=state.get("step_id"), batches_trained=done, epochs_trained=self._get_epoch_idx(done),
        ))

    # ---- validation ---------------------------------------------------------------------------------
    def _validation_step(self) -> Dict[str, Any]:
        """One validation as the loop runs it (reference _validate :941-1115 around the metric computation): the
        callbacks, the report, and a checkpoint if the policy asks for one."""
        if self.validation_loader is None:
            self.state.last_val = self.batches_trained
            return {}
        for cb in self.callbacks.values():
            cb.on_validation_start()
        metrics = self._validate()
        d = self.context.distributed
        if d.size > 1 and any(util.is_overridden(cb.on_validation_end, PyTorchCallback) for cb in self.callbacks.values()):
            metrics = d.broadcast(metrics)  # every rank's callbacks see the chief's metrics (:1040-1048)
        for cb in self.callbacks.values():
            cb.on_validation_end(metrics)
        self.state.last_val = n = self.batches_trained
        should_checkpoint = False
        if self.is_chief:
            self.validation_reports.append({"steps_completed": n, "validation_metrics": metrics})
            train = getattr(self.context._core, "train", None)
            if train is not None:
                train.report_validation_metrics(n, metrics)  # (:1075)
            if self.context.get_enable_tensorboard_logging():
                _metric_utils._log_tb_metrics(self.context.get_tensorboard_writer(), "val", n, metrics)
            if self.on_validation:
                self.on_validation(n, metrics)
            searcher_metric = self._check_searcher_metric(metrics) if self.searcher_metric_name else None
            if self.state.last_ckpt != n:
                if self.ckpt_policy == "all":
                    should_checkpoint = True
                elif self.ckpt_policy == "best":
                    assert self.searcher_metric_name, "checkpoint policy 'best' but searcher metric name not defined"
                    should_checkpoint = self._is_best_validation(searcher_metric, self._best_validation)
            if searcher_metric is not None and self._is_best_validation(searcher_metric, self._best_validation):
                self._best_validation = float(searcher_metric)
        if d.size > 1:
            should_checkpoint = d.broadcast(should_checkpoint)
        if should_checkpoint:
            self._checkpoint()
        return metrics

    def _is_best_validation(self, now: float, before: Optional[float]) -> bool:
        if before is None:
            return True
        return (now < before) if self.smaller_is_better else (now > before)

    def _check_searcher_metric(self, val_metrics: Dict) -> Any:
        if self.searcher_metric_name not in val_metrics:
            raise RuntimeError(
                f"Search method is configured to use metric '{self.searcher_metric_name}' but "
                f"model definition returned validation metrics {list(val_metrics.keys())}. The "
                f"metric used by the search method must be one of the validation "
                "metrics returned by the model definition."
            )
        searcher_metric = val_metrics[self.searcher_metric_name]
        if not util.is_numerical_scalar(searcher_metric):
            raise RuntimeError(
                f"Searcher validation metric '{self.searcher_metric_name}' returned "
                f"a non-scalar value: {searcher_metric}."
            )
        return searcher_metric

    @torch.no_grad()
    def _validate(self) -> Dict[str, Any]:
        """The metric computation of a validation: per-rank evaluate_batch loop, reduce over batches then ranks
        (weighted by batches per rank); or evaluate_full_dataset on the chief."""
        if self.validation_loader is None:
            raise RuntimeError("no validation data loader")
        for model in self.context.models:
            model.eval()
        self.context.reset_reducers()
        d = self.context.distributed
        if util.is_overridden(self.trial.evaluate_full_dataset, PyTorchTrial):
            if util.is_overridden(self.trial.evaluate_batch, PyTorchTrial):
                raise errors.InvalidExperimentException(
                    "Please define exactly one of: `evaluate_batch()` or `evaluate_full_dataset()`. "
                    "For most use cases `evaluate_batch()` is recommended because "
                    "it can be parallelized across all devices.",
                )
            metrics: Dict[str, Any] = {}
            if self.is_chief:
                metrics = self.trial.evaluate_full_dataset(data_loader=self.validation_loader)
                if not isinstance(metrics, dict):
                    raise TypeError(f"eval() must return a dictionary, got {type(metrics).__name__}.")
                metrics = _metric_utils._convert_metrics_to_numpy(metrics)
        else:
            metrics = self._evaluate_batches()
        metrics.update(_metric_utils._convert_metrics_to_numpy(self.context.reduce_metrics(for_training=False)))
        for model in self.context.models:
            model.train()
        return metrics if self.is_chief else {}

    def _evaluate_batches(self) -> Dict[str, Any]:
        batch_metrics: List[Dict[str, Any]] = []
        keys = None
        # a callback that wants every batch's results gets them as host values, as the reference produces them
        wants_outputs = any(util.is_overridden(cb.on_validation_epoch_end, PyTorchCallback) for cb in self.callbacks.values())
        device_buf = _metric_utils.DeviceMetricBuffer(self.context.device) if self._device_metrics and not wants_outputs else None
        takes_idx = util.has_param(self.trial.evaluate_batch, "batch_idx", 2)
        for cb in self.callbacks.values():
            cb.on_validation_epoch_start()
        for idx, batch in enumerate(self.validation_loader):
            if self.context.experimental._auto_to_device:
                batch = self.context.to_device(batch)
            vld = self.trial.evaluate_batch(batch=batch, batch_idx=idx) if takes_idx else self.trial.evaluate_batch(batch=batch)
            if not isinstance(vld, dict):
                raise TypeError(f"validation_metrics() must return a dictionary of string names to Tensor metrics; got {vld}.")
            if keys is None:
                keys = vld.keys()
            elif keys != vld.keys():
                raise ValueError(f"Validation metric names must match across all batches of data: {keys} != {vld.keys()}.")
            if device_buf is not None and not batch_metrics and device_buf.accepts(vld) and all(v is not None for v in vld.values()):
                device_buf.append(vld)  # stays on the GPU (the reference converts every batch: :999)
            else:
                if device_buf is not None and len(device_buf):
                    raise RuntimeError("evaluate_batch() must return the same kind of metrics for every batch")
                batch_metrics.append(_metric_utils._convert_metrics_to_numpy(vld))
            if self.test_mode:
                break
        for cb in self.callbacks.values():
            cb.on_validation_epoch_end(batch_metrics)
        d = self.context.distributed
        keys = d.broadcast(list(keys) if keys is not None else None) if d.size > 1 else keys
        reducers = _metric_utils._prepare_metrics_reducers(self.trial.evaluation_reducer(), keys=keys or [])
        if device_buf is not None and len(device_buf):
            return _metric_utils.reduce_validation_on_device(device_buf, d, reducers)
        return _metric_utils._reduce_metrics(d, batch_metrics=batch_metrics, keys=keys, metrics_reducers=reducers)
