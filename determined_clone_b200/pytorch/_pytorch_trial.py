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
import json
import logging
import pathlib
import pickle
import random
import time
from typing import Any, Callable, Dict, List, Optional, Union

import numpy as np
import torch

from determined_clone_b200 import util
from determined_clone_b200.pytorch import _metric_utils
from determined_clone_b200.pytorch._data import DataLoader, data_length
from determined_clone_b200.pytorch._lr_scheduler import LRScheduler
from determined_clone_b200.pytorch._pytorch_context import PyTorchTrialContext
from determined_clone_b200.pytorch._reducer import Reducer

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

    def build_callbacks(self) -> Dict[str, Any]:
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
    ) -> None:
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
        self._metric_buffer = _metric_utils.DeviceMetricBuffer(context.device, capacity=max(reporting_period, 1))
        self.training_loader: Any = None
        self.validation_loader: Any = None
        self.trial_id = trial_id
        self.state = _TrialState(trial_id=trial_id, batches_trained=steps_completed)
        self.val_from_previous_run = val_from_previous_run

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
        d = self.context.distributed
        train = self.trial.build_training_data_loader()
        self.context._epoch_len = (len(train) - d.rank + d.size - 1) // d.size if hasattr(train, "__len__") else None
        if isinstance(train, DataLoader):
            self.training_loader = train.get_data_loader(repeat=True, skip=self.batches_trained, num_replicas=d.size, rank=d.rank)
        else:
            self.training_loader = train
        if self.context._epoch_len is not None and d.size > 1:
            self.context._epoch_len = d.broadcast(self.context._epoch_len)  # ranks must agree on epoch boundaries
        val = self.trial.build_validation_data_loader()
        if isinstance(val, DataLoader):
            self.validation_loader = val.get_data_loader(repeat=False, skip=0, num_replicas=d.size, rank=d.rank)
        else:
            self.validation_loader = val

    # ---- training loop ------------------------------------------------------------------------------
    def run(self) -> None:
        if self.training_loader is None:
            self._set_data_loaders()
        it = iter(self.training_loader)
        self.context.reset_reducers()
        pending: List[Dict[str, Any]] = []
        previous: Any = None
        samples = 0
        t0 = time.time()
        with contextlib.ExitStack() as stack:
            if self.context.profiler:
                stack.enter_context(self.context.profiler)  # torch.profiler set through context.set_profiler()
            self._train_loop(it, pending, previous, samples, t0)

    def _train_loop(self, it: Any, pending: List[Dict[str, Any]], previous: Any, samples: int, t0: float) -> None:
        while self.batches_trained < self.max_length:
            batch_idx = self.batches_trained
            try:
                batch = next(it)
            except StopIteration:
                it = iter(self.training_loader)
                batch = next(it)
            self.context._current_batch_idx = batch_idx
            epoch_idx = batch_idx // self.context._epoch_len if self.context._epoch_len else 0
            metrics = self._train_batch(batch, epoch_idx, batch_idx)
            if self.context.profiler:
                self.context.profiler.step()
            samples += self.trial.get_batch_length(batch)
            if metrics is not None:
                pending.append(metrics)
            self.batches_trained += 1
            # context.set_stop_requested(True) from user code ends training after the current batch (reference
            # _pytorch_trial.py:521-525 raises ShouldExit there); what has been trained is still reported
            stopping = self._stop_requested()
            if self.batches_trained % self.reporting_period == 0 or self.batches_trained == self.max_length or stopping:
                # the reduction of this period is enqueued now and read one period later (or at the end), so
                # the host keeps launching batches instead of waiting for the GPU to drain
                finish = self._begin_report(pending)
                self._deliver(previous)
                previous = (finish, self.batches_trained, samples, time.time() - t0)
                pending, samples, t0 = [], 0, time.time()
                self.context.reset_reducers()
            if stopping:
                break
        self._deliver(previous)

    def _stop_requested(self) -> bool:
        """As in the reference, the flag is per rank: user code must make the same request on every rank (the next
        batch starts with collectives)."""
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
        if self.is_chief:
            report["samples_per_second"] = samples / dur * self.context.distributed.size if dur > 0 else 0.0
            report["steps_completed"] = steps
            self.reports.append(report)
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

    def _train_batch(self, batch: TorchData, epoch_idx: int, batch_idx: int) -> Optional[Dict[str, Any]]:
        self.context._loss_ids = {}
        if self.context.experimental._auto_to_device:
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
            "callbacks": {},
            "rng_state": rng_state,
        }
        if self.context._scaler:
            checkpoint["scaler_state_dict"] = self.context._scaler.state_dict()
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

        if (path / "trial_state.pkl").exists():
            with (path / "trial_state.pkl").open("rb") as f:
                self._load_state(pickle.load(f))
        elif (path / "workload_sequencer.pkl").exists():
            with (path / "workload_sequencer.pkl").open("rb") as f:
                self._load_wlsq_state(pickle.load(f))

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
    @torch.no_grad()
    def _validate(self) -> Dict[str, Any]:
        """Per-rank evaluate_batch loop, reduce over batches then ranks (weighted by batches per rank)."""
        if self.validation_loader is None:
            raise RuntimeError("no validation data loader")
        for model in self.context.models:
            model.eval()
        self.context.reset_reducers()
        batch_metrics: List[Dict[str, Any]] = []
        keys = None
        device_buf = _metric_utils.DeviceMetricBuffer(self.context.device) if self._device_metrics else None
        for idx, batch in enumerate(self.validation_loader):
            if self.context.experimental._auto_to_device:
                batch = self.context.to_device(batch)
            vld = self.trial.evaluate_batch(batch=batch, batch_idx=idx)
            if keys is None:
                keys = vld.keys()
            if device_buf is not None and not batch_metrics and device_buf.accepts(vld) and all(v is not None for v in vld.values()):
                device_buf.append(vld)  # stays on the GPU (the reference converts every batch: :999)
                continue
            if device_buf is not None and len(device_buf):
                raise RuntimeError("evaluate_batch() must return the same kind of metrics for every batch")
            batch_metrics.append(_metric_utils._convert_metrics_to_numpy(vld))
        d = self.context.distributed
        keys = d.broadcast(list(keys) if keys is not None else None) if d.size > 1 else keys
        reducers = _metric_utils._prepare_metrics_reducers(self.trial.evaluation_reducer(), keys=keys or [])
        if device_buf is not None and len(device_buf):
            metrics = _metric_utils.reduce_validation_on_device(device_buf, d, reducers)
        else:
            metrics = _metric_utils._reduce_metrics(d, batch_metrics=batch_metrics, keys=keys, metrics_reducers=reducers)
        metrics.update(_metric_utils._convert_metrics_to_numpy(self.context.reduce_metrics(for_training=False)))
        for model in self.context.models:
            model.train()
        return metrics if self.is_chief else {}
