"""pytorch.init() and pytorch.Trainer: the entry points a train.py uses
(reference harness/determined/pytorch/_trainer.py: Trainer.fit :83-213, _initialize_distributed_backend
:216-236, init :249-319; TrainUnit classes in _trainer_utils.py).  Searcher operations, checkpoints and
preemption belong to the control plane and are not rebuilt: fit() trains for max_length (or the
searcher.max_length of the experiment config) and returns the chief's reports."""
import contextlib
import os
from typing import Any, Dict, Iterator, List, Optional

from determined_clone_b200 import _info, core
from determined_clone_b200._trial_controller import _DistributedBackend
from determined_clone_b200.pytorch._pytorch_context import PyTorchTrialContext
from determined_clone_b200.pytorch._pytorch_trial import PyTorchTrial, _PyTorchTrialController


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


def _initialize_distributed_backend() -> Optional[core.DistributedContext]:
    """A DistributedContext from the launcher's environment (torchrun variables); the NCCL communicator
    itself is created lazily by the trial context, its id travelling over this context."""
    backend = _DistributedBackend()
    if backend.use_torch() or int(os.environ.get("WORLD_SIZE", "1")) > 1:
        for k, v in (("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")), ("GROUP_RANK", "0"), ("GROUP_WORLD_SIZE", "1"),
                     ("LOCAL_RANK", os.environ.get("RANK", "0"))):
            os.environ.setdefault(k, v)
        return core.DistributedContext.from_torch_distributed()
    return None


@contextlib.contextmanager
def init(
    *,
    hparams: Optional[Dict] = None,
    exp_conf: Optional[Dict[str, Any]] = None,
    distributed: Optional[core.DistributedContext] = None,
    aggregation_frequency: int = 1,
    enable_tensorboard_logging: bool = False,
) -> Iterator[PyTorchTrialContext]:
    """Creates the PyTorchTrialContext a PyTorchTrial is constructed with."""
    info = _info.get_cluster_info()
    trial_seed: Optional[int] = None
    if info is not None:
        exp_conf = info.trial._config
        hparams = hparams if hparams is not None else info.trial.hparams
        trial_seed = info.trial.trial_seed
        opts = exp_conf.get("optimizations", {}) or {}
        aggregation_frequency = int(opts.get("aggregation_frequency", aggregation_frequency))
    owned = distributed is None
    if distributed is None:
        distributed = _initialize_distributed_backend() or core.DummyDistributedContext()
    if trial_seed is not None:
        _PyTorchTrialController._set_random_seeds(trial_seed)
    context = PyTorchTrialContext(
        distributed=distributed, trial_seed=trial_seed, hparams=hparams, slots_per_trial=distributed.get_size(),
        exp_conf=exp_conf, aggregation_frequency=aggregation_frequency, managed_training=True,
        enable_tensorboard_logging=enable_tensorboard_logging,
    )
    if exp_conf:
        opts = exp_conf.get("optimizations", {}) or {}
        context._set_default_gradient_compression(bool(opts.get("gradient_compression", False)))
        context._set_default_average_aggregated_gradients(bool(opts.get("average_aggregated_gradients", True)))
    try:
        yield context
    finally:
        context.close()
        if owned:
            distributed.close()


class Trainer:
    def __init__(self, trial: PyTorchTrial, context: PyTorchTrialContext) -> None:
        self._trial, self._context = trial, context
        self.reports: List[Dict[str, Any]] = []

    def _to_batches(self, unit: TrainUnit, loader_len: Optional[int]) -> int:
        if isinstance(unit, Batch):
            return unit.value
        if isinstance(unit, Record):
            gbs = unit.global_batch_size or self._context.get_global_batch_size()
            return max(unit.value // gbs, 1)
        if loader_len is None:
            raise ValueError("max_length in epochs needs a training data loader with a length")
        return unit.value * loader_len

    def fit(self, max_length: Optional[TrainUnit] = None, reporting_period: TrainUnit = Batch(100),
            latest_checkpoint: Optional[str] = None, **_ignored: Any) -> List[Dict[str, Any]]:
        """Train for max_length; without it (on-cluster style) for searcher.max_length of the experiment
        config.  Returns the chief's per-period reports."""
        ctx = self._context
        if max_length is None:
            conf = ctx._exp_conf or {}
            ml = (conf.get("searcher", {}) or {}).get("max_length")
            if isinstance(ml, int):
                max_length = Batch(ml)
            elif isinstance(ml, dict) and len(ml) == 1:
                (unit, length), = ml.items()
                max_length = TrainUnit._from_searcher_unit(int(length), unit, ctx._global_batch_size)
            else:
                raise ValueError("fit() needs max_length, or an experiment config with searcher.max_length")
            if "scheduling_unit" in conf:
                reporting_period = Batch(int(conf["scheduling_unit"]))
        controller = _PyTorchTrialController(self._trial, ctx, max_length=0, reporting_period=1)
        controller._set_data_loaders()
        controller.max_length = self._to_batches(max_length, ctx._epoch_len)
        controller.reporting_period = max(self._to_batches(reporting_period, ctx._epoch_len), 1)
        controller.run()
        self.reports = controller.reports
        return controller.reports
