"""pytorch.init() and pytorch.Trainer: the entry points a train.py uses
(reference harness/determined/pytorch/_trainer.py: Trainer.fit :83-213, _initialize_distributed_backend
:216-236, init :249-319; TrainUnit classes in _trainer_utils.py).  Searcher operations, checkpoints and
preemption belong to the control plane and are not rebuilt: fit() trains for max_length (or the
searcher.max_length of the experiment config) and returns the chief's reports."""
import contextlib
import logging
import os
from typing import Any, Dict, Iterator, List, Optional

from determined_clone_b200 import _info, core
from determined_clone_b200._trial_controller import _DistributedBackend
from determined_clone_b200.pytorch._pytorch_context import PyTorchTrialContext
from determined_clone_b200.pytorch._pytorch_trial import PyTorchTrial, _PyTorchTrialController
from determined_clone_b200.pytorch._trainer_utils import Batch, Epoch, Record, TrainUnit, to_batches  # noqa: F401

logger = logging.getLogger("determined.pytorch")


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
        self.validation_reports: List[Dict[str, Any]] = []
        self.checkpoints: List[str] = []

    def _to_batches(self, unit: TrainUnit, loader_len: Optional[int]) -> int:
        return to_batches(unit, loader_len, self._context._global_batch_size)

    def fit(
        self,
        checkpoint_period: Optional[TrainUnit] = None,
        validation_period: Optional[TrainUnit] = None,
        max_length: Optional[TrainUnit] = None,
        reporting_period: TrainUnit = Batch(100),
        checkpoint_policy: str = "best",
        latest_checkpoint: Optional[str] = None,
        step_zero_validation: bool = False,
        test_mode: bool = False,
        checkpoint_dir: Optional[str] = None,
    ) -> List[Dict[str, Any]]:
        """Train, validate every validation_period, checkpoint every checkpoint_period (and after validations, per
        checkpoint_policy), report every reporting_period; validate and checkpoint once more at the end if the last
        batch was not covered (reference _trainer.py:83-213 -> _pytorch_trial.py:626-820).

        Launched with an experiment config (DET_EXPERIMENT_CONFIG): max_length, the searcher metric and
        smaller_is_better come from its `searcher` block, as on a cluster.  Otherwise ("local training") max_length
        is required and checkpoint_policy "best" falls back to "all", as in the reference.  Checkpoints are
        directories <checkpoint_dir>/<uuid>; latest_checkpoint is such a uuid, or a path.  Returns the chief's
        training reports; validation reports are in .validation_reports, checkpoint uuids in .checkpoints."""
        ctx = self._context
        conf = (ctx._exp_conf or {}) if _info.get_cluster_info() is not None else {}
        searcher = conf.get("searcher", {}) or {}
        searcher_metric_name, smaller_is_better = None, True
        if conf:
            if test_mode:
                raise ValueError("test_mode is only supported in local training mode.")
            if max_length is not None:
                logger.warning("max_length is ignored when training on-cluster. Please configure the searcher instead.")
            ml = searcher.get("max_length")
            if isinstance(ml, int):
                max_length = Batch(ml)
            elif isinstance(ml, dict) and len(ml) == 1:
                (unit, length), = ml.items()
                max_length = TrainUnit._from_searcher_unit(int(length), unit, ctx._global_batch_size)
            else:
                raise ValueError("the experiment config needs searcher.max_length")
            searcher_metric_name = searcher.get("metric")
            smaller_is_better = bool(searcher.get("smaller_is_better", True))
        else:
            if checkpoint_policy == "best":
                logger.warning("checkpoint_policy='best' is not supported in local training mode. Falling back to 'all'.")
                checkpoint_policy = "all"
            if max_length is None:
                raise ValueError("max_length must be defined in local training mode.")
            if not isinstance(max_length.value, int):
                raise TypeError("max_length must be configured in TrainUnit(int) types.")
        controller = _PyTorchTrialController(
            self._trial, ctx, max_length=0, reporting_period=1, test_mode=False,
            checkpoint_policy=checkpoint_policy, searcher_metric_name=searcher_metric_name,
            smaller_is_better=smaller_is_better, step_zero_validation=step_zero_validation, checkpoint_dir=checkpoint_dir,
            latest_checkpoint=latest_checkpoint, final_validation=True, final_checkpoint=True,
            trial_id=ctx.get_trial_id(),
        )
        if latest_checkpoint is not None:  # the loaders skip what the restored trial has trained; run() restores the rest
            controller.batches_trained = controller._peek_batches_trained(controller._checkpoint_path(latest_checkpoint))
        controller._set_data_loaders()
        epoch_len = ctx._epoch_len
        controller.max_length = controller.batches_trained + 1 if test_mode else self._to_batches(max_length, epoch_len)
        controller.test_mode = test_mode
        controller.reporting_period = max(self._to_batches(reporting_period, epoch_len), 1)
        controller.validation_period = self._to_batches(validation_period, epoch_len) if validation_period is not None else None
        controller.checkpoint_period = self._to_batches(checkpoint_period, epoch_len) if checkpoint_period is not None else None
        controller.run()
        self.reports = controller.reports
        self.validation_reports = controller.validation_reports
        self.checkpoints = controller.checkpoints
        return controller.reports
