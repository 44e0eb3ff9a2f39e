"""determined.pytorch-compatible surface of the B200 grad-sync path."""
from determined_clone_b200.pytorch._reducer import (
    MetricReducer,
    Reducer,
    _PyTorchReducerContext,
    _SimpleReducer,
    _simple_reduce_metrics,
)
from determined_clone_b200.pytorch._metric_utils import (
    DeviceMetricBuffer,
    _average_training_metrics,
    _log_tb_metrics,
    _combine_and_average_training_metrics,
    _combine_metrics_across_processes,
    _convert_metrics_to_numpy,
    _prepare_metrics_reducers,
    _process_combined_metrics_and_batches,
    _reduce_metrics,
)
from determined_clone_b200.pytorch._lr_scheduler import LRScheduler
from determined_clone_b200.pytorch._callback import PyTorchCallback
from determined_clone_b200.pytorch import samplers
from determined_clone_b200.pytorch._data import DataLoader, _dataset_repro_warning, adapt_batch_sampler, data_length, to_device
from determined_clone_b200.pytorch._comm import Comm
from determined_clone_b200.pytorch._gradsync import GradSyncEngine
from determined_clone_b200.pytorch._pytorch_context import PyTorchExperimentalContext, PyTorchTrialContext
from determined_clone_b200.pytorch._pytorch_trial import PyTorchTrial, TorchData, _PyTorchTrialController, _TrialState
from determined_clone_b200.pytorch._trainer import Batch, Epoch, Record, Trainer, TrainUnit, init
from determined_clone_b200.pytorch._load import CheckpointLoadContext, load_trial_from_checkpoint_path
