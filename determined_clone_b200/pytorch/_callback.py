"""PyTorchCallback: hooks a trial returns from build_callbacks() and the training loop calls around the grad-sync
path (reference harness/determined/pytorch/_callback.py; call sites in _pytorch_trial.py are cited per hook).

Every rank runs its own copy of every callback, except on_checkpoint_write_end / on_checkpoint_end, which only the
chief -- the rank that writes the files -- calls.  A callback that keeps state across batches overrides
state_dict() / load_state_dict(); that state travels in the checkpoint under the callback's name."""
from typing import Any, Dict, List, Optional


class PyTorchCallback:
    # ---- trial lifetime ---------------------------------------------------------------------------------------
    def on_trial_startup(self, first_batch_idx: int, checkpoint_uuid: Optional[str]) -> None:
        """First hook of a run, before data loaders are built (:573-581).  first_batch_idx: the batch training
        (re)starts at; checkpoint_uuid: the checkpoint about to be restored, or None."""

    def on_trial_shutdown(self) -> None:
        """Last hook of a run, also when it ends by an exception or a stop request (:581).  The trial may merely
        have been paused."""

    # ---- training ---------------------------------------------------------------------------------------------
    def on_training_start(self) -> None:
        """After the checkpoint (if any) is restored, before the first batch (:621-627)."""

    def on_training_epoch_start(self, epoch_idx: int) -> None:
        """Before the first batch of a training epoch (:371-386)."""

    def on_training_epoch_end(self, epoch_idx: int) -> None:
        """After the last batch of a training epoch (:388-393, :510-512)."""

    def on_training_workload_end(self, avg_metrics: Dict[str, Any], batch_metrics: Dict[str, Any]) -> None:
        """Whenever the loop stops training to report, validate or checkpoint: the metrics of the batches since the
        previous stop, averaged over ranks, on every rank (:766-777)."""

    # ---- validation -------------------------------------------------------------------------------------------
    def on_validation_start(self) -> None:
        """Before every validation (:956-957)."""

    def on_validation_epoch_start(self) -> None:
        """Before the first evaluate_batch of a validation (:969-970)."""

    def on_validation_epoch_end(self, outputs: List[Any]) -> None:
        """After the last evaluate_batch, with this rank's per-batch results (:1003-1004)."""

    def on_validation_end(self, metrics: Dict[str, Any]) -> None:
        """After the validation metrics are reduced; every rank gets the chief's metrics (:1040-1051)."""

    # ---- checkpoints ------------------------------------------------------------------------------------------
    def on_checkpoint_load_start(self, checkpoint: Dict[str, Any]) -> None:
        """With the loaded state_dict.pth contents, before anything is restored from them (:1137-1138)."""

    def on_checkpoint_save_start(self, checkpoint: Dict[str, Any]) -> None:
        """With the dictionary about to be written as state_dict.pth; may add to it (:1351-1352)."""

    def on_checkpoint_write_end(self, checkpoint_dir: str) -> None:
        """Chief only: all checkpoint files are in checkpoint_dir (:1385)."""

    def on_checkpoint_end(self, checkpoint_dir: str) -> None:
        """Deprecated spelling of on_checkpoint_write_end, still called (:1384)."""

    def on_checkpoint_upload_end(self, uuid: str) -> None:
        """Every rank: the checkpoint `uuid` is complete in checkpoint storage (:417-419)."""

    # ---- state ------------------------------------------------------------------------------------------------
    def state_dict(self) -> Dict[str, Any]:
        """Picklable state to keep across checkpoints."""
        return {}

    def load_state_dict(self, state_dict: Dict[str, Any]) -> None:
        """Restore what state_dict() returned."""
