"""core._dummy_init(): the off-cluster core.Context the reference builds when there is no master (reference
harness/determined/core/_context.py:142-176 and the Dummy* classes it assembles: DummyTrainContext _train.py:261-297,
DummyPreemptContext _preempt.py:260-312, DummyCheckpointContext _checkpoint.py:724-747).  The reference's own tests
and `pytorch.init()` in local mode hand such a context to PyTorchTrialContext(core_context=...); the training loop
then reports through `.train`, stores checkpoints through `.checkpoint` and asks `.preempt` whether to stop.

Nothing here talks to a master.  Reports are logged and kept in lists; checkpoints are directories
<base>/<uuid>/ on the local file system."""
import contextlib
import json
import logging
import os
import pathlib
import tempfile
import uuid
from typing import Any, Dict, Iterator, List, Optional, Tuple, Union

from determined_clone_b200.core._distributed import DistributedContext, DummyDistributedContext

logger = logging.getLogger("determined.core")


class DummyTrainContext:
    def __init__(self, tensorboard_path: Optional[pathlib.Path] = None) -> None:
        self._tbd_directory = tensorboard_path
        self._trial_id, self._exp_id = 0, 0
        self.training_reports: List[Tuple[int, Dict[str, Any], Optional[List[Dict[str, Any]]]]] = []
        self.validation_reports: List[Tuple[int, Dict[str, Any]]] = []

    def set_status(self, status: str) -> None:
        logger.info(f"status: {status}")

    def _get_last_validation(self) -> Optional[int]:
        return None

    def report_training_metrics(self, steps_completed: int, metrics: Dict[str, Any],
                                batch_metrics: Optional[List[Dict[str, Any]]] = None) -> None:
        logger.info(f"report_training_metrics(steps_completed={steps_completed}, metrics={metrics})")
        self.training_reports.append((steps_completed, metrics, batch_metrics))

    def report_validation_metrics(self, steps_completed: int, metrics: Dict[str, Any]) -> None:
        logger.info(f"report_validation_metrics(steps_completed={steps_completed}, metrics={metrics})")
        self.validation_reports.append((steps_completed, metrics))

    def upload_tensorboard_files(self, *args: Any, **kwargs: Any) -> None:
        logger.info("upload_tensorboard_files()")

    def report_early_exit(self, reason: Any) -> None:
        logger.info(f"report_early_exit({reason})")

    def get_experiment_best_validation(self) -> Optional[float]:
        return None

    def get_tensorboard_path(self) -> pathlib.Path:
        if self._tbd_directory is None:
            self._tbd_directory = pathlib.Path(tempfile.mkdtemp(prefix="determined-tensorboard-"))
        return pathlib.Path(self._tbd_directory)


class DummyPreemptContext:
    """Nobody preempts a local run."""

    def __init__(self, distributed: Any = None, preempt_mode: Any = None) -> None:
        self._dist = distributed

    def start(self) -> "DummyPreemptContext":
        return self

    def should_preempt(self, auto_ack: bool = True) -> bool:
        return False

    def acknowledge_preemption_signal(self) -> None:
        pass

    def close(self) -> None:
        pass


class DummyCheckpointContext:
    """store_path / restore_path over a local directory (the reference's shared_fs storage manager, no upload)."""

    def __init__(self, distributed: Any, base_path: Union[str, os.PathLike]) -> None:
        self._dist = distributed
        self._base = pathlib.Path(base_path)

    @contextlib.contextmanager
    def store_path(self, metadata: Optional[Dict[str, Any]] = None, shard: bool = False) -> Iterator[Tuple[pathlib.Path, str]]:
        storage_id = str(uuid.uuid4())
        path = self._base / storage_id
        path.mkdir(parents=True, exist_ok=True)
        yield path, storage_id
        with (path / "metadata.json").open("w") as f:
            json.dump(metadata or {}, f, indent=2)

    @contextlib.contextmanager
    def restore_path(self, storage_id: str, *args: Any, **kwargs: Any) -> Iterator[pathlib.Path]:
        path = self._base / storage_id
        if not path.is_dir():
            raise FileNotFoundError(f"checkpoint {storage_id} not found in {self._base}")
        yield path

    def get_metadata(self, storage_id: str) -> Dict[str, Any]:
        with (self._base / storage_id / "metadata.json").open() as f:
            return json.load(f)


class Context:
    """What core.init() / core._dummy_init() return: the sub-contexts as attributes; also a context manager."""

    def __init__(self, distributed: Any, checkpoint: Any = None, preempt: Any = None, train: Any = None, searcher: Any = None) -> None:
        self.distributed, self.checkpoint, self.preempt, self.train, self.searcher = distributed, checkpoint, preempt, train, searcher

    def __enter__(self) -> "Context":
        if self.preempt is not None:
            self.preempt.start()
        return self

    def __exit__(self, *exc: Any) -> None:
        self.close()

    def close(self) -> None:
        if self.preempt is not None:
            self.preempt.close()
        self.distributed.close()


def _dummy_init(*, distributed: Optional[DistributedContext] = None, checkpoint_storage: Optional[Union[str, Dict[str, Any]]] = None,
                tensorboard_path: Optional[pathlib.Path] = None, preempt_mode: Any = None) -> Context:
    distributed = distributed or DummyDistributedContext()
    if isinstance(checkpoint_storage, dict):
        if checkpoint_storage.get("type", "shared_fs") != "shared_fs":
            raise ValueError("only shared_fs checkpoint storage exists off-cluster")
        base: Any = os.path.join(checkpoint_storage["host_path"], checkpoint_storage.get("storage_path") or "")
    else:
        base = checkpoint_storage or os.path.join(tempfile.gettempdir(), "determined-checkpoints")
    return Context(
        distributed=distributed,
        checkpoint=DummyCheckpointContext(distributed, base),
        preempt=DummyPreemptContext(distributed, preempt_mode),
        train=DummyTrainContext(tensorboard_path),
    )
